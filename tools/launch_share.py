"""Developer tool: cut the timed step out of an ncu launch list of `bench.py --steps 1 --warmup 3 --no-cpu-baseline`.

    ncu --metrics gpu__time_duration.sum --clock-control none -c 1800 --csv --log-file all.csv python bench.py ...
    python tools/launch_share.py all.csv profiles/rNN_launches.csv > profiles/rNN_launches_summary.txt

The timed step is what follows the 256 MiB L2-flush fill (the only long FillFunctor launch) up to conv_post."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
H = rows[hdr]
ki, vi, ui = H.index("Kernel Name"), H.index("Metric Value"), H.index("Metric Unit")
L = []
for r in rows[hdr + 1:]:
    if len(r) <= vi:
        continue
    v = float(r[vi].replace(",", ""))
    L.append((r[ki], v / 1e3 if r[ui] in ("ns", "nsecond") else v, r))
fills = [i for i, (k, v, _) in enumerate(L) if "FillFunctor" in k and v > 30]
if not fills:
    sys.exit("no L2-flush fill found in the launch list")
seg = []
for k, v, r in L[fills[0] + 1:]:
    seg.append((k, v, r))
    if "conv_post" in k:
        break
agg, n = collections.Counter(), collections.Counter()
for k, v, _ in seg:
    name = k.split("(")[0][:60]
    agg[name] += v
    n[name] += 1
tot = sum(agg.values())
print("launches of the timed step (after the 256 MiB L2-flush fill); per-launch times are cold-cache and serialised:")
for k, v in agg.most_common(12):
    print(f"{k:62s} n={n[k]:4d} {v / 1e3:8.3f} ms {100 * v / tot:5.1f}%")
print(f"total {tot / 1e3:.3f} ms over {len(seg)} launches")
if len(sys.argv) > 2:
    with open(sys.argv[2], "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(H)
        for _, _, r in seg:
            w.writerow(r)
