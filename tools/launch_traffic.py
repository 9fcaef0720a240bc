"""Developer tool: DRAM traffic per launch over the timed step of `bench.py --steps 1 --warmup 3 --no-cpu-baseline`.

    ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none -c 1800 --csv \
        --log-file all.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-v1
    python tools/launch_traffic.py all.csv profiles/ncu_traffic.json > profiles/rNN_launch_traffic.txt

The timed step is what follows the 256 MiB L2-flush fill up to conv_post (same cut as tools/launch_share.py).  Updates the
`conv_umma_kernel_bytes_per_launch_avg` key of the JSON file bench.py reads for `roofline.traffic` (per-launch AVERAGE of
dram__bytes_read.sum + dram__bytes_write.sum over every conv_umma_kernel launch of the step -- the same aggregation as
`roofline.achieved`)."""
import collections
import csv
import json
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
H = rows[hdr]
idi, ki, mi, vi, ui = H.index("ID"), H.index("Kernel Name"), H.index("Metric Name"), H.index("Metric Value"), H.index("Metric Unit")
SCALE = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "nsecond": 1e-3, "us": 1.0, "usecond": 1.0,
         "ms": 1e3, "msecond": 1e3}
launch = collections.OrderedDict()
for r in rows[hdr + 1:]:
    if len(r) <= vi:
        continue
    d = launch.setdefault(r[idi], {"name": r[ki]})
    d[r[mi]] = float(r[vi].replace(",", "")) * SCALE.get(r[ui], 1.0)
L = list(launch.values())
fills = [i for i, d in enumerate(L) if "FillFunctor" in d["name"] and d.get("gpu__time_duration.sum", 0) > 30]
if not fills:
    sys.exit("no L2-flush fill found in the launch list")
seg = []
for d in L[fills[0] + 1:]:
    seg.append(d)
    if "conv_post" in d["name"]:
        break
agg = collections.OrderedDict()
for d in seg:
    name = d["name"].split("(")[0][:64]
    a = agg.setdefault(name, {"n": 0, "us": 0.0, "rd": 0.0, "wr": 0.0})
    a["n"] += 1
    a["us"] += d.get("gpu__time_duration.sum", 0.0)
    a["rd"] += d.get("dram__bytes_read.sum", 0.0)
    a["wr"] += d.get("dram__bytes_write.sum", 0.0)
print("DRAM traffic of the timed step per kernel (ncu launch list: serialised, cold L2 after the 256 MiB flush fill)")
print(f"{'kernel':66s} {'n':>4s} {'ms':>8s} {'read MB':>10s} {'write MB':>10s} {'MB/launch':>10s} {'GB/s':>8s}")
tot_b = 0.0
for k, a in agg.items():
    b = a["rd"] + a["wr"]
    tot_b += b
    print(f"{k:66s} {a['n']:4d} {a['us'] / 1e3:8.3f} {a['rd'] / 1e6:10.1f} {a['wr'] / 1e6:10.1f} {b / a['n'] / 1e6:10.1f} "
          f"{b / max(a['us'], 1e-9) / 1e3:8.1f}")
print(f"total {tot_b / 1e9:.3f} GB over {len(seg)} launches")
if len(sys.argv) > 2:
    try:
        j = json.load(open(sys.argv[2]))
    except Exception:
        j = {}
    for k, a in agg.items():
        if "conv_umma_kernel" in k:
            j["conv_umma_kernel_bytes_per_launch_avg"] = (a["rd"] + a["wr"]) / a["n"]
            j["conv_umma_kernel_launches"] = a["n"]
    j["step_dram_bytes"] = tot_b
    j["launch_list_source"] = ("ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum over the timed "
                               "step of bench.py --steps 1 --warmup 3 (tools/launch_traffic.py)")
    json.dump(j, open(sys.argv[2], "w"), indent=1)
