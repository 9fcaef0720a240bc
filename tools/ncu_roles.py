"""Developer tool: warp-stall samples of a conv_umma capture, listing the sync / TMA / UMMA instructions and hot spots.
    python tools/ncu_roles.py prof.ncu-rep [min_samples]"""
import csv, io, subprocess, sys
out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
body = rows[2:]
thr = int(sys.argv[2]) if len(sys.argv) > 2 else 150
tot = sum(int(r[4]) for r in body)
print("total samples", tot, "instructions", len(body))
for i, r in enumerate(body):
    t = r[1].strip()
    if int(r[4]) >= thr or any(k in t for k in ("TRYWAIT", "UTCHMMA", "UTMALDG", "UTMASTG", "LDTM", "UTCBAR")):
        print(f"{i:5d} S={int(r[4]):6d} E={r[5]:>9s}  {t[:100]}")
