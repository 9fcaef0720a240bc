"""Developer tool: per-kernel SASS opcode histogram of the shipped library (evidence that the tcgen05 / TMA path is what
the .so contains).

    python tools/sass_opcodes.py index-tts_b200/libbvg.so > profiles/rNN_libbvg_sass_opcodes.txt
"""
import collections
import re
import subprocess
import sys

path = sys.argv[1] if len(sys.argv) > 1 else "index-tts_b200/libbvg.so"
out = subprocess.run(["cuobjdump", "-sass", path], capture_output=True, text=True).stdout
kern, hist = None, collections.OrderedDict()
for line in out.splitlines():
    m = re.search(r"Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        kern = re.sub(r"\(.*", "", kern)[:110]
        hist[kern] = collections.Counter()
        continue
    m = re.match(r"\s+/\*[0-9a-f]{4,}\*/\s+(?:@!?U?P\d+\s+)?([A-Z][A-Z0-9_.]*)", line)
    if m and kern:
        hist[kern][m.group(1).split(".")[0]] += 1
KEY = ["UTCHMMA", "UTCQMMA", "UTCBAR", "UTCATOMSWS", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAPF", "UBLKCP", "SYNCS", "HMMA",
       "LDSM", "STSM", "FFMA2", "FMUL2", "FADD2", "FFMA", "MUFU", "LDGSTS", "ACQBULK", "ELECT", "NANOSLEEP"]
print(f"# cuobjdump -sass {path}: instruction counts per kernel (tcgen05.mma = UTCHMMA, tcgen05.ld/st = LDTM/STTM, "
      "tcgen05.commit = UTCBAR, TMA tensor load/store = UTMALDG/UTMASTG, mbarrier = SYNCS)")
for k, c in hist.items():
    tot = sum(c.values())
    if tot == 0:
        continue
    keys = " ".join(f"{n}={c[n]}" for n in KEY if c[n])
    top = " ".join(f"{n}={v}" for n, v in c.most_common(6))
    print(f"{k}\n    total={tot}  {keys}\n    top: {top}")
