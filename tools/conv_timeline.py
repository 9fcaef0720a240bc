"""Developer tool: per-tile timeline of the tcgen05 conv kernel's pipeline roles on CTA 0 (bvg_debug_conv_trace).

    python tools/conv_timeline.py C T k res [n_tiles]
Prints, per role event, the median cycle distance between consecutive tiles and the event-to-event latencies.
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from index_tts_b200 import ops, _lib
C, T, k, res = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3]), int(sys.argv[4])
N = int(sys.argv[5]) if len(sys.argv) > 5 else 96
B = 16
torch.manual_seed(0)
x = torch.randn(B, T, C, device="cuda").bfloat16()
w = ops.pack_conv1d_weight(torch.randn(C, C, k, device="cuda") / (C * k) ** 0.5, 1)
bias = torch.randn(C, device="cuda")
r = torch.randn(B, T, C, device="cuda") if res else None
fn = (lambda: ops.conv1d_nlc(x, w, bias, 1, 1, residual=r)) if res else (lambda: ops.conv1d_nlc(x, w, bias, 1, 1, out_f32=False, out_bf16=True))
fn(); fn(); torch.cuda.synchronize()
buf = torch.zeros(16 * N, dtype=torch.int64, device="cuda")
lib = _lib.load()
lib.bvg_debug_conv_trace(buf.data_ptr(), N)
fn(); torch.cuda.synchronize()
lib.bvg_debug_conv_trace(None, 0)
t = buf.cpu().view(16, N)
names = {0: "prod a_empty ok", 1: "mma acc_empty ok", 2: "mma a_full ok", 3: "mma committed", 4: "epi acc_full ok", 5: "epi tmem read", 6: "epi tile done",
         7: "mma1 acc_empty ok", 8: "mma1 a_full ok", 9: "mma1 committed", 10: "epi1 acc_full ok", 11: "epi1 tmem read", 12: "epi1 tile done"}
base = int(t[t > 0].min())
def med(v):
    v = sorted(v); return v[len(v) // 2] if v else -1
for s in range(13):
    row = t[s]
    n = int((row > 0).sum())
    if n < 8: continue
    d = [int(row[i + 1] - row[i]) for i in range(n // 2, n - 1)]
    print(f"{names[s]:18s} n={n:3d} first={int(row[0]) - base:7d} median period (2nd half) {med(d):6d} cyc")
def lat(a, b, shift=0, label=""):
    n = min(int((t[a] > 0).sum()), int((t[b] > 0).sum())) - abs(shift)
    if n < 8: return
    d = [int(t[b][i + max(shift, 0)] - t[a][i + max(-shift, 0)]) for i in range(n // 2, n)]
    print(f"  {label or (names[a] + ' -> ' + names[b])}: median {med(d)} cyc")
lat(2, 3, 0, "mma: a_full ok -> committed (issue time)")
lat(3, 4, 0, "mma committed -> epi sees acc_full")
lat(4, 5, 0, "epi acc_full -> tmem read done + acc_empty arrive")
lat(5, 6, 0, "epi after tmem read -> tile done")
lat(5, 1, 2, "epi acc_empty arrive (tile i) -> mma sees acc_empty (tile i+2)")
lat(1, 2, 0, "mma acc_empty ok -> a_full ok")
lat(0, 2, 0, "producer issued load -> mma sees a_full")
if N >= 148:
    st, en = t[13][:148], t[14][:148]
    t0 = int(st[st > 0].min())
    dur = sorted((int(e - s_), i) for i, (s_, e) in enumerate(zip(st, en)) if s_ > 0)
    print("CTA durations ns: min", dur[0], "median", dur[len(dur) // 2], "max", dur[-1], " kernel span", int(en.max()) - t0,
          " start skew", int(st.max()) - t0)
    d2 = [int(e - s_) // 1000 for s_, e in zip(st, en)]
    print("CTA durations us by blockIdx:", d2)
    print("CTA0 role end cycles:", {names[s]: int(t[s][(t[s] > 0).sum() - 1]) - base for s in (0, 3, 6, 9, 12) if (t[s] > 0).sum() > 0})
print("raw first 12 tiles (cycles from start):")
for s in range(13):
    if int((t[s] > 0).sum()) >= 8:
        print(f"{names[s]:18s}", [int(v) - base for v in t[s][:12]])
