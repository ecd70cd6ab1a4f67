"""Timeline of one extend_tc_kernel CTA from a -DB200_DEV -DB200_TRACE=<block> build:
    B200ATTN_LIB=benchmarks/_ab/lib_trace.so python benchmarks/_ab/trace_extend.py [page]
Prints, per kv block j (steady state), the clock deltas between the hand-offs of the S -> softmax -> P -> PV chain."""
import ctypes, os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch
from sweep import extend_case
from sglang_b200 import _C

page = int(sys.argv[1]) if len(sys.argv) > 1 else 1
D = int(sys.argv[2]) if len(sys.argv) > 2 else 128           # 256: the one-tile D 256 instance (Gemma-2 heads)
one_tile = len(sys.argv) > 3 and sys.argv[3] == "one"
if D == 256:
    print(extend_case("trace D256 8 req x (16K + 1K)", 8, 16384, 1024, 16, 8, 256, 3, page=page))
else:
    print(extend_case("trace 8 req x (16K + 1K)", 8, 16384, 1024, 32, 8, 128, 3, page=page, one_tile=one_tile))
torch.cuda.synchronize()
buf = np.zeros(16 * 512, dtype=np.int64)
rc = _C.lib.b200attn_dev_trace_read(buf.ctypes.data_as(ctypes.c_void_p))
assert rc == 0, rc
tr = buf.reshape(16, 512)
n = int((tr[0] > 0).sum())
print("kv blocks traced:", n)
names = {0: "S0rdy", 1: "S0ld", 2: "S0max", 3: "P0arr", 4: "S1rdy", 5: "S1ld", 6: "S1max", 7: "P1arr",
         8: "mmaKV", 9: "P0seen", 10: "t0iss", 11: "P1seen", 12: "t1iss", 13: "prK", 14: "prV"}
lo, hi = 20, min(n - 4, 120)
base = tr[0, lo]
print("j   " + " ".join(f"{names[s]:>7}" for s in range(15)))
for j in range(lo, lo + 12):
    print(f"{j:3d} " + " ".join(f"{tr[s, j] - base:7d}" for s in range(15)))
def avg(a, b, sa=0, sb=0):
    d = tr[b, lo + sb:hi + sb] - tr[a, lo + sa:hi + sa]
    return float(d.mean()), float(np.median(d)), int(d.min()), int(d.max())
print("period (S0rdy j -> j+1):", avg(0, 0, 0, 1))
for t in (0, 1):
    o = 4 * t
    print(f"tile {t}: S ready -> ld done      ", avg(o + 0, o + 1))
    print(f"tile {t}: ld done -> max/rescale  ", avg(o + 1, o + 2))
    print(f"tile {t}: exp+pack+st -> P arrive ", avg(o + 2, o + 3))
    print(f"tile {t}: P arrive -> mma sees P  ", avg(o + 3, 9 + 2 * t))
    print(f"tile {t}: mma sees P -> issued    ", avg(9 + 2 * t, 10 + 2 * t))
    print(f"tile {t}: issued(j) -> S ready(j+1)", avg(10 + 2 * t, o + 0, 0, 1))
    print(f"tile {t}: P arrive(j) -> S ready(j+1)", avg(o + 3, o + 0, 0, 1))
print("mma: t1 issued(j) -> KV seen(j+1)", avg(12, 8, 0, 1))
print("mma: KV seen -> P0 seen", avg(8, 9))
print("mma: t0 issued -> P1 seen", avg(10, 11))
