"""One extend case for profiling: python benchmarks/extend_one.py [bs P E Hq Hkv iters fp8 D]"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent))
from sweep import extend_case
a = [int(x) for x in sys.argv[1:]] + [None] * 8
bs, P, E, Hq, Hkv, iters = (a[0] or 8, a[1] if a[1] is not None else 16384, a[2] or 1024, a[3] or 32, a[4] or 8, a[5] or 3)
import torch
dt = torch.float8_e4m3fn if a[6] else torch.bfloat16
D = a[7] or 128
print(extend_case(f"{bs} req x ({P} + {E}) Hq{Hq} Hkv{Hkv} D{D} {'fp8' if a[6] else 'bf16'}", bs, P, E, Hq, Hkv, D, iters, pool_dtype=dt))
