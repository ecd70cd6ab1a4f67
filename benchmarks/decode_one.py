"""One decode case for profiling: python benchmarks/decode_one.py [bs L Hq Hkv fp8(0/1) page iters D]"""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent))
from sweep import decode_case
a = [int(x) for x in sys.argv[1:]] + [None] * 8
bs, L, Hq, Hkv, fp8, page, iters = (a[0] or 256, a[1] or 8192, a[2] or 32, a[3] or 8, a[4] if a[4] is not None else 1,
                                    a[5] or 1, a[6] or 10)
D = a[7] or 128
dt = torch.float8_e4m3fn if fp8 else torch.bfloat16
print(decode_case(f"bs{bs} ctx{L} Hq{Hq} Hkv{Hkv} {'fp8' if fp8 else 'bf16'} page{page} D{D}", bs, L, Hq, Hkv, D, dt, page, iters))
