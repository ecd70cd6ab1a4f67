"""cfg5-size KV store and fused RoPE+store, two launches each, for ncu: python benchmarks/store_one.py"""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent))
import sweep
bf, e4 = torch.bfloat16, torch.float8_e4m3fn
sweep.time_cuda = lambda fn, iters, warmup=5: (fn(), fn(), torch.cuda.synchronize(), 1.0)[-1]
print(sweep.store_case("cfg5 store bf16", 131072, 8, 128, bf, 2)["case"])
print(sweep.store_case("cfg5 store fp8", 131072, 8, 128, e4, 2)["case"])
print(sweep.rope_store_case("cfg5 rope+store bf16", 131072, 32, 8, 128, bf, 2)["case"])
print(sweep.rope_store_case("cfg5 rope+store fp8", 131072, 32, 8, 128, e4, 2)["case"])
