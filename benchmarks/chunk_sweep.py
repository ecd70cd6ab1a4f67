"""Fixed-chunk sweep of the split-KV decode (planner cost-model check): python benchmarks/chunk_sweep.py"""
import sys
from pathlib import Path
import torch
sys.path.insert(0, str(Path(__file__).resolve().parent.parent)); sys.path.insert(0, str(Path(__file__).resolve().parent))
from sglang_b200 import ops
from sweep import time_cuda, time_cuda_graph

def run(bs, L, Hq, Hkv, D=128, chunks=(0, 256, 512, 768, 1024, 1216, 1824, 2048, 2736, 3648, 4096, 8192)):
    dev = "cuda"; g = torch.Generator(device=dev).manual_seed(0)
    rows = bs * L + 1
    kp = torch.randn(rows, Hkv, D, generator=g, device=dev, dtype=torch.bfloat16); vp = torch.randn(rows, Hkv, D, generator=g, device=dev, dtype=torch.bfloat16)
    perm = torch.randperm(bs * L, generator=g, device=dev) + 1
    rtt = torch.zeros(bs + 1, L, dtype=torch.int32, device=dev); rtt[1:] = perm.view(bs, L).int()
    q = torch.randn(bs, Hq, D, generator=g, device=dev, dtype=torch.bfloat16); o = torch.empty_like(q)
    lens = torch.full((bs,), L, dtype=torch.int32, device=dev); rpi = torch.arange(1, bs + 1, device=dev)
    sm, _, _ = ops.device_info()
    nbytes = bs * L * Hkv * 2 * D * 2 + bs * L * 4 + bs * Hq * 2 * D * 2
    for cps in sorted({ops.decode_ctas_per_sm(D, torch.bfloat16), 2, 3}):
        ws = ops.DecodeWorkspace.create(bs, Hq, D, dev, num_ctas=sm * cps)
        for c in chunks:
            try:
                ops.decode_plan(ws, lens, Hkv, fixed_chunk_tokens=c,
                                items_per_chunk=ops.decode_items_per_chunk(Hq, Hkv, D, D, torch.bfloat16, torch.bfloat16))
                fn = lambda: ops.decode_attention(q, o, kp, vp, rtt, rpi, lens, ws, D**-0.5, kv_dtype=torch.bfloat16)
                t = time_cuda_graph(fn)
                print(f"bs{bs} L{L} Hq{Hq} Hkv{Hkv} ctas/sm {cps} fixed_chunk {c:5d} -> chunk {int(ws.plan[0]):5d} chunks {int(ws.plan[1]):5d}  {t*1e6:7.1f} us  {nbytes/t/1e9:7.1f} GB/s", flush=True)
            except Exception as e:
                print(f"chunk {c}: {type(e).__name__} {str(e)[:100]}")

if __name__ == "__main__":
    run(32, 32768, 8, 1)
    run(32, 32768, 16, 2)
    run(8, 32768, 8, 1)
