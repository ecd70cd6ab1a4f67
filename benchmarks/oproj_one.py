"""Single-GPU timing of the o_proj GEMM path of the fused kernel (TP 1: no exchange) vs cuBLAS, with the weights
rotated through copies larger than L2 so that every launch streams them from HBM (as a real layer stack does).

    python benchmarks/oproj_one.py [T N K copies]"""
import sys
from pathlib import Path

import torch

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
from sglang_b200.comm import OProjAllReduce  # noqa: E402


def graph_time(fns, iters=10):
    """us per call of the sequence `fns` captured once in a CUDA graph."""
    for f in fns:
        f()
    torch.cuda.synchronize()
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=side):
            for f in fns:
                f()
    torch.cuda.current_stream().wait_stream(side)
    for _ in range(3):
        g.replay()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        g.replay()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters / len(fns) * 1e3


def main():
    a = [int(v) for v in sys.argv[1:]]
    shapes = [(a[0], a[1], a[2])] if len(a) >= 3 else [(8, 8192, 4096), (64, 8192, 4096), (128, 8192, 4096),
                                                       (64, 8192, 1024), (64, 4096, 4096), (512, 8192, 4096)]
    for T, N, K in shapes:
        copies = a[3] if len(a) > 3 else max(2, (300 << 20) // (N * K * 2) + 1)
        x = (torch.randn(T, K, device="cuda") * 0.5).bfloat16()
        ws = [(torch.randn(N, K, device="cuda") * 0.5).bfloat16() for _ in range(copies)]
        out = torch.empty(T, N, dtype=torch.bfloat16, device="cuda")
        comm = OProjAllReduce(None, max(T, 128), N, torch.bfloat16, "cuda")
        t_fused = graph_time([(lambda w=w: comm.forward(x, w)) for w in ws])
        t_cublas = graph_time([(lambda w=w: torch.matmul(x, w.t(), out=out)) for w in ws])
        byt = N * K * 2 + T * K * 2 + T * N * 2
        print(f"T={T} N={N} K={K} copies={copies}: fused {t_fused:.2f} us = {byt / t_fused / 1e3:.0f} GB/s | "
              f"cuBLAS {t_cublas:.2f} us = {byt / t_cublas / 1e3:.0f} GB/s", flush=True)


if __name__ == "__main__":
    main()
