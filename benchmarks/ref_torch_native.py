"""CPU arm of bench.py: the REFERENCE's own `TorchNativeAttnBackend` (imported from the staged tree oracle/_ref,
never restated) on the host cores, driven through the reference's real `ReqToTokenPool`, `MHATokenToKVPool`,
`ForwardBatch`, `RadixAttention` and `ForwardContext` on device "cpu".

    python benchmarks/ref_torch_native.py --mode decode --bs 64 --ctx 4096 --hq 32 --hkv 8 --d 128 --steps 3 --warmup 1

Reference: srt/layers/attention/torch_native_backend.py:176-277 (decode), :61-174 (extend); SURVEY section 8d
"CPU baseline beside it".  One JSON object on the last stdout line: seconds per step and the thread count.
This file is bench/test infrastructure: it is the only place besides tests/ that executes anything under oracle/.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time
from pathlib import Path
from types import SimpleNamespace

REPO = Path(__file__).resolve().parent.parent
REF = REPO / "oracle" / "_ref"


def main() -> int:
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="decode", choices=["decode", "extend"])
    ap.add_argument("--bs", type=int, default=64)
    ap.add_argument("--ctx", type=int, default=4096, help="decode: tokens per request; extend: prefix length")
    ap.add_argument("--ext", type=int, default=128, help="extend: new tokens per request")
    ap.add_argument("--hq", type=int, default=32)
    ap.add_argument("--hkv", type=int, default=8)
    ap.add_argument("--d", type=int, default=128)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=1)
    ap.add_argument("--budget-s", type=float, default=0.0, help="stop after this many seconds (at least 1 step)")
    a = ap.parse_args()
    if not (REF / "MANIFEST.json").exists():
        print(json.dumps({"unavailable": "oracle/_ref is not staged"}))
        return 0
    sys.path[:0] = [str(REF / "stubs"), str(REF / "python")]
    import torch

    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    from sglang.srt.layers.attention.torch_native_backend import TorchNativeAttnBackend
    from sglang.srt.layers.radix_attention import RadixAttention
    from sglang.srt.mem_cache.memory_pool import MHATokenToKVPool, ReqToTokenPool
    from sglang.srt.model_executor.forward_batch_info import ForwardBatch, ForwardMode
    from sglang.srt.model_executor.forward_context import ForwardContext, forward_context

    bs, Hq, Hkv, D = a.bs, a.hq, a.hkv, a.d
    dt = torch.bfloat16
    g = torch.Generator().manual_seed(0)
    decode = a.mode == "decode"
    new = 1 if decode else a.ext
    L = a.ctx if decode else a.ctx + a.ext
    rows = bs * L + 1
    rtp = ReqToTokenPool(size=bs, max_context_len=L, device="cpu", enable_memory_saver=False)
    pool = MHATokenToKVPool(size=rows - 1, page_size=1, dtype=dt, head_num=Hkv, head_dim=D, layer_num=1,
                            device="cpu", enable_memory_saver=False, enable_alt_stream=False)
    pool.k_buffer[0].copy_(torch.randn(pool.k_buffer[0].shape, generator=g).to(dt))
    pool.v_buffer[0].copy_(torch.randn(pool.v_buffer[0].shape, generator=g).to(dt))
    perm = (torch.randperm(rows - 1, generator=g) + 1).view(bs, L).to(torch.int32)
    rtp.req_to_token[1 : bs + 1, :L] = perm
    rpi = torch.arange(1, bs + 1, dtype=torch.int64)
    T = bs * new
    q = torch.randn(T, Hq * D, generator=g).to(dt)
    k = torch.randn(T, Hkv, D, generator=g).to(dt)
    v = torch.randn(T, Hkv, D, generator=g).to(dt)
    loc = perm[:, L - new :].reshape(-1).to(torch.int64)
    seq = torch.full((bs,), L, dtype=torch.int32)
    fb = ForwardBatch(forward_mode=ForwardMode.DECODE if decode else ForwardMode.EXTEND, batch_size=bs,
                      input_ids=torch.arange(T, dtype=torch.int64), req_pool_indices=rpi, seq_lens=seq,
                      seq_lens_cpu=seq.clone(), out_cache_loc=loc, seq_lens_sum=bs * L,
                      positions=torch.zeros(T, dtype=torch.int64))
    if not decode:
        fb.extend_prefix_lens = torch.full((bs,), a.ctx, dtype=torch.int32)
        fb.extend_prefix_lens_cpu = [a.ctx] * bs
        fb.extend_seq_lens = torch.full((bs,), a.ext, dtype=torch.int32)
        fb.extend_seq_lens_cpu = [a.ext] * bs
        fb.extend_num_tokens = T
    backend = TorchNativeAttnBackend(SimpleNamespace(device="cpu", req_to_token_pool=rtp, token_to_kv_pool=pool))
    layer = RadixAttention(num_heads=Hq, head_dim=D, scaling=D ** -0.5, num_kv_heads=Hkv, layer_id=0)

    def step():
        with torch.no_grad(), forward_context(ForwardContext(attn_backend=backend)):
            backend.init_forward_metadata(fb)
            return backend.forward(q, k, v, layer, fb, save_kv_cache=True)

    for _ in range(a.warmup):
        step()
    t0, n = time.perf_counter(), 0
    while n < a.steps:
        step()
        n += 1
        if a.budget_s > 0 and time.perf_counter() - t0 > a.budget_s:
            break
    dt_s = (time.perf_counter() - t0) / n
    print(json.dumps({"seconds_per_step": dt_s, "steps": n, "cores": cores, "mode": a.mode, "bs": bs, "ctx": a.ctx,
                      "ext": a.ext, "backend": f"{TorchNativeAttnBackend.__module__}.TorchNativeAttnBackend",
                      "sglang_file": sys.modules["sglang"].__file__}))
    return 0


if __name__ == "__main__":
    sys.exit(main())
