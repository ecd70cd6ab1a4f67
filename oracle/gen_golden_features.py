"""Generate tests/golden/features.npz by running the REFERENCE itself on CPU (test infrastructure).

Run in the build container only (needs /root/reference; the GPU box uses the committed file):

    TRITON_INTERPRET=1 PYTHONPATH=oracle/stubs:/root/reference/python python oracle/gen_golden_features.py

What runs is the reference's own code, unmodified:
  - `extend_attention_fwd`  (kernels/ops/attention/extend_attention.py:664-812)  -- the production Triton extend
    kernel, executed by Triton's CPU interpreter in float32: attention sinks, tanh logit cap, sliding window,
    custom (tree) masks with and without SKIP_PREFIX_CUSTOM_MASK, LSE output
  - `decode_attention_fwd`  (kernels/ops/attention/decode_attention.py:968-1044) -- split-KV decode stage 1 + 2:
    sinks, logit cap, and a sliding window expressed the way triton_backend.py does it (window kv_indices)
  - `merge_state_triton`    (kernels/ops/attention/merge_state.py:69-100)
  - `TorchNativeAttnBackend.forward_decode / forward_extend` with sliding-window layers
    (srt/layers/attention/torch_native_backend.py:36-48, 150-157, 251-258)
Every case stores a pool + page table the way this repo's kernels see them (new tokens already stored), the
query, the options, and the reference output (float32).
"""
from __future__ import annotations

import os
import sys
from pathlib import Path
from types import SimpleNamespace

os.environ.setdefault("TRITON_INTERPRET", "1")

import numpy as np
import torch

REPO = Path(__file__).resolve().parent.parent
GOLDEN = REPO / "tests" / "golden"

from sglang.kernels.ops.attention.decode_attention import decode_attention_fwd  # noqa: E402
from sglang.kernels.ops.attention.extend_attention import extend_attention_fwd  # noqa: E402
from sglang.kernels.ops.attention.merge_state import merge_state_triton  # noqa: E402

OUT = {}
NAMES = []


def put(name, **arrays):
    NAMES.append(name)
    for k, v in arrays.items():
        if isinstance(v, torch.Tensor):
            v = v.detach().cpu().numpy()
        OUT[f"{name}.{k}"] = np.asarray(v)


def make_layout(gen, prefix_lens, extend_lens, Hkv, D, slots):
    """Random pool + page table: request r's tokens 0..P+E sit at random distinct slots (slot 0 unused)."""
    bs = len(prefix_lens)
    max_ctx = max(p + e for p, e in zip(prefix_lens, extend_lens)) + 3
    perm = (torch.randperm(slots - 1, generator=gen) + 1).to(torch.int32)
    rtt = torch.zeros(bs + 1, max_ctx, dtype=torch.int32)
    rpi = torch.arange(1, bs + 1, dtype=torch.int64)
    used = 0
    for r in range(bs):
        n = prefix_lens[r] + extend_lens[r]
        rtt[r + 1, :n] = perm[used : used + n]
        used += n
    assert used < slots
    kp = torch.randn(slots, Hkv, D, generator=gen)
    vp = torch.randn(slots, Hkv, D, generator=gen)
    return rtt, rpi, kp, vp


def triton_extend_case(name, seed, Hq, Hkv, D, prefix_lens, extend_lens, *, logit_cap=0.0, window=-1, sinks=False,
                       mask=None, skip_prefix_mask=True, causal=True):
    gen = torch.Generator().manual_seed(seed)
    bs = len(prefix_lens)
    rtt, rpi, kp, vp = make_layout(gen, prefix_lens, extend_lens, Hkv, D, 256)
    T = sum(extend_lens)
    q = torch.randn(T, Hq, D, generator=gen)
    scale = D ** -0.5
    # the reference kernel reads the prefix through kv_indices and the new tokens from contiguous k/v_extend
    qo_indptr = torch.zeros(bs + 1, dtype=torch.int32)
    kv_indptr = torch.zeros(bs + 1, dtype=torch.int32)
    qo_indptr[1:] = torch.tensor(extend_lens).cumsum(0)
    kv_indptr[1:] = torch.tensor(prefix_lens).cumsum(0)
    kv_indices = torch.cat([rtt[r + 1, : prefix_lens[r]] for r in range(bs)]).to(torch.int64)
    new_slots = torch.cat([rtt[r + 1, prefix_lens[r] : prefix_lens[r] + extend_lens[r]] for r in range(bs)]).long()
    k_ext, v_ext = kp[new_slots].contiguous(), vp[new_slots].contiguous()
    sink_t = torch.randn(Hq, generator=gen) if sinks else None
    custom_mask = mask_indptr = None
    if mask is not None:
        parts, offs = [], [0]
        for r in range(bs):
            P, E = prefix_lens[r], extend_lens[r]
            if mask == "tree":      # a draft tree: token i sees a random subset of earlier new tokens + itself
                m = torch.zeros(E, P + E, dtype=torch.bool)
                m[:, :P] = True
                for i in range(E):
                    m[i, P + i] = True
                    for j in range(i):
                        m[i, P + j] = bool(torch.rand((), generator=gen) < 0.5)
            else:                   # "full": random mask over prefix and new tokens, diagonal kept
                m = torch.rand(E, P + E, generator=gen) < 0.6
                for i in range(E):
                    m[i, P + i] = True
            parts.append(m.reshape(-1))
            offs.append(offs[-1] + E * (P + E))
        custom_mask = torch.cat(parts).to(torch.uint8)
        mask_indptr = torch.tensor(offs, dtype=torch.int64)
    o = torch.zeros(T, Hq, D)
    lse = torch.zeros(T, Hq)
    extend_attention_fwd(
        q, k_ext, v_ext, o, kp, vp, qo_indptr, kv_indptr, kv_indices,
        custom_mask.to(torch.bool) if custom_mask is not None else None, causal, mask_indptr, max(extend_lens),
        1.0, 1.0, scale, logit_cap=logit_cap, skip_prefix_custom_mask=skip_prefix_mask,
        sliding_window_size=window, sinks=sink_t, lse_extend=lse,
    )
    put(name, kind=np.int32(0), q=q, k_pool=kp, v_pool=vp, req_to_token=rtt, req_pool_indices=rpi,
        prefix_lens=np.array(prefix_lens, np.int32), extend_lens=np.array(extend_lens, np.int32),
        opts=np.array([logit_cap, window + 1 if window > -1 else 0, int(skip_prefix_mask), int(causal)], np.float32),
        sinks=sink_t if sinks else np.zeros(0, np.float32),
        custom_mask=custom_mask if custom_mask is not None else np.zeros(0, np.uint8),
        mask_indptr=mask_indptr if mask_indptr is not None else np.zeros(0, np.int64),
        out=o, lse=lse)
    print(name, tuple(o.shape), float(o.abs().max()))


def triton_decode_case(name, seed, Hq, Hkv, D, seq_lens, *, logit_cap=0.0, window_keys=0, sinks=False, splits=4):
    gen = torch.Generator().manual_seed(seed)
    bs = len(seq_lens)
    rtt, rpi, kp, vp = make_layout(gen, seq_lens, [0] * bs, Hkv, D, 512)
    q = torch.randn(bs, Hq, D, generator=gen)
    scale = D ** -0.5
    # sliding window the way triton_backend.py:686-699 feeds the kernel: kv_indices of the last tokens only
    lens = [min(L, window_keys) if window_keys > 0 else L for L in seq_lens]
    kv_indptr = torch.zeros(bs + 1, dtype=torch.int32)
    kv_indptr[1:] = torch.tensor(lens).cumsum(0)
    kv_indices = torch.cat([rtt[r + 1, seq_lens[r] - lens[r] : seq_lens[r]] for r in range(bs)]).to(torch.int64)
    sink_t = torch.randn(Hq, generator=gen) if sinks else None
    o = torch.zeros(bs, Hq, D)
    attn_logits = torch.zeros(bs, Hq, splits, D)
    attn_lse = torch.zeros(bs, Hq, splits)
    num_kv_splits = torch.full((bs,), splits, dtype=torch.int32)
    decode_attention_fwd(q, kp, vp, o, kv_indptr, kv_indices, attn_logits, attn_lse, num_kv_splits, splits,
                         scale, 1.0, 1.0, logit_cap=logit_cap, sinks=sink_t)
    put(name, kind=np.int32(1), q=q, k_pool=kp, v_pool=vp, req_to_token=rtt, req_pool_indices=rpi,
        prefix_lens=np.array(seq_lens, np.int32), extend_lens=np.zeros(bs, np.int32),
        opts=np.array([logit_cap, window_keys, 1, 1], np.float32),
        sinks=sink_t if sinks else np.zeros(0, np.float32),
        custom_mask=np.zeros(0, np.uint8), mask_indptr=np.zeros(0, np.int64), out=o, lse=np.zeros(0, np.float32))
    print(name, tuple(o.shape), float(o.abs().max()))


def merge_case(name, seed, T, H, D):
    gen = torch.Generator().manual_seed(seed)
    oa, ob = torch.randn(T, H, D, generator=gen), torch.randn(T, H, D, generator=gen)
    la, lb = torch.randn(T, H, generator=gen) * 3, torch.randn(T, H, generator=gen) * 3
    la[0, 0] = float("inf")        # the reference reads +inf as an empty state
    lb[1, 1] = float("-inf")
    out, lse = merge_state_triton(oa, la, ob, lb)
    put(name, kind=np.int32(2), o_a=oa, lse_a=la, o_b=ob, lse_b=lb, out=out, lse=lse)
    print(name, tuple(out.shape))


# ----------------------------------------------------------------------------- torch_native with sliding windows
def torch_native_swa_case(name, seed, Hq, Hkv, D, prefix_lens, extend_lens, window, is_decode):
    from sglang.srt.layers.attention.torch_native_backend import TorchNativeAttnBackend
    from sglang.srt.layers.radix_attention import RadixAttention
    from sglang.srt.mem_cache.memory_pool import MHATokenToKVPool, ReqToTokenPool
    from sglang.srt.model_executor.forward_batch_info import ForwardBatch, ForwardMode
    from sglang.srt.model_executor.forward_context import ForwardContext, forward_context

    gen = torch.Generator().manual_seed(seed)
    bs = len(prefix_lens)
    in_lens = [1] * bs if is_decode else list(extend_lens)
    rtt, rpi, kp, vp = make_layout(gen, prefix_lens, in_lens, Hkv, D, 256)
    T = sum(in_lens)
    q = torch.randn(T, Hq, D, generator=gen)
    pool = MHATokenToKVPool(size=255, page_size=1, dtype=torch.float32, head_num=Hkv, head_dim=D, layer_num=1,
                            device="cpu", enable_memory_saver=False, enable_alt_stream=False)
    pool.k_buffer[0].copy_(kp)
    pool.v_buffer[0].copy_(vp)
    rtp = ReqToTokenPool(size=bs + 1, max_context_len=rtt.shape[1], device="cpu", enable_memory_saver=False)
    rtp.req_to_token[: bs + 1].copy_(rtt)
    runner = SimpleNamespace(device="cpu", req_to_token_pool=rtp, token_to_kv_pool=pool)
    backend = TorchNativeAttnBackend(runner)
    layer = RadixAttention(num_heads=Hq, head_dim=D, scaling=D ** -0.5, num_kv_heads=Hkv, layer_id=0,
                           sliding_window_size=window)
    seq_lens = [p + e for p, e in zip(prefix_lens, in_lens)]
    loc = torch.cat([rtt[r + 1, prefix_lens[r] : seq_lens[r]] for r in range(bs)]).long()
    fb = ForwardBatch(
        forward_mode=ForwardMode.DECODE if is_decode else ForwardMode.EXTEND, batch_size=bs,
        input_ids=torch.zeros(T, dtype=torch.int64), req_pool_indices=rpi, seq_lens=torch.tensor(seq_lens),
        out_cache_loc=loc, seq_lens_sum=sum(seq_lens), seq_lens_cpu=torch.tensor(seq_lens),
    )
    if not is_decode:
        fb.extend_prefix_lens = torch.tensor(prefix_lens)
        fb.extend_seq_lens = torch.tensor(in_lens)
        fb.extend_prefix_lens_cpu, fb.extend_seq_lens_cpu = list(prefix_lens), list(in_lens)
        fb.extend_num_tokens = T
    backend.init_forward_metadata(fb)
    k_new, v_new = kp[loc].clone(), vp[loc].clone()       # storing them again leaves the pool unchanged
    fn = backend.forward_decode if is_decode else backend.forward_extend
    o = fn(q.reshape(T, -1), k_new, v_new, layer, fb, save_kv_cache=True)
    put(name, kind=np.int32(1 if is_decode else 0), q=q, k_pool=kp, v_pool=vp, req_to_token=rtt,
        req_pool_indices=rpi, prefix_lens=np.array(seq_lens if is_decode else prefix_lens, np.int32),
        extend_lens=np.array([0] * bs if is_decode else in_lens, np.int32),
        opts=np.array([0.0, window + 1, 1, 1], np.float32), sinks=np.zeros(0, np.float32),
        custom_mask=np.zeros(0, np.uint8), mask_indptr=np.zeros(0, np.int64), out=o.view(T, Hq, D),
        lse=np.zeros(0, np.float32))
    print(name, tuple(o.shape), float(o.abs().max()))


def main():
    torch.set_grad_enabled(False)
    # Triton extend kernel (interpreter)
    triton_extend_case("tri_ext_plain", 1, 4, 2, 16, [5, 0, 9], [4, 6, 1])
    triton_extend_case("tri_ext_sinks", 2, 4, 2, 16, [5, 3], [4, 2], sinks=True)
    triton_extend_case("tri_ext_cap", 3, 4, 4, 16, [7, 0], [3, 5], logit_cap=2.5)
    triton_extend_case("tri_ext_window", 4, 4, 2, 16, [3, 9, 0], [5, 4, 12], window=4)
    triton_extend_case("tri_ext_window_sinks_cap", 5, 8, 2, 32, [20, 2], [9, 17], window=6, sinks=True, logit_cap=4.0)
    triton_extend_case("tri_ext_tree_mask", 6, 4, 2, 16, [6, 11], [5, 5], mask="tree")
    triton_extend_case("tri_ext_full_mask", 7, 4, 1, 16, [6, 3], [4, 7], mask="full", skip_prefix_mask=False)
    triton_extend_case("tri_ext_noncausal", 8, 4, 2, 16, [4, 0], [3, 6], causal=False)
    # Triton decode kernels (interpreter)
    triton_decode_case("tri_dec_plain_gqa", 11, 8, 2, 16, [37, 5, 64])
    triton_decode_case("tri_dec_sinks_mha", 12, 4, 4, 16, [21, 48], sinks=True)
    triton_decode_case("tri_dec_cap_sinks", 13, 8, 2, 32, [70, 9, 33], logit_cap=3.0, sinks=True)
    triton_decode_case("tri_dec_window", 14, 8, 4, 16, [50, 7, 16], window_keys=16)
    merge_case("merge_basic", 21, 5, 3, 16)
    # the CPU backend itself with sliding windows
    torch_native_swa_case("tn_swa_extend", 31, 4, 4, 16, [3, 4, 5, 0], [2, 2, 2, 7], 4, False)
    torch_native_swa_case("tn_swa_decode", 32, 4, 2, 16, [2, 4, 5, 30], None, 4, True)
    OUT["names"] = np.array(NAMES)
    np.savez_compressed(GOLDEN / "features.npz", **OUT)
    print("features.npz", len(NAMES), "cases")


if __name__ == "__main__":
    sys.exit(main())
