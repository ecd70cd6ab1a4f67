"""ORACLE (test infrastructure) -- CPU restatement of the reference's rotary embedding, the step the
fused RoPE + KV-store kernel replaces.

Follows (reference python/sglang/...):
  srt/layers/rotary_embedding/base.py:143-180   _compute_inv_freq / _compute_cos_sin_cache
  srt/layers/rotary_embedding/base.py:234-275   forward_native (view, rotary_dim split, concat)
  srt/layers/rotary_embedding/utils.py:23-60    rotate pairing + apply_rotary_emb arithmetic
Pinned by tests/golden/rope.npz, which oracle/gen_golden_rope.py produced by running the reference's own
`apply_rotary_emb` on CPU (tests/test_oracle.py::test_rope_*).
"""
from __future__ import annotations

import torch


def cos_sin_cache(rotary_dim: int, max_position: int, base: float = 10000.0,
                  dtype: torch.dtype = torch.float32) -> torch.Tensor:
    """base.py:143-180: inv_freq = base ** -(arange(0, rot, 2) / rot); cache = cat(cos(t f), sin(t f))."""
    inv_freq = 1.0 / (base ** (torch.arange(0, rotary_dim, 2, dtype=torch.float) / rotary_dim))
    t = torch.arange(max_position, dtype=torch.float)
    freqs = torch.einsum("i,j -> ij", t, inv_freq)
    return torch.cat((freqs.cos(), freqs.sin()), dim=-1).to(dtype)


def _rotate(x: torch.Tensor, cos: torch.Tensor, sin: torch.Tensor, is_neox: bool) -> torch.Tensor:
    """utils.py:36-60 with x: [T, H, rot]; cos, sin: [T, rot/2] already in the arithmetic dtype."""
    cos, sin = cos.unsqueeze(-2), sin.unsqueeze(-2)
    if is_neox:
        x1, x2 = torch.chunk(x, 2, dim=-1)
    else:
        x1, x2 = x[..., ::2], x[..., 1::2]
    o1 = x1 * cos - x2 * sin
    o2 = x2 * cos + x1 * sin
    if is_neox:
        return torch.cat((o1, o2), dim=-1)
    return torch.stack((o1, o2), dim=-1).flatten(-2)


def rope_ref(positions: torch.Tensor, q: torch.Tensor, k: torch.Tensor, cache: torch.Tensor, head_dim: int,
             is_neox: bool, native_rounding: bool):
    """Returns (q_rot, k_rot) in the input dtype.

    native_rounding=True: forward_native -- the cache row is cast to the activation dtype and every product
    and sum is an op in that dtype (one rounding each).  False: the same formula evaluated in float32 with a
    single rounding at the end (what the reference's CUDA kernel does with its float32 cache)."""
    T = positions.shape[0]
    rot = cache.shape[1]
    cs = cache.index_select(0, positions.long())
    cos, sin = cs.chunk(2, dim=-1)
    outs = []
    for x in (q, k):
        shape = x.shape
        x3 = x.reshape(T, -1, head_dim)
        x_rot, x_pass = x3[..., :rot], x3[..., rot:]
        if native_rounding:
            r = _rotate(x_rot, cos.to(x.dtype), sin.to(x.dtype), is_neox)
        else:
            r = _rotate(x_rot.float(), cos.float(), sin.float(), is_neox).to(x.dtype)
        outs.append(torch.cat((r, x_pass), dim=-1).reshape(shape))
    return outs[0], outs[1]
