"""Generate tests/golden/rope.npz by running the REFERENCE's rotary arithmetic on CPU (test infrastructure).

    PYTHONPATH=oracle/stubs:/root/reference/python python oracle/gen_golden_rope.py

What runs from /root/reference/python, unmodified: `apply_rotary_emb`
(srt/layers/rotary_embedding/utils.py:36-60).  `RotaryEmbedding.__init__` itself cannot be constructed
outside a server process (it reads the published runtime context, base.py:156), so the dozen lines of
`forward_native` around that call (base.py:247-275: index_select of the cache row, chunk into cos | sin,
rotary_dim split, concat) and the cache formula (base.py:143-180) are replayed here verbatim in spirit; the
arithmetic whose rounding the CUDA kernel must reproduce is the reference's own function.
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

REPO = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(REPO))
from sglang.srt.layers.rotary_embedding.utils import apply_rotary_emb  # noqa: E402  (the reference)

from oracle.rope_ref import cos_sin_cache  # noqa: E402


def bits(t):
    return t.detach().contiguous().view(torch.int16).numpy().view(np.uint16).copy()


def forward_native(positions, query, key, cache, head_size, rotary_dim, is_neox):
    positions = positions.flatten()
    num_tokens = positions.shape[0]
    cos_sin = cache.index_select(0, positions)
    cos, sin = cos_sin.chunk(2, dim=-1)
    out = []
    for x in (query, key):
        shape = x.shape
        x = x.view(num_tokens, -1, head_size)
        x_rot, x_pass = x[..., :rotary_dim], x[..., rotary_dim:]
        x_rot = apply_rotary_emb(x_rot, cos, sin, is_neox)
        out.append(torch.cat((x_rot, x_pass), dim=-1).reshape(shape))
    return out


def main():
    g = torch.Generator().manual_seed(20261104)
    out, names = {}, []
    for dname, dt in (("bf16", torch.bfloat16), ("f16", torch.float16)):
        for neox in (True, False):
            for D, rot in ((64, 64), (128, 128), (128, 64)):
                name = f"{dname}_{'neox' if neox else 'gptj'}_D{D}_rot{rot}"
                T, Hq, Hkv = 7, 4, 2
                # CPU reference: the cache is cast to the activation dtype at construction (base.py:100-101)
                cache = cos_sin_cache(rot, 512, 10000.0, torch.float32)
                pos = torch.tensor([0, 1, 2, 17, 255, 300, 511])
                q = (torch.randn(T, Hq * D, generator=g) * 3).to(dt)
                k = (torch.randn(T, Hkv * D, generator=g) * 3).to(dt)
                q2, k2 = forward_native(pos, q.clone(), k.clone(), cache.to(dt), D, rot, neox)
                out[f"{name}.meta"] = np.array([0 if dt == torch.bfloat16 else 1, int(neox), T, Hq, Hkv, D, rot])
                # only the cache rows the case touches are kept: tests use positions 0..T-1 into these rows
                out[f"{name}.pos"] = pos.numpy()
                out[f"{name}.cache_rows_f32"] = cache.index_select(0, pos).numpy()
                out[f"{name}.q"], out[f"{name}.k"] = bits(q), bits(k)
                out[f"{name}.q_out"], out[f"{name}.k_out"] = bits(q2), bits(k2)
                names.append(name)
    out["names"] = np.array(names)
    np.savez_compressed(REPO / "tests" / "golden" / "rope.npz", **out)
    print("rope.npz", len(names), "cases")


if __name__ == "__main__":
    main()
