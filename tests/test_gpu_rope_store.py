"""GPU parity: fused rotary embedding + KV-store through the C ABI.

native rounding  -> bit-exact against the reference's apply_rotary_emb (tests/golden/rope.npz) and the oracle;
fp32 rounding    -> within one ulp of the model dtype of the oracle's float32 evaluation;
pool rows        -> bit-exact against the oracle's store of the rotated k (byte work)."""
import numpy as np
import pytest
import torch

from helpers import GOLDEN, from_bits, store_dtype
from oracle import rope_ref
from oracle import torch_native_ref as ref

pytestmark = pytest.mark.gpu


def _names():
    return [str(n) for n in np.load(GOLDEN / "rope.npz")["names"]]


@pytest.mark.parametrize("name", _names())
def test_rope_matches_reference_bit_exact(name):
    from sglang_b200 import ops

    z = np.load(GOLDEN / "rope.npz")
    dt_i, neox, T, Hq, Hkv, D, rot = (int(x) for x in z[f"{name}.meta"])
    dt = torch.bfloat16 if dt_i == 0 else torch.float16
    q = from_bits(z[f"{name}.q"], dt).view(T, Hq * D).cuda()
    k = from_bits(z[f"{name}.k"], dt).view(T, Hkv * D).cuda()
    cache = torch.from_numpy(z[f"{name}.cache_rows_f32"].copy()).cuda()
    pos = torch.arange(T, device="cuda")
    ops.rope_store_kv(pos, q, k, None, cache, D, is_neox=bool(neox), native_rounding=True)
    torch.cuda.synchronize()
    assert torch.equal(q.cpu().view(torch.int16), from_bits(z[f"{name}.q_out"], dt).view(T, Hq * D).view(torch.int16))
    assert torch.equal(k.cpu().view(torch.int16), from_bits(z[f"{name}.k_out"], dt).view(T, Hkv * D).view(torch.int16))


@pytest.mark.parametrize("dt", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("pool_dt,scale", [(None, None), (torch.float8_e4m3fn, None), (torch.float8_e4m3fn, 1.5),
                                           (torch.float8_e5m2, None)])
@pytest.mark.parametrize("neox,D,rot,cache_dt", [(True, 128, 128, torch.float32), (False, 128, 128, torch.float32),
                                                 (True, 128, 64, None), (True, 64, 64, torch.float32),
                                                 (False, 256, 128, None)])
@pytest.mark.parametrize("native", [True, False])
def test_rope_store_vs_oracle(dt, pool_dt, scale, neox, D, rot, cache_dt, native):
    from sglang_b200 import ops

    g = torch.Generator().manual_seed(D + rot + int(neox))
    T, Hq, Hkv, rows = 37, 8, 2, 64
    pdt = pool_dt or dt
    cache = rope_ref.cos_sin_cache(rot, 1024, 500000.0, cache_dt or dt)
    pos = torch.randint(0, 1024, (T,), generator=g)
    q = (torch.randn(T, Hq * D, generator=g) * 2).to(dt)
    k = (torch.randn(T, Hkv * D, generator=g) * 2).to(dt)
    v = (torch.randn(T, Hkv * D, generator=g) * 2).to(dt)
    loc = torch.randperm(rows - 1, generator=g)[:T] + 1
    loc[3] = 0                                                   # reserved sink slot: must not be written
    want_q, want_k = rope_ref.rope_ref(pos, q, k, cache, D, neox, native_rounding=native)

    qd, kd, vd = q.cuda(), k.cuda(), v.cuda()
    kc = torch.zeros(rows, Hkv, D, dtype=store_dtype(pdt), device="cuda")
    vc = torch.zeros_like(kc)
    ops.rope_store_kv(pos.cuda(), qd, kd, vd, cache.cuda(), D, is_neox=neox, k_cache=kc, v_cache=vc, loc=loc.cuda(),
                      cache_dtype=pdt, k_scale=scale, v_scale=scale, native_rounding=native)
    torch.cuda.synchronize()
    if native:
        assert torch.equal(qd.cpu().view(torch.int16), want_q.view(torch.int16))
        assert torch.equal(kd.cpu().view(torch.int16), want_k.view(torch.int16))
    else:
        ulp = 2.0 ** -7 if dt == torch.bfloat16 else 2.0 ** -10
        for got, want in ((qd, want_q), (kd, want_k)):
            err = (got.cpu().float() - want.float()).abs()
            assert bool((err <= ulp * want.float().abs().clamp_min(2.0 ** -6) + 1e-6).all()), float(err.max())
    # pool rows: the oracle's store applied to the k the kernel itself produced (byte work -> bit-exact)
    kc_ref = torch.zeros(rows, Hkv, D, dtype=store_dtype(pdt))
    vc_ref = torch.zeros_like(kc_ref)
    ref.store_kv_ref(kd.cpu().view(T, Hkv, D), v.view(T, Hkv, D), kc_ref, vc_ref, loc, cache_dtype=pdt,
                     k_scale=scale, v_scale=scale, skip_index=0)
    assert torch.equal(kc.cpu().view(torch.uint8), kc_ref.view(torch.uint8))
    assert torch.equal(vc.cpu().view(torch.uint8), vc_ref.view(torch.uint8))
    assert not bool(kc[0].any()) and not bool(vc[0].any())


def test_rope_store_rejects_bad_shapes():
    from sglang_b200 import _C, ops

    q = torch.zeros(4, 2 * 128, dtype=torch.bfloat16, device="cuda")
    k = torch.zeros(4, 1 * 128, dtype=torch.bfloat16, device="cuda")
    cache = torch.zeros(16, 120, device="cuda")                  # neox needs rotary_dim % 16 == 0
    with pytest.raises(_C.B200AttnError):
        ops.rope_store_kv(torch.arange(4, device="cuda"), q, k, None, cache, 128)
    with pytest.raises(ValueError):
        ops.rope_store_kv(torch.arange(4, device="cuda"), q, k, None, cache.double(), 128)


@pytest.mark.parametrize("T,Hq,Hkv,D,neox,pool_dt", [
    (5000, 32, 8, 128, True, None),                  # one warp per token (prefill regime), Llama-8B heads
    (5000, 32, 8, 128, True, torch.float8_e4m3fn),
    (600, 32, 8, 128, False, torch.float8_e4m3fn),   # a token shared by several warps, interleaved pairs
    (64, 32, 8, 128, True, None),                    # decode regime: one pass per warp
    (333, 14, 2, 64, True, None),                    # Qwen2.5-0.5B heads: partial passes (8 heads per pass)
    (97, 64, 8, 64, True, torch.float8_e4m3fn),      # gpt-oss heads
    (130, 8, 4, 256, False, None),                   # D 256: 2 heads per pass
    (130, 16, 16, 256, True, torch.float8_e5m2),
])
def test_rope_store_warp_per_token_shapes(T, Hq, Hkv, D, neox, pool_dt):
    """The full-rotary fast path at the shapes that decide its work split; native rounding -> bit-exact."""
    from sglang_b200 import ops

    g = torch.Generator().manual_seed(T + D)
    dt = torch.bfloat16
    pdt = pool_dt or dt
    rows = T + 9
    cache = rope_ref.cos_sin_cache(D, 4096, 10000.0, torch.float32)
    pos = torch.randint(0, 4096, (T,), generator=g)
    q = torch.randn(T, Hq * D, generator=g).to(dt)
    k = torch.randn(T, Hkv * D, generator=g).to(dt)
    v = torch.randn(T, Hkv * D, generator=g).to(dt)
    loc = torch.randperm(rows - 1, generator=g)[:T] + 1
    loc[T // 2] = 0
    want_q, want_k = rope_ref.rope_ref(pos, q, k, cache, D, neox, native_rounding=True)
    for write_k in (True, False):
        qd, kd, vd = q.cuda(), k.cuda(), v.cuda()
        kc = torch.zeros(rows, Hkv, D, dtype=store_dtype(pdt), device="cuda")
        vc = torch.zeros_like(kc)
        ops.rope_store_kv(pos.cuda(), qd, kd, vd, cache.cuda(), D, is_neox=neox, k_cache=kc, v_cache=vc,
                          loc=loc.cuda(), cache_dtype=pdt, native_rounding=True, write_k=write_k)
        torch.cuda.synchronize()
        assert torch.equal(qd.cpu().view(torch.int16), want_q.view(torch.int16))
        assert torch.equal(kd.cpu().view(torch.int16), (want_k if write_k else k).view(torch.int16))
        assert torch.equal(vd.cpu().view(torch.int16), v.view(torch.int16))
        kc_ref = torch.zeros(rows, Hkv, D, dtype=store_dtype(pdt))
        vc_ref = torch.zeros_like(kc_ref)
        ref.store_kv_ref(want_k.view(T, Hkv, D), v.view(T, Hkv, D), kc_ref, vc_ref, loc, cache_dtype=pdt,
                         skip_index=0)
        assert torch.equal(kc.cpu().view(torch.uint8), kc_ref.view(torch.uint8))
        assert torch.equal(vc.cpu().view(torch.uint8), vc_ref.view(torch.uint8))
    # rotate only (no pool): q and k still rotated
    qd, kd = q.cuda(), k.cuda()
    ops.rope_store_kv(pos.cuda(), qd, kd, None, cache.cuda(), D, is_neox=neox, native_rounding=True)
    torch.cuda.synchronize()
    assert torch.equal(qd.cpu().view(torch.int16), want_q.view(torch.int16))
    assert torch.equal(kd.cpu().view(torch.int16), want_k.view(torch.int16))
