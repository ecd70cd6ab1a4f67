"""GPU parity: KV-store scatter through the C ABI vs the reference pool (golden) and the oracle.
Bit-exact (byte work), slot 0 excluded where the reserved-slot skip applies."""
import numpy as np
import pytest
import torch

from helpers import DTYPES, GOLDEN, from_bits, store_dtype
from oracle import torch_native_ref as ref

pytestmark = pytest.mark.gpu


def _names():
    return [str(n) for n in np.load(GOLDEN / "kv_store.npz")["names"]]


@pytest.mark.parametrize("name", _names())
def test_store_matches_reference_pool_bit_exact(name):
    from sglang_b200 import ops

    z = np.load(GOLDEN / "kv_store.npz")
    meta = z[f"{name}.meta"]
    sdt, pdt = DTYPES[int(meta[0])], DTYPES[int(meta[1])]
    T, H, D, rows = (int(x) for x in meta[2:6])
    k = from_bits(z[f"{name}.k"], sdt).view(T, H, D).cuda()
    v = from_bits(z[f"{name}.v"], sdt).view(T, H, D).cuda()
    loc = torch.from_numpy(z[f"{name}.loc"].copy()).cuda()
    ks, vs = (float(x) for x in z[f"{name}.scales"])
    kc = torch.zeros(rows, H, D, dtype=store_dtype(pdt), device="cuda")
    vc = torch.zeros_like(kc)
    ops.store_kv(k, v, kc, vc, loc, cache_dtype=pdt, k_scale=ks, v_scale=vs)
    torch.cuda.synchronize()
    want_k = from_bits(z[f"{name}.k_cache"], pdt).view(rows, H, D)
    want_v = from_bits(z[f"{name}.v_cache"], pdt).view(rows, H, D)
    assert torch.equal(kc.cpu().view(torch.uint8), want_k.view(torch.uint8)), name
    assert torch.equal(vc.cpu().view(torch.uint8), want_v.view(torch.uint8)), name


@pytest.mark.parametrize("src,dst", [(torch.bfloat16, torch.float8_e4m3fn), (torch.float16, torch.float8_e4m3fn),
                                     (torch.bfloat16, torch.float8_e5m2), (torch.float16, torch.float8_e5m2)])
def test_fp8_cast_exhaustive_on_device(src, dst):
    """All 65536 source patterns through the CUDA kernel equal torch's (= the oracle's) cast."""
    from sglang_b200 import ops

    pat = torch.arange(65536, dtype=torch.int32).to(torch.int16).view(src).view(512, 128).cuda()
    kc = torch.zeros(513, 128, dtype=torch.uint8, device="cuda")
    vc = torch.zeros_like(kc)
    loc = torch.arange(1, 513, device="cuda")
    ops.store_kv(pat, pat, kc, vc, loc, cache_dtype=dst)
    want = torch.from_numpy(ref.cast_table(src, dst)).view(512, 128)
    assert torch.equal(kc[1:].cpu(), want)
    assert torch.equal(vc[1:].cpu(), want)


@pytest.mark.parametrize("T,H,D", [(1, 8, 128), (64, 8, 128), (257, 1, 128), (16399, 2, 64), (33, 3, 80), (5, 1, 13)])
@pytest.mark.parametrize("idx_dtype", [torch.int32, torch.int64])
def test_store_random_shapes_and_slot0_sink(T, H, D, idx_dtype):
    """Sweep modelled on the reference's test_store_cache.py:31-45, 91-120 (bs 1..16399, int32/int64
    indices, reserved index 0 skipped)."""
    from sglang_b200 import ops

    g = torch.Generator().manual_seed(T * 131 + D)
    rows = T + 7
    k = torch.randn(T, H, D, generator=g).to(torch.bfloat16)
    v = torch.randn(T, H, D, generator=g).to(torch.bfloat16)
    loc = (torch.randperm(rows - 1, generator=g)[:T] + 1).to(idx_dtype)
    if T >= 5:
        loc[::5] = 0  # padded rows point at the sink
    kc = torch.full((rows, H, D), 7.0, dtype=torch.bfloat16)
    vc = torch.full((rows, H, D), 7.0, dtype=torch.bfloat16)
    kc_g, vc_g = kc.cuda(), vc.cuda()
    ops.store_kv(k.cuda(), v.cuda(), kc_g, vc_g, loc.cuda())
    ref.store_kv_ref(k, v, kc, vc, loc, cache_dtype=torch.bfloat16, skip_index=0)
    assert torch.equal(kc_g.cpu().view(torch.int16), kc.view(torch.int16))
    assert torch.equal(vc_g.cpu().view(torch.int16), vc.view(torch.int16))
    assert torch.all(kc_g[0] == 7.0)  # sink untouched


def test_store_strided_source_and_asymmetric_v():
    """K/V are slices of a fused qkv projection in the model (non-contiguous rows); V may be narrower."""
    from sglang_b200 import ops

    g = torch.Generator().manual_seed(5)
    T, H, Dk, Dv = 40, 2, 128, 64
    qkv = torch.randn(T, H * (Dk + Dv) + 64, generator=g).to(torch.bfloat16).cuda()
    k = qkv[:, 64 : 64 + H * Dk].view(T, H, Dk)
    v = qkv[:, 64 + H * Dk :].view(T, H, Dv)
    kc = torch.zeros(64, H, Dk, dtype=torch.bfloat16, device="cuda")
    vc = torch.zeros(64, H, Dv, dtype=torch.bfloat16, device="cuda")
    loc = (torch.randperm(63, generator=g)[:T] + 1).cuda()
    ops.store_kv(k, v, kc, vc, loc)
    assert torch.equal(kc[loc], k) and torch.equal(vc[loc], v)


def test_full_size_extend_store_roundtrip():
    """cfg5-sized store (131072 rows x 2 KiB): scatter then gather is the identity; untouched rows stay zero."""
    from sglang_b200 import ops

    T, H, D = 131072, 8, 128
    g = torch.Generator(device="cuda").manual_seed(0)
    k = torch.randn(T, H, D, generator=g, device="cuda", dtype=torch.bfloat16)
    v = torch.randn(T, H, D, generator=g, device="cuda", dtype=torch.bfloat16)
    rows = T + 4096
    kc = torch.zeros(rows, H, D, dtype=torch.bfloat16, device="cuda")
    vc = torch.zeros_like(kc)
    loc = torch.randperm(rows - 1, device="cuda", generator=g)[:T] + 1
    ops.store_kv(k, v, kc, vc, loc)
    assert torch.equal(kc[loc], k) and torch.equal(vc[loc], v)
    mask = torch.ones(rows, dtype=torch.bool, device="cuda")
    mask[loc] = False
    assert int(kc[mask].abs().sum()) == 0
