"""Pins the CPU oracle (oracle/) to outputs of the reference itself (tests/golden/, made by
oracle/gen_golden.py from /root/reference).  Runs without a GPU."""
import numpy as np
import pytest
import torch

from helpers import DTYPES, GOLDEN, GoldenAttention, from_bits
from oracle import torch_native_ref as ref


@pytest.mark.parametrize("src,dst", [("bf16", "e4m3"), ("bf16", "e5m2"), ("f16", "e4m3"), ("f16", "e5m2")])
def test_fp8_cast_tables_exhaustive(src, dst):
    """All 65536 source bit patterns: the C restatement equals torch's cast byte for byte."""
    z = np.load(GOLDEN / "fp8_cast_tables.npz")
    sdt = {"bf16": torch.bfloat16, "f16": torch.float16}[src]
    ddt = {"e4m3": torch.float8_e4m3fn, "e5m2": torch.float8_e5m2}[dst]
    got = ref.cast_table(sdt, ddt)
    want = z[f"{src}_to_{dst}"]
    assert np.array_equal(got, want), f"{int((got != want).sum())} mismatching casts"


def test_fp8_cast_documented_edge_values():
    """SURVEY.md section 9 probe: [1000,448,449,464,465,480,-1000,inf,1e-4,2^-10] -> [nan,448,448,448,448,nan,nan,nan,0,0]."""
    x = torch.tensor([1000, 448, 449, 464, 465, 480, -1000, float("inf"), 1e-4, 2**-10]).to(torch.bfloat16)
    table = ref.cast_table(torch.bfloat16, torch.float8_e4m3fn)
    got = torch.from_numpy(table[x.view(torch.int16).numpy().view(np.uint16)]).view(torch.float8_e4m3fn).float()
    want = torch.tensor([float("nan"), 448, 448, 448, 448, float("nan"), float("nan"), float("nan"), 0, 0])
    assert torch.equal(torch.isnan(got), torch.isnan(want))
    assert torch.equal(got[~torch.isnan(got)], want[~torch.isnan(want)])


def _kv_case(z, name):
    meta = z[f"{name}.meta"]
    sdt, pdt = DTYPES[int(meta[0])], DTYPES[int(meta[1])]
    T, H, D, rows = (int(x) for x in meta[2:6])
    k = from_bits(z[f"{name}.k"], sdt).view(T, H, D)
    v = from_bits(z[f"{name}.v"], sdt).view(T, H, D)
    loc = torch.from_numpy(z[f"{name}.loc"].copy())
    ks, vs = (float(x) for x in z[f"{name}.scales"])
    return sdt, pdt, T, H, D, rows, k, v, loc, ks, vs


def test_store_kv_ref_matches_reference_pool():
    """C scatter + cast == MHATokenToKVPool.set_kv_buffer on CPU, bit for bit (whole pool)."""
    z = np.load(GOLDEN / "kv_store.npz")
    for name in (str(n) for n in z["names"]):
        sdt, pdt, T, H, D, rows, k, v, loc, ks, vs = _kv_case(z, name)
        sdtype = torch.uint8 if pdt in (torch.float8_e4m3fn, torch.float8_e5m2) else pdt
        kc = torch.zeros(rows, H, D, dtype=sdtype)
        vc = torch.zeros(rows, H, D, dtype=sdtype)
        ref.store_kv_ref(k, v, kc, vc, loc, cache_dtype=pdt, k_scale=ks, v_scale=vs)
        want_k = from_bits(z[f"{name}.k_cache"], pdt).view(rows, H, D)
        want_v = from_bits(z[f"{name}.v_cache"], pdt).view(rows, H, D)
        assert torch.equal(kc.view(torch.uint8), want_k.view(torch.uint8)), name
        assert torch.equal(vc.view(torch.uint8), want_v.view(torch.uint8)), name


def test_store_kv_ref_skip_and_bounds():
    k = torch.randn(4, 1, 8).to(torch.bfloat16)
    kc = torch.zeros(6, 1, 8, dtype=torch.bfloat16)
    vc = torch.zeros_like(kc)
    ref.store_kv_ref(k, k, kc, vc, torch.tensor([0, 3, 0, 5]), cache_dtype=torch.bfloat16, skip_index=0)
    assert torch.equal(kc[0], torch.zeros(1, 8, dtype=torch.bfloat16))  # reserved sink untouched
    assert torch.equal(kc[3], k[1]) and torch.equal(kc[5], k[3])
    with pytest.raises(IndexError):
        ref.store_kv_ref(k, k, kc, vc, torch.tensor([1, 2, 6, 3]), cache_dtype=torch.bfloat16)


GA = GoldenAttention()


def _oracle_on_case(c):
    """Oracle run on the golden inputs: store the new rows, then attend."""
    kc, vc = c.k_pool.clone(), c.v_pool.clone()
    ref.store_kv_ref(c.k_new, c.v_new, kc, vc, c.out_cache_loc, cache_dtype=c.pool_dtype)
    seq = torch.tensor(c.seq_lens)
    if c.is_decode:
        out = ref.decode_ref(c.q, kc, vc, c.req_to_token, c.req_pool_indices, seq, c.D**-0.5,
                             kv_dtype=c.pool_dtype)
    else:
        out = ref.extend_ref(c.q, kc, vc, c.req_to_token, c.req_pool_indices, torch.tensor(c.prefix_lens),
                             torch.tensor(c.extend_lens), c.D**-0.5, kv_dtype=c.pool_dtype)
    return out, kc, vc


@pytest.mark.parametrize("name", GA.names)
def test_attention_oracle_matches_reference(name):
    c = GA.case(name)
    out, kc, vc = _oracle_on_case(c)
    # stored rows: bit-exact against the reference pool
    loc = c.out_cache_loc
    assert torch.equal(kc[loc].view(torch.uint8), c.k_rows_after.view(torch.uint8))
    assert torch.equal(vc[loc].view(torch.uint8), c.v_rows_after.view(torch.uint8))
    T = c.q.shape[0]
    # the reference's own code run in float32 on the same pool contents: tight
    want32 = c.out_ref_fp32.view(T, c.Hq, c.D)
    assert torch.allclose(out, want32, atol=2e-5, rtol=1e-4), float((out - want32).abs().max())
    # the reference as it runs in production (model dtype, SDPA internals): dtype-rounding tolerance
    want = c.out_ref.float().view(T, c.Hq, c.D)
    assert float((out - want).abs().max()) <= 2e-2


# ----------------------------------------------------------------------------- rotary embedding
def _rope_names():
    return [str(n) for n in np.load(GOLDEN / "rope.npz")["names"]]


@pytest.mark.parametrize("name", _rope_names())
def test_rope_oracle_matches_reference_bit_exact(name):
    """oracle/rope_ref (native rounding) == the reference's apply_rotary_emb run on CPU (oracle/gen_golden_rope.py)."""
    from oracle import rope_ref

    z = np.load(GOLDEN / "rope.npz")
    dt_i, neox, T, Hq, Hkv, D, rot = (int(x) for x in z[f"{name}.meta"])
    dt = torch.bfloat16 if dt_i == 0 else torch.float16
    q = from_bits(z[f"{name}.q"], dt).view(T, Hq * D)
    k = from_bits(z[f"{name}.k"], dt).view(T, Hkv * D)
    cache = torch.from_numpy(z[f"{name}.cache_rows_f32"].copy())
    q2, k2 = rope_ref.rope_ref(torch.arange(T), q, k, cache, D, bool(neox), native_rounding=True)
    assert torch.equal(q2.view(torch.int16), from_bits(z[f"{name}.q_out"], dt).view(T, Hq * D).view(torch.int16))
    assert torch.equal(k2.view(torch.int16), from_bits(z[f"{name}.k_out"], dt).view(T, Hkv * D).view(torch.int16))
    # the float32 evaluation of the same formula differs only by the op-by-op roundings (|x| <= ~12 here)
    q3, _ = rope_ref.rope_ref(torch.arange(T), q, k, cache, D, bool(neox), native_rounding=False)
    assert torch.allclose(q3.float(), q2.float(), rtol=0, atol=0.13 if dt_i == 0 else 0.02)


def test_rope_cache_formula():
    from oracle import rope_ref

    c = rope_ref.cos_sin_cache(64, 32, 10000.0)
    assert c.shape == (32, 64) and torch.all(c[0, :32] == 1) and torch.all(c[0, 32:] == 0)
    assert abs(float(c[1, 0]) - np.cos(1.0)) < 1e-6 and abs(float(c[3, 32]) - np.sin(3.0)) < 1e-6


# ----------------------------------------------------------------------------- feature options
from helpers import GoldenFeatures  # noqa: E402

_FEAT = GoldenFeatures()


@pytest.mark.parametrize("name", _FEAT.attn_names())
def test_oracle_features_match_reference_kernels(name):
    """Sinks / logit cap / sliding window / custom masks / LSE: the oracle equals the reference's own Triton
    kernels (run by Triton's CPU interpreter) and TorchNativeAttnBackend with sliding windows, in float32."""
    c = _FEAT.case(name)
    if c.is_decode:
        got = ref.decode_ref(c.q, c.k_pool, c.v_pool, c.req_to_token, c.req_pool_indices, c.prefix_lens, c.scale,
                             logit_cap=c.logit_cap, window_keys=c.window_keys, sinks=c.sinks)
        lse = None
    else:
        got, lse = ref.extend_ref(c.q, c.k_pool, c.v_pool, c.req_to_token, c.req_pool_indices, c.prefix_lens,
                                  c.extend_lens, c.scale, logit_cap=c.logit_cap, causal=c.causal,
                                  window_keys=c.window_keys, sinks=c.sinks, custom_mask=c.custom_mask,
                                  mask_indptr=c.mask_indptr, mask_skip_prefix=c.mask_skip_prefix, want_lse=True)
    assert (got - c.out.view_as(got)).abs().max().item() < 2e-5, name
    if c.lse is not None:
        assert (lse - c.lse).abs().max().item() < 2e-5, name


def test_merge_state_ref_matches_reference_kernel():
    c = _FEAT.case("merge_basic")
    out, lse = ref.merge_state_ref(c.o_a, c.lse_a, c.o_b, c.lse_b)
    assert torch.allclose(out, c.out, atol=1e-6, equal_nan=True)
    assert torch.allclose(lse, c.lse, atol=1e-6, equal_nan=True)
