"""GPU parity: paged decode attention through B200AttnBackend / the C ABI.
Tolerance is BASELINE.json's: max-abs <= 2e-2 against the FP32-accumulated oracle."""
import pytest
import torch

from helpers import ATOL, GoldenAttention, run_backend_on_case, store_dtype, assert_close_scaled
from oracle import torch_native_ref as ref

pytestmark = pytest.mark.gpu

GA = GoldenAttention()
DECODE_CASES = [n for n in GA.names if GA.case(n).is_decode]
KERNELS = {"generic": 1, "auto": 0}


@pytest.mark.parametrize("kernel", list(KERNELS))
@pytest.mark.parametrize("name", DECODE_CASES)
def test_golden_decode(name, kernel):
    """Reference kit cases (12 x 3 slot layouts) + scaled BASELINE configs, run by the reference on CPU."""
    c = GA.case(name)
    out, pool = run_backend_on_case(c, decode_kernel=KERNELS[kernel])
    T = c.q.shape[0]
    got = out.float().cpu().view(T, c.Hq, c.D)
    assert_close_scaled(got, c.out_ref_fp32.view(T, c.Hq, c.D), name)
    # the rows this step stored: bit-exact against the reference pool
    loc = c.out_cache_loc
    assert torch.equal(pool.k_buffer[0][loc.cuda()].cpu().view(torch.uint8), c.k_rows_after.view(torch.uint8))
    assert torch.equal(pool.v_buffer[0][loc.cuda()].cpu().view(torch.uint8), c.v_rows_after.view(torch.uint8))


def _random_problem(bs, Hq, Hkv, D, lens, page_size, dtype, pool_dtype, seed, idx_dtype=torch.int64):
    g = torch.Generator().manual_seed(seed)
    n_pages = sum((L + page_size - 1) // page_size for L in lens) + 3
    rows = (n_pages + 1) * page_size
    perm = (torch.randperm(n_pages, generator=g) + 1).tolist()
    max_ctx = max(lens) + 1
    rtt = torch.zeros(bs + 1, max_ctx, dtype=torch.int32)
    for r, L in enumerate(lens):
        pos = 0
        while pos < L:
            pg = perm.pop()
            n = min(page_size, L - pos)
            rtt[r + 1, pos : pos + n] = torch.arange(pg * page_size, pg * page_size + n)
            pos += n
    q = torch.randn(bs, Hq, D, generator=g).to(dtype)
    kp = torch.randn(rows, Hkv, D, generator=g).to(dtype)
    vp = torch.randn(rows, Hkv, D, generator=g).to(dtype)
    if store_dtype(pool_dtype) == torch.uint8:
        kp, vp = kp.to(pool_dtype).view(torch.uint8), vp.to(pool_dtype).view(torch.uint8)
    return q, kp, vp, rtt, torch.arange(1, bs + 1, dtype=idx_dtype), torch.tensor(lens, dtype=torch.int32)


def _run_ops_decode(q, kp, vp, rtt, rpi, seq, scale, pool_dtype, kernel=0, fixed_chunk=0, k_scale=1.0, v_scale=1.0,
                    logit_cap=0.0):
    from sglang_b200 import ops

    sm, _, _ = ops.device_info()
    bs, Hq, D = q.shape
    ws = ops.DecodeWorkspace.create(bs, Hq, D, "cuda", num_ctas=sm * 2)
    seq_g = seq.cuda()
    ops.decode_plan(ws, seq_g, kp.shape[1], fixed_chunk_tokens=fixed_chunk)
    o = torch.empty_like(q, device="cuda")
    ops.decode_attention(q.cuda(), o, kp.cuda(), vp.cuda(), rtt.cuda(), rpi.cuda(), seq_g, ws, scale,
                         kv_dtype=pool_dtype, k_scale=k_scale, v_scale=v_scale, logit_cap=logit_cap, kernel=kernel)
    torch.cuda.synchronize()
    return o.float().cpu(), ws


SHAPES = [
    # bs, Hq, Hkv, D, lens, page
    (3, 32, 8, 128, [1, 17, 600], 1),          # Llama-8B heads, ragged incl. length 1
    (5, 8, 1, 128, [2048, 33, 700, 64, 65], 16),  # 70B TP8 shard
    (2, 14, 2, 64, [515, 575], 1),             # Qwen2.5-0.5B heads (group 7)
    (2, 16, 1, 128, [300, 129], 1),            # group 16 (upper MMA rows)
    (2, 4, 4, 256, [90, 257], 32),             # MHA, D 256
    (2, 4, 4, 80, [70, 33], 1),                # odd head dims -> generic path
    (2, 4, 2, 13, [40, 9], 1),
    (1, 64, 2, 64, [130], 1),                  # group 32 -> generic path
]


@pytest.mark.parametrize("kernel", list(KERNELS))
@pytest.mark.parametrize("dtype,pool_dtype", [(torch.bfloat16, torch.bfloat16), (torch.float16, torch.float16),
                                              (torch.bfloat16, torch.float8_e4m3fn)])
@pytest.mark.parametrize("shape", SHAPES, ids=[f"bs{s[0]}_h{s[1]}_{s[2]}_d{s[3]}_p{s[5]}" for s in SHAPES])
def test_decode_vs_oracle(shape, dtype, pool_dtype, kernel):
    """Configs follow the reference's Triton kernel test (B,Hq,Hkv,D) grid, test_triton_attention_kernels.py:563-573."""
    bs, Hq, Hkv, D, lens, page = shape
    q, kp, vp, rtt, rpi, seq = _random_problem(bs, Hq, Hkv, D, lens, page, dtype, pool_dtype, seed=D + bs)
    scale = D**-0.5
    got, _ = _run_ops_decode(q, kp, vp, rtt, rpi, seq, scale, pool_dtype, kernel=KERNELS[kernel])
    want = ref.decode_ref(q, kp, vp, rtt, rpi, seq, scale, kv_dtype=pool_dtype)
    assert_close_scaled(got, want)


@pytest.mark.parametrize("shape", [(3, 32, 8, 128, [1, 17, 600], 1), (5, 8, 1, 128, [2048, 33, 700, 64, 65], 16),
                                   (2, 16, 1, 128, [300, 129], 1)], ids=["hg_8B", "perhead_70Btp8", "group16"])
def test_decode_e5m2_fast_kernels(shape):
    """FP8-e5m2 pools (--kv-cache-dtype fp8_e5m2) on the tensor-core decode kernels (head-group and per-head)."""
    from sglang_b200 import _C

    bs, Hq, Hkv, D, lens, page = shape
    pool_dtype = torch.float8_e5m2
    assert _C.lib.b200attn_decode_fast_supported(D, D, Hq // Hkv, _C.B200_BF16, _C.B200_FP8_E5M2) == 1
    q, kp, vp, rtt, rpi, seq = _random_problem(bs, Hq, Hkv, D, lens, page, torch.bfloat16, pool_dtype, seed=7)
    want = ref.decode_ref(q, kp, vp, rtt, rpi, seq, D**-0.5, kv_dtype=pool_dtype)
    for kernel in (2, 0):                       # KERNEL_FAST (refuses shapes it does not cover), KERNEL_AUTO
        got, _ = _run_ops_decode(q, kp, vp, rtt, rpi, seq, D**-0.5, pool_dtype, kernel=kernel)
        assert_close_scaled(got, want, f"kernel {kernel}")
    gen, _ = _run_ops_decode(q, kp, vp, rtt, rpi, seq, D**-0.5, pool_dtype, kernel=KERNELS["generic"])
    assert float((got - gen).abs().max()) <= 2e-2


D64_FP8_SHAPES = [
    # bs, Hq, Hkv, D, lens, page
    (3, 64, 8, 64, [1, 17, 1300], 1),          # gpt-oss heads (group 8), ragged incl. length 1
    (2, 32, 8, 64, [515, 64], 16),             # Llama-3.2-1B heads (group 4), paged
    (2, 64, 4, 64, [300, 129], 1),             # group 16 (upper MMA rows)
    (4, 8, 8, 64, [33, 2, 700, 65], 1),        # MHA
]


@pytest.mark.parametrize("pool_dtype", [torch.float8_e4m3fn, torch.float8_e5m2], ids=["e4m3", "e5m2"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16], ids=["bf16", "f16"])
@pytest.mark.parametrize("shape", D64_FP8_SHAPES, ids=[f"bs{s[0]}_h{s[1]}_{s[2]}_p{s[5]}" for s in D64_FP8_SHAPES])
def test_decode_d64_fp8_head_group_kernel(shape, dtype, pool_dtype):
    """D = 64 FP8 pools (64-byte rows) on the tensor-core head-group kernel: half-warp token copies, swizzle across
    two heads of a 128-byte line.  KERNEL_FAST refuses shapes it has no instance for, so passing proves the kernel ran."""
    bs, Hq, Hkv, D, lens, page = shape
    q, kp, vp, rtt, rpi, seq = _random_problem(bs, Hq, Hkv, D, lens, page, dtype, pool_dtype, seed=3 * bs + Hkv)
    want = ref.decode_ref(q, kp, vp, rtt, rpi, seq, D**-0.5, kv_dtype=pool_dtype, k_scale=0.75, v_scale=1.5)
    for kernel, chunk in ((2, 0), (0, 0), (2, 64)):       # FAST, AUTO, FAST with 64-token chunks (merge of many partials)
        got, _ = _run_ops_decode(q, kp, vp, rtt, rpi, seq, D**-0.5, pool_dtype, kernel=kernel, fixed_chunk=chunk,
                                 k_scale=0.75, v_scale=1.5)
        assert_close_scaled(got, want, f"kernel {kernel} chunk {chunk}")
    gen, _ = _run_ops_decode(q, kp, vp, rtt, rpi, seq, D**-0.5, pool_dtype, kernel=KERNELS["generic"], k_scale=0.75,
                             v_scale=1.5)
    assert float((got - gen).abs().max()) <= 2e-2


ODD_D_SHAPES = [
    # bs, Hq, Hkv, D, lens, page
    (3, 32, 32, 96, [1, 17, 1300], 1),         # Phi-3-mini heads (MHA, D 96)
    (2, 32, 32, 80, [515, 64], 16),            # Phi-2 heads (D 80), paged
    (2, 16, 4, 112, [300, 129], 1),            # D 112, group 4
    (2, 64, 4, 96, [700, 33], 1),              # group 16 (upper MMA rows)
]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16], ids=["bf16", "f16"])
@pytest.mark.parametrize("shape", ODD_D_SHAPES, ids=[f"h{s[1]}_{s[2]}_d{s[3]}_p{s[5]}" for s in ODD_D_SHAPES])
def test_decode_d80_d96_d112_head_group_kernel(shape, dtype):
    """Head dims 80 / 96 / 112 (rows of 160 / 192 / 224 bytes) with 16-bit pools on the tensor-core head-group kernel:
    flat chunk gather, swizzle across heads, generic split merge.  KERNEL_FAST refuses shapes without an instance."""
    bs, Hq, Hkv, D, lens, page = shape
    q, kp, vp, rtt, rpi, seq = _random_problem(bs, Hq, Hkv, D, lens, page, dtype, dtype, seed=D + bs)
    want = ref.decode_ref(q, kp, vp, rtt, rpi, seq, D**-0.5, kv_dtype=dtype)
    for kernel, chunk in ((2, 0), (0, 0), (2, 64)):
        got, _ = _run_ops_decode(q, kp, vp, rtt, rpi, seq, D**-0.5, dtype, kernel=kernel, fixed_chunk=chunk)
        assert_close_scaled(got, want, f"kernel {kernel} chunk {chunk}")
    gen, _ = _run_ops_decode(q, kp, vp, rtt, rpi, seq, D**-0.5, dtype, kernel=KERNELS["generic"])
    assert float((got - gen).abs().max()) <= 2e-2


def test_decode_d64_fp8_needs_four_kv_heads_for_the_fast_path():
    """Hkv not a multiple of 4 (Qwen2.5-0.5B: 2 kv heads) keeps the CUDA-core kernel for D = 64 FP8 pools: AUTO works,
    FAST refuses loudly."""
    q, kp, vp, rtt, rpi, seq = _random_problem(2, 14, 2, 64, [515, 575], 1, torch.bfloat16, torch.float8_e4m3fn, seed=5)
    want = ref.decode_ref(q, kp, vp, rtt, rpi, seq, 64**-0.5, kv_dtype=torch.float8_e4m3fn)
    got, _ = _run_ops_decode(q, kp, vp, rtt, rpi, seq, 64**-0.5, torch.float8_e4m3fn, kernel=0)
    assert_close_scaled(got, want)
    with pytest.raises(Exception, match="fast kernel does not cover"):
        _run_ops_decode(q, kp, vp, rtt, rpi, seq, 64**-0.5, torch.float8_e4m3fn, kernel=2)


def test_decode_fp8_descale_and_logit_cap():
    q, kp, vp, rtt, rpi, seq = _random_problem(3, 8, 2, 128, [400, 3, 129], 1, torch.bfloat16,
                                               torch.float8_e4m3fn, seed=11)
    scale = 128**-0.5
    got, _ = _run_ops_decode(q, kp, vp, rtt, rpi, seq, scale, torch.float8_e4m3fn, k_scale=0.5, v_scale=1.75)
    want = ref.decode_ref(q, kp, vp, rtt, rpi, seq, scale, kv_dtype=torch.float8_e4m3fn, k_scale=0.5, v_scale=1.75)
    assert_close_scaled(got, want)
    got, _ = _run_ops_decode(q, kp, vp, rtt, rpi, seq, scale, torch.float8_e4m3fn, logit_cap=5.0)
    want = ref.decode_ref(q, kp, vp, rtt, rpi, seq, scale, kv_dtype=torch.float8_e4m3fn, logit_cap=5.0)
    assert_close_scaled(got, want)


@pytest.mark.parametrize("idx_dtype", [torch.int32, torch.int64])
def test_split_invariance_and_index_dtypes(idx_dtype):
    """The split-KV result must not depend on how the planner cuts the sequence (64-token vs one chunk)."""
    q, kp, vp, rtt, rpi, seq = _random_problem(4, 32, 8, 128, [1000, 64, 63, 2500], 1, torch.bfloat16,
                                               torch.bfloat16, seed=3, idx_dtype=idx_dtype)
    scale = 128**-0.5
    a, ws_a = _run_ops_decode(q, kp, vp, rtt, rpi, seq, scale, torch.bfloat16, fixed_chunk=64)
    b, ws_b = _run_ops_decode(q, kp, vp, rtt, rpi, seq, scale, torch.bfloat16, fixed_chunk=4096)
    assert int(ws_a.plan[0]) == 64 and int(ws_b.plan[0]) == 4096
    assert int(ws_a.plan[1]) == sum((L + 63) // 64 for L in [1000, 64, 63, 2500]) and int(ws_b.plan[1]) == 4
    assert float((a - b).abs().max()) <= 4e-3
    want = ref.decode_ref(q, kp, vp, rtt, rpi, seq, scale)
    assert_close_scaled(a, want)
