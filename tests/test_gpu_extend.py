"""GPU parity: prefix + causal extend attention through B200AttnBackend / the C ABI."""
import pytest
import torch

from helpers import ATOL, GoldenAttention, run_backend_on_case, store_dtype, assert_close_scaled
from oracle import torch_native_ref as ref

pytestmark = pytest.mark.gpu

GA = GoldenAttention()
EXTEND_CASES = [n for n in GA.names if not GA.case(n).is_decode]
KERNELS = {"generic": 1, "auto": 0}


@pytest.mark.parametrize("kernel", list(KERNELS))
@pytest.mark.parametrize("name", EXTEND_CASES)
def test_golden_extend(name, kernel):
    c = GA.case(name)
    out, pool = run_backend_on_case(c, extend_kernel=KERNELS[kernel])
    T = c.q.shape[0]
    got = out.float().cpu().view(T, c.Hq, c.D)
    assert_close_scaled(got, c.out_ref_fp32.view(T, c.Hq, c.D), name)
    loc = c.out_cache_loc
    assert torch.equal(pool.k_buffer[0][loc.cuda()].cpu().view(torch.uint8), c.k_rows_after.view(torch.uint8))
    assert torch.equal(pool.v_buffer[0][loc.cuda()].cpu().view(torch.uint8), c.v_rows_after.view(torch.uint8))


def make_extend_problem(Hq, Hkv, D, prefix, extend, page_size, dtype, pool_dtype, seed, shared_prefix=0):
    g = torch.Generator().manual_seed(seed)
    bs = len(prefix)
    lens = [p + e for p, e in zip(prefix, extend)]
    n_pages = sum((L + page_size - 1) // page_size for L in lens) + 3
    rows = (n_pages + 1) * page_size
    perm = (torch.randperm(n_pages, generator=g) + 1).tolist()
    rtt = torch.zeros(bs + 1, max(lens) + 1, dtype=torch.int32)
    shared = [perm.pop() for _ in range(shared_prefix // page_size)]
    for r, L in enumerate(lens):
        pos = 0
        for pg in shared:
            rtt[r + 1, pos : pos + page_size] = torch.arange(pg * page_size, (pg + 1) * page_size)
            pos += page_size
        while pos < L:
            pg = perm.pop()
            n = min(page_size, L - pos)
            rtt[r + 1, pos : pos + n] = torch.arange(pg * page_size, pg * page_size + n)
            pos += n
    T = sum(extend)
    q = torch.randn(T, Hq, D, generator=g).to(dtype)
    kp = torch.randn(rows, Hkv, D, generator=g).to(dtype)
    vp = torch.randn(rows, Hkv, D, generator=g).to(dtype)
    if store_dtype(pool_dtype) == torch.uint8:
        kp, vp = kp.to(pool_dtype).view(torch.uint8), vp.to(pool_dtype).view(torch.uint8)
    return q, kp, vp, rtt, torch.arange(1, bs + 1), torch.tensor(prefix, dtype=torch.int32), \
        torch.tensor(extend, dtype=torch.int32)


def run_ops_extend(q, kp, vp, rtt, rpi, pre, ext, scale, pool_dtype, kernel=0, causal=True, want_lse=False,
                   page_size=1, one_tile=False):
    from sglang_b200 import ops

    sm, _, _ = ops.device_info()
    T, Hq, D = q.shape
    group = Hq // kp.shape[1]
    bq = ops.extend_block_q(D, D, group, q.dtype, pool_dtype, kernel)
    if one_tile:                     # the tcgen05 kernel's one-tile instance (valid for any extend length)
        bq = ops.extend_block_q_one_tile(D, D, group, q.dtype, pool_dtype, kernel)
        assert bq > 0
    ws = ops.ExtendWorkspace.create(len(pre), T // bq + len(pre) + 1, "cuda", num_ctas=sm)
    ext_g, pre_g = ext.cuda(), pre.cuda()
    rtt_g, rpi_g = rtt.cuda(), rpi.cuda()
    ops.extend_plan(ws, ext_g, bq, page_size=page_size, req_to_token=rtt_g, req_pool_indices=rpi_g,
                    extend_prefix_lens=pre_g)
    o = torch.empty_like(q, device="cuda")
    lse = torch.empty(T, Hq, dtype=torch.float32, device="cuda") if want_lse else None
    ops.extend_attention(q.cuda(), o, kp.cuda(), vp.cuda(), rtt_g, rpi_g, pre_g, ext_g, ws, scale,
                         kv_dtype=pool_dtype, kernel=kernel, causal=causal, lse=lse, page_size=page_size)
    torch.cuda.synchronize()
    run_ops_extend.page_rows = int(ws.plan[4])     # rows per verified page box, 0 = row gather
    return o.float().cpu(), lse


SHAPES = [
    # Hq, Hkv, D, prefix, extend, page
    (32, 8, 128, [0, 130, 64], [257, 1, 128], 1),      # Llama-8B heads; no-prefix, single-token and exact tiles
    (8, 1, 128, [700], [300], 16),                     # 70B TP8 shard
    (14, 2, 64, [512, 512], [63, 5], 1),               # Qwen2.5-0.5B heads (group 7) -> kHD = 64 tcgen05 instance
    (16, 4, 64, [300, 0, 127], [260, 129, 1], 16),     # D 64, group 4: several q tiles and kv blocks, page 16
    (4, 4, 64, [200], [400], 1),                       # D 64 MHA: 128 tokens per q tile
    (4, 4, 80, [10, 0], [33, 17], 1),                  # D 80 -> padded onto the D 128 one-tile instance (16-bit pools)
    (4, 4, 256, [33], [70], 32),                       # D 256 MHA -> one-tile tcgen05 instance (16-bit and FP8 pools)
    (8, 2, 256, [300, 0, 127], [260, 129, 1], 1),      # D 256, group 4 (Gemma-like): several q tiles / kv blocks
    (16, 8, 256, [1000], [200], 16),                   # D 256, group 2, page 16
    (16, 1, 128, [100], [150], 1),                     # group 16
    # page-granular pools: K/V arrive by TMA page boxes (16-bit pools) -- ragged tails inside the last page,
    # requests shorter than one page, pages larger than a 128-row kv block
    (32, 8, 128, [0, 130, 64], [257, 1, 128], 64),
    (8, 1, 128, [1000, 3], [300, 5], 128),
    (8, 2, 128, [100, 700], [513, 20], 256),
    (8, 2, 64, [31, 500], [200, 77], 32),
]


@pytest.mark.parametrize("dtype,pool_dtype", [(torch.bfloat16, torch.bfloat16), (torch.float16, torch.float16),
                                              (torch.bfloat16, torch.float8_e4m3fn), (torch.float16, torch.float8_e5m2)])
@pytest.mark.parametrize("shape", SHAPES, ids=[f"h{s[0]}_{s[1]}_d{s[2]}_p{s[5]}" for s in SHAPES])
def test_extend_one_tile_instance_vs_oracle(shape, dtype, pool_dtype):
    """The one-tile instance (S double-buffered across kv blocks; the backend picks it for short extends) on every
    shape the tcgen05 kernel covers, incl. multi-tile requests, and bit-compared with the two-tile instance's LSE
    within float tolerance."""
    from sglang_b200 import ops

    Hq, Hkv, D, prefix, extend, page = shape
    if not ops.extend_fast_supported(D, D, Hq // Hkv, dtype, pool_dtype):
        pytest.skip("generic kernel shape")
    q, kp, vp, rtt, rpi, pre, ext = make_extend_problem(Hq, Hkv, D, prefix, extend, page, dtype, pool_dtype, seed=D + 1)
    scale = D**-0.5
    got, lse1 = run_ops_extend(q, kp, vp, rtt, rpi, pre, ext, scale, pool_dtype, page_size=page, one_tile=True,
                               want_lse=True)
    want = ref.extend_ref(q, kp, vp, rtt, rpi, pre, ext, scale, kv_dtype=pool_dtype)
    assert_close_scaled(got, want)
    got2, lse2 = run_ops_extend(q, kp, vp, rtt, rpi, pre, ext, scale, pool_dtype, page_size=page, want_lse=True)
    assert float((lse1 - lse2).abs().max()) <= 2e-3
    assert float((got - got2).abs().max()) <= 2e-2


@pytest.mark.parametrize("kernel", list(KERNELS))
@pytest.mark.parametrize("dtype,pool_dtype", [(torch.bfloat16, torch.bfloat16), (torch.float16, torch.float16),
                                              (torch.bfloat16, torch.float8_e4m3fn), (torch.float16, torch.float8_e5m2)])
@pytest.mark.parametrize("shape", SHAPES, ids=[f"h{s[0]}_{s[1]}_d{s[2]}_p{s[5]}" for s in SHAPES])
def test_extend_vs_oracle(shape, dtype, pool_dtype, kernel):
    Hq, Hkv, D, prefix, extend, page = shape
    q, kp, vp, rtt, rpi, pre, ext = make_extend_problem(Hq, Hkv, D, prefix, extend, page, dtype, pool_dtype, seed=D)
    scale = D**-0.5
    got, _ = run_ops_extend(q, kp, vp, rtt, rpi, pre, ext, scale, pool_dtype, kernel=KERNELS[kernel], page_size=page)
    want = ref.extend_ref(q, kp, vp, rtt, rpi, pre, ext, scale, kv_dtype=pool_dtype)
    assert_close_scaled(got, want)


@pytest.mark.parametrize("page", [16, 64, 256])
def test_page_box_fill_is_bit_identical_to_row_gather(page):
    """Same pool, same page table: the TMA page-box fill and the cp.async row gather feed the same bytes to the same
    MMAs, so the outputs are bit-equal -- including with NaN-free garbage in the unused tail of each last page."""
    q, kp, vp, rtt, rpi, pre, ext = make_extend_problem(32, 8, 128, [0, 777, 130], [300, 129, 1], page,
                                                        torch.bfloat16, torch.bfloat16, seed=page)
    scale = 128**-0.5
    a, _ = run_ops_extend(q, kp, vp, rtt, rpi, pre, ext, scale, torch.bfloat16, page_size=page)
    assert run_ops_extend.page_rows == min(page, 128)
    b, _ = run_ops_extend(q, kp, vp, rtt, rpi, pre, ext, scale, torch.bfloat16, page_size=1)
    assert run_ops_extend.page_rows == 0
    assert torch.equal(a, b)


def test_scattered_page_table_falls_back_to_row_gather():
    """req_to_token is token-granular: the reference kit's non_monotonic_extend layout permutes the new tokens'
    slots inside their pages (dense_attention.py:685-703).  The plan's page check must see that and the kernel
    must read row by row -- one swapped pair in one request is enough."""
    page = 16
    q, kp, vp, rtt, rpi, pre, ext = make_extend_problem(8, 2, 128, [40, 0], [200, 90], page, torch.bfloat16,
                                                        torch.bfloat16, seed=77)
    rtt = rtt.clone()
    rtt[2, [35, 36]] = rtt[2, [36, 35]]
    scale = 128**-0.5
    got, _ = run_ops_extend(q, kp, vp, rtt, rpi, pre, ext, scale, torch.bfloat16, page_size=page)
    assert run_ops_extend.page_rows == 0
    want = ref.extend_ref(q, kp, vp, rtt, rpi, pre, ext, scale)
    assert_close_scaled(got, want)


def test_fast_path_covers_d256_16bit_pools():
    from sglang_b200 import ops

    assert ops.extend_fast_supported(256, 256, 2, torch.bfloat16, torch.bfloat16)
    assert ops.extend_fast_supported(256, 256, 4, torch.float16, torch.float16)
    assert ops.extend_fast_supported(256, 256, 2, torch.bfloat16, torch.float8_e4m3fn)     # FP8 pools: same instance
    assert ops.extend_fast_supported(256, 256, 2, torch.float16, torch.float8_e5m2)
    assert ops.extend_block_q(256, 256, 2, torch.bfloat16, torch.bfloat16) == 64      # one 128-row tile per item


def test_fast_path_covers_d64_and_d128():
    from sglang_b200 import ops

    for D in (64, 128):
        for kv in (torch.bfloat16, torch.float8_e4m3fn):
            assert ops.extend_fast_supported(D, D, 4, torch.bfloat16, kv)
    assert ops.extend_fast_supported(80, 80, 4, torch.bfloat16, torch.bfloat16)           # padded onto the D 128 tile
    assert not ops.extend_fast_supported(80, 80, 4, torch.bfloat16, torch.float8_e4m3fn)  # ... 16-bit pools only
    assert not ops.extend_fast_supported(72, 72, 4, torch.bfloat16, torch.bfloat16)
    assert not ops.extend_fast_supported(128, 64, 4, torch.bfloat16, torch.bfloat16)


def test_extend_non_causal_and_lse():
    q, kp, vp, rtt, rpi, pre, ext = make_extend_problem(8, 2, 128, [40, 0], [50, 90], 1, torch.bfloat16,
                                                        torch.bfloat16, seed=21)
    scale = 128**-0.5
    got, _ = run_ops_extend(q, kp, vp, rtt, rpi, pre, ext, scale, torch.bfloat16, causal=False)
    want = ref.extend_ref(q, kp, vp, rtt, rpi, pre, ext, scale, causal=False)
    assert_close_scaled(got, want)
    # LSE (natural log) of the causal scores, as chunked-prefix merging consumes it
    _, lse = run_ops_extend(q, kp, vp, rtt, rpi, pre, ext, scale, torch.bfloat16, want_lse=True)
    t = 0
    for r in range(2):
        P, E = int(pre[r]), int(ext[r])
        tok = rtt[r + 1, : P + E].long()
        k = kp[tok].float().repeat_interleave(4, dim=1)
        for i in (0, E - 1):
            s = torch.einsum("hd,khd->hk", q[t + i].float(), k[: P + i + 1]) * scale
            assert torch.allclose(lse[t + i].cpu(), torch.logsumexp(s, dim=-1), atol=2e-2)
        t += E


def test_extend_equals_decode_on_last_token():
    """Consistency between the two paths: a 1-token extend is a decode step."""
    from test_gpu_decode import _run_ops_decode

    q, kp, vp, rtt, rpi, pre, ext = make_extend_problem(32, 8, 128, [333, 64], [1, 1], 1, torch.bfloat16,
                                                        torch.bfloat16, seed=4)
    scale = 128**-0.5
    a, _ = run_ops_extend(q, kp, vp, rtt, rpi, pre, ext, scale, torch.bfloat16)
    b, _ = _run_ops_decode(q, kp, vp, rtt, rpi, pre + ext, scale, torch.bfloat16)
    assert float((a - b).abs().max()) <= 8e-3
