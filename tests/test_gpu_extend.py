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


def run_ops_extend(q, kp, vp, rtt, rpi, pre, ext, scale, pool_dtype, kernel=0, causal=True, want_lse=False):
    """Runs the kernel the planner would pick for a short forward (one-CTA tcgen05 / generic) and -- where the
    CTA-pair kernel covers the shape (D = 128, 16-bit pool) -- the pair kernel too; both must agree, the pair
    kernel's output is returned."""
    from sglang_b200 import ops

    sm, _, _ = ops.device_info()
    T, Hq, D = q.shape
    group = Hq // kp.shape[1]
    ext_g, pre_g = ext.cuda(), pre.cuda()
    qg, kg, vg, rg, ig = q.cuda(), kp.cuda(), vp.cuda(), rtt.cuda(), rpi.cuda()

    def run(bq):
        ws = ops.ExtendWorkspace.create(len(pre), T // bq + len(pre) + 1, "cuda", num_ctas=sm)
        ops.extend_plan(ws, ext_g, bq)
        o = torch.empty_like(q, device="cuda")
        lse = torch.empty(T, Hq, dtype=torch.float32, device="cuda") if want_lse else None
        ops.extend_attention(qg, o, kg, vg, rg, ig, pre_g, ext_g, ws, scale, kv_dtype=pool_dtype, kernel=kernel,
                             causal=causal, lse=lse)
        torch.cuda.synchronize()
        return o.float().cpu(), lse

    bq1 = ops.extend_block_q(D, D, group, q.dtype, pool_dtype, kernel)
    bq2 = ops.extend_block_q(D, D, group, q.dtype, pool_dtype, kernel, num_tokens=1 << 30, batch_size=1)
    o, lse = run(bq1)
    if bq2 != bq1:
        o2, lse2 = run(bq2)
        assert float((o2 - o).abs().max()) <= 2e-3, "CTA-pair kernel disagrees with the one-CTA kernel"
        return o2, lse2
    return o, lse


SHAPES = [
    # Hq, Hkv, D, prefix, extend, page
    (32, 8, 128, [0, 130, 64], [257, 1, 128], 1),      # Llama-8B heads; no-prefix, single-token and exact tiles
    (8, 1, 128, [700], [300], 16),                     # 70B TP8 shard
    (14, 2, 64, [512, 512], [63, 5], 1),               # Qwen2.5-0.5B heads (group 7) -> kHD = 64 tcgen05 instance
    (16, 4, 64, [300, 0, 127], [260, 129, 1], 16),     # D 64, group 4: several q tiles and kv blocks, page 16
    (4, 4, 64, [200], [400], 1),                       # D 64 MHA: 128 tokens per q tile
    (4, 4, 80, [10, 0], [33, 17], 1),                  # odd head dim -> generic
    (4, 4, 256, [33], [70], 32),
    (16, 1, 128, [100], [150], 1),                     # group 16
]


@pytest.mark.parametrize("kernel", list(KERNELS))
@pytest.mark.parametrize("dtype,pool_dtype", [(torch.bfloat16, torch.bfloat16), (torch.float16, torch.float16),
                                              (torch.bfloat16, torch.float8_e4m3fn), (torch.float16, torch.float8_e5m2)])
@pytest.mark.parametrize("shape", SHAPES, ids=[f"h{s[0]}_{s[1]}_d{s[2]}_p{s[5]}" for s in SHAPES])
def test_extend_vs_oracle(shape, dtype, pool_dtype, kernel):
    Hq, Hkv, D, prefix, extend, page = shape
    q, kp, vp, rtt, rpi, pre, ext = make_extend_problem(Hq, Hkv, D, prefix, extend, page, dtype, pool_dtype, seed=D)
    scale = D**-0.5
    got, _ = run_ops_extend(q, kp, vp, rtt, rpi, pre, ext, scale, pool_dtype, kernel=KERNELS[kernel])
    want = ref.extend_ref(q, kp, vp, rtt, rpi, pre, ext, scale, kv_dtype=pool_dtype)
    assert_close_scaled(got, want)


def test_fast_path_covers_d64_and_d128():
    from sglang_b200 import ops

    for D in (64, 128):
        for kv in (torch.bfloat16, torch.float8_e4m3fn):
            assert ops.extend_fast_supported(D, D, 4, torch.bfloat16, kv)
    assert not ops.extend_fast_supported(80, 80, 4, torch.bfloat16, torch.bfloat16)
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
