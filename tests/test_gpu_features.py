"""GPU parity for the options around the hot path: sliding window, attention sinks, tanh logit
cap, custom (tree) masks, LSE output, TARGET_VERIFY, merge_state -- all through the C ABI.

Oracle: oracle/torch_native_ref.py, pinned by tests/golden/features.npz to the reference's own Triton kernels
(CPU interpreter) and to TorchNativeAttnBackend with sliding windows.  Tolerance: max-abs <= 2e-2 (BASELINE)."""
from types import SimpleNamespace

import pytest
import torch

from helpers import ATOL, GoldenFeatures, make_runner, assert_close_scaled
from oracle import torch_native_ref as ref
from test_gpu_decode import _random_problem
from test_gpu_extend import make_extend_problem

pytestmark = pytest.mark.gpu

GF = GoldenFeatures()
KERNELS = {"generic": 1, "auto": 0}


def _decode(q, kp, vp, rtt, rpi, seq, scale, pool_dtype, *, kernel=0, window_keys=0, sinks=None, logit_cap=0.0,
            fixed_chunk=0, want_lse=False):
    from sglang_b200 import ops

    sm, _, _ = ops.device_info()
    bs, Hq, D = q.shape
    Hkv = kp.shape[1]
    ctas = sm * ops.decode_ctas_per_sm(D, pool_dtype) if D in (64, 128, 256) else sm * 2
    ws = ops.DecodeWorkspace.create(bs, Hq, vp.shape[2], "cuda", num_ctas=ctas)
    seq_g = seq.cuda()
    items = ops.decode_items_per_chunk(Hq, Hkv, D, vp.shape[2], q.dtype, pool_dtype)
    ops.decode_plan(ws, seq_g, Hkv, fixed_chunk_tokens=fixed_chunk, window_keys=window_keys, items_per_chunk=items)
    o = torch.empty(bs, Hq, vp.shape[2], dtype=q.dtype, device="cuda")
    lse = torch.full((bs, Hq), float("nan"), device="cuda") if want_lse else None
    ops.decode_attention(q.cuda(), o, kp.cuda(), vp.cuda(), rtt.cuda(), rpi.cuda(), seq_g, ws, scale,
                         kv_dtype=pool_dtype, logit_cap=logit_cap, kernel=kernel, window_keys=window_keys,
                         sinks=None if sinks is None else sinks.cuda(), lse=lse)
    torch.cuda.synchronize()
    return o.float().cpu(), (lse.cpu() if want_lse else None)


def _extend(q, kp, vp, rtt, rpi, pre, ext, scale, pool_dtype, *, kernel=0, causal=True, window_keys=0, sinks=None,
            logit_cap=0.0, custom_mask=None, mask_indptr=None, mask_skip_prefix=True, want_lse=False):
    from sglang_b200 import ops

    sm, _, _ = ops.device_info()
    T, Hq, D = q.shape
    Hkv = kp.shape[1]
    bq = ops.extend_block_q(D, vp.shape[2], Hq // Hkv, q.dtype, pool_dtype, kernel)
    ws = ops.ExtendWorkspace.create(len(pre), T // bq + len(pre) + 1, "cuda", num_ctas=sm)
    ops.extend_plan(ws, ext.cuda(), bq)
    o = torch.empty(T, Hq, vp.shape[2], dtype=q.dtype, device="cuda")
    lse = torch.full((T, Hq), float("nan"), device="cuda") if want_lse else None
    ops.extend_attention(q.cuda(), o, kp.cuda(), vp.cuda(), rtt.cuda(), rpi.cuda(), pre.cuda(), ext.cuda(), ws, scale,
                         kv_dtype=pool_dtype, kernel=kernel, causal=causal, lse=lse, window_keys=window_keys,
                         sinks=None if sinks is None else sinks.cuda(), logit_cap=logit_cap,
                         custom_mask=None if custom_mask is None else custom_mask.cuda(),
                         mask_indptr=None if mask_indptr is None else mask_indptr.cuda(),
                         mask_skip_prefix=mask_skip_prefix)
    torch.cuda.synchronize()
    return o.float().cpu(), (lse.cpu() if want_lse else None)


# ----------------------------------------------------------------------------- reference-generated cases
@pytest.mark.parametrize("name", GF.attn_names())
def test_golden_features(name):
    """The inputs the reference's Triton kernels / torch_native ran on (float32 there, bf16 here)."""
    c = GF.case(name)
    dt = torch.bfloat16
    q, kp, vp = c.q.to(dt), c.k_pool.to(dt), c.v_pool.to(dt)
    if c.is_decode:
        got, _ = _decode(q, kp, vp, c.req_to_token, c.req_pool_indices, c.prefix_lens, c.scale, dt,
                         window_keys=c.window_keys, sinks=c.sinks, logit_cap=c.logit_cap)
        want = ref.decode_ref(q, kp, vp, c.req_to_token, c.req_pool_indices, c.prefix_lens, c.scale,
                              logit_cap=c.logit_cap, window_keys=c.window_keys, sinks=c.sinks)
    else:
        got, lse = _extend(q, kp, vp, c.req_to_token, c.req_pool_indices, c.prefix_lens, c.extend_lens, c.scale, dt,
                           causal=c.causal, window_keys=c.window_keys, sinks=c.sinks, logit_cap=c.logit_cap,
                           custom_mask=c.custom_mask, mask_indptr=c.mask_indptr,
                           mask_skip_prefix=c.mask_skip_prefix, want_lse=True)
        want, want_lse = ref.extend_ref(q, kp, vp, c.req_to_token, c.req_pool_indices, c.prefix_lens, c.extend_lens,
                                        c.scale, logit_cap=c.logit_cap, causal=c.causal, window_keys=c.window_keys,
                                        sinks=c.sinks, custom_mask=c.custom_mask, mask_indptr=c.mask_indptr,
                                        mask_skip_prefix=c.mask_skip_prefix, want_lse=True)
        assert float((lse - want_lse).abs().max()) <= ATOL, name
    assert_close_scaled(got, want, name)
    # and against what the reference itself produced in float32 (the reference kit's own tolerance, 3e-2)
    assert float((got - c.out.view_as(got)).abs().max()) <= 3e-2, name


# ----------------------------------------------------------------------------- decode: fast kernels
DEC_SHAPES = [
    # bs, Hq, Hkv, D, lens, page
    (4, 32, 8, 128, [1, 100, 700, 3000], 1),      # head-group kernel
    (3, 8, 1, 128, [2500, 63, 640], 16),          # per-head kernel (70B TP8 shard)
    (3, 14, 2, 64, [515, 40, 1300], 1),           # Qwen2.5-0.5B heads
    (2, 64, 8, 64, [900, 130], 1),                # gpt-oss heads (group 8, D 64)
    (2, 4, 4, 80, [70, 333], 1),                  # generic path
]


@pytest.mark.parametrize("pool_dtype", [torch.bfloat16, torch.float8_e4m3fn])
@pytest.mark.parametrize("opts", [dict(window_keys=128), dict(window_keys=129, sinks=True), dict(logit_cap=30.0),
                                  dict(sinks=True), dict(window_keys=1000, logit_cap=50.0, sinks=True),
                                  dict(window_keys=1)],
                         ids=["win128", "win129_sinks", "cap30", "sinks", "win1000_cap_sinks", "win1"])
@pytest.mark.parametrize("shape", DEC_SHAPES, ids=[f"h{s[1]}_{s[2]}_d{s[3]}_p{s[5]}" for s in DEC_SHAPES])
def test_decode_options(shape, opts, pool_dtype):
    bs, Hq, Hkv, D, lens, page = shape
    if pool_dtype == torch.float8_e4m3fn and D == 80:
        pytest.skip("one generic FP8 case is enough")
    q, kp, vp, rtt, rpi, seq = _random_problem(bs, Hq, Hkv, D, lens, page, torch.bfloat16, pool_dtype, seed=7 + D)
    sinks = torch.randn(Hq, generator=torch.Generator().manual_seed(5)) * 2 if opts.get("sinks") else None
    kw = dict(window_keys=opts.get("window_keys", 0), logit_cap=opts.get("logit_cap", 0.0), sinks=sinks)
    scale = D ** -0.5
    want, want_lse = ref.decode_ref(q, kp, vp, rtt, rpi, seq, scale, kv_dtype=pool_dtype, want_lse=True, **kw)
    for kernel in (0, 1):
        got, lse = _decode(q, kp, vp, rtt, rpi, seq, scale, pool_dtype, kernel=kernel, want_lse=True, **kw)
        assert_close_scaled(got, want, f"kernel {kernel}")
        assert float((lse - want_lse).abs().max()) <= ATOL, f"kernel {kernel} lse"
    # the split must not matter (windowed ranges start mid-request: chunk starts are not 16-aligned)
    got64, lse64 = _decode(q, kp, vp, rtt, rpi, seq, scale, pool_dtype, fixed_chunk=64, want_lse=True, **kw)
    assert_close_scaled(got64, want)
    assert float((lse64 - want_lse).abs().max()) <= ATOL


# ----------------------------------------------------------------------------- extend: tcgen05 kernel + generic
EXT_SHAPES = [
    # Hq, Hkv, D, prefix, extend, page
    (32, 8, 128, [0, 700, 130], [300, 260, 1], 1),       # Llama-8B heads: several q tiles, windows crossing blocks
    (8, 1, 128, [1500], [600], 16),                      # group 8, long prefix: whole leading blocks are skipped
    (16, 16, 128, [5, 0], [250, 129], 1),                # MHA: 128 tokens per q tile
    (14, 2, 64, [600, 3], [200, 140], 1),                # D 64 (Qwen2.5-0.5B heads): the kHD = 64 instance with options
    (4, 2, 64, [40, 0], [33, 70], 1),                    # D 64, group 2, short sequences
]


@pytest.mark.parametrize("pool_dtype", [torch.bfloat16, torch.float8_e4m3fn])
@pytest.mark.parametrize("opts", [dict(window_keys=128), dict(window_keys=100, sinks=True), dict(logit_cap=30.0),
                                  dict(sinks=True), dict(window_keys=513, logit_cap=50.0, sinks=True),
                                  dict(window_keys=1)],
                         ids=["win128", "win100_sinks", "cap30", "sinks", "win513_cap_sinks", "win1"])
@pytest.mark.parametrize("shape", EXT_SHAPES, ids=[f"h{s[0]}_{s[1]}_d{s[2]}_p{s[5]}" for s in EXT_SHAPES])
def test_extend_options(shape, opts, pool_dtype):
    Hq, Hkv, D, prefix, extend, page = shape
    q, kp, vp, rtt, rpi, pre, ext = make_extend_problem(Hq, Hkv, D, prefix, extend, page, torch.bfloat16, pool_dtype,
                                                        seed=3 + D + Hq)
    sinks = torch.randn(Hq, generator=torch.Generator().manual_seed(9)) * 2 if opts.get("sinks") else None
    kw = dict(window_keys=opts.get("window_keys", 0), logit_cap=opts.get("logit_cap", 0.0), sinks=sinks)
    scale = D ** -0.5
    want, want_lse = ref.extend_ref(q, kp, vp, rtt, rpi, pre, ext, scale, kv_dtype=pool_dtype, want_lse=True, **kw)
    for kernel in (0, 1):
        got, lse = _extend(q, kp, vp, rtt, rpi, pre, ext, scale, pool_dtype, kernel=kernel, want_lse=True, **kw)
        assert_close_scaled(got, want, f"kernel {kernel}")
        assert float((lse - want_lse).abs().max()) <= ATOL, f"kernel {kernel} lse"


@pytest.mark.parametrize("pool_dtype", [torch.bfloat16, torch.float8_e4m3fn])
@pytest.mark.parametrize("Hq,Hkv,D", [(8, 2, 128), (14, 2, 64), (4, 4, 80)])
def test_extend_custom_mask_large(Hq, Hkv, D, pool_dtype):
    """Tree mask over 64 draft tokens on top of a 900-token prefix, both SKIP_PREFIX_CUSTOM_MASK settings, with and
    without a sliding window; D 128 / 64 run the tcgen05 kernel (kOpts instance) and the generic one, D 80 generic."""
    prefix, extend = [900, 17, 0], [64, 64, 150]
    q, kp, vp, rtt, rpi, pre, ext = make_extend_problem(Hq, Hkv, D, prefix, extend, 1, torch.bfloat16, pool_dtype, 21)
    g = torch.Generator().manual_seed(4)
    parts, offs = [], [0]
    for P, E in zip(prefix, extend):
        m = torch.rand(E, P + E, generator=g) < 0.5
        m[:, P:] &= torch.tril(torch.ones(E, E, dtype=torch.bool))
        m[torch.arange(E), P + torch.arange(E)] = True
        parts.append(m.reshape(-1))
        offs.append(offs[-1] + E * (P + E))
    mask, indptr = torch.cat(parts).to(torch.uint8), torch.tensor(offs, dtype=torch.int64)
    for skip in (True, False):
        for window in (0, 200):
            kw = dict(custom_mask=mask, mask_indptr=indptr, mask_skip_prefix=skip, window_keys=window)
            want, want_lse = ref.extend_ref(q, kp, vp, rtt, rpi, pre, ext, D ** -0.5, kv_dtype=pool_dtype,
                                            want_lse=True, **kw)
            for kernel in (0, 1):
                got, lse = _extend(q, kp, vp, rtt, rpi, pre, ext, D ** -0.5, pool_dtype, kernel=kernel,
                                   want_lse=True, **kw)
                assert_close_scaled(got, want, f"skip_prefix={skip} window={window} kernel={kernel}")
                assert float((lse - want_lse).abs().max()) <= ATOL, f"lse skip_prefix={skip} kernel={kernel}"


# ----------------------------------------------------------------------------- merge_state
def test_merge_state_golden_and_random():
    from sglang_b200 import ops

    c = GF.case("merge_basic")
    for dt in (torch.bfloat16, torch.float16):
        oa, ob = c.o_a.to(dt), c.o_b.to(dt)
        out, lse = ops.merge_state(oa.cuda(), c.lse_a.cuda(), ob.cuda(), c.lse_b.cuda())
        want, want_lse = ref.merge_state_ref(oa, c.lse_a, ob, c.lse_b)
        assert torch.allclose(out.float().cpu(), want, atol=2e-2, equal_nan=True)
        assert torch.allclose(lse.cpu(), want_lse, atol=1e-5, equal_nan=True)
        assert torch.allclose(lse.cpu(), c.lse, atol=1e-5, equal_nan=True)     # the reference kernel's own LSE
    g = torch.Generator().manual_seed(0)
    T, H, D = 777, 32, 128
    oa, ob = torch.randn(T, H, D, generator=g).bfloat16(), torch.randn(T, H, D, generator=g).bfloat16()
    la, lb = torch.randn(T, H, generator=g) * 4, torch.randn(T, H, generator=g) * 4
    big = torch.empty(T, H + 3, D, dtype=torch.bfloat16, device="cuda")       # strided output view
    out, lse = ops.merge_state(oa.cuda(), la.cuda(), ob.cuda(), lb.cuda(), out=big[:, 1 : H + 1])
    want, want_lse = ref.merge_state_ref(oa, la, ob, lb)
    assert_close_scaled(out.float().cpu(), want)
    assert float((lse.cpu() - want_lse).abs().max()) <= 1e-5


def test_chunked_prefix_equals_one_pass():
    """Chunked-prefix MHA (radix_attention.py:244-277): attention over the prefix and over the new tokens computed
    separately with LSE and merged == one pass over everything."""
    from sglang_b200 import ops

    Hq, Hkv, D = 32, 8, 128
    q, kp, vp, rtt, rpi, pre, ext = make_extend_problem(Hq, Hkv, D, [640, 100], [200, 77], 1, torch.bfloat16,
                                                        torch.bfloat16, seed=8)
    scale = D ** -0.5
    full, full_lse = _extend(q, kp, vp, rtt, rpi, pre, ext, scale, torch.bfloat16, want_lse=True)
    # (b) new tokens only: causal extend with zero prefix over a page table that starts at the new tokens
    zero = torch.zeros_like(pre)
    rtt_new = torch.zeros_like(rtt)
    for r in range(len(pre)):
        P, E = int(pre[r]), int(ext[r])
        rtt_new[r + 1, :E] = rtt[r + 1, P : P + E]
    o_b, lse_b = _extend(q, kp, vp, rtt_new, rpi, zero, ext, scale, torch.bfloat16, want_lse=True)
    # (a) prefix only: every new token attends the whole prefix (decode-style, with LSE)
    o_a = torch.empty_like(o_b)
    lse_a = torch.empty_like(lse_b)
    t0 = 0
    for r in range(len(pre)):
        P, E = int(pre[r]), int(ext[r])
        qq = q[t0 : t0 + E]
        oo, ll = _decode(qq, kp, vp, rtt[r + 1 : r + 2].expand(E, -1).contiguous(),
                         torch.arange(E), torch.full((E,), P, dtype=torch.int32), scale, torch.bfloat16, want_lse=True)
        o_a[t0 : t0 + E], lse_a[t0 : t0 + E] = oo, ll
        t0 += E
    out, lse = ops.merge_state(o_a.bfloat16().cuda(), lse_a.cuda(), o_b.bfloat16().cuda(), lse_b.cuda())
    assert_close_scaled(out.float().cpu(), full)
    assert float((lse.cpu() - full_lse).abs().max()) <= ATOL


# ----------------------------------------------------------------------------- through B200AttnBackend
def _backend_problem(Hq, Hkv, D, prefix, new_lens, *, sliding_window=None, seed=0, pool_dtype=torch.bfloat16):
    """Mirror pools + backend with a random page table; returns everything a forward needs."""
    from sglang_b200.backend import B200AttnBackend

    g = torch.Generator().manual_seed(seed)
    bs = len(prefix)
    lens = [p + e for p, e in zip(prefix, new_lens)]
    rows = sum(lens) + 8
    runner = make_runner(device="cuda", dtype=torch.bfloat16, pool_dtype=pool_dtype, Hq=Hq, Hkv=Hkv, D=D, page_size=1,
                         pool_rows=rows + 1, max_bs=bs + 2, max_context_len=max(lens) + 4, layer_num=2)
    runner.sliding_window_size = sliding_window
    perm = (torch.randperm(rows - 1, generator=g) + 1).to(torch.int32)
    rtt = torch.zeros(bs + 3, max(lens) + 4, dtype=torch.int32)
    used = 0
    for r, L in enumerate(lens):
        rtt[r + 1, :L] = perm[used : used + L]
        used += L
    runner.req_to_token_pool.req_to_token.copy_(rtt.cuda())
    pool = runner.token_to_kv_pool
    kp = torch.randn(rows + 1, Hkv, D, generator=g).bfloat16()
    vp = torch.randn(rows + 1, Hkv, D, generator=g).bfloat16()
    for layer in range(2):
        pool.k_buffer[layer].copy_(kp.cuda())
        pool.v_buffer[layer].copy_(vp.cuda())
    T = sum(new_lens)
    q = torch.randn(T, Hq, D, generator=g).bfloat16()
    k_new = torch.randn(T, Hkv, D, generator=g).bfloat16()
    v_new = torch.randn(T, Hkv, D, generator=g).bfloat16()
    loc = torch.cat([rtt[r + 1, prefix[r] : lens[r]] for r in range(bs)]).long()
    kp[loc], vp[loc] = k_new, v_new          # what the pool holds after the store
    return SimpleNamespace(backend=B200AttnBackend(runner), runner=runner, rtt=rtt, rpi=torch.arange(1, bs + 1),
                           q=q, k_new=k_new, v_new=v_new, loc=loc, kp=kp, vp=vp, lens=lens, bs=bs, T=T)


def test_backend_short_extend_plans_one_tile_items():
    """A few new tokens behind cached prefixes (multi-turn chat): the scheduler's host copy of the extend lengths tells
    the backend that one 128-row tile per request is enough; a longer request in the batch switches back."""
    from sglang_b200.srt_mirror import ForwardBatch, ForwardMode, RadixAttention

    Hq, Hkv, D = 32, 8, 128
    layer = RadixAttention(Hq, D, D ** -0.5, Hkv, 0)
    for prefix, ext, want_bq in (([900, 0, 130], [17, 32, 1], 32), ([900, 0, 130], [17, 33, 1], 64)):
        pb = _backend_problem(Hq, Hkv, D, prefix, ext, seed=sum(ext))
        fb = ForwardBatch(forward_mode=ForwardMode.EXTEND, batch_size=pb.bs, req_pool_indices=pb.rpi.cuda(),
                          seq_lens=torch.tensor(pb.lens, dtype=torch.int32, device="cuda"), out_cache_loc=pb.loc.cuda(),
                          seq_lens_sum=sum(pb.lens),
                          extend_prefix_lens=torch.tensor(prefix, dtype=torch.int32, device="cuda"),
                          extend_seq_lens=torch.tensor(ext, dtype=torch.int32, device="cuda"),
                          extend_prefix_lens_cpu=prefix, extend_seq_lens_cpu=ext, extend_num_tokens=sum(ext))
        pb.backend.init_forward_metadata(fb)
        assert pb.backend.extend_ws.block_q == want_bq
        out = layer(pb.q.cuda().reshape(pb.T, -1), pb.k_new.cuda().reshape(pb.T, -1), pb.v_new.cuda().reshape(pb.T, -1),
                    fb, pb.backend)
        torch.cuda.synchronize()
        want = ref.extend_ref(pb.q, pb.kp, pb.vp, pb.rtt, pb.rpi, torch.tensor(prefix, dtype=torch.int32),
                              torch.tensor(ext, dtype=torch.int32), D ** -0.5)
        assert_close_scaled(out.float().cpu().view_as(want), want, f"extend lens {ext}")


@pytest.mark.parametrize("Hq,Hkv,D", [(32, 8, 128), (64, 8, 64), (8, 2, 128), (16, 8, 256), (32, 32, 96), (16, 4, 80)])
def test_backend_hybrid_sliding_window(Hq, Hkv, D):
    """A hybrid model: layer 0 sliding window (W = 127 -> 128 keys) with sinks, layer 1 full attention; decode and
    extend through the plugin API, one metadata call per forward (two decode plans)."""
    from sglang_b200.srt_mirror import ForwardBatch, ForwardMode, RadixAttention

    W = 127
    sinks = torch.randn(Hq, generator=torch.Generator().manual_seed(2))
    scale = D ** -0.5
    layers = [RadixAttention(Hq, D, scale, Hkv, 0, sliding_window_size=W), RadixAttention(Hq, D, scale, Hkv, 1)]
    # ---- decode
    prefix = [700, 50, 127, 128, 2000]
    pb = _backend_problem(Hq, Hkv, D, prefix, [1] * 5, sliding_window=W, seed=D)
    fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=pb.bs, req_pool_indices=pb.rpi.cuda(),
                      seq_lens=torch.tensor(pb.lens, dtype=torch.int32, device="cuda"), out_cache_loc=pb.loc.cuda(),
                      seq_lens_sum=sum(pb.lens))
    pb.backend.init_forward_metadata(fb)
    seq = torch.tensor(pb.lens, dtype=torch.int32)
    for layer, kw, fkw in ((layers[0], dict(window_keys=W + 1, sinks=sinks), dict(sinks=sinks.cuda())),
                           (layers[1], {}, {})):
        out = layer(pb.q.cuda().reshape(pb.T, -1), pb.k_new.cuda().reshape(pb.T, -1), pb.v_new.cuda().reshape(pb.T, -1),
                    fb, pb.backend, **fkw)
        torch.cuda.synchronize()
        want = ref.decode_ref(pb.q, pb.kp, pb.vp, pb.rtt, pb.rpi, seq, scale, **kw)
        assert_close_scaled(out.float().cpu().view_as(want), want, f"decode layer {layer.layer_id}")
    # ---- extend
    prefix, ext = [300, 0, 130], [200, 129, 3]
    pb = _backend_problem(Hq, Hkv, D, prefix, ext, sliding_window=W, seed=D + 1)
    fb = ForwardBatch(forward_mode=ForwardMode.EXTEND, batch_size=pb.bs, req_pool_indices=pb.rpi.cuda(),
                      seq_lens=torch.tensor(pb.lens, dtype=torch.int32, device="cuda"), out_cache_loc=pb.loc.cuda(),
                      seq_lens_sum=sum(pb.lens),
                      extend_prefix_lens=torch.tensor(prefix, dtype=torch.int32, device="cuda"),
                      extend_seq_lens=torch.tensor(ext, dtype=torch.int32, device="cuda"),
                      extend_prefix_lens_cpu=prefix, extend_seq_lens_cpu=ext, extend_num_tokens=sum(ext))
    pb.backend.init_forward_metadata(fb)
    pre_t, ext_t = torch.tensor(prefix, dtype=torch.int32), torch.tensor(ext, dtype=torch.int32)
    for layer, kw, fkw in ((layers[0], dict(window_keys=W + 1, sinks=sinks), dict(sinks=sinks.cuda())),
                           (layers[1], {}, {})):
        out = layer(pb.q.cuda().reshape(pb.T, -1), pb.k_new.cuda().reshape(pb.T, -1), pb.v_new.cuda().reshape(pb.T, -1),
                    fb, pb.backend, **fkw)
        torch.cuda.synchronize()
        want = ref.extend_ref(pb.q, pb.kp, pb.vp, pb.rtt, pb.rpi, pre_t, ext_t, scale, **kw)
        assert_close_scaled(out.float().cpu().view_as(want), want, f"extend layer {layer.layer_id}")


def test_backend_target_verify_and_lse():
    """TARGET_VERIFY: draft_token_num tokens per request under a tree mask (triton_backend.py:807-866), and
    `mha_return_lse` -> (o, lse)."""
    from sglang_b200.srt_mirror import ForwardBatch, ForwardMode, RadixAttention

    Hq, Hkv, D, n = 32, 8, 128, 8
    prefix = [333, 40, 1025]
    pb = _backend_problem(Hq, Hkv, D, prefix, [n] * 3, seed=77)
    g = torch.Generator().manual_seed(6)
    parts = []
    for P in prefix:
        m = torch.zeros(n, P + n, dtype=torch.bool)
        m[:, :P] = True
        tri = torch.tril(torch.rand(n, n, generator=g) < 0.5)
        tri[torch.arange(n), torch.arange(n)] = True
        m[:, P:] = tri
        parts.append(m.reshape(-1))
    mask = torch.cat(parts)
    spec = SimpleNamespace(draft_token_num=n, custom_mask=mask.cuda())
    fb = ForwardBatch(forward_mode=ForwardMode.TARGET_VERIFY, batch_size=pb.bs, req_pool_indices=pb.rpi.cuda(),
                      seq_lens=torch.tensor(prefix, dtype=torch.int32, device="cuda"), out_cache_loc=pb.loc.cuda(),
                      seq_lens_sum=sum(prefix), spec_info=spec, mha_return_lse=True)
    pb.backend.init_forward_metadata(fb)
    assert pb.backend.extend_ws.block_q == 128 // (Hq // Hkv)      # 8 draft tokens per request: one-tile items
    layer = RadixAttention(Hq, D, D ** -0.5, Hkv, 0)
    out, lse = layer(pb.q.cuda().reshape(pb.T, -1), pb.k_new.cuda().reshape(pb.T, -1),
                     pb.v_new.cuda().reshape(pb.T, -1), fb, pb.backend)
    torch.cuda.synchronize()
    indptr = torch.tensor([0] + [n * (P + n) for P in prefix]).cumsum(0)
    want, want_lse = ref.extend_ref(pb.q, pb.kp, pb.vp, pb.rtt, pb.rpi, torch.tensor(prefix), torch.tensor([n] * 3),
                                    D ** -0.5, custom_mask=mask.to(torch.uint8), mask_indptr=indptr, want_lse=True)
    assert tuple(out.shape) == (pb.T, Hq, D) and tuple(lse.shape) == (pb.T, Hq)
    assert_close_scaled(out.float().cpu(), want)
    assert float((lse.cpu() - want_lse).abs().max()) <= ATOL


def test_backend_swa_decode_cuda_graph():
    """CUDA-graph capture of a hybrid SWA decode step at a padded batch size; replays with new lengths."""
    from sglang_b200.srt_mirror import RadixAttention

    Hq, Hkv, D, W = 32, 8, 128, 63
    scale = D ** -0.5
    prefix = [500, 64, 9, 1200]
    pb = _backend_problem(Hq, Hkv, D, prefix, [1] * 4, sliding_window=W, seed=5)
    be = pb.backend
    max_bs = 6
    be.init_cuda_graph_state(max_bs, max_bs)
    seq_buf = torch.ones(max_bs, dtype=torch.int32, device="cuda")
    rpi_buf = torch.zeros(max_bs, dtype=torch.int64, device="cuda")
    loc_buf = torch.zeros(max_bs, dtype=torch.int64, device="cuda")
    q_buf = torch.zeros(max_bs, Hq * D, dtype=torch.bfloat16, device="cuda")
    k_buf = torch.zeros(max_bs, Hkv * D, dtype=torch.bfloat16, device="cuda")
    v_buf = torch.zeros(max_bs, Hkv * D, dtype=torch.bfloat16, device="cuda")
    from sglang_b200.srt_mirror import ForwardMode

    view = SimpleNamespace(forward_mode=ForwardMode.DECODE, batch_size=max_bs, req_pool_indices=rpi_buf,
                           seq_lens=seq_buf, seq_lens_cpu=None, out_cache_loc=loc_buf, seq_lens_sum=max_bs,
                           spec_info=None)
    layers = [RadixAttention(Hq, D, scale, Hkv, 0, sliding_window_size=W), RadixAttention(Hq, D, scale, Hkv, 1)]
    be.init_forward_metadata_out_graph(view, in_capture=True)
    outs = []
    graph = torch.cuda.CUDAGraph()
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        be.init_forward_metadata_in_graph(view)
        for layer in layers:
            layer(q_buf, k_buf, v_buf, view, be)
    torch.cuda.current_stream().wait_stream(s)
    with torch.cuda.graph(graph):
        be.init_forward_metadata_in_graph(view)
        outs = [layer(q_buf, k_buf, v_buf, view, be) for layer in layers]
    for it in range(2):
        lens = [L - 3 * it for L in pb.lens]
        loc = torch.tensor([int(pb.rtt[r + 1, lens[r] - 1]) for r in range(pb.bs)])
        seq_buf.fill_(1); rpi_buf.zero_(); loc_buf.zero_()
        seq_buf[: pb.bs] = torch.tensor(lens, dtype=torch.int32)
        rpi_buf[: pb.bs] = pb.rpi
        loc_buf[: pb.bs] = loc
        q_buf.zero_(); q_buf[: pb.bs] = pb.q.reshape(pb.bs, -1)
        kp, vp = pb.kp.clone(), pb.vp.clone()
        k_buf[: pb.bs], v_buf[: pb.bs] = kp[loc].reshape(pb.bs, -1), vp[loc].reshape(pb.bs, -1)
        be.init_forward_metadata_out_graph(view)
        graph.replay()
        torch.cuda.synchronize()
        seq = torch.tensor(lens, dtype=torch.int32)
        for layer, out, kw in ((layers[0], outs[0], dict(window_keys=W + 1)), (layers[1], outs[1], {})):
            want = ref.decode_ref(pb.q, kp, vp, pb.rtt, pb.rpi, seq, scale, **kw)
            got = out[: pb.bs].float().cpu().view_as(want)
            assert_close_scaled(got, want, f"replay {it} layer {layer.layer_id}")
