"""GPU parity: split-KV of the one-tile extend items (speculative verify / draft extend / a few new tokens behind a long
prefix) -- every kv range cut into parts that run as separate CTAs, merged by a second launch.
Tolerance is BASELINE.json's (max-abs <= 2e-2 vs the FP32-accumulated oracle) plus the scale-aware bound."""
import pytest
import torch

from helpers import ATOL, assert_close_scaled
from oracle import torch_native_ref as ref
from test_gpu_extend import make_extend_problem

pytestmark = pytest.mark.gpu


def _run(q, kp, vp, rtt, rpi, pre, ext, scale, pool_dtype, splits, *, page_size=1, want_lse=True, poison=None,
         tiles_hint=0, **kw):
    from sglang_b200 import ops

    sm, _, _ = ops.device_info()
    T, Hq, D = q.shape
    group = Hq // kp.shape[1]
    bq = ops.extend_block_q_one_tile(D, D, group, q.dtype, pool_dtype)
    assert bq > 0
    ws = ops.ExtendWorkspace.create(len(pre), T // bq + len(pre) + 1, "cuda", num_ctas=sm)
    ext_g, pre_g, rtt_g, rpi_g = ext.cuda(), pre.cuda(), rtt.cuda(), rpi.cuda()
    ops.extend_plan(ws, ext_g, bq, page_size=page_size, req_to_token=rtt_g, req_pool_indices=rpi_g,
                    extend_prefix_lens=pre_g)
    o = torch.full_like(q, float("nan"), device="cuda")
    lse = torch.full((T, Hq), float("nan"), device="cuda") if want_lse else None
    split_ws = None
    if splits > 1:
        # stale workspace contents must never leak into a result: fill it with NaN (or a caller-chosen poison)
        split_ws = torch.full((ops.extend_split_ws_floats(splits, T, Hq, D),), float("nan") if poison is None else poison,
                              dtype=torch.float32, device="cuda")
    for name in ("sinks", "custom_mask", "mask_indptr"):
        if kw.get(name) is not None:
            kw[name] = kw[name].cuda()
    ops.extend_attention(q.cuda(), o, kp.cuda(), vp.cuda(), rtt_g, rpi_g, pre_g, ext_g, ws, scale, kv_dtype=pool_dtype,
                         lse=lse, page_size=page_size, num_splits=splits, split_ws=split_ws, num_tiles_hint=tiles_hint,
                         **kw)
    torch.cuda.synchronize()
    return o.float().cpu(), (lse.cpu() if want_lse else None)


SHAPES = [
    # Hq, Hkv, D, prefix, extend, page
    (32, 8, 128, [4096, 700, 0, 130, 2500], [8, 8, 8, 8, 8], 1),       # verify-shaped: 8 draft tokens, ragged prefixes
    (32, 8, 128, [3000, 513], [32, 17], 16),                          # a full tile of 32 tokens / a ragged one, paged
    (8, 1, 128, [5000], [16], 1),                                     # 70B TP8 shard: one item, deep split
    (16, 16, 128, [1500, 2], [128, 100], 1),                          # MHA: 128 tokens per tile
    (14, 2, 64, [2100, 40], [18, 5], 1),                              # D 64, group 7 (126 of 128 MMA rows)
    (16, 8, 256, [2000, 300], [40, 64], 1),                           # D 256 instance (64-row kv blocks)
    (32, 8, 128, [1200, 0], [300, 70], 1),                            # multi-tile requests: every tile is split
]


@pytest.mark.parametrize("splits", [2, 3, 16])
@pytest.mark.parametrize("dtype,pool_dtype", [(torch.bfloat16, torch.bfloat16), (torch.float16, torch.float16),
                                              (torch.bfloat16, torch.float8_e4m3fn)])
@pytest.mark.parametrize("shape", SHAPES, ids=[f"h{s[0]}_{s[1]}_d{s[2]}_p{s[5]}_e{s[4][0]}" for s in SHAPES])
def test_split_kv_extend_vs_oracle(shape, dtype, pool_dtype, splits):
    from sglang_b200 import ops

    Hq, Hkv, D, prefix, extend, page = shape
    if not ops.extend_fast_supported(D, D, Hq // Hkv, dtype, pool_dtype):
        pytest.skip("no tcgen05 instance for this dtype pair")
    q, kp, vp, rtt, rpi, pre, ext = make_extend_problem(Hq, Hkv, D, prefix, extend, page, dtype, pool_dtype, seed=D + splits)
    scale = D**-0.5
    want, want_lse = ref.extend_ref(q, kp, vp, rtt, rpi, pre, ext, scale, kv_dtype=pool_dtype, want_lse=True)
    got, lse = _run(q, kp, vp, rtt, rpi, pre, ext, scale, pool_dtype, splits, page_size=page)
    assert_close_scaled(got, want, f"splits {splits}")
    assert float((lse - want_lse).abs().max()) <= ATOL
    one, lse1 = _run(q, kp, vp, rtt, rpi, pre, ext, scale, pool_dtype, 1, page_size=page)
    assert float((got - one).abs().max()) <= 2e-2
    assert float((lse - lse1).abs().max()) <= 2e-3


@pytest.mark.parametrize("pool_dtype", [torch.bfloat16, torch.float8_e4m3fn])
@pytest.mark.parametrize("opts", [dict(window_keys=700), dict(window_keys=129, sinks=True), dict(logit_cap=30.0),
                                  dict(sinks=True), dict(window_keys=2000, logit_cap=50.0, sinks=True)],
                         ids=["win700", "win129_sinks", "cap30", "sinks", "win2000_cap_sinks"])
def test_split_kv_extend_options(opts, pool_dtype):
    """Sliding window (the item's block range starts mid-prefix and is split from there), sinks (applied once, by the
    merge), logit cap."""
    Hq, Hkv, D = 32, 8, 128
    q, kp, vp, rtt, rpi, pre, ext = make_extend_problem(Hq, Hkv, D, [3000, 640, 100], [8, 32, 20], 1, torch.bfloat16,
                                                        pool_dtype, seed=17)
    sinks = torch.randn(Hq, generator=torch.Generator().manual_seed(9)) * 2 if opts.get("sinks") else None
    kw = dict(window_keys=opts.get("window_keys", 0), logit_cap=opts.get("logit_cap", 0.0), sinks=sinks)
    want, want_lse = ref.extend_ref(q, kp, vp, rtt, rpi, pre, ext, D**-0.5, kv_dtype=pool_dtype, want_lse=True, **kw)
    for splits in (2, 5):
        got, lse = _run(q, kp, vp, rtt, rpi, pre, ext, D**-0.5, pool_dtype, splits, **dict(kw))
        assert_close_scaled(got, want, f"splits {splits}")
        assert float((lse - want_lse).abs().max()) <= ATOL


@pytest.mark.parametrize("skip", [True, False])
def test_split_kv_extend_tree_mask(skip):
    """TARGET_VERIFY: a tree mask over the draft tokens on top of a long prefix; with SKIP_PREFIX_CUSTOM_MASK off the
    mask also hides random prefix positions, so a part of a row's prefix may be empty."""
    Hq, Hkv, D = 32, 8, 128
    prefix, extend = [2500, 600, 0], [16, 8, 24]
    q, kp, vp, rtt, rpi, pre, ext = make_extend_problem(Hq, Hkv, D, prefix, extend, 1, torch.bfloat16, torch.bfloat16, 23)
    g = torch.Generator().manual_seed(4)
    parts, offs = [], [0]
    for P, E in zip(prefix, extend):
        m = torch.rand(E, P + E, generator=g) < 0.5
        if P >= 1024:
            m[0, :1024] = False                      # row 0 sees nothing in the first parts of the prefix
        m[:, P:] &= torch.tril(torch.ones(E, E, dtype=torch.bool))
        m[torch.arange(E), P + torch.arange(E)] = True
        parts.append(m.reshape(-1))
        offs.append(offs[-1] + E * (P + E))
    mask, indptr = torch.cat(parts).to(torch.uint8), torch.tensor(offs, dtype=torch.int64)
    kw = dict(custom_mask=mask, mask_indptr=indptr, mask_skip_prefix=skip)
    want, want_lse = ref.extend_ref(q, kp, vp, rtt, rpi, pre, ext, D**-0.5, want_lse=True, **kw)
    for splits in (2, 4):
        got, lse = _run(q, kp, vp, rtt, rpi, pre, ext, D**-0.5, torch.bfloat16, splits, **dict(kw))
        assert_close_scaled(got, want, f"splits {splits} skip {skip}")
        assert float((lse - want_lse).abs().max()) <= ATOL


@pytest.mark.parametrize("bs,pool_dtype", [(24, torch.bfloat16), (40, torch.bfloat16), (21, torch.float8_e4m3fn)])
def test_split_kv_tail_only_for_batches_above_one_wave(bs, pool_dtype):
    """More work items than SMs and an exact tile count: the tiles that fill whole waves run unsplit, the tail is cut
    into parts (two attention launches + merge; rows of the unsplit tiles keep the NaN fill in the split-0 LSE block and
    are left alone by the merge).  Prefixes include short ones, which the device keeps unsplit inside the tail."""
    from sglang_b200 import ops

    Hq, Hkv, D, n = 32, 8, 128, 8
    g = torch.Generator().manual_seed(bs)
    prefix = torch.randint(200, 2600, (bs,), generator=g).tolist()
    prefix[-1], prefix[-2] = 300, 2500                   # a short and a long context inside the tail
    q, kp, vp, rtt, rpi, pre, ext = make_extend_problem(Hq, Hkv, D, prefix, [n] * bs, 1, torch.bfloat16, pool_dtype, seed=bs)
    sm, _, _ = ops.device_info()
    splits = max(ops.extend_pick_splits(bs * Hkv, sm), 2)       # (the heuristic may decline; the mechanism is under test)
    assert bs * Hkv > sm
    scale = D**-0.5
    want, want_lse = ref.extend_ref(q, kp, vp, rtt, rpi, pre, ext, scale, kv_dtype=pool_dtype, want_lse=True)
    got, lse = _run(q, kp, vp, rtt, rpi, pre, ext, scale, pool_dtype, splits, tiles_hint=bs)
    assert_close_scaled(got, want, f"tail split {splits}")
    assert float((lse - want_lse).abs().max()) <= ATOL
    one, lse1 = _run(q, kp, vp, rtt, rpi, pre, ext, scale, pool_dtype, 1)
    # the tiles of the full waves ran the unsplit kernel: bit-identical rows
    tiles_full = (bs * Hkv // sm) * sm // Hkv
    assert torch.equal(got[: tiles_full * n], one[: tiles_full * n])
    assert float((got - one).abs().max()) <= 2e-2 and float((lse - lse1).abs().max()) <= 2e-3


def test_split_kv_short_contexts_are_not_split():
    """The device trims the split count so that a part keeps >= 4 kv blocks: items below 8 blocks are finished by
    split 0 exactly like an unsplit launch (bit-identical), whatever the workspace holds."""
    Hq, Hkv, D = 32, 8, 128
    q, kp, vp, rtt, rpi, pre, ext = make_extend_problem(Hq, Hkv, D, [800, 0, 300], [8, 8, 8], 1, torch.bfloat16,
                                                        torch.bfloat16, seed=31)
    a, la = _run(q, kp, vp, rtt, rpi, pre, ext, D**-0.5, torch.bfloat16, 1)
    b, lb = _run(q, kp, vp, rtt, rpi, pre, ext, D**-0.5, torch.bfloat16, 4, poison=1e30)
    assert torch.equal(a, b) and torch.equal(la, lb)


def test_split_kv_argument_checks():
    from sglang_b200 import ops, _C

    Hq, Hkv, D = 32, 8, 128
    q, kp, vp, rtt, rpi, pre, ext = make_extend_problem(Hq, Hkv, D, [300], [8], 1, torch.bfloat16, torch.bfloat16, seed=1)
    sm, _, _ = ops.device_info()
    T = q.shape[0]
    ws = ops.ExtendWorkspace.create(1, T + 2, "cuda", num_ctas=sm)
    args = (q.cuda(), torch.empty_like(q, device="cuda"), kp.cuda(), vp.cuda(), rtt.cuda(), rpi.cuda(), pre.cuda(),
            ext.cuda(), ws, D**-0.5)
    small = torch.empty(16, dtype=torch.float32, device="cuda")
    ops.extend_plan(ws, ext.cuda(), ops.extend_block_q_one_tile(D, D, 4, torch.bfloat16, torch.bfloat16))
    with pytest.raises(_C.B200AttnError, match="split workspace"):
        ops.extend_attention(*args, num_splits=2, split_ws=small)
    big = torch.empty(ops.extend_split_ws_floats(32, T, Hq, D), dtype=torch.float32, device="cuda")
    with pytest.raises(_C.B200AttnError, match="num_splits"):
        ops.extend_attention(*args, num_splits=32, split_ws=big)
    ops.extend_plan(ws, ext.cuda(), ops.extend_block_q(D, D, 4, torch.bfloat16, torch.bfloat16))      # two-tile plan
    with pytest.raises(_C.B200AttnError, match="one-tile plan"):
        ops.extend_attention(*args, num_splits=2, split_ws=big)
    torch.cuda.synchronize()


def test_pick_splits_heuristic():
    from sglang_b200 import ops

    assert ops.extend_pick_splits(8, 148) == 16                # one request x 8 kv heads
    assert ops.extend_pick_splits(32, 148) == 4
    assert ops.extend_pick_splits(64, 148) == 2
    assert ops.extend_pick_splits(148, 148) == 1               # exactly one wave
    assert ops.extend_pick_splits(512, 148) == 1               # 3.46 waves: cutting 68 tail items in two does not pay
    assert ops.extend_pick_splits(320, 148) == 6               # 24 tail items behind two full waves (measured -9 %)
    assert ops.extend_pick_splits(192, 148) == 3               # 44 tail items behind one full wave (measured -11 %)
    assert ops.extend_pick_splits(600, 148, 8192) == 16        # 8 tail items would otherwise cost a whole wave
    assert ops.extend_pick_splits(592, 148) == 1               # exactly four waves
    assert ops.extend_pick_splits(64, 148, 256) == 1           # short contexts: fixed costs dominate
    assert ops.extend_pick_splits(4096, 148) == 1              # many waves: nothing to gain
    assert ops.extend_pick_splits(0, 148) == 1


def test_backend_verify_picks_split_kv_and_matches_oracle():
    """TARGET_VERIFY through the plugin API at a batch that cannot fill the SMs: the backend plans one-tile items, picks
    a split count from bs x Hkv (and, eagerly, the mean context length) and the result matches the oracle; then the
    same step captured in a CUDA graph and replayed with a new mask."""
    from types import SimpleNamespace

    from sglang_b200.srt_mirror import ForwardBatch, ForwardMode, RadixAttention
    from test_gpu_features import _backend_problem

    Hq, Hkv, D, n = 32, 8, 128, 8
    prefix = [3000, 5100]
    pb = _backend_problem(Hq, Hkv, D, prefix, [n] * 2, seed=78)
    g = torch.Generator().manual_seed(6)

    def tree_mask():
        parts = []
        for P in prefix:
            m = torch.zeros(n, P + n, dtype=torch.bool)
            m[:, :P] = True
            tri = torch.tril(torch.rand(n, n, generator=g) < 0.5)
            tri[torch.arange(n), torch.arange(n)] = True
            m[:, P:] = tri
            parts.append(m.reshape(-1))
        return torch.cat(parts)

    indptr = torch.tensor([0] + [n * (P + n) for P in prefix]).cumsum(0)
    mask = tree_mask()
    spec = SimpleNamespace(draft_token_num=n, custom_mask=mask.cuda())
    seq = torch.tensor(prefix, dtype=torch.int32, device="cuda")
    fb = ForwardBatch(forward_mode=ForwardMode.TARGET_VERIFY, batch_size=pb.bs, req_pool_indices=pb.rpi.cuda(),
                      seq_lens=seq, out_cache_loc=pb.loc.cuda(), seq_lens_sum=sum(prefix), spec_info=spec,
                      mha_return_lse=True)
    be = pb.backend
    be.init_forward_metadata(fb)
    assert be.extend_ws.block_q == 128 // (Hq // Hkv)
    assert be._ext_splits > 1 and be._ext_split_ws is not None
    layer = RadixAttention(Hq, D, D ** -0.5, Hkv, 0)
    qf, kf, vf = (x.cuda().reshape(pb.T, -1) for x in (pb.q, pb.k_new, pb.v_new))
    out, lse = layer(qf, kf, vf, fb, be)
    torch.cuda.synchronize()
    want, want_lse = ref.extend_ref(pb.q, pb.kp, pb.vp, pb.rtt, pb.rpi, torch.tensor(prefix), torch.tensor([n] * 2),
                                    D ** -0.5, custom_mask=mask.to(torch.uint8), mask_indptr=indptr, want_lse=True)
    assert_close_scaled(out.float().cpu(), want, "eager")
    assert float((lse.cpu() - want_lse).abs().max()) <= ATOL

    # the same step under a CUDA graph (the split count is chosen from the batch shape at capture)
    be.init_cuda_graph_state(pb.bs, pb.bs * n)
    fb.mha_return_lse = False
    be.init_forward_metadata_out_graph(fb, in_capture=True)
    assert be._ext_splits > 1
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        be.init_forward_metadata_in_graph(fb)
        layer(qf, kf, vf, fb, be)
    torch.cuda.current_stream().wait_stream(s)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        be.init_forward_metadata_in_graph(fb)
        out_g = layer(qf, kf, vf, fb, be)
    mask2 = tree_mask()
    fb.spec_info = SimpleNamespace(draft_token_num=n, custom_mask=mask2.cuda())
    be.init_forward_metadata_out_graph(fb)
    graph.replay()
    torch.cuda.synchronize()
    want2 = ref.extend_ref(pb.q, pb.kp, pb.vp, pb.rtt, pb.rpi, torch.tensor(prefix), torch.tensor([n] * 2), D ** -0.5,
                           custom_mask=mask2.to(torch.uint8), mask_indptr=indptr)
    assert_close_scaled(out_g.float().cpu().view(pb.T, Hq, D), want2, "graph replay with a new mask")

    # EAGLE's overlapped plan: the replay prep ran with a stale mask, the worker hands the real one over after the
    # draft through update_verify_buffers_to_fill_after_draft (eagle_worker_common.py:505-535)
    mask3 = tree_mask()
    fb.spec_info = SimpleNamespace(draft_token_num=n, custom_mask=torch.ones_like(mask3).cuda())      # stale
    be.init_forward_metadata_out_graph(fb)
    be.update_verify_buffers_to_fill_after_draft(SimpleNamespace(draft_token_num=n, custom_mask=mask3.cuda()), pb.bs)
    graph.replay()
    torch.cuda.synchronize()
    want3 = ref.extend_ref(pb.q, pb.kp, pb.vp, pb.rtt, pb.rpi, torch.tensor(prefix), torch.tensor([n] * 2), D ** -0.5,
                           custom_mask=mask3.to(torch.uint8), mask_indptr=indptr)
    assert_close_scaled(out_g.float().cpu().view(pb.T, Hq, D), want3, "mask refreshed after the draft")


PADDED_SHAPES = [
    # Hq, Hkv, D, prefix, extend, page
    (32, 32, 96, [0, 700, 130], [300, 260, 1], 1),       # Phi-3-mini heads (MHA): several q tiles
    (32, 32, 80, [1500], [600], 16),                     # Phi-2 heads, long prefix, paged
    (16, 4, 112, [5, 0], [250, 129], 1),                 # D 112, group 4
    (8, 1, 96, [2100, 40], [18, 5], 1),                  # group 8, short extends
]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16], ids=["bf16", "f16"])
@pytest.mark.parametrize("shape", PADDED_SHAPES, ids=[f"h{s[0]}_{s[1]}_d{s[2]}_p{s[5]}" for s in PADDED_SHAPES])
def test_extend_d80_d96_d112_on_the_padded_d128_tile(shape, dtype):
    """Head dims 80 / 96 / 112 (16-bit pools) run on the tcgen05 D = 128 one-tile instance with zero-filled padding;
    KERNEL_FAST (2) refuses shapes without a tensor-core instance.  Also with split-KV, options and LSE."""
    from sglang_b200 import ops
    from test_gpu_extend import run_ops_extend

    Hq, Hkv, D, prefix, extend, page = shape
    assert ops.extend_fast_supported(D, D, Hq // Hkv, dtype, dtype)
    assert ops.extend_block_q(D, D, Hq // Hkv, dtype, dtype) == 128 // (Hq // Hkv)
    q, kp, vp, rtt, rpi, pre, ext = make_extend_problem(Hq, Hkv, D, prefix, extend, page, dtype, dtype, seed=D)
    scale = D**-0.5
    want, want_lse = ref.extend_ref(q, kp, vp, rtt, rpi, pre, ext, scale, kv_dtype=dtype, want_lse=True)
    got, lse = run_ops_extend(q, kp, vp, rtt, rpi, pre, ext, scale, dtype, kernel=2, page_size=page, want_lse=True)
    assert_close_scaled(got, want, "fast")
    assert float((lse.cpu() - want_lse).abs().max()) <= ATOL
    got3, lse3 = _run(q, kp, vp, rtt, rpi, pre, ext, scale, dtype, 3, page_size=page)
    assert_close_scaled(got3, want, "3 splits")
    assert float((lse3 - want_lse).abs().max()) <= ATOL
    sinks = torch.randn(Hq, generator=torch.Generator().manual_seed(2)) * 2
    kw = dict(window_keys=200, logit_cap=40.0, sinks=sinks)
    want_o = ref.extend_ref(q, kp, vp, rtt, rpi, pre, ext, scale, kv_dtype=dtype, **kw)
    got_o, _ = _run(q, kp, vp, rtt, rpi, pre, ext, scale, dtype, 1, page_size=page, **dict(kw))
    assert_close_scaled(got_o, want_o, "window + cap + sinks")


def test_backend_verify_tail_split_above_one_wave():
    """TARGET_VERIFY through the plugin API with more work items than SMs (24 requests x 8 kv heads): the backend hands
    the exact tile count over and the library splits the tail only."""
    from types import SimpleNamespace

    from sglang_b200.srt_mirror import ForwardBatch, ForwardMode, RadixAttention
    from test_gpu_features import _backend_problem

    Hq, Hkv, D, n, bs = 32, 8, 128, 8, 24
    g = torch.Generator().manual_seed(3)
    prefix = torch.randint(2500, 4500, (bs,), generator=g).tolist()
    pb = _backend_problem(Hq, Hkv, D, prefix, [n] * bs, seed=79)
    parts = []
    for P in prefix:
        m = torch.zeros(n, P + n, dtype=torch.bool)
        m[:, :P] = True
        tri = torch.tril(torch.rand(n, n, generator=g) < 0.5)
        tri[torch.arange(n), torch.arange(n)] = True
        m[:, P:] = tri
        parts.append(m.reshape(-1))
    mask = torch.cat(parts)
    indptr = torch.tensor([0] + [n * (P + n) for P in prefix]).cumsum(0)
    fb = ForwardBatch(forward_mode=ForwardMode.TARGET_VERIFY, batch_size=bs, req_pool_indices=pb.rpi.cuda(),
                      seq_lens=torch.tensor(prefix, dtype=torch.int32, device="cuda"), out_cache_loc=pb.loc.cuda(),
                      seq_lens_sum=sum(prefix), spec_info=SimpleNamespace(draft_token_num=n, custom_mask=mask.cuda()))
    be = pb.backend
    be.init_forward_metadata(fb)
    assert be._ext_splits > 1 and be._ext_tiles == bs and bs * Hkv > be.sm_count
    layer = RadixAttention(Hq, D, D ** -0.5, Hkv, 0)
    out = layer(pb.q.cuda().reshape(pb.T, -1), pb.k_new.cuda().reshape(pb.T, -1), pb.v_new.cuda().reshape(pb.T, -1), fb, be)
    torch.cuda.synchronize()
    want = ref.extend_ref(pb.q, pb.kp, pb.vp, pb.rtt, pb.rpi, torch.tensor(prefix), torch.tensor([n] * bs), D ** -0.5,
                          custom_mask=mask.to(torch.uint8), mask_indptr=indptr)
    assert_close_scaled(out.float().cpu().view(pb.T, Hq, D), want, "backend verify, tail-only split")
