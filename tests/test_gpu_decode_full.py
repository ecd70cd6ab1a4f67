"""GPU tests at BASELINE.json's full decode sizes (size-independent properties), the planner,
and the CUDA-graph contract of the backend."""
import numpy as np
import pytest
import torch

from helpers import ATOL, make_runner, assert_close_scaled
from oracle import torch_native_ref as ref

pytestmark = pytest.mark.gpu


def _expected_plan(lens, G, num_ctas, waves, min_chunk, max_chunks):
    """numpy restatement of decode_plan_kernel: candidate m = 1..waves rounds, smallest 16-multiple chunk
    whose item count fits m * num_ctas (bisection), cost = rounds * (chunk + min_chunk / 4)."""
    total, bs = sum(lens), len(lens)

    def items(chunk):
        return G * sum((L + chunk - 1) // chunk for L in lens if L > 0)

    u_min = (min_chunk + 15) // 16
    u_max = max(u_min, (max(lens + [0]) + 15) // 16)
    best = None
    for m in range(1, min(waves, 32) + 1):
        lo, hi = u_min, u_max
        while lo < hi:
            mid = (lo + hi) >> 1
            if items(16 * mid) <= m * num_ctas:
                hi = mid
            else:
                lo = mid + 1
        chunk = 16 * lo
        cost = ((items(chunk) + num_ctas - 1) // num_ctas) * (chunk + min_chunk // 4)
        if best is None or cost < best[0]:
            best = (cost, chunk)
    chunk = best[1]
    while total // chunk + bs > max_chunks:
        chunk *= 2
    n = [(L + chunk - 1) // chunk for L in lens]
    off = np.concatenate([[0], np.cumsum(n)])
    req = np.repeat(np.arange(bs), n)
    return chunk, off, req


@pytest.mark.parametrize("lens", [[4096] * 64, [1], [5, 0, 7000, 64, 65], list(range(1, 1500, 7)), [32768] * 8,
                                  [100] * 3000])
@pytest.mark.parametrize("idx_dtype", [torch.int32, torch.int64])
def test_decode_planner_matches_host_model(lens, idx_dtype):
    """Page-index/plan arithmetic is integer work: bit-exact against a numpy restatement."""
    from sglang_b200 import ops

    bs = len(lens)
    ws = ops.DecodeWorkspace.create(bs, 8, 128, "cuda", num_ctas=296)
    seq = torch.tensor(lens, dtype=idx_dtype, device="cuda")
    ops.decode_plan(ws, seq, 8, waves=16, min_chunk_tokens=128)
    torch.cuda.synchronize()
    plan = ws.plan.cpu().numpy()
    chunk, off, req = _expected_plan(lens, 8, 296, 16, 128, ws.max_chunks)
    assert plan[0] == chunk and plan[1] == off[-1] and plan[2] == bs
    assert np.array_equal(plan[8 : 8 + bs + 1], off)
    assert np.array_equal(plan[8 + ws.max_bs + 1 : 8 + ws.max_bs + 1 + off[-1]], req)


def _full_problem(bs, L, Hq, Hkv, D, pool_dtype, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    rows = bs * L + 1
    kp = torch.randn(rows, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
    vp = torch.randn(rows, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
    if pool_dtype != torch.bfloat16:
        kp, vp = kp.to(pool_dtype).view(torch.uint8), vp.to(pool_dtype).view(torch.uint8)
    perm = torch.randperm(rows - 1, generator=g, device="cuda") + 1   # BASELINE cfg2: random slot permutation
    rtt = torch.zeros(bs + 1, L, dtype=torch.int32, device="cuda")
    rtt[1:] = perm.view(bs, L).to(torch.int32)
    q = torch.randn(bs, Hq, D, generator=g, device="cuda", dtype=torch.bfloat16)
    rpi = torch.arange(1, bs + 1, device="cuda")
    seq = torch.full((bs,), L, dtype=torch.int32, device="cuda")
    return q, kp, vp, rtt, rpi, seq


@pytest.mark.parametrize("pool_dtype", [torch.bfloat16, torch.float8_e4m3fn])
def test_cfg2_full_size_properties(pool_dtype):
    """BASELINE cfg2 (bs 64 x 4096 ctx, Hq 32 / Hkv 8 / D 128).
    1. oracle check on a request subset (the oracle handles 4 requests in seconds);
    2. permutation invariance: re-shuffling which request owns which rows of the batch permutes the output;
    3. convexity: with V == one fixed row everywhere the output is that row (softmax weights sum to 1)."""
    from sglang_b200 import ops

    bs, L, Hq, Hkv, D = 64, 4096, 32, 8, 128
    q, kp, vp, rtt, rpi, seq = _full_problem(bs, L, Hq, Hkv, D, pool_dtype)
    sm, _, _ = ops.device_info()
    ws = ops.DecodeWorkspace.create(bs, Hq, D, "cuda", num_ctas=2 * sm)
    scale = D**-0.5

    def run(rpi_, q_, vp_=vp):
        ops.decode_plan(ws, seq, Hkv)
        o = torch.empty_like(q_)
        ops.decode_attention(q_, o, kp, vp_, rtt, rpi_, seq, ws, scale, kv_dtype=pool_dtype)
        torch.cuda.synchronize()
        return o

    o = run(rpi, q)
    sub = [0, 17, 40, 63]
    want = ref.decode_ref(q[sub].cpu(), kp.cpu(), vp.cpu(), rtt.cpu(), rpi[sub].cpu(), seq[sub].cpu(), scale,
                          kv_dtype=pool_dtype)
    assert_close_scaled(o[sub].float().cpu(), want)
    p = torch.randperm(bs, device="cuda")
    o_perm = run(rpi[p], q[p])
    assert torch.equal(o_perm, o[p])
    row = torch.randn(Hkv, D, device="cuda").to(torch.bfloat16)
    v_const = row.expand(vp.shape[0], Hkv, D).contiguous()
    if pool_dtype != torch.bfloat16:
        v_const = v_const.to(pool_dtype).view(torch.uint8)
        row = v_const[0].view(pool_dtype).to(torch.bfloat16)
    o_c = run(rpi, q, v_const).view(bs, Hkv, Hq // Hkv, D)
    assert float((o_c.float() - row.float()[None, :, None, :]).abs().max()) <= 2e-2


def test_cuda_graph_capture_and_replay():
    """The contract of decode_cuda_graph_runner.py:946-953,1168: capture at a padded batch size (padded rows:
    pool index 0, seq_len = fill value, out_cache_loc 0), replay with new lengths -> equals eager and oracle."""
    from sglang_b200.backend import B200AttnBackend
    from sglang_b200.srt_mirror import ForwardBatch, ForwardMode, RadixAttention

    Hq, Hkv, D, max_bs, ctx = 8, 2, 128, 8, 700
    rows = max_bs * ctx + 1
    runner = make_runner(device="cuda", dtype=torch.bfloat16, pool_dtype=torch.bfloat16, Hq=Hq, Hkv=Hkv, D=D,
                         page_size=1, pool_rows=rows, max_bs=max_bs, max_context_len=ctx)
    g = torch.Generator().manual_seed(7)
    pool, rtp = runner.token_to_kv_pool, runner.req_to_token_pool
    pool.k_buffer[0].copy_(torch.randn(rows, Hkv, D, generator=g).to(torch.bfloat16))
    pool.v_buffer[0].copy_(torch.randn(rows, Hkv, D, generator=g).to(torch.bfloat16))
    pool.k_buffer[0][0].zero_(); pool.v_buffer[0][0].zero_()   # the reserved sink row
    rtp.req_to_token[1:].copy_((torch.randperm(rows - 1, generator=g) + 1).view(max_bs, ctx).to(torch.int32))
    backend = B200AttnBackend(runner)
    backend.init_cuda_graph_state(max_bs, max_bs)
    fill = backend.get_cuda_graph_seq_len_fill_value()
    layer = RadixAttention(num_heads=Hq, head_dim=D, scaling=D**-0.5, num_kv_heads=Hkv, layer_id=0)
    # static buffers owned by the graph runner
    s_rpi = torch.zeros(max_bs, dtype=torch.int64, device="cuda")
    s_seq = torch.full((max_bs,), fill, dtype=torch.int32, device="cuda")
    s_loc = torch.zeros(max_bs, dtype=torch.int64, device="cuda")
    s_q = torch.zeros(max_bs, Hq * D, dtype=torch.bfloat16, device="cuda")
    s_k = torch.zeros(max_bs, Hkv * D, dtype=torch.bfloat16, device="cuda")
    s_v = torch.zeros(max_bs, Hkv * D, dtype=torch.bfloat16, device="cuda")
    fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=max_bs, req_pool_indices=s_rpi, seq_lens=s_seq,
                      out_cache_loc=s_loc, seq_lens_sum=max_bs * fill)
    stream = torch.cuda.Stream()
    stream.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(stream):
        for _ in range(2):  # warm-up
            backend.init_forward_metadata(fb)
            layer(s_q, s_k, s_v, fb, backend)
    torch.cuda.current_stream().wait_stream(stream)
    graph = torch.cuda.CUDAGraph()
    backend.init_forward_metadata_out_graph(fb, in_capture=True)
    with torch.cuda.graph(graph, stream=stream):
        backend.init_forward_metadata_in_graph(fb)
        s_out = layer(s_q, s_k, s_v, fb, backend)

    for raw_bs, lens in [(3, [5, 300, 699]), (8, [64, 65, 1, 2, 128, 699, 333, 17]), (1, [450])]:
        q = torch.randn(raw_bs, Hq * D, generator=g).to(torch.bfloat16)
        k = torch.randn(raw_bs, Hkv * D, generator=g).to(torch.bfloat16)
        v = torch.randn(raw_bs, Hkv * D, generator=g).to(torch.bfloat16)
        s_rpi.zero_(); s_seq.fill_(fill); s_loc.zero_()
        s_rpi[:raw_bs] = torch.arange(1, raw_bs + 1)
        s_seq[:raw_bs] = torch.tensor(lens, dtype=torch.int32)
        loc = rtp.req_to_token[torch.arange(1, raw_bs + 1), torch.tensor(lens) - 1].long()
        s_loc[:raw_bs] = loc
        s_q[:raw_bs], s_k[:raw_bs], s_v[:raw_bs] = q.cuda(), k.cuda(), v.cuda()
        backend.init_forward_metadata_out_graph(fb)
        graph.replay()
        torch.cuda.synchronize()
        got = s_out[:raw_bs].float().cpu().view(raw_bs, Hq, D)
        want = ref.decode_ref(q.view(raw_bs, Hq, D), pool.k_buffer[0].cpu(), pool.v_buffer[0].cpu(),
                              rtp.req_to_token.cpu(), torch.arange(1, raw_bs + 1), torch.tensor(lens), D**-0.5)
        assert_close_scaled(got, want)
        assert torch.equal(pool.k_buffer[0][loc].cpu().view(raw_bs, -1), k)  # new rows stored by the replay
        assert int(pool.k_buffer[0][0].abs().sum()) == 0                     # slot-0 sink never written
