"""GPU parity: multi-step EAGLE draft decode (chain drafting, topk 1) through B200MultiStepDraftBackend -- step i of a
draft attends the request's prefix plus the i + 1 tokens drafted so far (triton_backend.py:1867-2042,
kernels/ops/speculative/cache_locs.py:56-130 for the per-step kv indices the reference materialises)."""
from types import SimpleNamespace

import pytest
import torch

from helpers import assert_close_scaled
from oracle import torch_native_ref as ref
from test_gpu_features import _backend_problem

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("Hq,Hkv,D", [(32, 8, 128), (64, 8, 64), (8, 1, 128)])
def test_multistep_draft_decode_eager_and_graph(Hq, Hkv, D):
    from sglang_b200.backend import B200MultiStepDraftBackend
    from sglang_b200.srt_mirror import ForwardBatch, ForwardMode, RadixAttention

    steps = 3                                      # speculative_num_steps = 4: three decode steps after the draft prefill
    prefix = [700, 1, 129, 2048]
    pb = _backend_problem(Hq, Hkv, D, prefix, [steps] * len(prefix), seed=D + Hkv)
    bs = pb.bs
    ms = B200MultiStepDraftBackend(pb.runner, 1, steps + 1)
    assert len(ms.attn_backends) == steps and [b.draft_step_offset for b in ms.attn_backends] == [1, 2, 3]
    layer = RadixAttention(Hq, D, D ** -0.5, Hkv, 0)
    loc = pb.loc.view(bs, steps)
    rows = torch.arange(bs) * steps
    seq = torch.tensor(prefix, dtype=torch.int32)

    loc_dev = loc.cuda()

    def step_inputs(i):
        idx = rows + i            # the slot column is a STRIDED view, as the draft worker's per-step out_cache_loc is
        return (pb.q[idx].cuda().reshape(bs, -1), pb.k_new[idx].cuda().reshape(bs, -1),
                pb.v_new[idx].cuda().reshape(bs, -1), loc_dev[:, i])

    def want(i):
        return ref.decode_ref(pb.q[rows + i], pb.kp, pb.vp, pb.rtt, pb.rpi, seq + i + 1, D ** -0.5)

    # ---- eager: one metadata call per draft, then step i under attn_backends[i]
    fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=bs, req_pool_indices=pb.rpi.cuda(),
                      seq_lens=seq.cuda(), out_cache_loc=loc[:, 0].contiguous().cuda(), seq_lens_sum=sum(prefix))
    ms.init_forward_metadata(fb)
    for i in range(steps):
        q, k, v, fb.out_cache_loc = step_inputs(i)
        out = layer(q, k, v, fb, ms.attn_backends[i])
        torch.cuda.synchronize()
        assert_close_scaled(out.float().cpu().view(bs, Hq, D), want(i), f"eager step {i}")

    # ---- CUDA graph at a padded batch size, replayed (EAGLEDraftCudaGraphRunner's call sequence)
    max_bs = bs + 2
    ms.init_cuda_graph_state(max_bs, max_bs)
    seq_buf = torch.ones(max_bs, dtype=torch.int32, device="cuda")
    rpi_buf = torch.zeros(max_bs, dtype=torch.int64, device="cuda")
    loc_buf = torch.zeros(steps, max_bs, dtype=torch.int64, device="cuda")
    q_buf = torch.zeros(steps, max_bs, Hq * D, dtype=torch.bfloat16, device="cuda")
    k_buf = torch.zeros(steps, max_bs, Hkv * D, dtype=torch.bfloat16, device="cuda")
    v_buf = torch.zeros(steps, max_bs, Hkv * D, dtype=torch.bfloat16, device="cuda")
    view = SimpleNamespace(forward_mode=ForwardMode.DECODE, batch_size=max_bs, req_pool_indices=rpi_buf, seq_lens=seq_buf,
                           seq_lens_cpu=None, out_cache_loc=loc_buf[0], seq_lens_sum=max_bs, spec_info=None)

    def run_steps():
        ms.init_forward_metadata_in_graph(view)
        outs = []
        for i in range(steps):
            view.out_cache_loc = loc_buf[i]
            outs.append(layer(q_buf[i], k_buf[i], v_buf[i], view, ms.attn_backends[i]))
        return outs

    ms.init_forward_metadata_out_graph(view, in_capture=True)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        run_steps()
    torch.cuda.current_stream().wait_stream(s)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        outs = run_steps()
    # the pool rows the eager pass stored are rewritten by the replay with the same values
    seq_buf[:bs].copy_(seq.cuda())
    rpi_buf[:bs].copy_(pb.rpi.cuda())
    for i in range(steps):
        q, k, v, lo = step_inputs(i)
        q_buf[i, :bs], k_buf[i, :bs], v_buf[i, :bs], loc_buf[i, :bs] = q, k, v, lo
    ms.init_forward_metadata_out_graph(view)
    graph.replay()
    torch.cuda.synchronize()
    for i in range(steps):
        assert_close_scaled(outs[i][:bs].float().cpu().view(bs, Hq, D), want(i), f"graph step {i}")


@pytest.mark.parametrize("Hq,Hkv,D,topk", [(32, 8, 128, 2), (8, 1, 128, 4), (64, 8, 64, 3)])
def test_multistep_tree_draft_decode(Hq, Hkv, D, topk):
    """Tree drafting (topk > 1): branch t of a request attends the prefix plus the slots drafted on its own branch,
    which the worker keeps at req_to_token[req, seq_len + t * num_steps + j] (cache_locs.py:104-108).  The wrapper
    builds one page-table row per (request, branch) and every step is a plain decode over num_seqs x topk rows."""
    from sglang_b200.backend import B200MultiStepDraftBackend
    from sglang_b200.srt_mirror import ForwardBatch, ForwardMode, RadixAttention

    num_steps = 4
    steps = num_steps - 1
    prefix = [700, 1, 129, 1500]
    per_req = topk * num_steps
    pb = _backend_problem(Hq, Hkv, D, prefix, [per_req] * len(prefix), seed=D + topk)
    bs = pb.bs
    ms = B200MultiStepDraftBackend(pb.runner, topk, num_steps)
    layer = RadixAttention(Hq, D, D ** -0.5, Hkv, 0)
    # the oracle's page table: one row per (request, branch) = prefix ++ the branch's slots
    W = pb.rtt.shape[1]
    vrtt = torch.zeros(bs * topk + 1, W, dtype=torch.int32)
    for b, P in enumerate(prefix):
        for t in range(topk):
            vrtt[b * topk + t + 1, :P] = pb.rtt[b + 1, :P]
            vrtt[b * topk + t + 1, P : P + num_steps] = pb.rtt[b + 1, P + t * num_steps : P + (t + 1) * num_steps]
    vrpi = torch.arange(1, bs * topk + 1)
    vseq = torch.tensor(prefix, dtype=torch.int32).repeat_interleave(topk)
    # row (b, t) of step i is new token index b * per_req + t * num_steps + i
    base = (torch.arange(bs).view(-1, 1) * per_req + torch.arange(topk).view(1, -1) * num_steps).reshape(-1)
    loc_all = pb.loc.cuda()

    def step_inputs(i):
        idx = base + i
        return (pb.q[idx].cuda().reshape(bs * topk, -1), pb.k_new[idx].cuda().reshape(bs * topk, -1),
                pb.v_new[idx].cuda().reshape(bs * topk, -1), loc_all[idx.cuda()])

    def want(i):
        return ref.decode_ref(pb.q[base + i], pb.kp, pb.vp, vrtt, vrpi, vseq + i + 1, D ** -0.5)

    fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=bs, req_pool_indices=pb.rpi.cuda(),
                      seq_lens=torch.tensor(prefix, dtype=torch.int32, device="cuda"), out_cache_loc=loc_all[:bs * topk],
                      seq_lens_sum=sum(prefix))
    ms.init_forward_metadata(fb)
    for i in range(steps):
        q, k, v, fb.out_cache_loc = step_inputs(i)
        out = layer(q, k, v, fb, ms.attn_backends[i])
        torch.cuda.synchronize()
        assert_close_scaled(out.float().cpu().view(bs * topk, Hq, D), want(i), f"eager step {i}")

    # CUDA graph at a padded batch, replayed
    max_bs = bs + 1
    ms.init_cuda_graph_state(max_bs, max_bs * topk)
    rows = max_bs * topk
    seq_buf = torch.ones(max_bs, dtype=torch.int32, device="cuda")
    rpi_buf = torch.zeros(max_bs, dtype=torch.int64, device="cuda")
    loc_buf = torch.zeros(steps, rows, dtype=torch.int64, device="cuda")
    q_buf = torch.zeros(steps, rows, Hq * D, dtype=torch.bfloat16, device="cuda")
    k_buf = torch.zeros(steps, rows, Hkv * D, dtype=torch.bfloat16, device="cuda")
    v_buf = torch.zeros(steps, rows, Hkv * D, dtype=torch.bfloat16, device="cuda")
    view = SimpleNamespace(forward_mode=ForwardMode.DECODE, batch_size=max_bs, req_pool_indices=rpi_buf, seq_lens=seq_buf,
                           seq_lens_cpu=None, out_cache_loc=loc_buf[0], seq_lens_sum=max_bs, spec_info=None)

    def run_steps():
        ms.init_forward_metadata_in_graph(view)
        outs = []
        for i in range(steps):
            view.out_cache_loc = loc_buf[i]
            outs.append(layer(q_buf[i], k_buf[i], v_buf[i], view, ms.attn_backends[i]))
        return outs

    ms.init_forward_metadata_out_graph(view, in_capture=True)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        run_steps()
    torch.cuda.current_stream().wait_stream(s)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        outs = run_steps()
    seq_buf[:bs].copy_(torch.tensor(prefix, dtype=torch.int32))
    rpi_buf[:bs].copy_(pb.rpi.cuda())
    n = bs * topk
    for i in range(steps):
        q, k, v, lo = step_inputs(i)
        q_buf[i, :n], k_buf[i, :n], v_buf[i, :n], loc_buf[i, :n] = q, k, v, lo
    ms.init_forward_metadata_out_graph(view)
    graph.replay()
    torch.cuda.synchronize()
    for i in range(steps):
        assert_close_scaled(outs[i][:n].float().cpu().view(n, Hq, D), want(i), f"graph step {i}")
