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


def test_multistep_draft_backend_refuses_tree_drafting():
    from sglang_b200.backend import B200MultiStepDraftBackend

    pb = _backend_problem(8, 2, 128, [10], [1], seed=1)
    with pytest.raises(ValueError, match="topk 1"):
        B200MultiStepDraftBackend(pb.runner, 2, 3)
