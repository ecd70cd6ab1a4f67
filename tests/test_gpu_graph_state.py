"""CUDA-graph state of the backend (round-1 advisor findings).

* `init_cuda_graph_state(max_bs, ...)` is called by the reference's graph runner with the largest CAPTURE batch
  (decode_cuda_graph_runner.py:318,444), usually far below the pool size: the decode workspace must not shrink,
  and an eager decode batch above the captured maximum must still run.
* TARGET_VERIFY under a real CUDA graph: the replay's tree mask is a different tensor from the capture-time one
  (triton_backend.py:507-514 copies it into a persistent buffer), and later, larger eager prefills must not
  invalidate what the captured graph points at.
"""
from types import SimpleNamespace

import pytest
import torch

from helpers import ATOL, assert_close_scaled
from oracle import torch_native_ref as ref
from test_gpu_features import _backend_problem

pytestmark = pytest.mark.gpu


def test_graph_state_smaller_than_eager_batch():
    from sglang_b200.srt_mirror import ForwardBatch, ForwardMode, RadixAttention

    Hq, Hkv, D = 32, 8, 128
    prefix = [97, 300, 12, 555, 64, 1, 2047, 130]
    pb = _backend_problem(Hq, Hkv, D, prefix, [1] * 8, seed=11)
    be = pb.backend
    ws_before = be.decode_ws
    be.init_cuda_graph_state(4, 4)             # capture sizes stop at 4 ...
    be.init_cuda_graph_state(4, 4)             # ... and the runner calls it twice (:318 and :444)
    assert be.decode_ws is ws_before, "the decode workspace must only grow"
    fb = ForwardBatch(forward_mode=ForwardMode.DECODE, batch_size=pb.bs, req_pool_indices=pb.rpi.cuda(),
                      seq_lens=torch.tensor(pb.lens, dtype=torch.int32, device="cuda"), out_cache_loc=pb.loc.cuda(),
                      seq_lens_sum=sum(pb.lens))
    be.init_forward_metadata(fb)               # ... but an eager batch of 8 arrives
    layer = RadixAttention(Hq, D, D ** -0.5, Hkv, 0)
    out = layer(pb.q.cuda().reshape(pb.T, -1), pb.k_new.cuda().reshape(pb.T, -1), pb.v_new.cuda().reshape(pb.T, -1),
                fb, be)
    torch.cuda.synchronize()
    want = ref.decode_ref(pb.q, pb.kp, pb.vp, pb.rtt, pb.rpi, torch.tensor(pb.lens, dtype=torch.int32), D ** -0.5)
    assert_close_scaled(out.float().cpu().view_as(want), want)


def _tree_mask(prefix, n, seed):
    g = torch.Generator().manual_seed(seed)
    parts = []
    for P in prefix:
        m = torch.zeros(n, P + n, dtype=torch.bool)
        m[:, :P] = True
        tri = torch.tril(torch.rand(n, n, generator=g) < 0.5)
        tri[torch.arange(n), torch.arange(n)] = True
        m[:, P:] = tri
        parts.append(m.reshape(-1))
    return torch.cat(parts)


def test_target_verify_graph_replay_with_new_mask_and_interleaved_prefill():
    from sglang_b200.srt_mirror import ForwardBatch, ForwardMode, RadixAttention

    Hq, Hkv, D, n = 32, 8, 128, 4
    scale = D ** -0.5
    prefix = [333, 40, 700]
    bs, cap_bs = len(prefix), 4
    pb = _backend_problem(Hq, Hkv, D, prefix, [n] * bs, seed=21)
    be = pb.backend
    be.num_draft_tokens = n
    be.init_cuda_graph_state(cap_bs, cap_bs * n)
    T = cap_bs * n
    seq_buf = torch.ones(cap_bs, dtype=torch.int32, device="cuda")
    rpi_buf = torch.zeros(cap_bs, dtype=torch.int64, device="cuda")
    loc_buf = torch.zeros(T, dtype=torch.int64, device="cuda")
    q_buf = torch.zeros(T, Hq * D, dtype=torch.bfloat16, device="cuda")
    k_buf = torch.zeros(T, Hkv * D, dtype=torch.bfloat16, device="cuda")
    v_buf = torch.zeros(T, Hkv * D, dtype=torch.bfloat16, device="cuda")
    layer = RadixAttention(Hq, D, scale, Hkv, 0)

    def view(mask):
        return SimpleNamespace(forward_mode=ForwardMode.TARGET_VERIFY, batch_size=cap_bs, req_pool_indices=rpi_buf,
                               seq_lens=seq_buf, seq_lens_cpu=None, out_cache_loc=loc_buf, seq_lens_sum=cap_bs,
                               spec_info=SimpleNamespace(draft_token_num=n, custom_mask=mask))

    # ---- capture with an all-ones mask of the capture shape (seq_len fill value 1 everywhere)
    cap_mask = torch.ones(cap_bs * n * (1 + n), dtype=torch.bool, device="cuda")
    cap_view = view(cap_mask)
    be.init_forward_metadata_out_graph(cap_view, in_capture=True)
    s = torch.cuda.Stream()
    s.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(s):
        be.init_forward_metadata_in_graph(cap_view)
        layer(q_buf, k_buf, v_buf, cap_view, be)
    torch.cuda.current_stream().wait_stream(s)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        be.init_forward_metadata_in_graph(cap_view)
        out_buf = layer(q_buf, k_buf, v_buf, cap_view, be)
    del cap_mask, cap_view                      # the capture-time mask is gone, as in serving

    # ---- an eager prefill far larger than the graph state runs in between (must not touch the graph's buffers)
    big_prefix, big_ext = [0, 100], [3000, 2500]
    pb2 = _backend_problem(Hq, Hkv, D, big_prefix, big_ext, seed=22)
    # same backend object, second problem's pools: swap them in for the eager call
    be.req_to_token_pool, be.token_to_kv_pool = pb2.runner.req_to_token_pool, pb2.runner.token_to_kv_pool
    fb2 = ForwardBatch(forward_mode=ForwardMode.EXTEND, batch_size=pb2.bs, req_pool_indices=pb2.rpi.cuda(),
                       seq_lens=torch.tensor(pb2.lens, dtype=torch.int32, device="cuda"),
                       out_cache_loc=pb2.loc.cuda(), seq_lens_sum=sum(pb2.lens),
                       extend_prefix_lens=torch.tensor(big_prefix, dtype=torch.int32, device="cuda"),
                       extend_seq_lens=torch.tensor(big_ext, dtype=torch.int32, device="cuda"),
                       extend_prefix_lens_cpu=big_prefix, extend_seq_lens_cpu=big_ext, extend_num_tokens=sum(big_ext))
    graph_state = be._ext_graph
    be.init_forward_metadata(fb2)
    assert be._ext is not graph_state and be._ext_graph is graph_state
    out2 = layer(pb2.q.cuda().reshape(pb2.T, -1), pb2.k_new.cuda().reshape(pb2.T, -1),
                 pb2.v_new.cuda().reshape(pb2.T, -1), fb2, be)
    torch.cuda.synchronize()
    want2 = ref.extend_ref(pb2.q, pb2.kp, pb2.vp, pb2.rtt, pb2.rpi, torch.tensor(big_prefix), torch.tensor(big_ext),
                           scale)
    assert_close_scaled(out2.float().cpu().view_as(want2), want2)
    be.req_to_token_pool, be.token_to_kv_pool = pb.runner.req_to_token_pool, pb.runner.token_to_kv_pool

    # ---- replays: 3 live requests in a batch padded to 4, a different tree mask every time
    for it in range(2):
        mask = _tree_mask(prefix, n, seed=100 + it)
        pad = torch.ones(n * (1 + n), dtype=torch.bool)         # padded row: seq_len 1 + n draft tokens
        seq_buf.fill_(1); rpi_buf.zero_(); loc_buf.zero_()
        seq_buf[:bs] = torch.tensor(prefix, dtype=torch.int32)
        rpi_buf[:bs] = pb.rpi
        loc_buf[: bs * n] = pb.loc
        q_buf.zero_(); k_buf.zero_(); v_buf.zero_()
        q_buf[: bs * n] = pb.q.reshape(bs * n, -1)
        k_buf[: bs * n] = pb.k_new.reshape(bs * n, -1)
        v_buf[: bs * n] = pb.v_new.reshape(bs * n, -1)
        be.init_forward_metadata_out_graph(view(torch.cat([mask, pad]).cuda()))
        graph.replay()
        torch.cuda.synchronize()
        indptr = torch.tensor([0] + [n * (P + n) for P in prefix]).cumsum(0)
        want = ref.extend_ref(pb.q, pb.kp, pb.vp, pb.rtt, pb.rpi, torch.tensor(prefix), torch.tensor([n] * bs),
                              scale, custom_mask=mask.to(torch.uint8), mask_indptr=indptr)
        got = out_buf[: bs * n].float().cpu().view_as(want)
        assert_close_scaled(got, want, f"replay {it}")
