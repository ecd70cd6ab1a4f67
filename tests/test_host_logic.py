"""Host-side logic that needs no GPU: the interface mirror, plugin registration, TP head sharding
(incl. a 2-rank gloo run), and the backend's refusal to run without CUDA."""
import os
import socket
import sys
from pathlib import Path
from types import SimpleNamespace

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import torch_native_ref as ref
from sglang_b200.tp import shard_heads

REPO = Path(__file__).resolve().parent.parent


def test_forward_mode_semantics():
    from sglang_b200.srt_mirror import ForwardMode as M

    assert M.DECODE.is_decode() and not M.DECODE.is_extend()
    assert M.EXTEND.is_extend() and M.MIXED.is_extend() and M.TARGET_VERIFY.is_extend()
    assert M.IDLE.is_idle() and M.IDLE.is_decode_or_idle()
    assert M.DRAFT_EXTEND_V2.is_extend(include_draft_extend_v2=True) and not M.DRAFT_EXTEND_V2.is_extend()


def test_mirror_pool_layout_matches_reference_contract():
    from sglang_b200.srt_mirror import MHATokenToKVPool, ReqToTokenPool

    rtp = ReqToTokenPool(size=4, max_context_len=16, device="cpu")
    assert rtp.req_to_token.shape == (5, 16) and rtp.req_to_token.dtype == torch.int32  # row 0 = dummy
    assert rtp.alloc(2) == [1, 2] and rtp.available_size() == 2
    pool = MHATokenToKVPool(size=64, page_size=16, dtype=torch.float8_e4m3fn, head_num=2, head_dim=8,
                            layer_num=3, device="cpu", start_layer=2)
    assert pool.store_dtype == torch.uint8 and pool.k_buffer[0].shape == (80, 2, 8)   # size + page_size rows
    assert pool.get_key_buffer(2).dtype == torch.float8_e4m3fn                          # start_layer offset
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        pool.set_kv_buffer(SimpleNamespace(layer_id=2), torch.tensor([1]), torch.zeros(1, 2, 8, dtype=torch.bfloat16),
                           torch.zeros(1, 2, 8, dtype=torch.bfloat16))


def test_plugin_registers_backend_name():
    from sglang_b200 import BACKEND_NAME, plugin
    from sglang_b200.srt_mirror import ATTENTION_BACKENDS

    plugin.register()
    plugin.register()  # idempotent: load_plugins() runs in every scheduler process
    assert BACKEND_NAME in ATTENTION_BACKENDS
    runner = SimpleNamespace(device="cpu")
    with pytest.raises(RuntimeError, match="CUDA"):
        ATTENTION_BACKENDS[BACKEND_NAME](runner)


def test_backend_implements_the_plugin_api():
    from sglang_b200.backend import B200AttnBackend
    from sglang_b200.srt_mirror import AttentionBackend

    assert issubclass(B200AttnBackend, AttentionBackend)
    for name in ("init_forward_metadata", "init_forward_metadata_out_graph", "init_forward_metadata_in_graph",
                 "init_cuda_graph_state", "get_cuda_graph_seq_len_fill_value", "forward", "forward_decode",
                 "forward_extend"):
        assert callable(getattr(B200AttnBackend, name))
    assert B200AttnBackend.forward_decode is not AttentionBackend.forward_decode


@pytest.mark.parametrize("tp,want", [(1, (64, 8, 1)), (2, (32, 4, 1)), (4, (16, 2, 1)), (8, (8, 1, 1)),
                                     (16, (4, 1, 2))])
def test_llama70b_head_shards(tp, want):
    """per-GPU (Hq, Hkv) = (64,8)/(32,4)/(16,2)/(8,1); KV heads replicate beyond tp 8."""
    seen_q = []
    for rank in range(tp):
        s = shard_heads(64, 8, tp, rank)
        assert (s.q_heads, s.kv_heads, s.kv_replicas) == want
        # every local q head maps onto a local kv head
        group = 64 // 8
        for h in range(s.q_head_start, s.q_head_start + s.q_heads):
            assert s.kv_head_start <= h // group < s.kv_head_start + s.kv_heads
        seen_q += list(range(s.q_head_start, s.q_head_start + s.q_heads))
    assert seen_q == list(range(64))
    with pytest.raises(ValueError):
        shard_heads(14, 2, 4, 0)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _tp_worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        g = torch.Generator().manual_seed(0)   # same scheduler decisions on every rank
        Hq, Hkv, D, rows, lens = 8, 4, 16, 400, [70, 31, 99]
        q = torch.randn(3, Hq, D, generator=g)
        kc, vc = torch.randn(rows, Hkv, D, generator=g), torch.randn(rows, Hkv, D, generator=g)
        rtt = torch.zeros(4, 100, dtype=torch.int32)
        rtt[1:] = (torch.randperm(rows - 1, generator=g)[:300] + 1).view(3, 100).to(torch.int32)
        rpi, seq = torch.arange(1, 4), torch.tensor(lens)
        s = shard_heads(Hq, Hkv, world, rank)
        qs = q[:, s.q_head_start : s.q_head_start + s.q_heads]
        ks = kc[:, s.kv_head_start : s.kv_head_start + s.kv_heads].contiguous()  # this rank's pool slice
        vs = vc[:, s.kv_head_start : s.kv_head_start + s.kv_heads].contiguous()
        local = ref.decode_ref(qs, ks, vs, rtt, rpi, seq, D**-0.5)
        parts = [torch.empty_like(local) for _ in range(world)]
        dist.all_gather(parts, local)          # test-only gather; the product path has no collective here
        full = ref.decode_ref(q, kc, vc, rtt, rpi, seq, D**-0.5)
        ok = torch.allclose(torch.cat(parts, dim=1), full, atol=1e-6)
        if rank == 0:
            ret.put(bool(ok))
    finally:
        dist.destroy_process_group()


def test_two_rank_head_sharding_needs_no_collective():
    """world_size 2 over gloo: attention on each rank's KV-head shard, concatenated, equals the unsharded result."""
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_tp_worker, args=(r, 2, port, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
        assert p.exitcode == 0
    assert ret.get(timeout=10) is True


# ----------------------------------------------------------------------------- o_proj + all-reduce host logic (8f-4)
def test_oproj_tile_plan_and_symmetric_layout():
    from sglang_b200 import _C
    from sglang_b200.comm import num_tiles, symmetric_layout, tile_owner

    assert num_tiles(64, 8192) == 128 and num_tiles(129, 8192) == 256 and num_tiles(1, 64) == 1
    assert num_tiles(64, 8192) == _C.lib.b200attn_oproj_allreduce_tiles(64, 8192)
    assert [tile_owner(t, 8) for t in range(10)] == [0, 1, 2, 3, 4, 5, 6, 7, 0, 1]
    lay = symmetric_layout(128, 8192, 2, num_tiles(128, 8192))
    one = 128 * 8192 * 2
    # two staging buffers (alternating by launch), the output, the flag area; 1 KiB-aligned regions, no overlap
    assert lay["staging"] == 0 and lay["staging_bytes"] >= one and lay["out"] >= 2 * lay["staging_bytes"]
    assert lay["flags"] >= lay["out"] + one and lay["total"] >= lay["flags"] + 4 * _C.lib.b200attn_oproj_allreduce_flag_words(128)
    assert all(v % 1024 == 0 for v in lay.values())
    assert _C.lib.b200attn_oproj_allreduce_flag_words(128) == 128 * 9 + 2


def _oproj_worker(rank, world, port, ret):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from sglang_b200.comm import oproj_allreduce_ref

        g = torch.Generator().manual_seed(3)
        T, N, K = 9, 64, 128 * world
        x = torch.randn(T, K, generator=g).bfloat16()            # the attention output, all heads
        w = torch.randn(N, K, generator=g).bfloat16()            # o_proj weight [hidden, Hq * D]
        ks = K // world
        xs = [x[:, r * ks : (r + 1) * ks] for r in range(world)]  # RowParallelLinear: input split along K
        ws = [w[:, r * ks : (r + 1) * ks] for r in range(world)]
        # the reference's data flow on this rank: local GEMM in the model dtype, then all-reduce
        part = (xs[rank].double() @ ws[rank].double().t()).float().bfloat16().float()
        dist.all_reduce(part)                                      # gloo sums in float32, rank order 0 + 1
        want = oproj_allreduce_ref(xs, ws)
        ok = torch.equal(part.bfloat16(), want)
        # and the sharded sum approximates the unsharded projection
        full = (x.double() @ w.double().t()).float()
        ok = ok and float((want.float() - full).abs().max()) <= 2.0 ** -7 * float(full.abs().max()) * 2
        if rank == 0:
            ret.put(bool(ok))
    finally:
        dist.destroy_process_group()


def test_two_rank_oproj_allreduce_reference_flow():
    """world_size 2 over gloo: the definition the fused kernel is held to (per-rank partial in the model dtype,
    fp32 sum, one rounding) equals RowParallelLinear + all-reduce as the reference runs it."""
    ctx = mp.get_context("spawn")
    ret = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_oproj_worker, args=(r, 2, port, ret)) for r in range(2)]
    for p in procs:
        p.start()
    for p in procs:
        p.join(timeout=180)
        assert p.exitcode == 0
    assert ret.get(timeout=10) is True
