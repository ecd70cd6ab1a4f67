"""Full-size parity: BASELINE.json configs[2..4] at their real context lengths, and a layer buffer larger than 4 GiB.

The oracle (oracle/torch_native_ref, pinned to the reference by tests/test_oracle.py) runs on a request / query-row
SUBSET so the CPU side finishes in seconds; every comparison asserts the contractual max-abs <= 2e-2 and the
scale-aware rms(err)/rms(ref) <= 1e-2 (helpers.assert_close_scaled).  SURVEY section 8d sizes:
  cfg3  8B heads, FP8-e4m3 pool, bs 256 x 8192 tokens
  cfg4  70B shards (32,4) / (16,2) / (8,1) at 32768 tokens (bs 8 here; the kernels are batch-agnostic)
  cfg5  16384-token shared prefix + 1024 new tokens (first / last 64 query rows of two requests checked)
  > 4 GiB: a 2.2 M-slot bf16 pool (4.5 GB per K / V buffer), requests living in the top slots -- the kernels'
          slot * row_bytes arithmetic must be 64-bit (a production 8B pool on a 180 GB B200 exceeds 4 GiB per layer)
"""
import pytest
import torch

from helpers import assert_close_scaled
from oracle import torch_native_ref as ref

pytestmark = pytest.mark.gpu


def _compact(rtt_rows, lens, kp_dev, vp_dev):
    """CPU copy of only the pool rows the given requests touch, with a page table renumbered onto it."""
    used = torch.cat([rtt_rows[i, :L] for i, L in enumerate(lens)]).long()
    uniq, inv = torch.unique(used, return_inverse=True)
    kp = kp_dev[uniq.cuda()].cpu()
    vp = vp_dev[uniq.cuda()].cpu()
    rtt = torch.zeros(len(lens), max(lens), dtype=torch.int32)
    off = 0
    for i, L in enumerate(lens):
        rtt[i, :L] = inv[off : off + L].int()
        off += L
    return kp, vp, rtt


def _decode_problem(bs, L, Hq, Hkv, D, pool_dtype, seed, slot_base=0, extra_rows=0):
    from sglang_b200 import ops

    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(seed)
    rows = slot_base + bs * L + 1 + extra_rows
    fp8 = pool_dtype in (torch.float8_e4m3fn, torch.float8_e5m2)
    if fp8:
        kp = torch.empty(rows, Hkv, D, device=dev, dtype=torch.uint8)
        vp = torch.empty(rows, Hkv, D, device=dev, dtype=torch.uint8)
        step = 1 << 18
        for r0 in range(0, rows, step):                     # bounded temporaries
            r1 = min(rows, r0 + step)
            kp[r0:r1] = torch.randn(r1 - r0, Hkv, D, generator=g, device=dev).to(pool_dtype).view(torch.uint8)
            vp[r0:r1] = torch.randn(r1 - r0, Hkv, D, generator=g, device=dev).to(pool_dtype).view(torch.uint8)
    else:
        kp = torch.empty(rows, Hkv, D, device=dev, dtype=torch.bfloat16).normal_(generator=g)
        vp = torch.empty(rows, Hkv, D, device=dev, dtype=torch.bfloat16).normal_(generator=g)
    perm = torch.randperm(bs * L, generator=g, device=dev) + 1 + slot_base
    rtt = torch.zeros(bs + 1, L, dtype=torch.int32, device=dev)
    rtt[1:] = perm.view(bs, L).int()
    q = torch.randn(bs, Hq, D, generator=g, device=dev, dtype=torch.bfloat16)
    rpi = torch.arange(1, bs + 1, device=dev)
    lens = torch.full((bs,), L, dtype=torch.int32, device=dev)
    sm, _, _ = ops.device_info()
    ws = ops.DecodeWorkspace.create(bs, Hq, D, dev, num_ctas=sm * ops.decode_ctas_per_sm(D, pool_dtype))
    ops.decode_plan(ws, lens, Hkv,
                    items_per_chunk=ops.decode_items_per_chunk(Hq, Hkv, D, D, torch.bfloat16, pool_dtype))
    o = torch.empty_like(q)
    ops.decode_attention(q, o, kp, vp, rtt, rpi, lens, ws, D ** -0.5, kv_dtype=pool_dtype)
    torch.cuda.synchronize()
    return q, o, kp, vp, rtt, lens


def _check_decode_subset(q, o, kp, vp, rtt, L, sub, pool_dtype, what):
    rows = rtt[1:].cpu()[sub]
    kc, vc, rt = _compact(rows, [L] * len(sub), kp, vp)
    want = ref.decode_ref(q[sub].cpu(), kc, vc, rt, torch.arange(len(sub)), torch.full((len(sub),), L), q.shape[2] ** -0.5,
                          kv_dtype=pool_dtype)
    assert_close_scaled(o[sub], want, what)


def test_cfg3_full_size_fp8_decode():
    bs, L, Hq, Hkv, D = 256, 8192, 32, 8, 128
    q, o, kp, vp, rtt, _ = _decode_problem(bs, L, Hq, Hkv, D, torch.float8_e4m3fn, seed=3)
    _check_decode_subset(q, o, kp, vp, rtt, L, [0, 101, 255], torch.float8_e4m3fn, "cfg3")
    # size-independent property: a constant V makes every output equal that constant whatever the keys are
    vp.fill_(torch.tensor(0.5).to(torch.float8_e4m3fn).view(torch.uint8).item())
    from sglang_b200 import ops

    sm, _, _ = ops.device_info()
    ws = ops.DecodeWorkspace.create(bs, Hq, D, "cuda", num_ctas=sm * ops.decode_ctas_per_sm(D, torch.float8_e4m3fn))
    lens = torch.full((bs,), L, dtype=torch.int32, device="cuda")
    ops.decode_plan(ws, lens, Hkv, items_per_chunk=ops.decode_items_per_chunk(Hq, Hkv, D, D, torch.bfloat16,
                                                                              torch.float8_e4m3fn))
    o2 = torch.empty_like(q)
    ops.decode_attention(q, o2, kp, vp, rtt, torch.arange(1, bs + 1, device="cuda"), lens, ws, D ** -0.5,
                         kv_dtype=torch.float8_e4m3fn)
    torch.cuda.synchronize()
    assert float((o2.float() - 0.5).abs().max()) <= 2e-3


@pytest.mark.parametrize("Hq,Hkv", [(32, 4), (16, 2), (8, 1)], ids=["tp2", "tp4", "tp8"])
def test_cfg4_shards_at_32k(Hq, Hkv):
    bs, L, D = 8, 32768, 128
    q, o, kp, vp, rtt, _ = _decode_problem(bs, L, Hq, Hkv, D, torch.bfloat16, seed=Hq)
    _check_decode_subset(q, o, kp, vp, rtt, L, [0, 7], torch.bfloat16, f"cfg4 ({Hq},{Hkv})")


def test_cfg4_tp_shards_concatenate_to_the_unsharded_result():
    """One 70B-headed problem (64 q / 8 kv heads), sharded with tp.shard_heads for TP 2 / 4 / 8: the CUDA kernels run
    per shard on that rank's head slice of the pool; the concatenation must equal the UNSHARDED oracle."""
    from sglang_b200 import ops
    from sglang_b200.tp import shard_heads

    bs, L, Hq, Hkv, D = 3, 4500, 64, 8, 128
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(9)
    rows = bs * L + 1
    kp = torch.randn(rows, Hkv, D, generator=g, device=dev, dtype=torch.bfloat16)
    vp = torch.randn(rows, Hkv, D, generator=g, device=dev, dtype=torch.bfloat16)
    rtt = torch.zeros(bs + 1, L, dtype=torch.int32, device=dev)
    rtt[1:] = (torch.randperm(bs * L, generator=g, device=dev) + 1).view(bs, L).int()
    q = torch.randn(bs, Hq, D, generator=g, device=dev, dtype=torch.bfloat16)
    rpi = torch.arange(1, bs + 1, device=dev)
    lens = torch.tensor([L, L - 1777, 33], dtype=torch.int32, device=dev)
    want = ref.decode_ref(q.cpu(), kp.cpu(), vp.cpu(), rtt.cpu(), rpi.cpu(), lens.cpu(), D ** -0.5)
    sm, _, _ = ops.device_info()
    for tp in (2, 4, 8):
        parts = []
        for rank in range(tp):
            sh = shard_heads(Hq, Hkv, tp, rank)
            # the rank's pool holds only its kv-head slice of every token (model_config.py:1121-1128)
            kps = kp[:, sh.kv_head_start : sh.kv_head_start + sh.kv_heads].contiguous()
            vps = vp[:, sh.kv_head_start : sh.kv_head_start + sh.kv_heads].contiguous()
            qs = q[:, sh.q_head_start : sh.q_head_start + sh.q_heads].contiguous()
            ws = ops.DecodeWorkspace.create(bs, sh.q_heads, D, dev, num_ctas=sm * ops.decode_ctas_per_sm(D, torch.bfloat16))
            ops.decode_plan(ws, lens, sh.kv_heads, items_per_chunk=ops.decode_items_per_chunk(
                sh.q_heads, sh.kv_heads, D, D, torch.bfloat16, torch.bfloat16))
            os_ = torch.empty_like(qs)
            ops.decode_attention(qs, os_, kps, vps, rtt, rpi, lens, ws, D ** -0.5)
            parts.append(os_)
        torch.cuda.synchronize()
        assert_close_scaled(torch.cat(parts, dim=1), want, f"tp{tp}")


def test_cfg5_full_size_shared_prefix_extend():
    """16384 shared-prefix tokens (identical slots in every request) + 1024 private new tokens; 4 of the 128 requests
    are run (the kernel treats requests independently), the first and last 64 query rows of two of them checked."""
    from sglang_b200 import ops

    bs, P, E, Hq, Hkv, D = 4, 16384, 1024, 32, 8, 128
    dev = "cuda"
    g = torch.Generator(device=dev).manual_seed(5)
    rows = P + bs * E + 1
    kp = torch.randn(rows, Hkv, D, generator=g, device=dev, dtype=torch.bfloat16)
    vp = torch.randn(rows, Hkv, D, generator=g, device=dev, dtype=torch.bfloat16)
    perm = torch.randperm(rows - 1, generator=g, device=dev).int() + 1
    rtt = torch.zeros(bs + 1, P + E, dtype=torch.int32, device=dev)
    rtt[1:, :P] = perm[:P]
    rtt[1:, P:] = perm[P : P + bs * E].view(bs, E)
    q = torch.randn(bs * E, Hq, D, generator=g, device=dev, dtype=torch.bfloat16)
    o = torch.empty_like(q)
    rpi = torch.arange(1, bs + 1, device=dev)
    pre = torch.full((bs,), P, dtype=torch.int32, device=dev)
    ext = torch.full((bs,), E, dtype=torch.int32, device=dev)
    sm, _, _ = ops.device_info()
    bq = ops.extend_block_q(D, D, Hq // Hkv, torch.bfloat16, torch.bfloat16)
    ws = ops.ExtendWorkspace.create(bs, bs * E // bq + bs + 1, dev, num_ctas=sm)
    ops.extend_plan(ws, ext, bq)
    ops.extend_attention(q, o, kp, vp, rtt, rpi, pre, ext, ws, D ** -0.5)
    torch.cuda.synchronize()
    kc, vc, rt = kp.cpu(), vp.cpu(), rtt.cpu()
    n = 64
    for r in (0, bs - 1):
        rows_q = q[r * E : (r + 1) * E].cpu()
        # first n rows == an extend of n tokens over P; last n rows == an extend of n tokens over P + E - n
        for name, pre_len, sl in (("first", P, slice(0, n)), ("last", P + E - n, slice(E - n, E))):
            want = ref.extend_ref(rows_q[sl], kc, vc, rt, torch.tensor([r + 1]), torch.tensor([pre_len]),
                                  torch.tensor([n]), D ** -0.5)
            assert_close_scaled(o[r * E : (r + 1) * E][sl], want, f"req {r} {name} {n} rows")


def test_layer_buffer_larger_than_4gib():
    """Store + decode + extend against slots whose byte offsets exceed 2^32 inside one K / V layer buffer."""
    from sglang_b200 import ops

    Hq, Hkv, D = 32, 8, 128                      # 2 KiB per row: offset 2^32 is slot 2 097 152
    base = (1 << 21) + 4096                      # every slot this test touches lies beyond 4 GiB
    bs, L = 4, 2048
    q, o, kp, vp, rtt, lens = _decode_problem(bs, L, Hq, Hkv, D, torch.bfloat16, seed=41, slot_base=base)
    assert kp.numel() * 2 > (1 << 32) and int(rtt[1:].min()) * Hkv * D * 2 > (1 << 32)
    _check_decode_subset(q, o, kp, vp, rtt, L, [0, 3], torch.bfloat16, "decode > 4 GiB")
    # store: rewrite the last token of every request, bit-exact
    g = torch.Generator(device="cuda").manual_seed(42)
    loc = rtt[1:, L - 1].long().contiguous()
    k_new = torch.randn(bs, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
    v_new = torch.randn(bs, Hkv, D, generator=g, device="cuda", dtype=torch.bfloat16)
    ops.store_kv(k_new, v_new, kp, vp, loc)
    torch.cuda.synchronize()
    assert torch.equal(kp[loc], k_new) and torch.equal(vp[loc], v_new)
    # extend: prefix 1500 + 548 new tokens per request over the same high slots
    P, E = 1500, L - 1500
    qe = torch.randn(bs * E, Hq, D, generator=g, device="cuda", dtype=torch.bfloat16)
    oe = torch.empty_like(qe)
    pre = torch.full((bs,), P, dtype=torch.int32, device="cuda")
    ext = torch.full((bs,), E, dtype=torch.int32, device="cuda")
    sm, _, _ = ops.device_info()
    bq = ops.extend_block_q(D, D, Hq // Hkv, torch.bfloat16, torch.bfloat16)
    ws = ops.ExtendWorkspace.create(bs, bs * E // bq + bs + 1, "cuda", num_ctas=sm)
    ops.extend_plan(ws, ext, bq)
    ops.extend_attention(qe, oe, kp, vp, rtt, torch.arange(1, bs + 1, device="cuda"), pre, ext, ws, D ** -0.5)
    torch.cuda.synchronize()
    sub = [1]
    kc, vc, rt = _compact(rtt[1:].cpu()[sub], [L], kp, vp)
    want = ref.extend_ref(qe[E : 2 * E].cpu(), kc, vc, rt, torch.tensor([0]), torch.tensor([P]), torch.tensor([E]),
                          D ** -0.5)
    assert_close_scaled(oe[E : 2 * E], want, "extend > 4 GiB")
