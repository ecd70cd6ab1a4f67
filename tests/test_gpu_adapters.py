"""GPU parity of the reference-signature adapters (sglang_b200/adapters.py) over the fused kernels."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("neox", [True, False])
@pytest.mark.parametrize("T,Hq,Hkv,D", [(37, 32, 8, 128), (5, 14, 2, 64)])
def test_fused_rope_adapter_equals_rotary_then_set_kv_buffer(T, Hq, Hkv, D, neox):
    from sglang_b200 import adapters, ops

    g = torch.Generator(device="cuda").manual_seed(3)
    dt = torch.bfloat16
    q = torch.randn(T, Hq, D, generator=g, device="cuda", dtype=dt)
    k = torch.randn(T, Hkv, D, generator=g, device="cuda", dtype=dt)
    v = torch.randn(T, Hkv * D, generator=g, device="cuda", dtype=dt)
    slots = 4 * T + 1
    kbuf = torch.zeros(slots, Hkv, D, device="cuda", dtype=dt)
    vbuf = torch.zeros_like(kbuf)
    loc = torch.randperm(slots - 1, generator=g, device="cuda")[:T] + 1
    pos = torch.randint(0, 4096, (T,), generator=g, device="cuda")
    inv = 1.0 / (10000.0 ** (torch.arange(0, D, 2, device="cuda").float() / D))
    fr = torch.arange(4096, device="cuda").float()[:, None] * inv[None]
    cache = torch.cat((fr.cos(), fr.sin()), dim=-1).contiguous()

    # the unfused sequence: rotate, then the pool's set_kv_buffer scatter
    q_ref, k_ref = q.clone(), k.clone()
    adapters.apply_rope_with_cos_sin_cache_inplace(q_ref, k_ref, cache, pos, is_neox=neox)
    k_pool_ref, v_pool_ref = torch.zeros_like(kbuf), torch.zeros_like(vbuf)
    ops.store_kv(k_ref, v.view(T, Hkv, D), k_pool_ref, v_pool_ref, loc)

    class Layer:
        layer_id, k_scale, v_scale = 0, None, None

    class Pool:
        def get_key_buffer(self, i): return kbuf
        def get_value_buffer(self, i): return vbuf

    class Batch:
        token_to_kv_pool, out_cache_loc = Pool(), loc

    fused = adapters.create_fused_set_kv_buffer_arg(v, Layer(), Batch())
    assert fused.k_buffer.shape == (slots, Hkv * D) and fused.k_buffer.data_ptr() == kbuf.data_ptr()
    adapters.apply_rope_with_cos_sin_cache_inplace(q, k, cache, pos, is_neox=neox, fused_args=fused)
    torch.cuda.synchronize()
    assert torch.equal(q, q_ref) and torch.equal(k, k_ref)
    assert torch.equal(kbuf, k_pool_ref) and torch.equal(vbuf, v_pool_ref)
    assert not torch.equal(q, torch.zeros_like(q)) and kbuf[loc].abs().sum() > 0


@pytest.mark.parametrize("T", [1, 64, 300])
def test_row_parallel_linear_tp1(T):
    from sglang_b200 import adapters

    g = torch.Generator(device="cuda").manual_seed(5)
    lin = adapters.B200RowParallelLinear(1024, 512, bias=True, tp_rank=0, tp_size=1, fused_max_tokens=128,
                                         max_tokens=128)
    lin.weight.data.copy_(torch.randn(512, 1024, generator=g, device="cuda", dtype=torch.bfloat16) * 0.05)
    lin.bias.data.copy_(torch.randn(512, generator=g, device="cuda", dtype=torch.bfloat16))
    x = torch.randn(T, 1024, generator=g, device="cuda", dtype=torch.bfloat16)
    out, out_bias = lin(x)
    ref = (x.double() @ lin.weight.double().t()).float().to(torch.bfloat16).float() + lin.bias.float()
    assert out_bias is None and out.shape == (T, 512)
    assert (out.float() - ref).abs().max().item() <= 2e-2 * max(1.0, ref.abs().max().item())
    lin.skip_bias_add = True
    out2, out_bias2 = lin(x)
    assert out_bias2 is lin.bias
    assert (out2.float() + lin.bias.float() - ref).abs().max().item() <= 2e-2 * max(1.0, ref.abs().max().item())
