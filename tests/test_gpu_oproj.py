"""GPU parity: the o_proj GEMM of the fused GEMM + all-reduce kernel on ONE GPU (TP 1), through
the C ABI.  The multi-GPU exchange is covered by benchmarks/multi/oproj_ar.py under torchrun (2+ GPUs) and by the
world-size-2 gloo test of the host logic in tests/test_host_logic.py."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float16])
@pytest.mark.parametrize("T,N,K", [(1, 64, 128), (64, 4096, 4096), (37, 8192, 1024), (128, 512, 8192),
                                   (129, 256, 384), (300, 19 * 64, 640), (1000, 4096, 512)])
def test_oproj_single_gpu_matches_fp64(T, N, K, dtype):
    from sglang_b200.comm import OProjAllReduce, oproj_allreduce_ref

    g = torch.Generator().manual_seed(T + N + K)
    x = (torch.randn(T, K, generator=g) * 0.5).to(dtype)
    w = (torch.randn(N, K, generator=g) * 0.5).to(dtype)
    comm = OProjAllReduce(None, max_tokens=max(T, 128), out_features=N, dtype=dtype, device="cuda")
    got = comm.forward(x.cuda(), w.cuda())
    torch.cuda.synchronize()
    want, mag = oproj_allreduce_ref([x], [w], with_scale=True)
    # fp32 accumulation in a different order than the float64 reference: one ulp of the model dtype
    ulp = 2.0 ** -8 if dtype == torch.bfloat16 else 2.0 ** -11
    err = (got.float().cpu() - want.float()).abs()
    tol = ulp * mag.clamp_min(1.0) * 2
    assert bool((err <= tol).all()), f"max err {float(err.max())}"
    # strided input rows (x is a column slice of a wider activation) and a second call on the same communicator
    wide = torch.zeros(T, K + 64, dtype=dtype, device="cuda")
    wide[:, 32 : 32 + K] = x.cuda()
    if (32 * 2) % 16 == 0:
        got2 = comm.forward(wide[:, 32 : 32 + K], w.cuda())
        torch.cuda.synchronize()
        assert torch.equal(got2.cpu(), got.cpu())


def test_oproj_rejects_bad_shapes():
    from sglang_b200 import _C
    from sglang_b200.comm import OProjAllReduce

    with pytest.raises(ValueError):
        OProjAllReduce(None, 128, 100, torch.bfloat16, "cuda")
    comm = OProjAllReduce(None, 128, 128, torch.bfloat16, "cuda")
    x = torch.zeros(4, 96, dtype=torch.bfloat16, device="cuda")
    w = torch.zeros(128, 96, dtype=torch.bfloat16, device="cuda")
    with pytest.raises(_C.B200AttnError):
        comm.forward(x, w)                       # K must be a multiple of 128
    with pytest.raises(RuntimeError):
        comm.forward(x.cpu(), w.cpu())
