"""GPU parity: the o_proj GEMM of the fused GEMM + all-reduce kernel on ONE GPU (TP 1), through
the C ABI, and -- when the box has two or more GPUs -- the exchange itself: benchmarks/multi/oproj_ar.py under
torchrun with 2 ranks (one-shot / two-shot, multimem / peer loads, repeated calls, CUDA-graph replays), every rank's
output within an ulp of the fp64 reference and bit-identical across ranks.  bench.py --gpus N re-checks the fused
kernel against cuBLAS + NCCL inside every scaling run."""
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch

REPO = Path(__file__).resolve().parent.parent

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


def test_oproj_allreduce_two_ranks(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (runs in the multi-GPU tiers; bench.py --gpus N checks the exchange as well)")
    out = tmp_path / "oproj_tp2.json"
    env = {k: v for k, v in os.environ.items() if k not in ("RANK", "WORLD_SIZE", "LOCAL_RANK")}
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", "29541", str(REPO / "benchmarks" / "multi" / "oproj_ar.py"), "--json",
           str(out), "--parity-only"]
    proc = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=900)
    assert proc.returncode == 0, proc.stdout[-3000:] + proc.stderr[-3000:]
    res = json.loads(out.read_text())
    assert res["ok"] and res["world"] == 2
    assert len(res["parity"]) >= 20 and all(r["within_ulp"] and r["identical_on_all_ranks"] for r in res["parity"])
    assert any(r.get("graph_replay") for r in res["parity"])
