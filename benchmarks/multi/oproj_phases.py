"""Phase breakdown of the fused o_proj + all-reduce kernel from in-kernel %globaltimer stamps (tuning aid).
torchrun --nproc-per-node N benchmarks/multi/oproj_phases.py [T N K]"""
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

sys.path.insert(0, str(Path(__file__).resolve().parent.parent.parent))
from sglang_b200.comm import OProjAllReduce  # noqa: E402

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl", device_id=torch.device("cuda", torch.cuda.current_device()))
a = [int(v) for v in sys.argv[1:]] + [None] * 3
T, N, Kf = a[0] or 64, a[1] or 8192, a[2] or 8192
K = Kf // world
x = (torch.randn(T, K, device="cuda") * 0.5).bfloat16()
ws = [(torch.randn(N, K, device="cuda") * 0.5).bfloat16() for _ in range(6)]
for mc, mode in ((True, 1), (True, 2), (False, 1), (False, 2)):
    comm = OProjAllReduce(dist.group.WORLD, max(T, 128), N, torch.bfloat16, "cuda", use_multicast=mc, mode=mode)
    comm.debug_times = torch.zeros(comm.max_tiles, 8, dtype=torch.int64, device="cuda")
    for i in range(12):
        dist.barrier()
        comm.forward(x, ws[i % 6])
    torch.cuda.synchronize()
    t = comm.debug_times.cpu()
    nt = (T + 127) // 128 * (N // 64)
    t = t[:nt]
    t0 = t[:, 0].min()
    own = torch.arange(nt) % world == rank
    names = ["start", "gemm done", "flag sent", "peers arrived", "reduced", "fenced", "done seen", "partial stored"]
    if rank == 0:
        print(f"--- multicast={mc and comm.multicast} mode={'one-shot' if mode == 1 else 'two-shot'} T={T} N={N} K_shard={K} (us since the first CTA started, mean / max over tiles)")
        for i, n in enumerate(names):
            if mode == 1 and i in (4, 5):
                continue
            sel = t[own] if (i in (3, 4, 5) and mode == 2) else t
            v = (sel[:, i] - t0).float() / 1e3
            print(f"  {n:14s} mean {v.mean():7.2f}  max {v.max():7.2f}")
    del comm
dist.barrier()
dist.destroy_process_group()
