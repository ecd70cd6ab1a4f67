import torch, sys
sys.path.insert(0, '.')
from sglang_b200 import ops
sm,_,_ = ops.device_info()
for bs, L in ((64, 4096), (1, 4096), (256, 8192), (32, 32768)):
    ws = ops.DecodeWorkspace.create(bs, 32, 128, "cuda", num_ctas=sm*2)
    lens = torch.full((bs,), L, dtype=torch.int32, device="cuda")
    for _ in range(5): ops.decode_plan(ws, lens, 8, items_per_chunk=2)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(20): ops.decode_plan(ws, lens, 8, items_per_chunk=2)
    g.replay(); torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10): g.replay()
    b.record(); torch.cuda.synchronize()
    print(f"plan bs {bs} L {L}: {a.elapsed_time(b)/200*1e3:.2f} us per launch (graph replay of 20), chunk {int(ws.plan[0])} chunks {int(ws.plan[1])}")
