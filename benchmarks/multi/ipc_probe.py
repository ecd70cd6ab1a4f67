"""Probe: can two ranks on one box map each other's device memory with CUDA IPC?  (torchrun --nproc-per-node 2)"""
import ctypes as C
import os

import torch
import torch.distributed as dist

rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
dist.init_process_group("nccl")
rt = C.CDLL("libcudart.so.12")
ptr = C.c_void_p()
assert rt.cudaMalloc(C.byref(ptr), 1 << 20) == 0


class Handle(C.Structure):
    _fields_ = [("reserved", C.c_char * 64)]


handle = Handle()
rc = rt.cudaIpcGetMemHandle(C.byref(handle), ptr)
print(rank, "cudaIpcGetMemHandle rc", rc, flush=True)
assert rt.cudaMemset(ptr, rank + 1, 1 << 20) == 0
torch.cuda.synchronize()
handles = [None] * world
dist.all_gather_object(handles, bytes(C.string_at(C.byref(handle), 64)))
peer = (rank + 1) % world
pp = C.c_void_p()
h = Handle.from_buffer_copy(handles[peer])
rt.cudaIpcOpenMemHandle.argtypes = [C.POINTER(C.c_void_p), Handle, C.c_uint]
rc = rt.cudaIpcOpenMemHandle(C.byref(pp), h, 1)
print(rank, "cudaIpcOpenMemHandle rc", rc, hex(pp.value or 0), flush=True)
host = (C.c_ubyte * 16)()
rc = rt.cudaMemcpy(host, pp, 16, 2)
print(rank, "peer bytes", list(host)[:4], "rc", rc, flush=True)
can = C.c_int()
rt.cudaDeviceCanAccessPeer(C.byref(can), rank, peer)
print(rank, "canAccessPeer", can.value, flush=True)
t = torch.ones(4, device="cuda") * (rank + 1)
dist.all_reduce(t)
print(rank, "nccl allreduce", t[0].item(), flush=True)
try:
    import torch.distributed._symmetric_memory as symm
    b = symm.empty(1024, dtype=torch.float32, device="cuda")
    hdl = symm.rendezvous(b, dist.group.WORLD)
    print(rank, "symm_mem ok: multicast_ptr", hex(getattr(hdl, "multicast_ptr", 0) or 0), flush=True)
except Exception as e:  # noqa: BLE001
    print(rank, "symm_mem failed:", repr(e)[:300], flush=True)
dist.barrier()
dist.destroy_process_group()
