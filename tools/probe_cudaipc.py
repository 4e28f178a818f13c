"""Probe: can rank A open rank B's cudaMalloc allocation with cudaIpc and write into it (peer window for the fused
scatter-exchange)? torchrun --nproc-per-node 2 tools/probe_cudaipc.py"""
import ctypes as C
import os

import torch
import torch.distributed as dist

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
torch.zeros(1, device="cuda")
libs = sorted({l.split()[-1] for l in open("/proc/self/maps") if "libcudart" in l})
print(rank, libs, flush=True)
rt = C.CDLL(libs[0])
N = 1 << 20
p = C.c_void_p()
assert rt.cudaMalloc(C.byref(p), N) == 0
assert rt.cudaMemset(p, rank + 1, N) == 0
class Handle(C.Structure):
    _fields_ = [("reserved", C.c_char * 64)]


h = Handle()
rc = rt.cudaIpcGetMemHandle(C.byref(h), p)
print(rank, "get handle rc", rc, flush=True)
handles = [None] * world
dist.all_gather_object(handles, bytes(h))
peer = (rank + 1) % world
q = C.c_void_p()
hb = Handle.from_buffer_copy(handles[peer])
rt.cudaIpcOpenMemHandle.argtypes = [C.POINTER(C.c_void_p), Handle, C.c_uint]
rc = rt.cudaIpcOpenMemHandle(C.byref(q), hb, 1)
print(rank, "open rc", rc, hex(q.value or 0), flush=True)
if rc == 0:
    src = torch.full((N,), 100 + rank, dtype=torch.uint8, device="cuda")
    rc = rt.cudaMemcpy(q, C.c_void_p(src.data_ptr()), N // 2, 3)
    print(rank, "peer write rc", rc, flush=True)
torch.cuda.synchronize()
dist.barrier()
torch.cuda.synchronize()
out = torch.empty(N, dtype=torch.uint8, device="cuda")
rt.cudaMemcpy(C.c_void_p(out.data_ptr()), p, N, 3)
print(rank, "own window now:", int(out[0]), int(out[-1]), "expected", 100 + (rank - 1) % world, rank + 1, flush=True)
dist.barrier()
if rc == 0:
    rt.cudaIpcCloseMemHandle(q)
dist.barrier()
rt.cudaFree(p)
dist.destroy_process_group()
