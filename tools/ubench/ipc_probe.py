"""Probe: does CUDA IPC (cudaIpcGetMemHandle / cudaIpcOpenMemHandle) work between two torchrun ranks on this box, and how
fast are remote reads over NVLink from a plain copy kernel (torch copy from the mapped pointer)?"""
import ctypes as C
import os
import time

import torch
import torch.distributed as dist

rank = int(os.environ["RANK"])
torch.cuda.set_device(rank)
dist.init_process_group("nccl", device_id=torch.device("cuda", rank))
rt = C.CDLL("libcudart.so")


class Handle(C.Structure):
    _fields_ = [("reserved", C.c_ubyte * 64)]


rt.cudaIpcOpenMemHandle.argtypes = [C.POINTER(C.c_void_p), Handle, C.c_uint]
n = 1 << 30
raw = C.c_void_p()
assert rt.cudaMalloc(C.byref(raw), C.c_size_t(n)) == 0
assert rt.cudaMemset(raw, rank + 1, C.c_size_t(n)) == 0
h = Handle()
rc = rt.cudaIpcGetMemHandle(C.byref(h), raw)
print(rank, "get handle rc", rc, flush=True)
handles = [None, None]
dist.all_gather_object(handles, bytes(h))
peer = 1 - rank
ph = Handle.from_buffer_copy(handles[peer])
p = C.c_void_p()
rc = rt.cudaIpcOpenMemHandle(C.byref(p), ph, C.c_uint(1))
print(rank, "open rc", rc, hex(p.value or 0), flush=True)
if rc == 0:
    class V:
        pass
    v = V()
    v.__cuda_array_interface__ = {"shape": (n,), "typestr": "|u1", "data": (p.value, False), "version": 2}
    remote = torch.as_tensor(v, device="cuda")
    local = torch.empty(n, dtype=torch.uint8, device="cuda")
    dist.barrier()
    for _ in range(3):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        local.copy_(remote)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
    print(rank, "remote read", n / dt / 1e9, "GB/s; first byte", int(local[0]), flush=True)
dist.barrier()
