#!/usr/bin/env python
"""trace_pass.py under torchrun with the communicator attached (GT_TRACE build through GT_LIB)."""
import ctypes as C
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import _pkg
gt = _pkg.load_package()
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
N = 100000
lv, ng, fr, _ = gt.synth_cluster(4, 0x5EED0004, rank * N, N)
with gt.Allocator(N, device=lr, first_node=rank * N) as a:
    a.load_nodes(lv, ng, fr)
    uid = torch.from_numpy(gt.comm_unique_id() if rank == 0 else np.zeros(128, np.uint8)).cuda()
    dist.broadcast(uid, 0)
    a.comm_init(uid.cpu().numpy(), rank, world)
    L = gt._lib()
    L.gt_debug_ktrace.argtypes = [C.c_void_p, C.POINTER(C.c_uint64), C.POINTER(C.c_uint32)]
    buf = np.zeros((64, 8), np.uint64)
    ps = C.c_uint32(0)
    for _ in range(40):
        a.score_all_async()
    L.gt_debug_ktrace(a._h, buf.ctypes.data_as(C.POINTER(C.c_uint64)), C.byref(ps))
    p0 = ps.value
    dist.barrier()
    for _ in range(40):
        a.score_all_async()
    L.gt_debug_ktrace(a._h, buf.ctypes.data_as(C.POINTER(C.c_uint64)), C.byref(ps))
    rows = [buf[(p0 + i) & 63].astype(np.int64) for i in range(40)]
    if rank == 0:
        print("pass | K0 end-go | K1 go-K0end  K1 len | next K0 go - K1 end | period")
        for i in range(5, 20):
            r, nx = rows[i], rows[i + 1]
            print("%4d | %7d | %8d %8d | %8d | %8d" % (i, r[1] - r[5], r[3] - r[1], r[4] - r[3], nx[5] - r[4], nx[5] - r[5]))
    dist.barrier()
dist.destroy_process_group()
