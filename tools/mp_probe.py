#!/usr/bin/env python
"""Multi-GPU pass: host enqueue time vs completion time (run under torchrun on the GPU box)."""
import os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import _pkg
gt = _pkg.load_package()
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr)
dist.init_process_group("nccl", device_id=torch.device("cuda", lr))
for N in (1000, 100000):
    lv, ng, fr, _ = gt.synth_cluster(4, 0x5EED0004, rank * N, N)
    with gt.Allocator(N, device=lr, first_node=rank * N) as a:
        a.load_nodes(lv, ng, fr)
        if os.environ.get("NO_COMM") is None:
            uid = torch.from_numpy(gt.comm_unique_id() if rank == 0 else np.zeros(128, np.uint8)).cuda()
            dist.broadcast(uid, 0)
            a.comm_init(uid.cpu().numpy(), rank, world)
        for _ in range(100):
            a.score_all_async()
        a.sync()
        dist.barrier()
        T = 400       # short enough not to fill the launch queue
        t0 = time.perf_counter()
        for _ in range(T):
            a.score_all_async()
        t1 = time.perf_counter()
        a.sync()
        t2 = time.perf_counter()
        if rank == 0:
            print("world %d N %7d  enqueue %.2f us/pass   until done %.2f us/pass" % (world, N, (t1 - t0) / T * 1e6, (t2 - t0) / T * 1e6), flush=True)
        dist.barrier()
dist.destroy_process_group()
