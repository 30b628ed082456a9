"""Run under torchrun on N GPUs: every rank scores its shard, the library's NCCL min-exchange must give
every rank the placement keys of the whole cluster (checked against the oracle on rank 0)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _pkg
import oracle_lib


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    gt = _pkg.load_package()
    total = 50000
    lo, hi = gt.shard_bounds(total, rank, world)
    lv, n, fr, _ = gt.synth_cluster(4, 0x5EED0004, lo, hi - lo)
    with gt.Allocator(hi - lo, device=local, first_node=lo) as a:
        a.load_nodes(lv, n, fr)
        uid = torch.from_numpy(gt.comm_unique_id() if rank == 0 else np.zeros(128, np.uint8)).cuda()
        dist.broadcast(uid, 0)
        a.comm_init(uid.cpu().numpy(), rank, world)
        ks = np.arange(1, 17, dtype=np.uint8)
        for _ in range(3):
            a.score_all(fetch=False)
            got = a.best_node(ks)
    if rank == 0:
        oracle = oracle_lib.load()
        flv, fn, ffr, _ = gt.synth_cluster(4, 0x5EED0004, 0, total)
        want = oracle.cluster_argmin(oracle.exhaustive_cluster(flv, fn, ffr, oracle.lut(0), 8)[0])
        assert (got == want).all(), (got, want)
    t = torch.from_numpy(got.view(np.int64).copy()).cuda()
    ref = t.clone()
    dist.broadcast(ref, 0)
    assert bool((t == ref).all())
    dist.barrier()
    if rank == 0:
        print("mp_invariance ok: world=%d" % world)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
