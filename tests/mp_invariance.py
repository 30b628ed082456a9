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
        oracle = oracle_lib.load() if rank == 0 else None
        flv, fn, ffr, _ = gt.synth_cluster(4, 0x5EED0004, 0, total)

        def check(free_all, what):
            got = a.best_node(ks)
            if rank == 0:
                want = oracle.cluster_argmin(oracle.exhaustive_cluster(flv, fn, free_all, oracle.lut(0), 8)[0])
                assert (got == want).all(), (what, got, want)
            return got

        for _ in range(3):
            a.score_all(fetch=False)
            got = check(ffr, "full pass")
        # many passes back to back without a consumer in between: both key buffers are reused again and again and
        # each exchange must have let go of its buffer before the pass two later resets it
        for _ in range(40):
            a.score_all_async()
        got = check(ffr, "40 passes back to back")
        # the free masks change between passes (every rank draws the whole cluster's masks from the same seed)
        rng = np.random.RandomState(7)
        for it in range(4):
            cur = (ffr & rng.randint(0, 1 << 16, total).astype(np.uint16)).astype(np.uint16)
            a.update_free(cur[lo:hi])
            a.score_all_async()
            a.score_all_async()
            got = check(cur, "new masks %d" % it)
        # dirty path (K0 + K1 over the touched nodes, K2, event-ordered exchange), then full passes again
        ids = np.arange(0, hi - lo, 97, dtype=np.uint32)
        cur = cur.copy()
        cur[lo + ids] = ffr[lo + ids]
        gcur = torch.from_numpy(cur.astype(np.int32)).cuda()
        dist.all_reduce(gcur, op=dist.ReduceOp.MAX)      # every rank restored masks in its own shard only
        cur = gcur.cpu().numpy().astype(np.uint16)
        a.update_free(cur[lo + ids], node_ids=ids)
        a.score_dirty()
        got = check(cur, "dirty pass")
        for _ in range(5):
            a.score_all_async()
        got = check(cur, "full passes after a dirty pass")
    t = torch.from_numpy(got.view(np.int64).copy()).cuda()
    ref = t.clone()
    dist.broadcast(ref, 0)
    assert bool((t == ref).all())
    dist.barrier()
    if rank == 0:
        print("mp_invariance ok: world=%d exchange=%s" % (world, "nccl" if os.environ.get("GT_NO_P2P") else "peer memory"))
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
