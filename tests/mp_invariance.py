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


def stage(rank, what):
    if rank == 0:
        print("mp_invariance: " + what, flush=True)


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
        oracle = oracle_lib.load()
        flv, fn, ffr, _ = gt.synth_cluster(4, 0x5EED0004, 0, total)

        def check(free_all, what):
            got = a.best_node(ks)
            if rank == 0:
                want = oracle.cluster_argmin(oracle.exhaustive_cluster(flv, fn, free_all, oracle.lut(0), 8)[0])
                assert (got == want).all(), (what, got, want)
            return got

        stage(rank, "communicator up, mode %d" % a.comm_mode())
        for _ in range(3):
            a.score_all(fetch=False)
            got = check(ffr, "full pass")
        stage(rank, "full passes ok")
        # many passes back to back without a consumer in between: both key buffers are reused again and again and
        # each exchange must have let go of its buffer before the pass two later resets it
        for _ in range(40):
            a.score_all_async()
        got = check(ffr, "40 passes back to back")
        stage(rank, "40 passes back to back ok")
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
        stage(rank, "new masks ok")
        got = check(cur, "dirty pass")
        stage(rank, "dirty pass ok")
        for _ in range(5):
            a.score_all_async()
        got = check(cur, "full passes after a dirty pass")
        # what every exchange delivered is the min over the ranks' own keys
        mine = a.best_node_local()
        tk = torch.from_numpy(mine.view(np.int64).copy()).cuda()
        allk = [torch.empty_like(tk) for _ in range(world)]
        dist.all_gather(allk, tk)
        assert (got == gt.merge_place([x.cpu().numpy().view(np.uint64) for x in allk])).all()
    t = torch.from_numpy(got.view(np.int64).copy()).cuda()
    ref = t.clone()
    dist.broadcast(ref, 0)
    assert bool((t == ref).all())
    dist.barrier()
    stage(rank, "single context ok")
    schedule_sharded(gt, rank, world, local)
    stage(rank, "sharded gt_schedule ok")
    shared_communicator(gt, rank, world, local)
    stage(rank, "shared communicator ok")
    if not os.environ.get("GT_NO_P2P"):
        different_pass_orders(gt, rank, world, local)
        stage(rank, "different pass orders ok")
    if rank == 0:
        print("mp_invariance ok: world=%d exchange=%s" % (world, "nccl" if os.environ.get("GT_NO_P2P") else "peer memory"))
    dist.destroy_process_group()


def new_comm(gt, a, rank, world):
    uid = torch.from_numpy(gt.comm_unique_id() if rank == 0 else np.zeros(128, np.uint8)).cuda()
    dist.broadcast(uid, 0)
    a.comm_init(uid.cpu().numpy(), rank, world)


def schedule_sharded(gt, rank, world, local):
    """gt_schedule on a sharded cluster: the config-2 trace (arrivals) and a config-5 trace (arrivals and departures)
    bit-exact against the oracle's sequential replay on every rank; the device tables follow."""
    o = oracle_lib.load()
    total = 1000
    lo, hi = gt.shard_bounds(total, rank, world)
    flv, fn, ffr, _ = gt.synth_cluster(2, 0x5EED0002, 0, total)
    for config, seed, n_ev in ((2, 0x5EED0012, 10000), (5, 0x5EED0005, 60000)):
        ek, ea = gt.synth_events(config, seed, n_ev)
        with gt.Allocator(hi - lo, device=local, first_node=lo) as a:
            a.load_nodes(flv[lo:hi], fn[lo:hi], ffr[lo:hi])
            new_comm(gt, a, rank, world)
            placed, on, om = a.schedule(ek, ea)
            wp, won, wom, wfree = o.replay(flv, fn, ffr, o.lut(0), ek, ea)
            assert placed == wp and (on == won).all() and (om == wom).all(), (config, placed, wp)
            # the shard's tables followed the loop: a pass now scores the final masks
            got = a.score_all()
            want, _ = o.exhaustive_cluster(flv[lo:hi], fn[lo:hi], wfree[lo:hi], o.lut(0), 2)
            assert (got == want).all()
            keys = a.best_node(np.arange(1, 17, dtype=np.uint8))
            wall, _ = o.exhaustive_cluster(flv, fn, wfree, o.lut(0), 4)
            assert (keys == o.cluster_argmin(wall)).all()
    dist.barrier()


def shared_communicator(gt, rank, world, local):
    """Several contexts of a rank on one communicator (gt_comm_share): passes interleaved in the same order on
    every rank, each context's keys are its own cluster's."""
    o = oracle_lib.load()
    total = 6000
    lo, hi = gt.shard_bounds(total, rank, world)
    ctxs, wants = [], []
    for r in range(3):
        lv, n, fr, _ = gt.synth_cluster(4, 0xC0FFEE + r, lo, hi - lo)
        a = gt.Allocator(hi - lo, device=local, first_node=lo)
        a.load_nodes(lv, n, fr)
        if r == 0:
            new_comm(gt, a, rank, world)
        else:
            a.comm_share(ctxs[0])
        ctxs.append(a)
        flv, fn, ffr, _ = gt.synth_cluster(4, 0xC0FFEE + r, 0, total)
        wants.append(o.cluster_argmin(o.exhaustive_cluster(flv, fn, ffr, o.lut(0), 4)[0]))
    ks = np.arange(1, 17, dtype=np.uint8)
    for it in range(12):
        for a in ctxs:
            a.score_all_async()
        for a, want in zip(ctxs, wants):
            assert (a.best_node(ks) == want).all(), it
    for a in reversed(ctxs):
        a.close()
    dist.barrier()


def different_pass_orders(gt, rank, world, local):
    """Two contexts with a communicator each; the ranks issue their passes in different orders.  Nothing may hang:
    every wait of the exchange is bounded, and a call that could not get the cluster's keys fails with 'call out of
    order'."""
    ctxs = []
    for r in range(2):
        lo, hi = gt.shard_bounds(2000, rank, world)
        lv, n, fr, _ = gt.synth_cluster(4, 0xBAD0 + r, lo, hi - lo)
        a = gt.Allocator(hi - lo, device=local, first_node=lo)
        a.load_nodes(lv, n, fr)
        new_comm(gt, a, rank, world)
        ctxs.append(a)
    if ctxs[0].comm_mode() != 2:      # NCCL fallback: its collectives have no time limit, do not provoke them
        for a in ctxs:
            a.close()
        return
    order = ctxs if rank % 2 == 0 else ctxs[::-1]
    for a in order:
        a.score_all_async()
    outcomes = []
    for a in ctxs:
        try:
            a.best_node(np.arange(1, 17, dtype=np.uint8))
            outcomes.append("ok")
        except gt.GtError as e:
            assert e.code == -8, e
            outcomes.append("state")
    # it returned: that is the point.  (With the exchange riding on the passes of its own context, mismatched orders
    # across contexts do not even block; a rank that never shows up would end in "state" after the time limit.)
    assert all(o in ("ok", "state") for o in outcomes), outcomes
    for a in ctxs:
        a.close()
    dist.barrier()
    if rank == 0:
        print("different pass orders: %s" % outcomes)


if __name__ == "__main__":
    main()
