"""World-size-2 gloo test of the multi-GPU host logic: contiguous node shards,
global node indices inside the placement keys, min-exchange of 16 keys.  The per-shard tables come from
the oracle here (no GPU in this container); the GPU version of the same check is in test_gpu_parity.py."""
import os
import socket
import sys

import numpy as np
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, total, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import _pkg
    import oracle_lib
    gt = _pkg.load_package()
    oracle = oracle_lib.load()
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    lo, hi = gt.shard_bounds(total, rank, world)
    lv, n, fr, _ = gt.synth_cluster(4, 0x5EED0004, lo, hi - lo)
    best, _ = oracle.exhaustive_cluster(lv, n, fr, oracle.lut(0), 2)
    local = oracle.cluster_argmin(best, first=lo)
    merged = gt.allreduce_place_torch(local)
    np.save(os.path.join(out_dir, "merged_%d.npy" % rank), merged)
    dist.destroy_process_group()


def test_two_rank_min_exchange_matches_single_shard(gt, oracle, tmp_path):
    total = 3001          # odd on purpose: uneven shards
    port = _free_port()
    mp.spawn(_worker, args=(2, port, total, str(tmp_path)), nprocs=2, join=True)
    lv, n, fr, _ = gt.synth_cluster(4, 0x5EED0004, 0, total)
    best, _ = oracle.exhaustive_cluster(lv, n, fr, oracle.lut(0), 4)
    want = oracle.cluster_argmin(best)
    for r in range(2):
        got = np.load(os.path.join(str(tmp_path), "merged_%d.npy" % r))
        assert (got == want).all()
    # k = 9..16 can only be served by 16-GPU nodes; some k have no feasible node at all
    assert (want[:8] != gt.NO_PLACE).all()


def test_shard_bounds_cover_everything(gt):
    for total in (1, 7, 100000, 100003):
        for world in (1, 2, 4, 8):
            edges = [gt.shard_bounds(total, r, world) for r in range(world)]
            assert edges[0][0] == 0 and edges[-1][1] == total
            assert all(edges[i][1] == edges[i + 1][0] for i in range(world - 1))


def test_merge_place_keeps_no_place(gt):
    a = np.full(16, gt.NO_PLACE, np.uint64)
    b = a.copy()
    b[3] = (5 << 48) | (7 << 16) | 0x3
    m = gt.merge_place([a, b])
    assert int(m[3]) == int(b[3]) and int(m[0]) == gt.NO_PLACE
