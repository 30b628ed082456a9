"""gt_score_nodes_static: the scoring pass for callers that cannot allocate device memory.

An extender that lives inside a process whose memory pool already owns the GPU, or a scheduling loop
that must not trigger the device-wide synchronisation of cudaMalloc/cudaFree, hands the library
buffers it already owns; the kernels' scratch sits in a fixed array inside libgputopo.so.  This test
plays such a caller: the device buffers come from torch, the library allocates nothing, and the result
must equal what the regular entry point (gt_score_all) computes for the same nodes -- bit for bit, as
in the other parity tests -- on two calls in a row (the second sees the scratch the first left behind).

Collected last on purpose (file name), so a run with -x has already executed every other test."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def reference_best(gt, levels, n_gpus, free):
    with gt.Allocator(len(n_gpus)) as a:
        a.load_nodes(levels, n_gpus, free)
        return a.score_all()


def test_static_scratch_entry_point_matches_score_all_twice(gt):
    import torch
    N = gt.STATIC_MAX_NODES
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = torch.cuda.Stream()
    results, wants = [], []
    inputs = []
    for seed, n_nodes in ((0x5EED0004, N), (0xABCDEF, N - 1357)):
        lv, n, fr, _ = gt.synth_cluster(4, seed, 0, n_nodes)
        # default cost table: cost == level code, so the level tiles are the cost tiles
        d_cost = torch.from_numpy(lv).to(dev)
        d_n = torch.from_numpy(n).to(dev)
        d_free = torch.from_numpy(fr.view(np.int16)).to(dev)
        d_best = torch.zeros((n_nodes, 16), dtype=torch.int32, device=dev)
        inputs.append((d_cost, d_n, d_free, d_best, n_nodes))
        wants.append(reference_best(gt, lv, n, fr))
    torch.cuda.synchronize()
    # two calls in a row on one stream, no host synchronisation in between
    for d_cost, d_n, d_free, d_best, n_nodes in inputs:
        gt.score_nodes_static(d_cost.data_ptr(), d_n.data_ptr(), d_free.data_ptr(), n_nodes, d_best.data_ptr(),
                              stream.cuda_stream)
    stream.synchronize()
    for (d_cost, d_n, d_free, d_best, n_nodes), want in zip(inputs, wants):
        got = d_best.cpu().numpy().view(np.uint32)
        assert got.shape == want.shape
        assert (got == want).all()
    # and once more on the first input: same answer again
    d_cost, d_n, d_free, d_best, n_nodes = inputs[0]
    d_best.zero_()
    torch.cuda.synchronize()
    gt.score_nodes_static(d_cost.data_ptr(), d_n.data_ptr(), d_free.data_ptr(), n_nodes, d_best.data_ptr(),
                          stream.cuda_stream)
    stream.synchronize()
    assert (d_best.cpu().numpy().view(np.uint32) == wants[0]).all()


def test_static_scratch_entry_point_rejects_oversize(gt):
    import torch
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    n_nodes = gt.STATIC_MAX_NODES + 1
    d_cost = torch.zeros((n_nodes, 256), dtype=torch.uint8, device=dev)
    d_n = torch.zeros(n_nodes, dtype=torch.uint8, device=dev)
    d_free = torch.zeros(n_nodes, dtype=torch.int16, device=dev)
    d_best = torch.zeros((n_nodes, 16), dtype=torch.int32, device=dev)
    with pytest.raises(gt.GtError) as e:
        gt.score_nodes_static(d_cost.data_ptr(), d_n.data_ptr(), d_free.data_ptr(), n_nodes, d_best.data_ptr())
    assert e.value.code == -5 and str(gt.STATIC_MAX_NODES) in str(e.value)
    # host memory is refused as output
    host = np.zeros((4, 16), np.uint32)
    with pytest.raises(gt.GtError):
        gt.score_nodes_static(d_cost.data_ptr(), d_n.data_ptr(), d_free.data_ptr(), 4, host.ctypes.data)
