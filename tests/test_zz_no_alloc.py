"""gt_score_nodes_static: the scoring pass for callers that cannot allocate device memory.

An extender that lives inside a process whose memory pool already owns the GPU, or a scheduling loop
that must not trigger the device-wide synchronisation of cudaMalloc/cudaFree, hands the library
buffers it already owns; the kernels' scratch sits in a fixed array inside libgputopo.so.  This test
plays such a caller: the device buffers come from torch, the library allocates nothing, and the result
must equal what the regular entry point (gt_score_all) computes for the same nodes AND what the CPU
oracle computes -- bit for bit, as in the other parity tests -- on two calls in a row (the second sees
the scratch the first left behind).  The first call on a device may reserve memory (the runtime loads
the library's device code then); from the second call on the device's free-memory figure must not move.

Collected last on purpose (file name), so a run with -x has already executed every other test."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def reference_best(gt, levels, n_gpus, free):
    with gt.Allocator(len(n_gpus)) as a:
        a.load_nodes(levels, n_gpus, free)
        return a.score_all()


def free_bytes():
    import torch
    return torch.cuda.mem_get_info()[0]


def test_static_scratch_entry_point_matches_score_all_twice(gt, oracle):
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
        want = reference_best(gt, lv, n, fr)
        # the checker of last resort is the CPU oracle, not another CUDA entry point
        assert (want == oracle.exhaustive_cluster(lv, n, fr, oracle.lut(0), 4)[0]).all()
        wants.append(want)
    torch.cuda.synchronize()
    # two calls in a row on one stream, no host synchronisation in between
    frees = []
    for d_cost, d_n, d_free, d_best, n_nodes in inputs:
        gt.score_nodes_static(d_cost.data_ptr(), d_n.data_ptr(), d_free.data_ptr(), n_nodes, d_best.data_ptr(),
                              stream.cuda_stream)
        frees.append(free_bytes())      # after the 1st call (set-up done), after the 2nd
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
    frees.append(free_bytes())          # after the 3rd
    stream.synchronize()
    assert (d_best.cpu().numpy().view(np.uint32) == wants[0]).all()
    # nothing was reserved on the device by the 2nd and 3rd calls
    assert frees[1] == frees[0] and frees[2] == frees[0], frees


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
    # host memory is refused, as output and as any of the inputs, before anything is launched
    host = np.zeros((4, 16), np.uint32)
    hin = np.zeros(4 * 256, np.uint8)
    ptrs = [d_cost.data_ptr(), d_n.data_ptr(), d_free.data_ptr(), d_best.data_ptr()]
    for i, bad in enumerate((hin.ctypes.data, hin.ctypes.data, hin.ctypes.data, host.ctypes.data)):
        args = list(ptrs)
        args[i] = bad
        with pytest.raises(gt.GtError) as e:
            gt.score_nodes_static(args[0], args[1], args[2], 4, args[3])
        assert e.value.code == -1 and "not device memory" in str(e.value)
    torch.cuda.synchronize()
