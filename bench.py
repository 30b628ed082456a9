#!/usr/bin/env python
"""bench.py -- candidate GPU subsets scored per second on the topology-aware allocation path.

A step is one Prioritize pass (design.md:118) over this rank's shard of a synthetic cluster:
K0 k_bin (bin nodes by popcount(free)) -> K1 k_score (score every k-subset of every node's free set, per-node
argmin, cluster argmin folded in) (+ a 128-byte NCCL min across ranks when N > 1).  Workload: BASELINE.json config 4 (heterogeneous 100k-node
cluster), 100 000 nodes PER GPU (weak scaling: nodes are independent).

    python bench.py [--gpus N] [--steps K] [--warmup W]        (torchrun for N > 1)
    python bench.py --impl reference ...                       (CPU oracle on the host cores)

Prints ONE JSON line on rank 0.
"""
import argparse
import json
import os
import statistics
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "candidate_gpu_subsets_scored_per_sec"
UNIT = "subsets/s"
SEED = 0x5EED0004
WORKLOAD = "config4: heterogeneous cluster (40% DGX-1, 20% NV3-ring, 20% PCIe tree, 10% 4-GPU tree, 10% 16-GPU NVSwitch), p_used=0.35, all k"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20000)
    ap.add_argument("--warmup", type=int, default=100)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--nodes-per-gpu", type=int, default=100000)
    ap.add_argument("--replicas", type=int, default=8,
                    help="distinct cluster replicas rotated between steps so inputs exceed the 126 MB L2")
    ap.add_argument("--cpu-sample-nodes", type=int, default=0, help="nodes in the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    return ap.parse_args()


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons of one GPU with NVML while the measurement runs."""

    def __init__(self, index, period=0.02):
        super().__init__(daemon=True)
        self.index, self.period = index, period
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._halt = threading.Event()
        self.ok = False
        try:
            import pynvml
            self.nv = pynvml
            pynvml.nvmlInit()
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
            self.ok = True
        except Exception as e:   # noqa
            self.err = str(e)

    def run(self):
        if not self.ok:
            return
        nv = self.nv
        names = {
            getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8): "hw_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
            getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4): "sw_power_cap",
        }
        while not self._halt.is_set():
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            time.sleep(self.period)

    def stop(self):
        self._halt.set()
        if self.is_alive():
            self.join(timeout=2)
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": self.max_mhz, "reasons": [], "samples": 0}
        return {"sm_mhz": statistics.median(self.samples), "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(self.samples)}


def load_oracle():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib
    return oracle_lib.load()


def cpu_sample(gt, oracle, n_nodes, threads, repeats=1):
    """Times the CPU oracle (exhaustive, the contract) on the first n_nodes of the same workload."""
    lv, n, fr, _ = gt.synth_cluster(4, SEED, 0, n_nodes)
    lut = oracle.lut(0)
    best = None
    for _ in range(repeats):
        t = time.perf_counter()
        _, cnt = oracle.exhaustive_cluster(lv, n, fr, lut, threads)
        dt = time.perf_counter() - t
        best = dt if best is None else min(best, dt)
    return cnt, best


def run_reference(args, rank, world):
    """--impl reference: the reference ships no allocator, so the CPU arm is the
    in-repo oracle port (oracle/gt_oracle.c) run with every host thread."""
    if rank != 0:
        return
    import _pkg
    gt = _pkg.load_package()
    oracle = load_oracle()
    threads = os.cpu_count() or 1
    sample = args.cpu_sample_nodes or min(args.nodes_per_gpu, 100000)
    lv, n, fr, _ = gt.synth_cluster(4, SEED, 0, sample)
    lut = oracle.lut(0)
    steps = max(1, min(args.steps, 20))
    warm = max(1, min(args.warmup, 2))
    for _ in range(warm):
        oracle.exhaustive_cluster(lv, n, fr, lut, threads)
    total = 0
    t0 = time.perf_counter()
    for _ in range(steps):
        _, cnt = oracle.exhaustive_cluster(lv, n, fr, lut, threads)
        total += cnt
    dt = time.perf_counter() - t0
    val = total / dt
    sample_desc = "first %d nodes of the config-4 cluster (seed %#x), %d subsets per step" % (sample, SEED, cnt)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "nodes_per_gpu": args.nodes_per_gpu, "sample_nodes": sample},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample_desc},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def main():
    args = parse()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return

    import torch
    import torch.distributed as dist
    import _pkg
    gt = _pkg.load_package()
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))

    N = args.nodes_per_gpu
    R = max(1, args.replicas)
    stream = torch.cuda.Stream()
    ctxs, subsets, adds = [], [], []
    first = rank * N
    host_free = []
    for r in range(R):
        lv, n, fr, _ = gt.synth_cluster(4, SEED + r, first, N)
        a = gt.Allocator(N, device=local_rank, first_node=first, stream=stream.cuda_stream)
        a.load_nodes(lv, n, fr)
        if world > 1:
            uid = torch.from_numpy(gt.comm_unique_id() if rank == 0 else np.zeros(128, np.uint8)).cuda()
            dist.broadcast(uid, 0)
            a.comm_init(uid.cpu().numpy(), rank, world)
        s, ad = a.count_subsets()
        ctxs.append(a)
        subsets.append(s)
        adds.append(ad)
        pin = torch.empty(N, dtype=torch.int16).pin_memory()
        pin.numpy().view(np.uint16)[:] = fr
        host_free.append(pin)
    del lv
    read_mb = R * (N * 256 + N * 64) / 1e6

    sampler = ClockSampler(local_rank)
    sampler.start()

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput (`value`): inputs already in HBM
    for i in range(args.warmup):
        ctxs[i % R].score_all_async()
    barrier()
    launches0 = sum(c.launch_count for c in ctxs)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for i in range(args.steps):
        ctxs[i % R].score_all_async()
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    launches = sum(c.launch_count for c in ctxs) - launches0
    my_subsets = sum(subsets[i % R] for i in range(args.steps))
    t = torch.tensor([ms, float(my_subsets)], dtype=torch.float64, device="cuda")
    if world > 1:
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        ms, total_subsets = float(tmax[0]), float(t[1])
    else:
        total_subsets = float(my_subsets)
    value = total_subsets / (ms * 1e-3)

    # ---- end to end through the reference-facing calls with HOST buffers, copies inside the timed region.
    # One scheduling cycle as the extender sees it: the informer's free masks go up (pinned, 2 B/node), the
    # `sort` verb answers with one score per node for the pod's request size (1 B/node, design.md:118) and
    # kube-scheduler's host choice comes back as 16 placement keys.
    ks = np.arange(1, 17, dtype=np.uint8)
    score_pin = torch.empty(N, dtype=torch.uint8).pin_memory()
    scores_np = score_pin.numpy()
    e2e_steps = max(3, min(args.steps, 500))

    free_np = [h.numpy().view(np.uint16) for h in host_free]
    place_np = np.zeros(16, np.uint64)

    def cycle(i):
        c = ctxs[i % R]
        c.update_free(free_np[i % R])
        c.ext_sort_into(1 + i % 8, scores_np, place_np)

    for i in range(3):
        cycle(i)
    barrier()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        cycle(i)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    e2e_subsets = sum(subsets[i % R] for i in range(e2e_steps))
    t = torch.tensor([dt, float(e2e_subsets)], dtype=torch.float64, device="cuda")
    if world > 1:
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        dt, e2e_subsets = float(tmax[0]), float(t[1])
    e2e = {"value": e2e_subsets / dt, "unit": UNIT, "h2d_bytes_per_step": N * 2, "d2h_bytes_per_step": N + 128,
           "ms_per_step": dt / e2e_steps * 1e3, "steps": e2e_steps,
           "call": "gt_update_free(all masks, pinned) + gt_ext_sort(m, scores + 16 placement keys to host) [full K0+K1 pass inside]"}

    # the same cycle when the caller wants the whole best table (all 16 request sizes, 64 B/node) on the host
    out_pin = torch.empty((N, 16), dtype=torch.int32).pin_memory()
    full_steps = max(3, min(args.steps, 100))

    def cycle_full(i):
        c = ctxs[i % R]
        c.update_free(free_np[i % R])
        c.score_all_into(out_pin.data_ptr())
        c.best_node(ks)

    cycle_full(0)
    barrier()
    t0 = time.perf_counter()
    for i in range(full_steps):
        cycle_full(i)
    torch.cuda.synchronize()
    dtf = time.perf_counter() - t0
    e2e["full_table"] = {"value": sum(subsets[i % R] for i in range(full_steps)) * world / dtf, "unit": UNIT,
                         "ms_per_step": dtf / full_steps * 1e3, "d2h_bytes_per_step": N * 64 + 128,
                         "call": "gt_update_free + gt_score_all(best table to pinned host) + gt_best_node"}

    # ---- roofline of the dominant kernel (K1 k_score), CUDA events on the launching stream, live
    roofline, kernels = None, {}
    if rank == 0:
        c = ctxs[0]
        for name in ("step", "step_nograph", "bin", "score", "argmin"):
            kernels[name + "_us"] = c.time_kernel(name, 50) * 1e3
        # not the measured path: what a pass costs when the rows come out of the GPU-built (class, free mask)
        # table instead of being enumerated again (GT_FLAG_MEMO_PASS; SURVEY.md section 7 asks not to hide it)
        lv0, n0, fr0, _ = gt.synth_cluster(4, SEED, first, N)
        with gt.Allocator(N, device=local_rank, first_node=first, flags=gt.FLAG_MEMO_PASS) as am:
            am.load_nodes(lv0, n0, fr0)
            am.score_all(fetch=False)
            kernels["memo_pass_us"] = am.time_kernel("step", 50) * 1e3
        del lv0
        peak_ms = c.time_kernel("int32_peak", 5)
        sms = torch.cuda.get_device_properties(local_rank).multi_processor_count
        peak_adds = sms * 8 * 256 * 2048 * 128 / (peak_ms * 1e-3)
        t_score = kernels["score_us"] * 1e-6
        alg_adds = adds[0]
        n_np = None
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        traffic = None
        try:   # dram__bytes_read.sum + dram__bytes_write.sum of one k_score launch, from the committed ncu capture
            tr = json.load(open(os.path.join(ROOT, "profiles", "r1_traffic.json")))
            if tr.get("nodes") == N:
                traffic = tr["dram_bytes_read"] + tr["dram_bytes_write"]
        except Exception:
            pass
        alg_bytes = N * (256 + 2 + 64)      # BASELINE.md section 5 with n = 16 storage: matrix + mask + best row
        roofline = {
            "kernel": "k_score (K1 + folded cluster argmin)", "bound": "alu", "achieved": alg_adds / t_score / 1e9, "peak": peak_adds / 1e9,
            "unit": "Gadd/s", "frac": (alg_adds / t_score) / peak_adds, "traffic": traffic,
            "traffic_note": "DRAM bytes of one k_score launch from profiles/r1_traffic.json (ncu); algorithmic bytes per launch = %d" % alg_bytes,
            "note": "ALU-bound path (BASELINE.md section 5): achieved = algorithmic pair-cost additions sum C(f,2)*2^(f-2) "
                    "per launch / kernel duration; peak = INT32 add rate measured live (k_int32_peak). The kernel does "
                    "fewer real instructions than the algorithmic count (incremental tree), so frac may exceed 1.",
            "subsets_per_s_kernel": subsets[0] / t_score,
            "hbm": {"bound": "hbm", "achieved": alg_bytes / t_score / 1e9, "peak": hbm_peak, "unit": "GB/s",
                    "frac": alg_bytes / t_score / 1e9 / hbm_peak,
                    "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"},
        }

    clocks = sampler.stop()

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        oracle = load_oracle()
        sample = args.cpu_sample_nodes or min(N, 100000)
        cores = os.cpu_count() or 1
        cnt, dt1 = cpu_sample(gt, oracle, sample, 1)
        cntm, dtm = cpu_sample(gt, oracle, sample, cores, repeats=3)
        # the reference design's own greedy (design.md:153-186) on the same nodes, every m per node: it answers
        # the same questions without scoring every subset, so its rate is given in (node, m) decisions per second
        lvg, ng_, frg, _ = gt.synth_cluster(4, SEED, 0, sample)
        t0 = time.perf_counter()
        gbest, gdec = oracle.greedy_cluster(lvg, ng_, frg, oracle.lut(0))
        gdt = time.perf_counter() - t0
        xbest, _ = oracle.exhaustive_cluster(lvg, ng_, frg, oracle.lut(0), cores)
        feas = xbest != 0xFFFF0000
        worse = int(((gbest >> 16) > (xbest >> 16))[feas].sum())
        del lvg
        cpu_baseline = {"value": cnt / dt1, "unit": UNIT, "cores": 1, "kind": "port",
                        "greedy_design_md": {"decisions_per_s": gdec / gdt, "cores": 1, "decisions": gdec,
                                             "costlier_than_optimum": worse,
                                             "note": "node x request-size decisions/s of the literal design.md heuristic; "
                                                     "`costlier_than_optimum` of them pick a strictly costlier GPU set than the exhaustive path"},
                        "sample": "first %d nodes of the config-4 cluster (%d subsets), oracle/gt_oracle.c -O2" % (sample, cnt),
                        "all_cores": {"value": cntm / dtm, "cores": cores}}

    placements = None
    if rank == 0 and world == 1:
        # pod placements/s: BASELINE config 5 (1M-event arrival/departure trace) on the config-2 cluster, through
        # gt_schedule (GPU-built memo table + host commit loop, sequential semantics)
        lv2, n2, fr2, _ = gt.synth_cluster(2, 0x5EED0002, 0, 1000)
        ek, ea = gt.synth_events(5, 0x5EED0005, 1000000)
        with gt.Allocator(1000, device=local_rank) as a:
            a.load_nodes(lv2, n2, fr2)
            a.schedule(ek[:1000], ea[:1000])          # warm-up: builds the memo table on the GPU
            a.update_free(fr2)
            t0 = time.perf_counter()
            placed, _, _ = a.schedule(ek, ea)
            dt = time.perf_counter() - t0
        placements = {"value": placed / dt, "unit": "placements/s", "events": len(ek), "events_per_s": len(ek) / dt,
                      "placed": placed, "seconds": dt,
                      "workload": "config5: 1M arrival/departure events on 1000 DGX-1 nodes (config 2 cluster), sequential"}
        if not args.no_cpu_baseline:
            oracle = load_oracle()
            sub = 100000
            t0 = time.perf_counter()
            op, _, _, _ = oracle.replay(lv2, n2, fr2, oracle.lut(0), ek[:sub], ea[:sub])
            odt = time.perf_counter() - t0
            placements["cpu_baseline"] = {"value": op / odt, "unit": "placements/s", "cores": 1, "kind": "port",
                                          "sample": "first %d events of the same trace, oracle gto_replay" % sub}

    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": {"workload": WORKLOAD, "nodes_per_gpu": N, "total_nodes": N * world, "seed": hex(SEED),
                       "subsets_per_step_per_gpu": subsets[0], "parallelism": "nodes sharded, %d rank(s)" % world,
                       "exchange": {0: "none (one rank)", 1: "16 placement keys, NCCL min all-reduce on a side stream",
                                    2: "16 placement keys, own kernel over NVLink peer memory (k_exchange) on a side stream"}[ctxs[0].comm_mode()],
                       "l2": "%d cluster replicas rotated between steps (%.0f MB of tables > 126 MB L2)" % (R, read_mb)},
            "e2e": e2e, "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "kernels_us": kernels,
            "cpu_baseline": cpu_baseline, "placements": placements,
        }))
    for c in ctxs:
        c.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
