#!/usr/bin/env python
"""bench.py -- candidate GPU subsets scored per second on the topology-aware allocation path.

A step is one Prioritize pass (design.md:118) over this rank's shard of a synthetic cluster:
K0 k_bin (bin nodes by popcount(free)) -> K1 k_score (score every k-subset of every node's free set, per-node
argmin, cluster argmin folded in) (+ a 128-byte min-exchange of the placement keys across ranks when N > 1).
Workload: BASELINE.json config 4 (heterogeneous 100k-node cluster).  `value` is the weak-scaling figure
(100 000 nodes PER GPU: nodes are independent); `strong` beside it is BASELINE's config 4
as written -- ONE 100 000-node cluster sharded over the N ranks.

    python bench.py [--gpus N] [--steps K] [--warmup W]        (torchrun for N > 1)
    python bench.py --impl reference ...                       (CPU oracle on the host cores)

Prints ONE JSON line on rank 0.
"""
import argparse
import gc
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
L2_BYTES = 126e6


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20000)
    ap.add_argument("--warmup", type=int, default=100)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--nodes-per-gpu", type=int, default=100000)
    ap.add_argument("--total-nodes", type=int, default=100000,
                    help="size of the ONE cluster that the strong-scaling line shards over the ranks")
    ap.add_argument("--replicas", type=int, default=8,
                    help="distinct cluster replicas rotated between steps so inputs exceed the 126 MB L2")
    ap.add_argument("--cpu-sample-nodes", type=int, default=0, help="nodes in the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the config 1-3 / placements / strong-scaling extras")
    return ap.parse_args()


def workload_config(args, world, subsets_per_step):
    """The keys that name the workload: identical in both arms."""
    return {"workload": WORKLOAD, "nodes_per_gpu": args.nodes_per_gpu, "total_nodes": args.nodes_per_gpu * world,
            "seed": hex(SEED), "subsets_per_step_per_gpu": subsets_per_step,
            "parallelism": "nodes sharded, %d rank(s)" % world}


class ClockSampler(threading.Thread):
    """Samples SM clock + throttle reasons of one GPU with NVML while the measurement runs."""

    def __init__(self, index, period=0.02, start_delay=0.0):
        super().__init__(daemon=True)
        self.index, self.period, self.start_delay = index, period, start_delay
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._halt = threading.Event()
        self.active = threading.Event()    # samples count from here on (before: set-up, the GPU idles at low clocks)
        self.quiet = threading.Event()     # set around timed regions of under ~5 ms: an NVML query can hold up kernel launches
        self.ok = False                    # for 0.1-1 ms, which is most of such a region; the samples just before and after stand for it
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
        if self.start_delay:
            time.sleep(self.start_delay)
        while not self._halt.is_set():
            if self.quiet.is_set():
                time.sleep(0.005)          # (no busy polling: every wake-up of this thread takes the interpreter lock)
                continue
            try:
                mhz = nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                if self.active.is_set():
                    self.samples.append(mhz)
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
                "reasons": sorted(self.reasons), "samples": len(self.samples),
                "sampling": "NVML every %d ms from the first warm-up pass to the end of the run, paused inside timed regions "
                            "shorter than about 5 ms (a query can hold up kernel launches for most of such a region)" % int(self.period * 1e3)}


def load_oracle():
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    import oracle_lib
    return oracle_lib.load()


def cpu_sample(oracle, n_nodes, threads, repeats=1):
    """Times the CPU oracle (exhaustive, the contract) on the first n_nodes of the same workload."""
    lv, n, fr, _ = oracle.synth_cluster(4, SEED, 0, n_nodes)
    lut = oracle.lut(0)
    best = None
    for _ in range(repeats):
        t = time.perf_counter()
        _, cnt = oracle.exhaustive_cluster(lv, n, fr, lut, threads)
        dt = time.perf_counter() - t
        best = dt if best is None else min(best, dt)
    return cnt, best


def subsets_of(oracle, n_nodes):
    _, n, fr, _ = oracle.synth_cluster(4, SEED, 0, n_nodes)
    p = np.array([bin(int(f) & ((1 << int(g)) - 1)).count("1") for f, g in zip(fr, n)], np.int64)
    return int(((1 << p) - 1).sum())


def run_reference(args, rank, world):
    """--impl reference: the reference ships no allocator, so the CPU arm is the in-repo
    oracle port (oracle/gt_oracle.c) run with every host thread.  Nothing of the product is imported here: the
    cluster comes from the oracle's own generator (bit-identical to the library's, tests/test_host_abi.py)."""
    if rank != 0:
        return
    oracle = load_oracle()
    threads = os.cpu_count() or 1
    sample = args.cpu_sample_nodes or min(args.nodes_per_gpu, 100000)
    lv, n, fr, _ = oracle.synth_cluster(4, SEED, 0, sample)
    lut = oracle.lut(0)
    steps, warm = max(1, args.steps), max(0, args.warmup)
    cnt = 0
    for _ in range(warm):
        oracle.exhaustive_cluster(lv, n, fr, lut, threads)
    total = 0
    t0 = time.perf_counter()
    for _ in range(steps):
        _, cnt = oracle.exhaustive_cluster(lv, n, fr, lut, threads)
        total += cnt
    dt = time.perf_counter() - t0
    val = total / dt
    full = cnt if sample == args.nodes_per_gpu else subsets_of(oracle, args.nodes_per_gpu)
    sample_desc = "first %d nodes of the config-4 cluster (seed %#x), %d subsets per step" % (sample, SEED, cnt)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": warm, "ms_per_step": dt / steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32", "data": "synthetic",
        "config": workload_config(args, world, full),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample_desc},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


def algorithmic_bytes(n_gpus):
    """(n^2 + 2 + 4n) bytes per node -- matrix, free mask, best entries."""
    n = n_gpus.astype(np.int64)
    return int((n * n + 2 + 4 * n).sum())


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

    stream = torch.cuda.Stream()

    def make_contexts(config, seed, first, count, replicas):
        """`replicas` contexts holding different draws of the same cluster shard, all on one stream."""
        out = []
        for r in range(replicas):
            lv, n, fr, _ = gt.synth_cluster(config, seed + r, first, count)
            a = gt.Allocator(count, device=local_rank, first_node=first, stream=stream.cuda_stream)
            a.load_nodes(lv, n, fr)
            if world > 1 and r == 0:
                uid = torch.from_numpy(gt.comm_unique_id() if rank == 0 else np.zeros(128, np.uint8)).cuda()
                dist.broadcast(uid, 0)
                a.comm_init(uid.cpu().numpy(), rank, world)
            elif world > 1:
                a.comm_share(out[0]["ctx"])       # one communicator and one set of peer inboxes for all replicas
            a.score_all(fetch=False)          # set-up, not warm-up: first pass of the context (all ranks: it carries an exchange)
            s, ad = a.count_subsets()
            out.append({"ctx": a, "subsets": s, "adds": ad, "free": fr, "n_gpus": n, "levels": lv if r == 0 else None})
        return out

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def over_ranks(ms, units):
        """max over ranks of the time, sum over ranks of the units"""
        if world == 1:
            return ms, float(units)
        t = torch.tensor([ms, float(units)], dtype=torch.float64, device="cuda")
        tmax = t.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(tmax[0]), float(t[1])

    def timed_passes(reps, steps, warmup):
        """device-resident passes back to back on the rotated replicas, CUDA events on the launching stream"""
        R = len(reps)
        cs = [r["ctx"] for r in reps]
        gt.score_all_async_many(cs, warmup)
        l0 = sum(r["ctx"].launch_count for r in reps)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if steps < 250:
            sampler.quiet.set()
        e0.record(stream)                       # (creates the handles; recorded again below)
        e1.record(stream)
        barrier()
        # pass i on replica i % R; one call, so the host pays two launches per pass, and the library records the two events
        # itself: no interpreter time between the start event and the first launch
        gt.score_all_async_many_timed(cs, steps, e0.cuda_event, e1.cuda_event)
        barrier()
        sampler.quiet.clear()
        ms, total = over_ranks(e0.elapsed_time(e1), sum(reps[i % R]["subsets"] for i in range(steps)))
        return ms, total, sum(r["ctx"].launch_count for r in reps) - l0

    N = args.nodes_per_gpu
    R = max(1, args.replicas)
    first = rank * N
    # started before the contexts are built: its first NVML calls (the slow ones) are over when the first clock starts
    sampler = ClockSampler(local_rank, start_delay=float(os.environ.get("GT_BENCH_SAMPLER_DELAY", "0")))
    if not os.environ.get("GT_BENCH_NO_SAMPLER"):
        sampler.start()
    reps = make_contexts(4, SEED, first, N, R)
    sampler.active.set()
    ctxs = [r["ctx"] for r in reps]
    subsets = [r["subsets"] for r in reps]
    read_mb = R * (N * 256 + N * 64) / 1e6


    # ---- device-resident throughput (`value`): inputs already in HBM
    ms, total_subsets, launches = timed_passes(reps, args.steps, args.warmup)
    value = total_subsets / (ms * 1e-3)

    # ---- what was computed is what the oracle computes (outside the timed regions)
    ks = np.arange(1, 17, dtype=np.uint8)
    checks = {}
    oracle = load_oracle() if rank == 0 else None

    def check_exchange(rep, tag):
        """every rank: the keys the library's exchange delivered == min over ranks of the ranks' own keys"""
        c = rep["ctx"]
        got = c.best_node(ks)
        mine = c.best_node_local()
        if world > 1:
            t = torch.from_numpy(mine.view(np.int64).copy()).cuda()
            allk = [torch.empty_like(t) for _ in range(world)]
            dist.all_gather(allk, t)
            want = gt.merge_place([a.cpu().numpy().view(np.uint64) for a in allk])
        else:
            want = mine
        ok = bool((got == want).all())
        flag = torch.tensor([0 if ok else 1], device="cuda")
        if world > 1:
            dist.all_reduce(flag)
        if int(flag[0]) != 0:
            if not ok:
                sys.stderr.write("bench.py rank %d: exchanged placement keys differ (%s)\ngot  %s\nwant %s\n" % (rank, tag, got, want))
            raise SystemExit(3)
        return True

    rlast = (args.steps - 1) % R
    checks["exchange_checked"] = check_exchange(reps[rlast], "after the value loop")
    checks["exchange_mode"] = ctxs[0].comm_mode()
    got_all = reps[rlast]["ctx"].score_all()      # every rank: a pass carries an exchange, no rank may skip it
    if rank == 0:
        sample = min(4000, N)
        lvs, ns, frs, _ = oracle.synth_cluster(4, SEED + rlast, first, sample)
        want, _ = oracle.exhaustive_cluster(lvs, ns, frs, oracle.lut(0), os.cpu_count() or 1)
        if not (got_all[:sample] == want).all():
            sys.stderr.write("bench.py: best table differs from the oracle on the %d-node sample\n" % sample)
            raise SystemExit(3)
        checks["oracle_sample_nodes"] = sample
    del got_all
    barrier()

    # ---- end to end through the reference-facing calls with HOST buffers, copies inside the timed region.
    # One scheduling cycle as the extender sees it: the informer's free masks go up (pinned, 2 B/node), the
    # `sort` verb answers with one score per node for the pod's request size (1 B/node, design.md:118) and
    # kube-scheduler's host choice comes back as 16 placement keys.
    e2e_steps = max(3, min(args.steps, 500))
    score_pin = torch.empty(N, dtype=torch.uint8).pin_memory()
    scores_np = score_pin.numpy()
    free_pins = []
    for r in reps:
        pin = torch.empty(N, dtype=torch.int16).pin_memory()
        pin.numpy().view(np.uint16)[:] = r["free"]
        free_pins.append(pin)
    free_np = [h.numpy().view(np.uint16) for h in free_pins]
    place_np = np.zeros(16, np.uint64)

    def run_cycles(cycle, steps):
        gc.collect()
        gc.disable()                     # no collector pause inside the timed cycles (collected before the warm-up: it empties the caches)
        if steps < 100:
            sampler.quiet.set()
        for i in range(max(3, R)):       # untimed: every replica has answered one call of this kind before the clock starts
            cycle(i)
        barrier()
        t0 = time.perf_counter()
        for i in range(steps):
            cycle(i)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        gc.enable()
        sampler.quiet.clear()
        return over_ranks(dt, sum(subsets[i % R] for i in range(steps)))

    # the informer's masks live in the context's own pinned mirror (gt_host_free) and the scores are read in place
    # from the mapped reply buffer (gt_host_scores): every cycle uploads the 2 B/node from pinned memory and gets
    # 1 B/node + 128 B back, with no host-side memcpy on either side
    for r, c in zip(reps, ctxs):
        c.host_free()[:N] = r["free"]
    reply = [c.host_scores() for c in ctxs]          # numpy views of the mapped reply buffers
    seen = [0]

    def cycle(i):
        c = ctxs[i % R]
        c.host_free_changed()
        c.ext_sort_inplace(1 + i % 8, place_np)
        seen[0] += int(reply[i % R][i % N])          # the caller looks at the reply

    dt, e2e_subsets = run_cycles(cycle, e2e_steps)
    e2e = {"value": e2e_subsets / dt, "unit": UNIT, "h2d_bytes_per_step": N * 2, "d2h_bytes_per_step": N + 128,
           "ms_per_step": dt / e2e_steps * 1e3, "steps": e2e_steps,
           "call": "gt_host_free_changed(whole shard: masks uploaded from the pinned mirror) + gt_ext_sort(m, scores to the "
                   "mapped reply buffer + 16 placement keys) [full K0+K1 pass inside]"}
    elast = (e2e_steps - 1) % R
    mlast = 1 + (e2e_steps - 1) % 8
    if rank == 0:    # the reply of the last cycle is the score formula on the oracle's table
        sample = min(4000, N)
        lvs, ns, frs, _ = oracle.synth_cluster(4, SEED + elast, first, sample)
        wb, _ = oracle.exhaustive_cluster(lvs, ns, frs, oracle.lut(0), os.cpu_count() or 1)
        cmax, ln = oracle.cmax(oracle.lut(0)), mlast * (mlast - 1) // 2
        want = np.array([0 if b == 0xFFFF0000 else oracle.score_int(int(b) >> 16, ln, cmax) for b in wb[:, mlast - 1]], np.uint8)
        if not (reply[elast][:sample] == want).all():
            sys.stderr.write("bench.py: sort scores differ from the oracle on the %d-node sample\n" % sample)
            raise SystemExit(3)
        checks["oracle_sample_scores"] = sample
    keys_from_sort = place_np.copy()
    checks["exchange_checked_e2e"] = check_exchange(reps[elast], "after the e2e loop")
    if not (keys_from_sort == ctxs[elast].best_node(ks)).all():      # what the sort call itself handed back
        sys.stderr.write("bench.py rank %d: placement keys returned by gt_ext_sort differ from gt_best_node\n" % rank)
        raise SystemExit(3)

    # the same cycle for a caller with ordinary buffers: masks memcpy'd into the mirror, scores memcpy'd out
    def cycle_copying(i):
        c = ctxs[i % R]
        c.update_free(free_np[i % R])
        c.ext_sort_into(1 + i % 8, scores_np, place_np)

    dtc, subc = run_cycles(cycle_copying, e2e_steps)
    e2e["copying"] = {"value": subc / dtc, "unit": UNIT, "ms_per_step": dtc / e2e_steps * 1e3,
                      "call": "gt_update_free(all masks) + gt_ext_sort(m, scores + keys copied into caller buffers)"}
    if world > 1:
        # the same cycle when the caller does not ask for the cluster-wide keys: no cross-rank wait per call
        def cycle_scores_only(i):
            c = ctxs[i % R]
            c.host_free_changed()
            c.ext_sort_inplace(1 + i % 8, None)
        dts, subs = run_cycles(cycle_scores_only, e2e_steps)
        e2e["scores_only"] = {"value": subs / dts, "unit": UNIT, "ms_per_step": dts / e2e_steps * 1e3, "d2h_bytes_per_step": N,
                              "call": "gt_host_free_changed + gt_ext_sort(m, scores) without the cluster-wide placement keys: no cross-rank wait"}

    # the same cycle when the caller wants the whole best table (all 16 request sizes, 64 B/node) on the host
    out_pin = torch.empty((N, 16), dtype=torch.int32).pin_memory()
    full_steps = max(3, min(args.steps, 100))

    def cycle_full(i):
        c = ctxs[i % R]
        c.host_free_changed()
        c.score_all_into(out_pin.data_ptr())
        c.best_node(ks)

    dtf, subf = run_cycles(cycle_full, full_steps)
    e2e["full_table"] = {"value": subf / dtf, "unit": UNIT, "ms_per_step": dtf / full_steps * 1e3,
                         "d2h_bytes_per_step": N * 64 + 128,
                         "call": "gt_host_free_changed + gt_score_all(best table to pinned host) + gt_best_node"}

    # ---- roofline of the dominant kernel (K1): CUDA events around each launch, replicas taking turns so
    # that every launch reads its tables from DRAM like the launches of the `value` loop do
    roofline, kernels, peak_adds = None, {}, None
    if world > 1:       # what follows is rank-local: no context may still owe its peers an exchange
        for c in ctxs:
            c.best_node(ks)
        barrier()
    if rank == 0:
        c = ctxs[0]
        rest_us, k1_us, pass_us = gt.time_pass_rotated(ctxs, 256)
        kernels.update({"k0_and_handovers_cold_us": rest_us, "score_cold_us": k1_us, "pass_cold_us": pass_us,
                        "cold_note": "passes of the stamping mode: K1 needs its done counter for the end stamp and K0 does not run "
                                     "ahead of it, so pass_cold_us and k0_and_handovers_cold_us are about 1.5 us above the production "
                                     "pass (ms_per_step); score_cold_us is K1 itself, the same kernel in both modes"})
        for name in ("step", "bin", "score", "argmin"):      # one replica back to back: L2-resident tables
            kernels[name + "_l2warm_us"] = c.time_kernel(name, 50) * 1e3
        peak_ms = c.time_kernel("int32_peak", 5)
        sms = torch.cuda.get_device_properties(local_rank).multi_processor_count
        peak_adds = sms * 8 * 256 * 2048 * 128 / (peak_ms * 1e-3)
        alg_adds = sum(r["adds"] for r in reps) / R
        alg_bytes = algorithmic_bytes(reps[0]["n_gpus"])
        t_cold, t_warm = k1_us * 1e-6, kernels["score_l2warm_us"] * 1e-6
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        hw = {}
        try:   # hardware counters of one K1 launch from the committed ncu capture of this same command
            hw = json.load(open(os.path.join(ROOT, "profiles", "r2_hw.json")))
            if hw.get("nodes") != N:
                hw = {}
        except Exception:
            pass
        traffic = (hw["dram_bytes_read"] + hw["dram_bytes_write"]) if hw else None
        roofline = {
            "kernel": "k_score (K1 + folded cluster argmin)", "bound": "alu",
            "achieved": alg_adds / t_cold / 1e9, "peak": peak_adds / 1e9, "unit": "Gadd/s",
            "frac": (alg_adds / t_cold) / peak_adds,
            "frac_l2warm": (alg_adds / t_warm) / peak_adds,
            "timing": "K1's own global-timer stamps (first CTA past the dependency wait -> last CTA done) inside passes launched "
                      "back to back, %d replicas taking turns (%.0f MB of tables > L2): cold L2 and overlapped launches as in "
                      "the value loop; frac_l2warm = CUDA events around back-to-back K1 launches on one replica" % (R, read_mb),
            "note": "ALU-bound path (BASELINE.md section 5): achieved = ALGORITHMIC pair-cost additions sum C(f,2)*2^(f-2) per "
                    "launch / kernel duration; peak = INT32 add rate measured live (k_int32_peak, 128 lanes/clk/SM). The "
                    "incremental tree executes ~3 instructions per subset instead of C(k,2) additions, so this is an "
                    "algorithmic rate, not a distance from the hardware limit: that is `hw` below (ncu counters).",
            "subsets_per_s_kernel": subsets[0] / t_cold,
            "traffic": traffic, "alg_bytes": alg_bytes,
            "traffic_over_alg_bytes": (traffic / alg_bytes) if traffic else None,
            "hw": {k: hw.get(k) for k in ("kernel", "alu_pipe_pct", "fma_pipe_pct", "issue_active_pct", "inst_executed_pct_of_peak",
                                           "warp_instructions", "warps_active_pct", "dram_gbs_measured", "duration_us_under_ncu",
                                           "source")} if hw else None,
            "hbm": {"bound": "hbm", "achieved": alg_bytes / t_cold / 1e9, "peak": hbm_peak, "unit": "GB/s",
                    "frac": alg_bytes / t_cold / 1e9 / hbm_peak,
                    "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"},
        }
        if not args.no_extras:
            # not the measured path: what a pass costs when the rows come out of the GPU-built (class, free mask)
            # table instead of being enumerated again (GT_FLAG_MEMO_PASS; SURVEY.md section 7 asks not to hide it)
            with gt.Allocator(N, device=local_rank, first_node=first, flags=gt.FLAG_MEMO_PASS) as am:
                am.load_nodes(reps[0]["levels"], reps[0]["n_gpus"], reps[0]["free"])
                am.score_all(fetch=False)
                kernels["memo_pass_us"] = am.time_kernel("step", 50) * 1e3

    # ---- strong scaling: BASELINE config 4 as written, ONE cluster of --total-nodes sharded over the ranks
    strong = None
    if world > 1 and not args.no_extras:
        lo, hi = gt.shard_bounds(args.total_nodes, rank, world)
        per = max(1, hi - lo)
        Rs = int(min(64, max(R, np.ceil(1.3 * L2_BYTES / (per * 320.0)))))
        for c in ctxs:      # the weak-scaling replicas are done: their exchanges must not interleave with the new ones
            c.sync()
        barrier()
        sreps = make_contexts(4, SEED, lo, hi - lo, Rs)
        sms_, stotal, _ = timed_passes(sreps, args.steps, args.warmup)
        ok = check_exchange(sreps[(args.steps - 1) % Rs], "strong-scaling loop")
        strong = {"total_nodes": args.total_nodes, "nodes_per_gpu": hi - lo, "value": stotal / (sms_ * 1e-3), "unit": UNIT,
                  "ms_per_step": sms_ / args.steps, "replicas": Rs, "exchange_checked": ok,
                  "note": "config 4 as BASELINE.json words it: one 100k-node cluster, node-sharded; a %d-node shard is "
                          "launch-latency bound (two kernel hand-overs per pass), see DESIGN.md" % (hi - lo)}
        for r in reversed(sreps):      # the lender of the communicator goes last
            r["ctx"].close()

    clocks = sampler.stop()

    cpu_baseline = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        sample = args.cpu_sample_nodes or min(N, 100000)
        cores = os.cpu_count() or 1
        cnt, dt1 = cpu_sample(oracle, sample, 1)
        cntm, dtm = cpu_sample(oracle, sample, cores, repeats=3)
        # the reference design's own greedy (design.md:153-186) on the same nodes, every m per node: it answers
        # the same questions without scoring every subset, so its rate is given in (node, m) decisions per second
        lvg, ng_, frg, _ = oracle.synth_cluster(4, SEED, 0, sample)
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

    extras = None
    if rank == 0 and world == 1 and not args.no_extras:
        extras = other_configs(gt, local_rank, peak_adds)
    placements = None
    if not args.no_extras:
        placements = placements_line(gt, args, rank, world, local_rank, dist if world > 1 else None)

    if rank == 0:
        print(json.dumps({
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "u32", "data": "synthetic",
            "config": workload_config(args, world, subsets[0]),
            "run": {"exchange": {0: "none (one rank)", 1: "16 placement keys, NCCL min all-reduce on a side stream",
                                 2: "16 placement keys over NVLink peer memory by one warp of the next pass's K1 (exchange_step), no kernel or stream of its own"}[ctxs[0].comm_mode()],
                    "l2": "%d cluster replicas rotated between steps (%.0f MB of tables > 126 MB L2)" % (R, read_mb)},
            "e2e": e2e, "gpu_launches": launches, "clocks": clocks, "roofline": roofline, "kernels_us": kernels,
            "cpu_baseline": cpu_baseline, "placements": placements, "strong": strong, "checks": checks,
            "exchange_checked": checks["exchange_checked"], "exchange_mode": checks["exchange_mode"],
            "other_configs": extras,
        }))
    for c in reversed(ctxs):
        c.close()
    if world > 1:
        dist.destroy_process_group()


def other_configs(gt, device, peak_adds):
    """Driver-visible figures for BASELINE.json configs 1-3 (extra keys, not new metrics)."""
    out = {}
    # config 3: 16-GPU NVSwitch nodes with fragmented free masks (uniform costs: every subset ties, pure tie-break),
    # and the variant whose tree does real work: two 8-GPU boards, NVSwitch inside a board, SYS across
    for name, cfg in (("config3_nvswitch16", 3), ("config3_dual8", 33)):
        n3 = 10000
        lv, ng, fr, _ = gt.synth_cluster(cfg, 0x5EED0003, 0, n3)
        with gt.Allocator(n3, device=device) as a:
            a.load_nodes(lv, ng, fr)
            s3, ad3 = a.count_subsets()
            a.score_all(fetch=False)
            step = a.time_kernel("step", 50) * 1e-3
            k1 = a.time_kernel("score", 50) * 1e-3
        out[name] = {"nodes": n3, "subsets_per_pass": s3, "value": s3 / step, "unit": UNIT, "pass_us": step * 1e6,
                     "roofline": {"kernel": "K1", "bound": "alu", "achieved": ad3 / k1 / 1e9, "peak": peak_adds / 1e9,
                                  "unit": "Gadd/s", "frac": ad3 / k1 / peak_adds, "score_us": k1 * 1e6,
                                  "adds_per_subset": ad3 / s3,
                                  "note": "10 000 x 256 B tables stay in L2 between launches (one replica)"}}
    # config 2: 1 000 DGX-1 nodes, 10 000 pending pods with mixed 1-8 GPU requests, placed one after the other
    lv2, n2, fr2, _ = gt.synth_cluster(2, 0x5EED0002, 0, 1000)
    ek, ea = gt.synth_events(2, 0x5EED0012, 10000)
    with gt.Allocator(1000, device=device) as a:
        a.load_nodes(lv2, n2, fr2)
        s2, _ = a.count_subsets()
        a.score_all(fetch=False)
        step2 = a.time_kernel("step", 50) * 1e-3
        a.schedule(ek[:100], ea[:100])
        a.update_free(fr2)
        t0 = time.perf_counter()
        placed, _, _ = a.schedule(ek, ea)
        dt = time.perf_counter() - t0
    out["config2_dgx1_1000"] = {"nodes": 1000, "subsets_per_pass": s2, "value": s2 / step2, "unit": UNIT, "pass_us": step2 * 1e6,
                                "pods": len(ek), "placed": placed, "placements_per_s": placed / dt,
                                "note": "a 1 000-node pass is two kernel launches of latency; placements: host loop over a GPU-built table"}
    # config 1: one DGX-1 node, requests of 1/2/4/8 GPUs: latency of a `sort` + `bind` round trip on the HTTP bodies
    lv1, n1, fr1, _ = gt.synth_cluster(1, 0x5EED0001, 0, 1)
    lat = {}
    with gt.Allocator(1, device=device) as a:
        a.load_nodes(lv1, n1, fr1)
        http = gt.ExtenderHTTP(a, ["dgx1-0"])
        for m in (1, 2, 4, 8):
            ts = []
            for it in range(60):
                uid = "pod-%d-%d" % (m, it)
                pod = {"metadata": {"name": uid, "namespace": "default", "uid": uid},
                       "spec": {"containers": [{"resources": {"limits": {gt.RESOURCE_NAME: str(m)}}}]}}
                sort_body = json.dumps({"Pod": pod, "NodeNames": ["dgx1-0"]})
                bind_body = json.dumps({"PodName": uid, "PodNamespace": "default", "PodUID": uid, "Node": "dgx1-0"})
                t0 = time.perf_counter()
                http.handle("sort", sort_body)
                http.handle("bind", bind_body, now=1)
                ts.append(time.perf_counter() - t0)
                http.pod_deleted(uid)
            lat["m%d_us" % m] = statistics.median(ts[10:]) * 1e6
    out["config1_single_dgx1"] = {"sort_plus_bind_latency_us": lat,
                                  "note": "median of 50 sort+bind round trips through gt_ext_sort_json / gt_ext_bind_json "
                                          "(JSON in, GPU pass, JSON out; Python ctypes overhead included)"}
    return out


def placements_line(gt, args, rank, world, device, dist):
    """pod placements/s: BASELINE config 5 (1M-event arrival/departure trace) on the config-2 cluster through
    gt_schedule -- a host loop over a table the GPU filled once (GPU-built memo, sequential semantics).  With N ranks
    the 1000 nodes are sharded, the table is built in N slices and gathered, and the loop -- sequential by definition --
    runs replicated on every rank (include/gputopo.h): the figure does not grow with N, and is not meant to."""
    import torch
    total = 1000
    lo, hi = gt.shard_bounds(total, rank, world)
    lv2, n2, fr2, _ = gt.synth_cluster(2, 0x5EED0002, 0, total)
    ek, ea = gt.synth_events(5, 0x5EED0005, 1000000)
    with gt.Allocator(max(1, hi - lo), device=device, first_node=lo) as a:
        a.load_nodes(lv2[lo:hi], n2[lo:hi], fr2[lo:hi])
        if world > 1:
            uid = torch.from_numpy(gt.comm_unique_id() if rank == 0 else np.zeros(128, np.uint8)).cuda()
            dist.broadcast(uid, 0)
            a.comm_init(uid.cpu().numpy(), rank, world)
        t0 = time.perf_counter()
        a.schedule(ek[:1000], ea[:1000])          # warm-up: builds the table on the GPU(s)
        build_s = time.perf_counter() - t0
        a.update_free(fr2[lo:hi])
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        placed, on, om = a.schedule(ek, ea)
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt = float(t[0])
            h = torch.tensor([int(on.astype(np.uint64).sum() & 0x7FFFFFFFFFFF), int(om.astype(np.uint64).sum()), placed],
                             dtype=torch.int64, device="cuda")
            ref = h.clone()
            dist.broadcast(ref, 0)
            if not bool((h == ref).all()):      # every rank must have arrived at the same placements
                sys.stderr.write("bench.py rank %d: gt_schedule placements differ between ranks\n" % rank)
                raise SystemExit(3)
    if rank != 0:
        return None
    out = {"value": placed / dt, "unit": "placements/s", "events": len(ek), "events_per_s": len(ek) / dt,
           "placed": placed, "seconds": dt, "table_build_seconds": build_s, "ranks": world,
           "what": "host loop (tournament tree per request size) over a GPU-built (topology class, free mask) table; the "
                   "GPU is idle while the loop runs; with N ranks the table build is sharded and the loop replicated",
           "workload": "config5: 1M arrival/departure events on 1000 DGX-1 nodes (config 2 cluster), sequential"}
    if not args.no_cpu_baseline and world == 1:
        oracle = load_oracle()
        sub = 100000
        t0 = time.perf_counter()
        op, won, wom, _ = oracle.replay(lv2, n2, fr2, oracle.lut(0), ek[:sub], ea[:sub])
        odt = time.perf_counter() - t0
        out["matches_oracle_on_first_events"] = bool((on[:sub] == won).all() and (om[:sub] == wom).all())
        out["cpu_baseline"] = {"value": op / odt, "unit": "placements/s", "cores": 1, "kind": "port",
                               "sample": "first %d events of the same trace, oracle gto_replay (linear scan over the nodes "
                                         "per event: a simple restatement, not a tuned scheduler)" % sub}
    return out


if __name__ == "__main__":
    main()
