#!/usr/bin/env python
"""bench.py — decode-GEMV throughput of the quantized-matmul hot path on B200.

Workload at every N (BASELINE.json configs[1]): Q4_0 decode GEMVs at the three Llama-2-7B
shapes (4096x4096, 11008x4096, 4096x11008), batch 1.  One step = LAYERS x 3 GEMVs, every GEMV
on its own weight copy (LAYERS x 60.2 MB = 1.9 GB >> the 126 MB L2, so each launch streams cold
HBM), replayed as one captured op sequence (GGMLCudaBackend.graphCompute -> CUDA graph).

  value     algorithmic GB/s, weights/activations resident in HBM (CUDA events on the ctx stream)
  e2e       same step through the backend-buffer API with HOST activations and results:
            setTensor (H2D from pinned memory) + graphCompute + getTensor (D2H) every step
  roofline  the GEMV kernel: algorithmic bytes per launch / mean launch time, vs measured HBM peak
  cpu_baseline / --impl reference
            the oracle's restatement of the reference CPU loops (oracle/, the Kotlin reference
            cannot be built here: no JVM / Konan) on a bounded row sample of the same workload

N > 1 (torchrun, one rank per GPU): BASELINE.json configs[4] — the FFN of a synthetic Llama-2-70B layer
(gate/up 28672x8192, down 8192x28672, Q4_0) with the weight rows sharded over the N ranks, 8 chained FFNs
per step, both exchanges (h, y) on the critical path: STRONG scaling, value = total weight bytes / time
(SURVEY §8e).  Exchanges: peer stores from the GEMV epilogue over NVLink with the arrival / wait folded into
the launches (--gather sync, default), with separate flag-barrier launches (--gather barrier), or
ncclAllGather for h (--gather nccl).  The weak-scaled config-2 step of round 1 (every rank a full 7B-shape
shard, one exchange point per step) is kept under config.other_paths.weak_config2.
"""
from __future__ import annotations

import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

SHAPES = [(4096, 4096), (11008, 4096), (4096, 11008)]     # (M, K)
BLOCK_BYTES = {"Q4_0": 18, "Q4_1": 20, "Q8_0": 34}
TYPE_ID = {"Q4_0": 2, "Q4_1": 3, "Q8_0": 6}
METRIC = "Q4_0 decode GEMV algorithmic GB/s at Llama-2-7B shapes (batch 1)"


def algo_bytes(M: int, K: int, bb: int) -> int:
    """M*K/32*B + 4*K + 4*M."""
    return M * K // 32 * bb + 4 * K + 4 * M


def synth_blocks(rng, M: int, K: int, qtype: str) -> np.ndarray:
    """Synthetic packed blocks with the statistics of quantized N(0, 0.02^2) weights: scale d in
    the normal-half range the reference quantizer produces (amax/8 ~ 5-8e-3 for Q4_0), uniform
    quants.  Generated directly as bytes (no oracle on the product path)."""
    bb = BLOCK_BYTES[qtype]
    nb = M * K // 32
    raw = rng.integers(0, 256, (nb, bb), dtype=np.uint8)
    lo, hi = (3e-4, 5e-4) if qtype == "Q8_0" else (5e-3, 8e-3)
    raw[:, 0:2] = rng.uniform(lo, hi, nb).astype(np.float16).view(np.uint8).reshape(nb, 2)
    if qtype == "Q4_1":
        raw[:, 2:4] = rng.uniform(-0.06, -0.04, nb).astype(np.float16).view(np.uint8).reshape(nb, 2)
    return raw.reshape(-1)


# ---------------------------------------------------------------------------------------------
# clocks: nvidia-smi sampled while the timed regions run
# ---------------------------------------------------------------------------------------------
class ClockSampler:
    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.samples = []
        self.proc = None
        self.windows = []
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits", "-lms", "50",
                 "-i", str(index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            p = [x.strip() for x in line.split(",")]
            if len(p) >= 7:
                self.samples.append((time.time(), p))

    def window(self, t0, t1):
        self.windows.append((t0, t1))

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        inwin = [p for (t, p) in self.samples if any(a - 0.05 <= t <= b + 0.05 for a, b in self.windows)]
        use = inwin if inwin else [p for _, p in self.samples]
        if not use:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}

        def num(s):
            try:
                return float(s)
            except ValueError:
                return float("nan")
        sm = sorted(num(p[0]) for p in use)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for i, n in enumerate(names) if any(p[3 + i].lower().startswith("active") for p in use)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": num(use[0][1]), "reasons": reasons,
                "samples": len(use), "in_timed_region": bool(inwin),
                "power_w_max": max(num(p[2]) for p in use)}


# ---------------------------------------------------------------------------------------------
# CPU arm: the oracle (restatement of the reference's Kotlin loops) on the host cores
# ---------------------------------------------------------------------------------------------
def cpu_arm(qtype: str, seconds_budget: float, steps: int, warmup: int, threads: int, gpu_check=None) -> dict:
    """Times oracle matmul on a bounded row sample of every shape.  One CPU 'step' = the first R rows
    of each of the three GEMVs; R is calibrated so warmup+steps fit `seconds_budget`."""
    from oracle import oracle as O
    O.build()
    rng = np.random.default_rng(4242)
    bb = BLOCK_BYTES[qtype]
    tid = TYPE_ID[qtype]
    mats = []
    for (M, K) in SHAPES:
        rows = M
        mats.append((rows, K, synth_blocks(rng, rows, K, qtype), rng.normal(0, 1, K).astype(np.float32)))
    # calibrate: rows/s on a small slice
    t0 = time.perf_counter()
    for rows, K, a, x in mats:
        O.matmul(tid, a, K, rows, x, 1, nthreads=threads, rows=(0, min(rows, 8 * threads)))
    cal = time.perf_counter() - t0
    cal_rows = sum(min(r, 8 * threads) * K for r, K, _, _ in mats)
    macs_per_s = cal_rows / max(cal, 1e-6)
    per_step = seconds_budget / max(steps + warmup, 1)
    total_k = sum(K for _, K, _, _ in mats)
    R = int(max(threads, min(4096, per_step * macs_per_s / total_k)))
    for _ in range(warmup):
        for rows, K, a, x in mats:
            O.matmul(tid, a, K, rows, x, 1, nthreads=threads, rows=(0, min(R, rows)))
    t0 = time.perf_counter()
    for _ in range(steps):
        for rows, K, a, x in mats:
            O.matmul(tid, a, K, rows, x, 1, nthreads=threads, rows=(0, min(R, rows)))
    dt = time.perf_counter() - t0
    nbytes = steps * sum(min(R, rows) * K // 32 * bb + 4 * K + 4 * min(R, rows) for rows, K, _, _ in mats)
    # the reference's real execution mode: ONE thread (core/GGMLScheduler.kt:165-171 runs nodes serially, computeMatMul has
    # no threading), same loops, on a smaller row sample (~3 s)
    R1 = int(max(1, min(R, 3.0 * macs_per_s / max(threads, 1) / total_k)))
    t1 = time.perf_counter()
    for rows, K, a, x in mats:
        O.matmul(tid, a, K, rows, x, 1, nthreads=1, rows=(0, min(R1, rows)))
    dt1 = time.perf_counter() - t1
    nbytes1 = sum(min(R1, rows) * K // 32 * bb + 4 * K + 4 * min(R1, rows) for rows, K, _, _ in mats)
    return {"value": nbytes / dt / 1e9, "unit": "GB/s", "cores": threads, "kind": "port",
            "sample": f"first {R} rows of each of the 3 {qtype} GEMV shapes per step (a row sample of the 96-GEMV step: the rate "
                      f"per byte is what is compared), {steps} steps, oracle/lk_oracle.c (C restatement of computeMatMul; "
                      f"Kotlin/Native not buildable here); weights of the sample ({nbytes // max(steps, 1) // (1 << 20)} MiB) partly live in "
                      f"the host's last-level cache",
            "single_thread": {"value": nbytes1 / dt1 / 1e9, "unit": "GB/s", "cores": 1,
                              "sample": f"first {R1} rows of each shape, one pass: the reference itself is single-threaded"},
            "ms_per_step": dt / steps * 1e3}


# ---------------------------------------------------------------------------------------------
# GPU arm
# ---------------------------------------------------------------------------------------------
def pinned_array(lkctx, nbytes: int) -> np.ndarray:
    from llama_kotlin_b200 import _binding as B
    p = C.c_void_p()
    B.check(B.lib().lk_host_alloc(lkctx.handle, nbytes, C.byref(p)), lkctx.handle)
    return np.ctypeslib.as_array((C.c_uint8 * nbytes).from_address(p.value))


def gpu_arm(args, rank: int, world: int, local_rank: int) -> dict:
    """N = 1: the config-2 step (headline) with the other measured paths beside it.  N > 1: config 5 (70B FFN, rows
    sharded, strong scaling) as the line, the weak-scaled config-2 step under config.other_paths."""
    import torch
    import torch.distributed as dist

    import llama_kotlin_b200 as lk

    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group(backend="cpu:gloo,cuda:nccl", rank=rank, world_size=world,
                                device_id=torch.device("cuda", local_rank))
    be = lk.GGMLCudaBackend(local_rank)
    for opt in args.opt or []:
        k, v = opt.split("=")
        be.lk.set_option(k, int(v))
    if world == 1:
        res = config2_arm(args, rank, world, local_rank, lk, be, dist, torch)
    else:
        res = config5_arm(args, rank, world, local_rank, lk, be, dist, torch)
    be.free()
    if world > 1:
        dist.destroy_process_group()
    return res


def config5_arm(args, rank, world, local_rank, lk, be, dist, torch) -> dict:
    """BASELINE configs[4] on `world` GPUs; see the module docstring."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import ffn70b_bench
    peaks = load_peaks()
    peak = float(peaks.get("hbm_gbs", 6650.0))
    sampler = ClockSampler(local_rank) if rank == 0 else None
    w0 = time.time()
    l0 = be.lk.launch_count()
    r = ffn70b_bench.run(lk, be, dist, rank, world, layers=8, steps=args.steps, warmup=max(args.warmup, 3),
                         mode=args.gather, e2e_steps=min(args.steps, 100))
    launches = r["launches_per_step"] * args.steps
    if sampler:
        sampler.window(w0, time.time())
    clocks = sampler.stop() if sampler else None
    # the same step with the round-1 exchanges, for the record (short)
    alt = {}
    for mode in ("barrier", "nccl"):
        if mode != args.gather:
            a = ffn70b_bench.run(lk, be, dist, rank, world, layers=8, steps=max(10, args.steps // 4), warmup=3, mode=mode)
            alt[mode] = {k: a[k] for k in ("us_per_ffn", "aggregate_weight_GBps", "launches_per_step", "all_ranks_identical_y",
                                           "barrier_timed_out")}
    weak = None
    if not args.no_extras:
        wargs = argparse.Namespace(**vars(args))
        wargs.steps = max(20, args.steps // 4)
        weak = config2_arm(wargs, rank, world, local_rank, lk, be, dist, torch, brief=True)
    value = r["aggregate_weight_GBps"]
    per_rank = r["per_gpu_weight_GBps"]
    res = {
        "metric": "Q4_0 matmul aggregate GB/s, Llama-2-70B FFN rows sharded over N GPUs (weight bytes / time incl. both exchanges)",
        "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
        "ms_per_step": r["ms_per_step"], "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f32 (int4 x 24-bit block-fixed-point activations on DP4A, exact int32 block sums, F32 scales and sums)",
        "data": "synthetic",
        "config": {"workload": r["workload"], "baseline_config": "BASELINE.json configs[4]", "exchange": r["exchanges"],
                   "gather": args.gather, "us_per_ffn": r["us_per_ffn"], "weight_bytes_per_ffn": r["weight_bytes_per_ffn"],
                   "per_gpu_weight_GBps": per_rank, "launches_per_step": r["launches_per_step"],
                   "l2": f"inputs larger than L2: every rank streams {8 * r['weight_bytes_per_ffn'] // world} B of distinct weights per step",
                   "parallelism": f"weight rows sharded over {world} ranks, replicated activations, outputs stored into every rank's "
                                  f"vectors over NVLink from the GEMV epilogue",
                   "single_gpu_base": "the N = 1 line of this script reports the same 8-FFN step on one GPU under "
                                      "config.other_paths.ffn70b_1gpu (strong-scaling efficiency = value(N) / (N x that))",
                   "check": {"spot_check_rel_err": r["spot_check_rel_err"], "all_ranks_identical_y": r["all_ranks_identical_y"],
                             "barrier_timed_out": r["barrier_timed_out"]},
                   "other_exchanges": alt,
                   "other_paths": {"weak_config2": weak}},
        "e2e": {"value": r["e2e"]["aggregate_weight_GBps"], "unit": "GB/s", "h2d_bytes_per_step": r["e2e"]["h2d_bytes_per_step"],
                "d2h_bytes_per_step": r["e2e"]["d2h_bytes_per_step"], "ms_per_step": r["e2e"]["ms_per_step"], "api": r["e2e"]["api"]},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": per_rank, "peak": peak, "unit": "GB/s", "frac": per_rank / peak, "traffic": None,
                     "peak_source": "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6650 GB/s (B200_PROFILING.md)",
                     "frac_of_nominal_8000": per_rank / 8000.0, "kernel": "lk::gemv_kernel<Q4_0, fused> (per rank, incl. the exchanges)",
                     "avg_launch_us": r["ms_per_step"] * 1e3 / max(r["launches_per_step"], 1),
                     "algorithmic_bytes_per_launch": 8 * r["weight_bytes_per_ffn"] / world / max(r["launches_per_step"], 1)},
        "clocks": clocks,
        "check": {"max_err_vs_f64_rows": r["spot_check_rel_err"],
                  "ok": bool(r["spot_check_rel_err"] <= 1e-3 and r["all_ranks_identical_y"] and not r["barrier_timed_out"])},
    }
    return res


def load_peaks() -> dict:
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)
    except OSError:
        return {}


def config2_arm(args, rank, world, local_rank, lk, be, dist, torch, brief: bool = False) -> dict:
    """The config-2 step (see the module docstring).  brief: device-resident timing and the check only (the weak-scaled
    line under other_paths at N > 1)."""
    from llama_kotlin_b200 import GGMLCGraph, GGMLOp, GGMLTensor, GGMLType
    qtype = args.qtype
    gt = {"Q4_0": GGMLType.Q4_0, "Q4_1": GGMLType.Q4_1, "Q8_0": GGMLType.Q8_0}[qtype]
    bb = BLOCK_BYTES[qtype]
    layers = args.layers
    lkc = be.lk

    # ---- device layout: [weights: layers x 3 matrices][acts: layers x sum K][dst: layers x sum M x world]
    wbytes = [M * K // 32 * bb for M, K in SHAPES]
    wtotal = layers * sum(((w + 255) // 256 * 256) for w in wbytes)
    sumK = sum(K for _, K in SHAPES)
    sumM = sum(M for M, _ in SHAPES) * world
    ga = lk.GGMLGraphAllocator(be, size=wtotal + layers * (sumK + sumM) * 4 + (1 << 20))
    rng = np.random.default_rng(1234 + rank)
    protos = [synth_blocks(rng, M, K, qtype) for M, K in SHAPES]
    acts_all = ga.allocateTensor(GGMLTensor(GGMLType.F32, [layers * sumK]))
    dst_all = ga.allocateTensor(GGMLTensor(GGMLType.F32, [layers * sumM]))
    x_host = pinned_array(lkc, layers * sumK * 4).view(np.float32)
    y_host = pinned_array(lkc, layers * sumM * 4).view(np.float32)
    x_host[:] = np.random.default_rng(99).normal(0, 1, layers * sumK).astype(np.float32)
    ga.setBytes(acts_all, x_host)

    group = peers = None
    if world > 1:
        # weight rows sharded over ranks: every rank allocated dst_all with the same layout, so one
        # peer mapping of the allocator's device buffer serves every GEMV (sharded.py)
        def exchange(obj):
            out = [None] * world
            dist.all_gather_object(out, obj)
            return out
        group = lk.ShardGroup(be, rank, world, exchange)
        peers = group.map_dst_buffer(ga.deviceBuffer)
        group.barrier()          # first call outside capture: uploads the peer flag table
        lkc.sync()
    nodes, weights = [], []
    koff = moff = 0
    first = {}
    for layer in range(layers):
        for si, (M, K) in enumerate(SHAPES):
            a = ga.allocateTensor(GGMLTensor(gt, [K, M]))
            if si in first:      # distinct HBM copy of the same bytes (device-to-device)
                ga.deviceBuffer.copyTensor(first[si], a)
            else:
                ga.setBytes(a, protos[si])
                first[si] = a
            b = GGMLTensor(GGMLType.F32, [1, K], bufferId=0, dataOffset=acts_all.dataOffset + koff * 4,
                           buffer=ga.deviceBuffer)
            extra = {"shard": (group, rank * M, M * world, peers)} if world > 1 else None
            d = GGMLTensor(GGMLType.F32, [1, M * world], op=GGMLOp.MUL_MAT, src=[a, b] + [None] * 8, bufferId=0,
                           dataOffset=dst_all.dataOffset + moff * 4, buffer=ga.deviceBuffer, extra=extra)
            koff += K
            moff += M * world
            nodes.append(d)
            weights.append(a)
    if world > 1:   # one exchange point per step: all peer stores of the step are visible after it
        nodes.append(GGMLTensor(GGMLType.F32, [1], op=GGMLOp.NONE, bufferId=0, dataOffset=dst_all.dataOffset,
                                buffer=ga.deviceBuffer, extra={"shard_barrier": group}))
    graph = GGMLCGraph(nodes=nodes, allocator=ga)

    def step_device():
        st = be.graphCompute(graph)
        if st != lk.GGMLStatus.SUCCESS:
            raise RuntimeError("graphCompute FAILED")

    def step_e2e_serial():     # round 1's form: both copies in full around the step
        ga.deviceBuffer.setTensor(acts_all, x_host, 0, x_host.nbytes)          # H2D, pinned
        step_device()
        ga.deviceBuffer.getTensor(dst_all, y_host.view(np.uint8), 0, y_host.nbytes)  # D2H (synchronises)

    # The same bytes, copied per layer inside the step (graphCompute(hostIO=...) -> lk_graph_compute_io): layer l's
    # activations go up right before its first GEMV, its results come down right after its last one, on copy streams
    # beside the compute stream; the step ends when the last download has (synchronize()).
    graph_io = GGMLCGraph(nodes=nodes, allocator=ga)
    host_io = []
    sched = [int(v) for v in args.io_schedule.split(",")] if args.io_schedule else []
    if sum(sched) != layers or min(sched, default=0) < 1:      # other layer counts: 1, 1, 2, 4, ...
        sched, left, c = [], layers, 1
        while left > 0:
            sched.append(min(c, left))
            left -= sched[-1]
            c = c * 2 if len(sched) > 1 else 1
    # uploads: small chunk first (the step starts after 1 layer's activations), growing; downloads: the mirror image
    # (the step ends one small chunk after its last launch)
    l0 = 0
    for nl in sched:
        xv = GGMLTensor(GGMLType.F32, [nl * sumK], bufferId=0, dataOffset=acts_all.dataOffset + l0 * sumK * 4, buffer=ga.deviceBuffer)
        host_io.append(("in", xv, x_host[l0 * sumK:(l0 + nl) * sumK]))
        l0 += nl
    l0 = 0
    for nl in reversed(sched):
        yv = GGMLTensor(GGMLType.F32, [nl * sumM], bufferId=0, dataOffset=dst_all.dataOffset + l0 * sumM * 4, buffer=ga.deviceBuffer)
        host_io.append(("out", yv, y_host[l0 * sumM:(l0 + nl) * sumM]))
        l0 += nl

    def step_e2e():
        if be.graphCompute(graph_io, hostIO=host_io) != lk.GGMLStatus.SUCCESS:
            raise RuntimeError("graphCompute FAILED")
        be.synchronize()

    bytes_step = layers * sum(algo_bytes(M, K, bb) for M, K in SHAPES) * world
    sampler = ClockSampler(local_rank) if rank == 0 else None

    def barrier():
        if world > 1:
            dist.barrier()
        lkc.sync()
        torch.cuda.synchronize()

    # ---- device-resident timing ------------------------------------------------------------
    for _ in range(max(args.warmup, 3)):
        step_device()
    barrier()
    l0 = lkc.launch_count()
    w0 = time.time()
    lkc.timer_start()
    for _ in range(args.steps):
        step_device()
    ms = lkc.timer_stop()
    barrier()
    w1 = time.time()
    launches = lkc.launch_count() - l0
    if sampler:
        sampler.window(w0, w1)

    # ---- spread of single steps (median and p10 / p90): each step timed alone, a synchronize on both sides,
    # so nothing overlaps across steps; not part of `value`
    step_spread = None
    if world == 1 and not brief:
        singles = []
        for _ in range(min(args.steps, 50)):
            lkc.timer_start()
            step_device()
            singles.append(lkc.timer_stop())
        q = np.percentile(np.array(singles), [10, 50, 90])
        step_spread = {"n": len(singles), "p10_ms": float(q[0]), "median_ms": float(q[1]), "p90_ms": float(q[2]),
                       "how": "one graph replay between two CUDA events with a synchronize after each (no overlap between steps)"}

    ms_dep = e2e_s = e2e_serial_s = e2e_nocopy_s = float("nan")
    if not brief:
        # ---- same step with every GEMV forced to wait for its predecessor (a fully dependent chain) ----
        lkc.set_option("gemv_overlap", 0)
        graph_dep = GGMLCGraph(nodes=nodes, allocator=ga)

        def step_dep():
            if be.graphCompute(graph_dep) != lk.GGMLStatus.SUCCESS:
                raise RuntimeError("graphCompute FAILED")
        for _ in range(3):
            step_dep()
        barrier()
        lkc.timer_start()
        for _ in range(args.steps):
            step_dep()
        ms_dep = lkc.timer_stop()
        barrier()
        lkc.set_option("gemv_overlap", 1)

        # ---- end to end: host activations in, host results out, every step -----------------------
        for _ in range(3):
            step_e2e()
        barrier()
        e0 = time.time()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step_e2e()
        barrier()
        e2e_s = time.perf_counter() - t0
        if sampler:
            sampler.window(e0, time.time())
        y_pipelined = y_host.copy()
        y_host[:] = 0
        for _ in range(3):
            step_e2e_serial()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step_e2e_serial()
        barrier()
        e2e_serial_s = time.perf_counter() - t0
        # what the call sequence costs with no copies at all (launch + synchronize per step): the floor of any e2e form
        for _ in range(3):
            step_device(); be.synchronize()
        barrier()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            step_device()
            be.synchronize()
        barrier()
        e2e_nocopy_s = time.perf_counter() - t0
        # the two forms must deliver the same bytes (then the check below covers both)
        if not np.array_equal(y_pipelined, y_host):
            raise RuntimeError("pipelined and serial host copies disagree")
    else:
        ga.deviceBuffer.getTensor(dst_all, y_host.view(np.uint8), 0, y_host.nbytes)     # for the check below

    if world > 1:
        t = torch.tensor([ms, e2e_s, ms_dep], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_s, ms_dep = float(t[0]), float(t[1]), float(t[2])
    clocks = sampler.stop() if sampler else None

    # ---- spot check of the timed outputs against exact f64 arithmetic on the host (numpy, not the oracle)
    chk = check_outputs(protos, x_host, y_host, qtype, world, sumM, rank)

    peaks = load_peaks()
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if peaks else "fallback 6650 GB/s (B200_PROFILING.md)"
    value = bytes_step * args.steps / (ms * 1e-3) / 1e9
    n_gemv = layers * len(SHAPES)
    res = {
        "metric": METRIC, "value": value, "unit": "GB/s", "n_gpus": world, "steps": args.steps,
        "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": DTYPE, "data": "synthetic",
        "config": {"workload": workload_name(qtype, layers),
                   "qtype": qtype, "layers": layers, "gemv_per_step": n_gemv,
                   "weight_bytes_per_step_per_gpu": layers * sum(wbytes),
                   "l2": "inputs larger than L2 (each launch reads a weight copy not touched for a whole step)",
                   "launch_order": "the step's GEMVs are independent ops (own x, own dst): the backend's dependency "
                                   "tracking lets consecutive launches overlap (PDL); dependent_chain_GBps is the same "
                                   "step with every GEMV made to wait for its predecessor before reading x",
                   "dependent_chain_GBps": bytes_step * args.steps / (ms_dep * 1e-3) / 1e9,
                   "activation_bits": int(8 * lkc.get_option("gemv_act_planes")),
                   "single_step_spread": step_spread,
                   "parallelism": "single GPU" if world == 1 else
                                  f"weight rows sharded over {world} ranks (each rank a full 7B-shape shard), outputs stored "
                                  f"into every rank's dst over NVLink from the GEMV epilogue + one flag barrier per step",
                   "tflops": 2 * layers * sum(M * K for M, K in SHAPES) * world * args.steps / (ms * 1e-3) / 1e12},
        "e2e": {"value": bytes_step * args.steps / e2e_s / 1e9, "unit": "GB/s",
                "h2d_bytes_per_step": int(x_host.nbytes), "d2h_bytes_per_step": int(y_host.nbytes),
                "ms_per_step": e2e_s / args.steps * 1e3,
                "api": "GGMLCudaBackend.graphCompute(graph, hostIO = per-layer pinned host slices) + synchronize(): "
                       "lk_graph_compute_io issues each layer's upload before its first GEMV and its download after its last",
                "launch_and_sync_only_ms_per_step": e2e_nocopy_s / args.steps * 1e3,
                "serial_copies": {"value": bytes_step * args.steps / e2e_serial_s / 1e9, "ms_per_step": e2e_serial_s / args.steps * 1e3,
                                  "api": "GGMLCudaBuffer.setTensor + GGMLCudaBackend.graphCompute + GGMLCudaBuffer.getTensor "
                                         "(both copies in full around the step: round 1's form)"}},
        "gpu_launches": int(launches),
        "roofline": {"bound": "hbm", "achieved": value / world, "peak": peak, "unit": "GB/s",
                     "frac": value / world / peak, "traffic": None, "peak_source": peak_src,
                     "frac_of_nominal_8000": value / world / 8000.0,
                     "kernel": "lk::gemv_kernel<Q4_0>" if qtype == "Q4_0" else f"lk::gemv_kernel<{qtype}>",
                     "avg_launch_us": ms * 1e3 / max(launches, 1),
                     "algorithmic_bytes_per_launch": bytes_step / world / n_gemv},
        "clocks": clocks, "check": chk,
    }
    res["roofline"].update(measured_traffic(qtype))
    if group is not None:
        res["config"]["barrier_timed_out"] = group.timed_out()
        barrier()
        be.releaseGraphs()
        group.close()
    if brief:
        return {k: res[k] for k in ("metric", "value", "unit", "ms_per_step", "scaling", "gpu_launches", "check")} | {
            "workload": res["config"]["workload"], "parallelism": res["config"]["parallelism"],
            "barrier_timed_out": res["config"].get("barrier_timed_out"), "roofline_frac_per_gpu": res["roofline"]["frac"]}
    if world == 1 and not args.no_extras:
        # the other two measured paths of the scope table, reported beside the headline (not part of `value`)
        res["config"]["other_paths"] = {"prefill_gemm": prefill_gemm_line(lk, be, peaks),
                                        "layer_graph": layer_graph_line(lk, be, peak),
                                        "layer_graph_ctx2048": layer_graph_line(lk, be, peak, ctx_len=2048),
                                        "layer_graph_q8_0": layer_graph_line(lk, be, peak, qtype="Q8_0"),
                                        "gemv_act16": act16_line(lk, be, graph, step_device, args, bytes_step, peak,
                                                                 protos, x_host, y_host, ga, dst_all, qtype),
                                        "gemv_other_types": gemv_types_line(lk, be, peak, args.layers),
                                        "elementwise": elementwise_lines(lk, be, peak),
                                        "ffn70b_1gpu": ffn70b_line(lk, be, dist, peak)}
    return res


DTYPE = ("f32 output; inner products int4/int8 x 24-bit block-fixed-point activations on DP4A with exact int32 block sums, "
         "F32 scales and F32 sums across blocks (K1); tcgen05 GEMM: FP16 hi+lo operand pairs, F32 accumulation (K2)")


def kernel_sources_sha256() -> str:
    """sha256 over the CUDA sources the library is built from (csrc/*.cu, *.cuh, *.h, Makefile, sorted by name).  The built
    .so itself is not reproducible byte for byte (two builds of identical sources differ), so a record keyed on its hash
    would be lost with every rebuild; identical sources give the same kernels."""
    import glob
    import hashlib
    h = hashlib.sha256()
    d = os.path.join(ROOT, "llama.kotlin_b200", "csrc")
    for path in sorted(glob.glob(os.path.join(d, "*.cu")) + glob.glob(os.path.join(d, "*.cuh")) + glob.glob(os.path.join(d, "*.h")) + [os.path.join(d, "Makefile")]):
        h.update(os.path.basename(path).encode())
        with open(path, "rb") as f:
            h.update(f.read())
    return h.hexdigest()


def measured_traffic(qtype: str) -> dict:
    """roofline.traffic = DRAM bytes per launch from an ncu capture of THIS library (tools/measure_traffic.py writes
    profiles/gemv_traffic.json with the hash of the kernel sources); a record taken from other sources is not reported."""
    path = os.path.join(ROOT, "profiles", "gemv_traffic.json")
    out = {"traffic": None, "traffic_source": "no ncu capture of these kernel sources (profiles/gemv_traffic.json is from another build)"}
    try:
        with open(path) as f:
            rec = json.load(f)
        if rec.get("sources_sha256") == kernel_sources_sha256() and rec.get(qtype) is not None:
            out = {"traffic": rec[qtype], "traffic_source": rec.get("how", "ncu dram__bytes_read.sum + dram__bytes_write.sum, mean per launch")}
    except (OSError, ValueError):
        pass
    return out


def gemv_types_line(lk, be, peak, layers: int) -> dict:
    """BASELINE configs 1 and 2 name Q8_0 and Q4_1 beside Q4_0: the headline step (layers x the three 7B shapes, every GEMV
    with its own weights, activations and output, weights > L2) with those block types, device-timed like `value`; four rows
    of the first GEMV of every shape are checked against f64 arithmetic on the same bytes."""
    from llama_kotlin_b200 import GGMLCGraph, GGMLOp, GGMLTensor, GGMLType
    out = {}
    for qtype in ("Q4_1", "Q8_0"):
        gt = getattr(GGMLType, qtype)
        bb = BLOCK_BYTES[qtype]
        rng = np.random.default_rng(97)
        wbytes = sum(M * K // 32 * bb + 256 for M, K in SHAPES)
        vec = sum((M + K + 128) * 4 for M, K in SHAPES)
        ga = lk.GGMLGraphAllocator(be, size=layers * (wbytes + vec) + (4 << 20))
        protos = [synth_blocks(rng, M, K, qtype) for M, K in SHAPES]
        firsts, nodes, xs, chk = [None] * len(SHAPES), [], [], []
        for li in range(layers):
            for si, (M, K) in enumerate(SHAPES):
                a = ga.allocateTensor(GGMLTensor(gt, [K, M]))
                if firsts[si] is None:
                    ga.setBytes(a, protos[si])
                    firsts[si] = a
                else:
                    ga.deviceBuffer.copyTensor(firsts[si], a)
                b = ga.allocateTensor(GGMLTensor(GGMLType.F32, [1, K]))
                x = rng.normal(0, 1, K).astype(np.float32)
                ga.setBytes(b, x)
                d = ga.allocateTensor(GGMLTensor(GGMLType.F32, [1, M], op=GGMLOp.MUL_MAT, src=[a, b] + [None] * 8))
                nodes.append(d)
                if li == 0:
                    chk.append((si, M, K, x, d))
        graph = GGMLCGraph(nodes=nodes, allocator=ga)
        ok = True
        for _ in range(3):
            ok = ok and be.graphCompute(graph) == lk.GGMLStatus.SUCCESS
        be.lk.sync()
        reps = 20
        be.lk.timer_start()
        for _ in range(reps):
            be.graphCompute(graph)
        ms = be.lk.timer_stop() / reps
        err = 0.0
        for si, M, K, x, d in chk:                     # rows 0..3 in f64 from the packed bytes (the reference's accessors)
            blk = protos[si].reshape(M, K // 32, bb)[:4]
            dsc = blk[:, :, 0:2].copy().view(np.float16).astype(np.float64)[:, :, 0]
            if qtype == "Q8_0":
                w = blk[:, :, 2:].view(np.int8).astype(np.float64) * dsc[:, :, None]
            else:
                mn = blk[:, :, 2:4].copy().view(np.float16).astype(np.float64)[:, :, 0]
                q = blk[:, :, 4:]
                nib = np.stack([q & 15, q >> 4], axis=-1).reshape(4, K // 32, 32).astype(np.float64)    # byte j = elements 2j, 2j + 1
                w = nib * dsc[:, :, None] + mn[:, :, None]
            ref = w.reshape(4, K) @ x.astype(np.float64)
            got = ga.getFloats(d)[:4].astype(np.float64)
            err = max(err, float(np.max(np.abs(got - ref)) / max(np.max(np.abs(ref)), 1e-30)))
        nbytes = layers * sum(algo_bytes(M, K, bb) for M, K in SHAPES)
        gbs = nbytes / (ms * 1e-3) / 1e9
        out[qtype] = {"GBps": gbs, "ms_per_step": ms, "gemv_per_step": len(nodes), "frac_of_measured_hbm": gbs / peak if peak else None,
                      "max_err_vs_f64_rows": err, "ok": bool(ok and err <= 1e-3)}
        be.forgetGraph(graph)
        ga.deviceBuffer.free()
    return out


def elementwise_lines(lk, be, peak) -> dict:
    """RMSNorm / SiLU*up / softmax stand-alone kernels against the HBM roofline (SURVEY §8d algorithmic bytes per
    element: 12 / 12 / 8), on vectors larger than L2 so that every launch streams HBM."""
    from llama_kotlin_b200 import GGMLTensor, GGMLType
    lkc = be.lk
    ne0, rows = 4096, 16384                      # 268 MB per F32 tensor
    n = ne0 * rows
    ga = lk.GGMLGraphAllocator(be, size=3 * n * 4 + ne0 * 4 + (1 << 20))
    x = ga.allocateTensor(GGMLTensor(GGMLType.F32, [ne0, rows]))
    y = ga.allocateTensor(GGMLTensor(GGMLType.F32, [ne0, rows]))
    d = ga.allocateTensor(GGMLTensor(GGMLType.F32, [ne0, rows]))
    w = ga.allocateTensor(GGMLTensor(GGMLType.F32, [ne0]))
    rng = np.random.default_rng(8)
    chunk = rng.normal(0, 1, ne0 * 256).astype(np.float32)
    for t in (x, y):
        for r0 in range(0, rows, 256):
            ga.deviceBuffer.setTensor(t, chunk, r0 * ne0 * 4, chunk.nbytes)
    ga.setBytes(w, np.ones(ne0, np.float32))
    out = {}
    for name, fn, bpe in (("rms_norm", lambda: lk.computeRmsNorm(ga, None, x, w, 1e-6, d), 8 + 4.0 / rows),
                          ("silu_mul", lambda: lk.computeSiluMul(ga, None, x, y, d), 12),
                          ("softmax", lambda: lk.computeSoftMax(ga, None, x, d), 8)):
        for _ in range(2):
            fn()
        lkc.sync()
        lkc.timer_start()
        for _ in range(5):
            fn()
        ms = lkc.timer_stop() / 5
        gbps = n * bpe / (ms * 1e-3) / 1e9
        out[name] = {"elements": n, "row": ne0, "algorithmic_bytes_per_element": round(bpe, 3), "ms": ms, "GBps": gbps,
                     "roofline": {"bound": "hbm", "achieved": gbps, "peak": peak, "unit": "GB/s", "frac": gbps / peak}}
    return out


def ffn70b_line(lk, be, dist, peak) -> dict:
    """BASELINE configs[4] on ONE GPU: the strong-scaling base of the --gpus N > 1 lines (same 8-FFN step)."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import ffn70b_bench
    r = ffn70b_bench.run(lk, be, dist, 0, 1, layers=8, steps=20, warmup=3)
    return {k: r[k] for k in ("workload", "us_per_ffn", "ms_per_step", "aggregate_weight_GBps", "launches_per_step",
                              "spot_check_rel_err")} | {"hbm_frac": r["aggregate_weight_GBps"] / peak}


def workload_name(qtype: str, layers: int) -> str:
    return (f"{qtype} decode GEMV x F32, Llama-2-7B shapes 4096x4096 + 11008x4096 + 4096x11008, "
            f"N=1, {layers} distinct weight copies per shape per step")


def prefill_gemm_line(lk, be, peaks) -> dict:
    """BASELINE config 3: Q8_0 11008x4096 x 4096x512 F32 on the tcgen05 GEMM (pack + GEMM kernels), weights
    rotating over 5 copies (240 MB > L2), CUDA events around 10 replays of the 5-op captured sequence."""
    from llama_kotlin_b200 import GGMLCGraph, GGMLOp, GGMLTensor, GGMLType
    M, K, N, copies, reps = 11008, 4096, 512, 5, 10
    wb = M * K // 32 * 34
    ga = lk.GGMLGraphAllocator(be, size=copies * (wb + 256) + (K * N + M * N) * 4 + (1 << 20))
    rng = np.random.default_rng(3001)
    b = ga.allocateTensor(GGMLTensor(GGMLType.F32, [N, K]))
    ga.setBytes(b, rng.normal(0, 1, K * N).astype(np.float32))
    dst = ga.allocateTensor(GGMLTensor(GGMLType.F32, [N, M]))
    nodes, first = [], None
    for _ in range(copies):
        a = ga.allocateTensor(GGMLTensor(GGMLType.Q8_0, [K, M]))
        if first is None:
            ga.setBytes(a, synth_blocks(rng, M, K, "Q8_0"))
            first = a
        else:
            ga.deviceBuffer.copyTensor(first, a)
        nodes.append(GGMLTensor(GGMLType.F32, [N, M], op=GGMLOp.MUL_MAT, src=[a, b] + [None] * 8, bufferId=0,
                                dataOffset=dst.dataOffset, buffer=ga.deviceBuffer))
    graph = GGMLCGraph(nodes=nodes, allocator=ga)
    before = be.lk.get_option("gemm_tc_launches")
    for _ in range(3):
        assert be.graphCompute(graph) == lk.GGMLStatus.SUCCESS
    took_tc = be.lk.get_option("gemm_tc_launches") - before == copies
    be.lk.sync()
    be.lk.timer_start()
    for _ in range(reps):
        be.graphCompute(graph)
    ms = be.lk.timer_stop() / (reps * copies)
    tf = 2.0 * M * K * N / (ms * 1e-3) / 1e12
    pk = float(peaks.get("bf16_tflops_sustained", 1400.0))
    return {"workload": "Q8_0 11008x4096 x F32 4096x512 (BASELINE config 3), device-resident, 5 rotating weight copies",
            "us": ms * 1e3, "TFLOPs": tf, "tcgen05_path": bool(took_tc),
            "roofline": {"bound": "tensor", "achieved": tf, "peak": pk, "unit": "TFLOP/s", "frac": tf / pk,
                         "peak_source": "MEASURED_PEAKS.json bf16_tflops_sustained (cuBLAS bf16, same 16-bit tensor pipe)"}}


def layer_graph_line(lk, be, hbm_peak, ctx_len: int = 0, qtype: str = "Q4_0") -> dict:
    """BASELINE config 4: 32-layer synthetic Llama-2-7B decode step as one GGMLCGraph (tools/layer_bench.py); with
    ctx_len > 0 the layer also carries RoPE, the KV-cache append and decode attention over that many positions."""
    sys.path.insert(0, os.path.join(ROOT, "tools"))
    import layer_bench
    return layer_bench.run(lk, be, qtype, layers=32, steps=30, warmup=3, hbm_peak_gbs=hbm_peak, ctx_len=ctx_len)


def act16_line(lk, be, graph, step_device, args, bytes_step, peak, protos, x_host, y_host, ga, dst_all, qtype) -> dict:
    """The headline step again with the GEMV's opt-in 16-bit activation planes (gemv_act_planes = 2: two byte planes,
    2/3 of the DP4As, |x - x'| <= 2^-15 of the block maximum).  NOT the default and not `value`: the default keeps 24-bit
    planes (error ~1e-6, below the reference's own F32 summation error); this line shows what the looser mode buys and
    what it costs, against the same f64 check."""
    from llama_kotlin_b200 import GGMLCGraph
    lkc = be.lk
    lkc.set_option("gemv_act_planes", 2)
    g16 = GGMLCGraph(nodes=graph.nodes, allocator=graph.allocator)     # fresh object => fresh capture with the option
    try:
        for _ in range(3):
            assert be.graphCompute(g16) == lk.GGMLStatus.SUCCESS
        lkc.sync()
        lkc.timer_start()
        for _ in range(args.steps):
            be.graphCompute(g16)
        ms = lkc.timer_stop() / args.steps
        ga.deviceBuffer.getTensor(dst_all, y_host.view(np.uint8), 0, y_host.nbytes)
        chk = _check_slice(protos, x_host, y_host, qtype, 1, 0)
    finally:
        lkc.set_option("gemv_act_planes", 3)
    gbps = bytes_step / (ms * 1e-3) / 1e9
    return {"activation_bits": 16, "GBps": gbps, "frac_of_measured_hbm": gbps / peak if peak else None,
            "ms_per_step": ms, "max_err_vs_f64_rows": chk.get("max_err_vs_f64_rows"), "default": False}


def check_outputs(protos, x_host, y_host, qtype, world, sumM, rank) -> dict:
    """f64 dot products (numpy) of a few rows of layer 0 against the timed run's output: this rank's
    own slice and, for world > 1, the slice the next rank stored here over NVLink."""
    res = _check_slice(protos, x_host, y_host, qtype, world, rank)
    if world > 1:
        peer = (rank + 1) % world
        prng = np.random.default_rng(1234 + peer)
        pprotos = [synth_blocks(prng, M, K, qtype) for M, K in SHAPES]
        r2 = _check_slice(pprotos, x_host, y_host, qtype, world, peer)
        res = {"max_err_vs_f64_rows": max(res["max_err_vs_f64_rows"], r2["max_err_vs_f64_rows"]),
               "peer_slice_err": r2["max_err_vs_f64_rows"], "ok": bool(res["ok"] and r2["ok"])}
    return res


def _check_slice(protos, x_host, y_host, qtype, world, rank) -> dict:
    bb = BLOCK_BYTES[qtype]
    worst = 0.0
    moff = koff = 0
    for si, (M, K) in enumerate(SHAPES):
        blocks = protos[si].reshape(-1, bb)
        x = x_host[koff:koff + K].astype(np.float64)
        rows = [0, 1, M // 2, M - 1]
        for r in rows:
            blk = blocks[r * (K // 32):(r + 1) * (K // 32)]
            d = blk[:, 0:2].copy().view(np.float16).astype(np.float64)
            if qtype == "Q8_0":
                w = d * blk[:, 2:].view(np.int8).astype(np.float64)
            else:
                q = blk[:, (2 if qtype == "Q4_0" else 4):]
                nib = np.empty((q.shape[0], 32), np.float64)
                nib[:, 0::2] = q & 15          # Kotlin nibble order: byte j -> elements 2j (low), 2j+1 (high)
                nib[:, 1::2] = q >> 4
                if qtype == "Q4_0":
                    w = d * (nib - 8.0)
                else:
                    w = d * nib + blk[:, 2:4].copy().view(np.float16).astype(np.float64)
            ref = float(np.dot(w.reshape(-1), x))
            got = float(y_host[moff + rank * M + r])
            worst = max(worst, abs(got - ref) / max(abs(ref), 1.0))
        moff += M * world
        koff += K
    return {"max_err_vs_f64_rows": worst, "ok": bool(worst <= 1e-3)}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--qtype", default="Q4_0", choices=list(BLOCK_BYTES))
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--opt", action="append", help="lk_set_option name=value (tuning experiments)")
    ap.add_argument("--gather", default="sync", choices=["sync", "barrier", "nccl"],
                    help="N > 1: how the two exchanges of the sharded FFN are made (see the module docstring)")
    ap.add_argument("--io-schedule", default="1,3,12,16",
                    help="e2e: layers per upload chunk, in order (downloads use the reverse); must add up to --layers")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the config-3 / config-4 lines under config.other_paths")
    ap.add_argument("--cpu-seconds", type=float, default=15.0)
    args = ap.parse_args()
    # stdout carries exactly one JSON line: anything a library prints there (NCCL's version banner does) goes to stderr
    json_fd = os.dup(1)
    sys.stdout.flush()
    os.dup2(2, 1)

    def emit(obj):
        os.write(json_fd, (json.dumps(obj) + "\n").encode())

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    threads = os.cpu_count() or 1

    if args.impl == "reference":
        if rank != 0:
            return
        r = cpu_arm(args.qtype, 90.0, args.steps, max(args.warmup, 1), threads)
        out = {"impl": "reference", "metric": METRIC, "value": r["value"], "unit": "GB/s", "n_gpus": args.gpus,
               "steps": args.steps, "warmup": max(args.warmup, 1), "ms_per_step": r["ms_per_step"],
               "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
               "config": {"workload": workload_name(args.qtype, args.layers), "qtype": args.qtype, "layers": args.layers,
                          "gemv_per_step": args.layers * len(SHAPES),
                          "sample": "each step is a bounded row sample of that workload, see cpu_baseline.sample"},
               "cpu_baseline": {k: r[k] for k in ("value", "unit", "cores", "kind", "sample", "single_thread")},
               "e2e": {"value": r["value"], "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
               "gpu_launches": 0}
        emit(out)
        return

    res = gpu_arm(args, rank, world, local_rank)
    if rank == 0:
        if not args.no_cpu_baseline and world == 1:
            r = cpu_arm(args.qtype, args.cpu_seconds, 5, 1, threads)
            res["cpu_baseline"] = {k: r[k] for k in ("value", "unit", "cores", "kind", "sample", "single_thread")}
        emit(res)


if __name__ == "__main__":
    main()
