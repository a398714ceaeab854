#!/usr/bin/env python
"""BASELINE config 5: the FFN of a synthetic Llama-2-70B layer (hidden 8192, intermediate 28672, Q4_0) with
the weight rows sharded over the ranks of one box; graph built by
llama_kotlin_b200.ffn_sharded.build_sharded_ffn (modes "sync" / "barrier" / "nccl", see there).

`--layers` FFNs with distinct weights are chained (x of the next = RMSNorm(y), computed by every rank from its
full copy of y) so that a step streams more than L2 per rank and every exchange is on the critical path.  One process per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 \
        tools/ffn70b_bench.py
    python tools/ffn70b_bench.py            # G = 1, the strong-scaling base

Rank 0 prints one JSON line (time per FFN = max over ranks, aggregate weight GB/s, spot check).  `run()` is what
bench.py calls for --gpus N > 1.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import synth_blocks  # noqa: E402

HIDDEN, FFN, BB = 8192, 28672, 18


def deq_rows_q4_0(blocks: np.ndarray, K: int, rows) -> np.ndarray:
    """float64 rows of a packed Q4_0 matrix (reference nibble order: byte j -> elements 2j, 2j+1)."""
    rb = K // 32 * BB
    out = np.empty((len(rows), K))
    for i, r in enumerate(rows):
        raw = blocks[r * rb:(r + 1) * rb].reshape(-1, BB)
        d = raw[:, :2].copy().view(np.float16).astype(np.float64)
        q = raw[:, 2:]
        w = np.empty((raw.shape[0], 32))
        w[:, 0::2] = (q & 15).astype(np.float64) - 8
        w[:, 1::2] = (q >> 4).astype(np.float64) - 8
        out[i] = (w * d).reshape(-1)
    return out


def run(lk, be, dist, rank: int, world: int, layers: int = 8, steps: int = 50, warmup: int = 5, mode: str = "sync",
        group=None, hidden: int = HIDDEN, ffn: int = FFN, e2e_steps: int = 0) -> dict:
    """Builds and times the chained sharded FFNs on this rank.  Returns (on every rank) a dict with the device time per
    FFN (max over ranks), aggregate weight GB/s, the spot check and, with e2e_steps > 0, the end-to-end line (host x in
    on every rank, host y out on rank 0: the 32 KiB the caller needs, once).  `group`: an existing ShardGroup (bench.py
    shares one); else one is made here and closed again."""
    import torch
    from llama_kotlin_b200 import GGMLType
    from llama_kotlin_b200.ffn_sharded import build_sharded_ffn, ffn_buffer_bytes
    lkc = be.lk
    ga = lk.GGMLGraphAllocator(be, size=ffn_buffer_bytes(world, hidden, ffn, layers, BB))
    own_group = False
    peers = None
    if world > 1:
        if group is None:
            def exchange(obj):
                out = [None] * world
                dist.all_gather_object(out, obj)
                return out
            group = lk.ShardGroup(be, rank, world, exchange)
            own_group = True
        peers = group.map_dst_buffer(ga.deviceBuffer)
        if mode == "nccl":
            group.init_nccl()
        group.barrier()          # first call outside capture: uploads the peer flag table
        lkc.sync()
    rng = np.random.default_rng(5001 + rank)      # rank r holds ITS rows: different bytes per rank
    protos = {}

    def weights(li, name, rows, K):
        if name in protos:
            return None                           # distinct memory per layer; the bytes repeat (timing only)
        protos[name] = synth_blocks(rng, rows, K, "Q4_0")
        return protos[name]
    S = build_sharded_ffn(ga, group, peers, rank, world, hidden, ffn, layers, GGMLType.Q4_0, weights, mode)
    x0 = np.random.default_rng(5500).normal(0, 1, hidden).astype(np.float32)     # same on every rank
    ga.setBytes(S.x, x0)

    def sync_all():
        if world > 1:
            dist.barrier()
        lkc.sync()
        torch.cuda.synchronize()

    l0 = lkc.launch_count()
    for _ in range(max(warmup, 3)):
        assert be.graphCompute(S.graph) == lk.GGMLStatus.SUCCESS
    sync_all()
    launches_per_step = (lkc.launch_count() - l0) // max(warmup, 3)
    lkc.timer_start()
    for _ in range(steps):
        be.graphCompute(S.graph)
    ms = lkc.timer_stop()
    sync_all()
    # end to end: x from pinned host memory on every rank (the input is replicated), y of the last FFN back to the host
    # on rank 0 only — the caller needs the 32 KiB once, not once per rank
    e2e_s = None
    h2d = d2h = 0
    if e2e_steps > 0:
        from bench import pinned_array
        xh = pinned_array(lkc, hidden * 4).view(np.float32)
        yh = pinned_array(lkc, hidden * 4).view(np.float32)
        xh[:] = x0
        h2d, d2h = int(xh.nbytes), int(yh.nbytes)

        def step_e2e():
            ga.deviceBuffer.setTensor(S.x, xh, 0, xh.nbytes)
            if be.graphCompute(S.graph) != lk.GGMLStatus.SUCCESS:
                raise RuntimeError("graphCompute FAILED")
            if rank == 0:
                ga.deviceBuffer.getTensor(S.out, yh.view(np.uint8), 0, yh.nbytes)
        for _ in range(3):
            step_e2e()
        sync_all()
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            step_e2e()
        sync_all()
        e2e_s = time.perf_counter() - t0
    t = torch.tensor([ms, e2e_s or 0.0], dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, e2e_s = float(t[0]), (float(t[1]) if e2e_s is not None else None)

    # spot check on layer 0 (every rank: its own rows; y of the last layer must be identical everywhere)
    f0, fn_, h0, hn = S.f0, S.fn, S.h0, S.hn
    L0 = S.layers[0]
    g, u, h, y = (ga.getFloats(L0[k]) for k in ("g_full", "u_full", "h_full", "y_full"))
    rows = [0, fn_ // 2, fn_ - 1]
    gr = deq_rows_q4_0(protos["gate"], hidden, rows) @ x0.astype(np.float64)
    ur = deq_rows_q4_0(protos["up"], hidden, rows) @ x0.astype(np.float64)
    err = max(np.max(np.abs(g[f0 + np.array(rows)] - gr)) / max(np.max(np.abs(gr)), 1e-9),
              np.max(np.abs(u[f0 + np.array(rows)] - ur)) / max(np.max(np.abs(ur)), 1e-9))
    rows_d = [0, hn - 1]
    yr = deq_rows_q4_0(protos["down"], ffn, rows_d) @ h.astype(np.float64)
    err = max(err, np.max(np.abs(y[h0 + np.array(rows_d)] - yr)) / max(np.max(np.abs(yr)), 1e-9))
    ylast = ga.getFloats(S.layers[-1]["y_full"])
    digest = float(np.sum(ylast.astype(np.float64)))
    same, timed_out = True, False
    if world > 1:
        ds = [None] * world
        dist.all_gather_object(ds, (digest, bool(np.all(np.isfinite(ylast))), float(err), bool(group.timed_out())))
        same = all(d[0] == ds[0][0] and d[1] for d in ds)
        err = max(d[2] for d in ds)
        timed_out = any(d[3] for d in ds)
    wbytes_ffn = 3 * ffn * hidden // 32 * BB
    per_ffn_us = ms * 1e3 / (steps * layers)
    exch = {"sync": "folded into the launches: peer-store epilogues, the producer's last CTA arrives on every rank's counter, "
                    "the consumer waits in its prologue; one flag-barrier launch closes the step",
            "barrier": "peer-store epilogues + a separate flag-barrier launch after each half",
            "nccl": "ncclAllGather of h; peer-store epilogue + flag-barrier launch for y"}
    out = {"workload": f"Llama-2-70B FFN (gate/up {ffn}x{hidden}, down {hidden}x{ffn}, Q4_0), rows sharded over {world} rank(s), "
                       f"{layers} chained FFNs with distinct weights per step",
           "n_gpus": world, "mode": mode if world > 1 else "single GPU", "us_per_ffn": per_ffn_us, "ms_per_step": ms / steps,
           "weight_bytes_per_ffn": wbytes_ffn, "aggregate_weight_GBps": wbytes_ffn / (per_ffn_us * 1e-6) / 1e9,
           "per_gpu_weight_GBps": wbytes_ffn / world / (per_ffn_us * 1e-6) / 1e9,
           "launches_per_step": int(launches_per_step), "layers": layers, "steps": steps,
           "exchanges": "none" if world == 1 else exch[mode] + f" (h: {ffn * 4 // 1024} KiB, y: {hidden * 4 // 1024} KiB per FFN)",
           "spot_check_rel_err": float(err), "all_ranks_identical_y": bool(same), "barrier_timed_out": bool(timed_out)}
    if e2e_s is not None:
        out["e2e"] = {"ms_per_step": e2e_s / e2e_steps * 1e3, "aggregate_weight_GBps": wbytes_ffn * layers * e2e_steps / e2e_s / 1e9,
                      "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                      "api": "GGMLCudaBuffer.setTensor(x) on every rank + GGMLCudaBackend.graphCompute + GGMLCudaBuffer.getTensor(y) on rank 0"}
    sync_all()
    be.releaseGraphs()          # the captured sequence may hold NCCL kernels: it goes before the communicator
    if own_group:
        sync_all()
        group.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=8)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--mode", default="sync", choices=["sync", "barrier", "nccl"])
    ap.add_argument("--e2e-steps", type=int, default=0)
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    json_fd = os.dup(1)
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist
    import llama_kotlin_b200 as lk
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group(backend="cpu:gloo,cuda:nccl", rank=rank, world_size=world,
                                device_id=torch.device("cuda", local_rank))
    be = lk.GGMLCudaBackend(local_rank)
    out = run(lk, be, dist, rank, world, layers=args.layers, steps=args.steps, warmup=args.warmup, mode=args.mode,
              e2e_steps=args.e2e_steps)
    if rank == 0:
        os.write(json_fd, (json.dumps(out) + "\n").encode())
    be.free()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
