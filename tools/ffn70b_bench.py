#!/usr/bin/env python
"""BASELINE config 5: the FFN of a synthetic Llama-2-70B layer (hidden 8192, intermediate 28672, Q4_0) with
the weight rows sharded over the ranks of one box:

    gate_r, up_r : rows [r*28672/G, ...) of the 28672 x 8192 matrices   -> g_r, u_r   (local)
    h_r = silu(g_r) * u_r                                               -> h_full[r]  (local slice)
    exchange 1   : all-gather of h (28672 F32 = 112 KiB), ncclAllGather in place
    down_r       : rows [r*8192/G, ...) of the 8192 x 28672 matrix      -> y, stored into EVERY rank's y
                   from the GEMV epilogue over NVLink (peer mappings), then one flag barrier (exchange 2)

`--layers` FFNs with distinct weights are chained (x of the next = RMSNorm(y), computed by every rank from its
full copy of y) so that a step streams more than L2 per rank and every exchange is on the critical path.  One process per GPU:

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 \
        tools/ffn70b_bench.py
    python tools/ffn70b_bench.py            # G = 1, the strong-scaling base

Rank 0 prints one JSON line (time per FFN = max over ranks, aggregate weight GB/s, spot check).
"""
from __future__ import annotations

import argparse
import json
import os
import sys

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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--layers", type=int, default=8)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--fused", type=int, default=1, help="world > 1 with --exchange push: gate/up + SwiGLU (+ the RMSNorm before them) "
                    "+ the push of h as ONE launch (lk_compute_swiglu_sharded); 0 = the separate ops")
    ap.add_argument("--exchange", default="push", choices=["push", "nccl"],
                    help="exchange 1 (h): peer stores + flag barrier, or ncclAllGather")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    json_fd = os.dup(1)
    os.dup2(2, 1)

    import torch
    import torch.distributed as dist
    import llama_kotlin_b200 as lk
    from llama_kotlin_b200 import GGMLCGraph, GGMLOp, GGMLTensor, GGMLType
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group(backend="cpu:gloo,cuda:nccl", rank=rank, world_size=world,
                                device_id=torch.device("cuda", local_rank))
    be = lk.GGMLCudaBackend(local_rank)
    lkc = be.lk
    L = args.layers
    f0, fn_ = lk.shard_rows(FFN, world, rank)
    h0, hn = lk.shard_rows(HIDDEN, world, rank)
    assert fn_ * world == FFN and hn * world == HIDDEN
    gu_bytes, dn_bytes = fn_ * HIDDEN // 32 * BB, hn * FFN // 32 * BB
    ga = lk.GGMLGraphAllocator(be, size=L * (2 * gu_bytes + dn_bytes + 1024 + (4 * FFN + 2 * HIDDEN) * 4 + 1024) + HIDDEN * 4 + (1 << 20))
    group = peers = None
    if world > 1:
        def exchange(obj):
            out = [None] * world
            dist.all_gather_object(out, obj)
            return out
        group = lk.ShardGroup(be, rank, world, exchange)
        peers = group.map_dst_buffer(ga.deviceBuffer)
        if args.exchange == "nccl":
            group.init_nccl()
        group.barrier()
        lkc.sync()

    def vec(n):
        return ga.allocateTensor(GGMLTensor(GGMLType.F32, [n]))

    def view(t, ne, off_elems=0):
        return GGMLTensor(GGMLType.F32, ne, bufferId=0, dataOffset=t.dataOffset + off_elems * 4, buffer=ga.deviceBuffer)

    def node(op, ne, src, at, extra=None):
        return GGMLTensor(GGMLType.F32, ne, op=op, src=list(src) + [None] * (10 - len(src)), bufferId=0,
                          dataOffset=at.dataOffset, buffer=ga.deviceBuffer, extra=extra)

    rng = np.random.default_rng(5001 + rank)      # rank r holds ITS rows: different bytes per rank
    protos = {"gate": synth_blocks(rng, fn_, HIDDEN, "Q4_0"), "up": synth_blocks(rng, fn_, HIDDEN, "Q4_0"),
              "down": synth_blocks(rng, hn, FFN, "Q4_0")}
    x0 = np.random.default_rng(5500).normal(0, 1, HIDDEN).astype(np.float32)     # same on every rank
    x = vec(HIDDEN)
    ga.setBytes(x, x0)
    ones = vec(HIDDEN)
    ga.setBytes(ones, np.ones(HIDDEN, np.float32))
    nodes, first, keep = [], {}, []
    cur = x
    for li in range(L):
        w = {}
        for name, (rows, K) in (("gate", (fn_, HIDDEN)), ("up", (fn_, HIDDEN)), ("down", (hn, FFN))):
            t = ga.allocateTensor(GGMLTensor(GGMLType.Q4_0, [K, rows]))
            if name in first:
                ga.deviceBuffer.copyTensor(first[name], t)
            else:
                ga.setBytes(t, protos[name])
                first[name] = t
            w[name] = t
        g_full, u_full, a_full, h_full, y_full = vec(FFN), vec(FFN), vec(FFN), vec(FFN), vec(HIDDEN)
        keep.append((g_full, u_full, h_full, y_full))
        xin = view(cur, [1, HIDDEN])
        fused = world > 1 and args.exchange == "push" and args.fused
        if fused:
            # [RMSNorm ->] gate/up -> SwiGLU -> h slice into every peer, one launch; then the flag barrier
            d = {"group": group, "gate": w["gate"], "up": w["up"], "x": xin, "g_full": view(g_full, [1, FFN]),
                 "u_full": view(u_full, [1, FFN]), "silu_full": view(a_full, [1, FFN]), "row0": f0, "M_total": FFN, "peers": peers}
            if li > 0:          # the RMS_NORM that closes the previous FFN folds into this launch's prologue
                d.update(x=view(prev_y, [1, HIDDEN]), norm_w=ones, eps=1e-6, normed_dst=cur)
            nodes.append(node(GGMLOp.NONE, [1, FFN], [], h_full, {"shard_swiglu": d}))
            nodes.append(node(GGMLOp.NONE, [1], [], h_full, {"shard_barrier": group}))
            shd = {"shard": (group, h0, HIDDEN, peers)}
            nodes.append(node(GGMLOp.MUL_MAT, [1, HIDDEN], [w["down"], view(h_full, [1, FFN])], y_full, shd))
            nodes.append(node(GGMLOp.NONE, [1], [], y_full, {"shard_barrier": group}))
            xn = vec(HIDDEN)
            if li == L - 1:     # nothing follows: the closing RMS_NORM on its own
                nodes.append(node(GGMLOp.RMS_NORM, [HIDDEN], [y_full, ones], xn))
                nodes[-1].opParams[0] = 0x358637BD
            prev_y, cur = y_full, xn
            continue
        sh = {"shard": (group, f0, FFN, None)} if world > 1 else None
        nodes.append(node(GGMLOp.MUL_MAT, [1, FFN if world > 1 else fn_], [w["gate"], xin], g_full, sh))
        nodes.append(node(GGMLOp.MUL_MAT, [1, FFN if world > 1 else fn_], [w["up"], xin], u_full, sh))
        a_sl = view(a_full, [fn_], f0)
        nodes.append(node(GGMLOp.SILU, [fn_], [view(g_full, [fn_], f0)], a_sl))
        nodes.append(node(GGMLOp.MUL, [fn_], [a_sl, view(u_full, [fn_], f0)], view(h_full, [fn_], f0)))
        if world > 1 and args.exchange == "nccl":
            nodes.append(node(GGMLOp.NONE, [FFN], [], h_full, {"shard_allgather": (group, fn_)}))
        elif world > 1:     # this rank's h slice pushed into every peer's h_full, then the flag barrier
            nodes.append(node(GGMLOp.NONE, [fn_], [], view(h_full, [fn_], f0), {"shard_push": (group, fn_, peers)}))
            nodes.append(node(GGMLOp.NONE, [1], [], h_full, {"shard_barrier": group}))
        shd = {"shard": (group, h0, HIDDEN, peers)} if world > 1 else None
        nodes.append(node(GGMLOp.MUL_MAT, [1, HIDDEN], [w["down"], view(h_full, [1, FFN])], y_full, shd))
        if world > 1:
            nodes.append(node(GGMLOp.NONE, [1], [], y_full, {"shard_barrier": group}))
        xn = vec(HIDDEN)
        nodes.append(node(GGMLOp.RMS_NORM, [HIDDEN], [y_full, ones], xn))
        nodes[-1].opParams[0] = 0x358637BD      # eps = 1e-6 as F32 bits
        cur = xn
    graph = GGMLCGraph(nodes=nodes, allocator=ga)

    def sync_all():
        if world > 1:
            dist.barrier()
        lkc.sync()
        torch.cuda.synchronize()

    for _ in range(max(args.warmup, 3)):
        assert be.graphCompute(graph) == lk.GGMLStatus.SUCCESS
    sync_all()
    lkc.timer_start()
    for _ in range(args.steps):
        be.graphCompute(graph)
    ms = lkc.timer_stop()
    sync_all()
    t = torch.tensor([ms], dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms = float(t.item())

    # spot check on layer 0 (every rank: its own rows; y of the last layer must be identical everywhere)
    g_full, u_full, h_full, y_full = keep[0]
    g, u, h, y = (ga.getFloats(v) for v in (g_full, u_full, h_full, y_full))
    rows = [0, fn_ // 2, fn_ - 1]
    gr = deq_rows_q4_0(protos["gate"], HIDDEN, rows) @ x0.astype(np.float64)
    ur = deq_rows_q4_0(protos["up"], HIDDEN, rows) @ x0.astype(np.float64)
    err = max(np.max(np.abs(g[f0 + np.array(rows)] - gr)) / max(np.max(np.abs(gr)), 1e-9),
              np.max(np.abs(u[f0 + np.array(rows)] - ur)) / max(np.max(np.abs(ur)), 1e-9))
    rows_d = [0, hn - 1]
    yr = deq_rows_q4_0(protos["down"], FFN, rows_d) @ h.astype(np.float64)
    err = max(err, np.max(np.abs(y[h0 + np.array(rows_d)] - yr)) / max(np.max(np.abs(yr)), 1e-9))
    ylast = ga.getFloats(keep[-1][3])
    digest = float(np.sum(ylast.astype(np.float64)))
    same = True
    if world > 1:
        ds = [None] * world
        dist.all_gather_object(ds, (digest, bool(np.all(np.isfinite(ylast))), float(err)))
        same = all(d[0] == ds[0][0] and d[1] for d in ds)
        err = max(d[2] for d in ds)
        timed_out = group.timed_out()
    else:
        timed_out = False
    wbytes_ffn = 3 * FFN * HIDDEN // 32 * BB
    per_ffn_us = ms * 1e3 / (args.steps * L)
    if rank == 0:
        out = {"workload": f"Llama-2-70B FFN (gate/up 28672x8192, down 8192x28672, Q4_0), rows sharded over {world} rank(s), "
                           f"{L} chained FFNs with distinct weights per step",
               "n_gpus": world, "us_per_ffn": per_ffn_us, "weight_bytes_per_ffn": wbytes_ffn,
               "aggregate_weight_GBps": wbytes_ffn / (per_ffn_us * 1e-6) / 1e9,
               "per_gpu_weight_GBps": wbytes_ffn / world / (per_ffn_us * 1e-6) / 1e9,
               "exchanges": "none" if world == 1 else
               ("ncclAllGather" if args.exchange == "nccl" else
                "fused gate/up + SwiGLU + RMSNorm launch with peer-store epilogue + flag barrier" if args.fused else
                "peer-store push + flag barrier") +
               " (h, 112 KiB) + GEMV peer-store epilogue and flag barrier (y, 32 KiB)",
               "spot_check_rel_err": float(err), "all_ranks_identical_y": bool(same), "barrier_timed_out": bool(timed_out)}
        os.write(json_fd, (json.dumps(out) + "\n").encode())
    sync_all()
    be.releaseGraphs()          # the captured sequence holds NCCL kernels: it goes before the communicator
    if group is not None:
        sync_all()
        group.close()
    be.free()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
