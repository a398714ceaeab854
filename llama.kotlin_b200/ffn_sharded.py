"""BASELINE config 5 as a graph: the FFN of a Llama-2-70B-shaped layer with the weight rows sharded over the
ranks of one box, as a GGMLCGraph that GGMLCudaBackend.graphCompute runs.

    gate_r, up_r : rows [r*ffn/G, ...) of the ffn x hidden matrices      -> g_r, u_r   (local)
    h_r = silu(g_r) * u_r                                                -> h_full[r]  (this rank's slice, on EVERY rank)
    exchange 1   : every rank holds all of h                              (ffn floats)
    down_r       : rows [r*hidden/G, ...) of the hidden x ffn matrix     -> y_full[r]  (on EVERY rank)
    exchange 2   : every rank holds all of y                              (hidden floats)

`layers` FFNs with distinct weights are chained: x of the next FFN = RMSNorm(y) (gain = ones), which every rank computes
itself from its full copy of y — the op order of LlamaMLP.forward inside LlamaDecoderLayer.forward
(model/LlamaModel.kt:159-176, :253-285).  Modes of the two exchanges:

    "sync"    (default) one launch per half, the exchange folded into the launches (lk_compute_swiglu_sharded_sync /
              lk_compute_matmul_sharded_sync): peer stores from the GEMV epilogue, the producer's last CTA arrives on every
              rank's counter, the consumer waits for the arrivals in its prologue after its weight prefetch was issued;
              one flag-barrier launch closes the step (the host reads y afterwards)
    "barrier" round 1: the same two launches, each followed by a separate lk_peer_barrier launch
    "nccl"    baseline: separate matmul / SILU / MUL ops, ncclAllGather of h, peer stores + flag barrier for y
    world = 1 the unsharded ops (the strong-scaling base)

Used by bench.py (--gpus N > 1), tools/ffn70b_bench.py and tests/test_gpu_multi.py.
"""
from __future__ import annotations

from typing import Callable, Dict, List, Optional

import numpy as np

from .ggml_types import GGMLCGraph, GGMLOp, GGMLTensor, GGMLType
from .sharded import shard_rows

EPS_BITS = 0x358637BD      # 1e-6 as F32 bits (LlamaConfig.rmsNormEps, model/LlamaModel.kt:18)
SLOT_H, SLOT_Y = 0, 1      # exchange slots of mode "sync"


class ShardedFFN:
    """Handles of one built graph: graph, x (input vector), per-layer vectors, shard geometry."""

    def __init__(self):
        self.graph: Optional[GGMLCGraph] = None
        self.x: Optional[GGMLTensor] = None
        self.out: Optional[GGMLTensor] = None          # RMSNorm(y) of the last layer (hidden floats)
        self.layers: List[Dict[str, GGMLTensor]] = []  # per layer: gate, up, down, g_full, u_full, a_full, h_full, y_full
        self.f0 = self.fn = self.h0 = self.hn = 0
        self.launches_per_step = 0


def build_sharded_ffn(ga, group, peers, rank: int, world: int, hidden: int, ffn: int, layers: int, qtype: GGMLType,
                      weights: Callable[[int, str, int, int], Optional[np.ndarray]], mode: str = "sync") -> ShardedFFN:
    """weights(layer, name, rows, K) -> packed blocks of THIS rank's rows of matrix `name` ("gate" | "up" | "down") of
    `layer`, or None to reuse (device-to-device copy) the bytes of the previous layer that returned an array."""
    if mode not in ("sync", "barrier", "nccl"):
        raise ValueError(mode)
    S = ShardedFFN()
    S.f0, S.fn = shard_rows(ffn, world, rank)
    S.h0, S.hn = shard_rows(hidden, world, rank)
    if S.fn * world != ffn or S.hn * world != hidden:
        raise ValueError("ffn and hidden must divide evenly over the ranks")
    f0, fn_, h0, hn = S.f0, S.fn, S.h0, S.hn

    def vec(n):
        return ga.allocateTensor(GGMLTensor(GGMLType.F32, [n]))

    def view(t, ne, off_elems=0):
        return GGMLTensor(GGMLType.F32, ne, bufferId=0, dataOffset=t.dataOffset + off_elems * 4, buffer=ga.deviceBuffer)

    def node(op, ne, src, at, extra=None):
        return GGMLTensor(GGMLType.F32, ne, op=op, src=list(src) + [None] * (10 - len(src)), bufferId=0,
                          dataOffset=at.dataOffset, buffer=ga.deviceBuffer, extra=extra)

    S.x = vec(hidden)
    ones = vec(hidden)
    ga.setBytes(ones, np.ones(hidden, np.float32))
    nodes, last = [], {}
    cur, prev_y = S.x, None
    fused = world > 1 and mode in ("sync", "barrier")
    for li in range(layers):
        w = {}
        for name, (rows, K) in (("gate", (fn_, hidden)), ("up", (fn_, hidden)), ("down", (hn, ffn))):
            t = ga.allocateTensor(GGMLTensor(qtype, [K, rows]))
            data = weights(li, name, rows, K)
            if data is None:
                ga.deviceBuffer.copyTensor(last[name], t)
            else:
                ga.setBytes(t, data)
                last[name] = t
            w[name] = t
        g_full, u_full, a_full, h_full, y_full = vec(ffn), vec(ffn), vec(ffn), vec(ffn), vec(hidden)
        S.layers.append(dict(w, g_full=g_full, u_full=u_full, a_full=a_full, h_full=h_full, y_full=y_full))
        xin = view(cur, [1, hidden])
        if fused:
            sync_up = sync_dn = None
            if mode == "sync":
                # SwiGLU half: waits for every rank's y of the previous FFN (slot Y), arrives on slot H;
                # down half: waits on slot H, arrives on slot Y — except in the last FFN, whose y the closing barrier covers
                sync_up = (SLOT_Y if li > 0 else None, SLOT_H)
                sync_dn = (SLOT_H, SLOT_Y if li < layers - 1 else None)
            d = {"group": group, "gate": w["gate"], "up": w["up"], "x": xin, "g_full": view(g_full, [1, ffn]),
                 "u_full": view(u_full, [1, ffn]), "silu_full": view(a_full, [1, ffn]), "row0": f0, "M_total": ffn,
                 "peers": peers, "sync": sync_up}
            if li > 0:          # the RMS_NORM that closes the previous FFN folds into this launch's prologue
                d.update(x=view(prev_y, [1, hidden]), norm_w=ones, eps=1e-6, normed_dst=cur)
            nodes.append(node(GGMLOp.NONE, [1, ffn], [], h_full, {"shard_swiglu": d}))
            if mode == "barrier":
                nodes.append(node(GGMLOp.NONE, [1], [], h_full, {"shard_barrier": group}))
            nodes.append(node(GGMLOp.MUL_MAT, [1, hidden], [w["down"], view(h_full, [1, ffn])], y_full,
                              {"shard": (group, h0, hidden, peers, sync_dn)}))
            if mode == "barrier" or li == layers - 1:
                nodes.append(node(GGMLOp.NONE, [1], [], y_full, {"shard_barrier": group}))
            xn = vec(hidden)
            if li == layers - 1:     # nothing follows: the closing RMS_NORM on its own
                nodes.append(node(GGMLOp.RMS_NORM, [hidden], [y_full, ones], xn))
                nodes[-1].opParams[0] = EPS_BITS
            prev_y, cur = y_full, xn
            continue
        sh = {"shard": (group, f0, ffn, None)} if world > 1 else None
        nodes.append(node(GGMLOp.MUL_MAT, [1, ffn if world > 1 else fn_], [w["gate"], xin], g_full, sh))
        nodes.append(node(GGMLOp.MUL_MAT, [1, ffn if world > 1 else fn_], [w["up"], xin], u_full, sh))
        a_sl = view(a_full, [fn_], f0)
        nodes.append(node(GGMLOp.SILU, [fn_], [view(g_full, [fn_], f0)], a_sl))
        nodes.append(node(GGMLOp.MUL, [fn_], [a_sl, view(u_full, [fn_], f0)], view(h_full, [fn_], f0)))
        if world > 1:
            nodes.append(node(GGMLOp.NONE, [ffn], [], h_full, {"shard_allgather": (group, fn_)}))
        shd = {"shard": (group, h0, hidden, peers)} if world > 1 else None
        nodes.append(node(GGMLOp.MUL_MAT, [1, hidden], [w["down"], view(h_full, [1, ffn])], y_full, shd))
        if world > 1:
            nodes.append(node(GGMLOp.NONE, [1], [], y_full, {"shard_barrier": group}))
        xn = vec(hidden)
        nodes.append(node(GGMLOp.RMS_NORM, [hidden], [y_full, ones], xn))
        nodes[-1].opParams[0] = EPS_BITS
        cur = xn
    S.out = cur
    S.graph = GGMLCGraph(nodes=nodes, allocator=ga)
    return S


def ffn_buffer_bytes(world: int, hidden: int, ffn: int, layers: int, block_bytes: int) -> int:
    """Device bytes build_sharded_ffn allocates from the graph allocator (weights of this rank + vectors)."""
    fn_, hn = ffn // world, hidden // world
    per_layer = 2 * (fn_ * hidden // 32 * block_bytes) + hn * ffn // 32 * block_bytes + 3 * 256 + (4 * ffn + 2 * hidden) * 4 + 8 * 256
    return layers * per_layer + 2 * hidden * 4 + (1 << 20)
