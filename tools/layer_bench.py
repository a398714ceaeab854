#!/usr/bin/env python
"""BASELINE config 4: decode step of a synthetic Llama-2-7B decoder stack (Q4_0 weights, random init) as one
GGMLCGraph on the CUDA backend: 13 nodes per layer (2 RMS_NORM, 7 MUL_MAT, SILU, MUL, 2 ADD), every layer with
its own weights (32 layers = 3.6 GB, so every replay streams cold HBM), captured once and replayed.

Prints one JSON line: layer-tokens/s (layers processed per second), tok/s for the stack, the HBM roofline
of the layer's weight bytes, and the share of the step spent in the 7 GEMVs.  `run()` is also what
bench.py calls for its `config.layer` entry.
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import BLOCK_BYTES, synth_blocks  # noqa: E402


def run(lk, be, qtype: str = "Q4_0", layers: int = 32, steps: int = 50, warmup: int = 5, hidden: int = 4096,
        ffn: int = 11008, hbm_peak_gbs: float = 0.0, drop_ops=(), ctx_len: int = 0, heads: int = 32, max_seq: int = 0) -> dict:
    """ctx_len > 0: the 18-node layer with RoPE, KV-cache append and decode attention over `ctx_len` cached positions
    (the token being decoded sits at position ctx_len - 1; the cache rows hold synthetic N(0,1) keys / values)."""
    from llama_kotlin_b200 import GGMLTensor, GGMLType
    from llama_kotlin_b200.llama_layer import alloc_attention, alloc_layer, build_decoder_graph, weight_shapes
    gt = getattr(GGMLType, qtype)
    bb = BLOCK_BYTES[qtype]
    shapes = weight_shapes(hidden, ffn)
    wbytes = sum(M * K // 32 * bb for M, K in shapes.values())
    act_bytes = 13 * ffn * 4 * 2
    max_seq = max(max_seq, ctx_len)
    kv_bytes = 2 * max_seq * hidden * 4 + 1024 if ctx_len else 0
    ga = lk.GGMLGraphAllocator(be, size=layers * (wbytes + 8 * 256 + 2 * hidden * 4 + 2 * act_bytes + kv_bytes) + (4 << 20))
    rng = np.random.default_rng(4001)
    x = ga.allocateTensor(GGMLTensor(GGMLType.F32, [hidden]))
    ga.setBytes(x, rng.normal(0, 1, hidden).astype(np.float32))
    stack, first = [], None
    for _ in range(layers):
        L = alloc_layer(ga, gt, hidden, ffn)
        if first is None:
            for n, (M, K) in shapes.items():
                ga.setBytes(L.weights[n], synth_blocks(rng, M, K, qtype))
            first = L
        else:                                   # distinct memory per layer; the bytes repeat (timing only)
            for n in shapes:
                ga.deviceBuffer.copyTensor(first.weights[n], L.weights[n])
        ga.setBytes(L.attn_norm, np.ones(hidden, np.float32))
        ga.setBytes(L.ffn_norm, np.ones(hidden, np.float32))
        stack.append(L)
    attn = None
    if ctx_len:
        attn = alloc_attention(ga, heads, hidden // heads, max_seq, layers)
        fill = rng.normal(0, 1, ctx_len * hidden).astype(np.float32)
        rows = GGMLTensor(GGMLType.F32, [ctx_len * hidden])
        for la in attn:                              # the same synthetic rows in every layer's cache (distinct memory)
            for c in (la.keys, la.values):
                rows.bufferId, rows.dataOffset, rows.buffer = c.bufferId, c.dataOffset, c.buffer
                ga.setBytes(rows, fill)
        ga.setBytes(attn[0].position, np.array([ctx_len - 1, 0, 0, 0], np.int32))
    graph, out = build_decoder_graph(ga, stack, x, hidden, ffn, attn=attn)
    if drop_ops:     # timing experiments only (the numbers are then meaningless): how much of the step do these ops cost?
        graph.nodes = [n for n in graph.nodes if n.op.name not in drop_ops]
    lkc = be.lk
    launches0, chains0, mma0 = lkc.launch_count(), int(lkc.get_option("gemv_chain_launches")), int(lkc.get_option("gemv_chain_mma_launches"))
    assert be.graphCompute(graph) == lk.GGMLStatus.SUCCESS          # captures: the launch counters move here only
    launches, chains = lkc.launch_count() - launches0, int(lkc.get_option("gemv_chain_launches")) - chains0
    mma_chains = int(lkc.get_option("gemv_chain_mma_launches")) - mma0        # of those, taken by K1m (lk_gemv_mma.cu)
    for _ in range(warmup):
        assert be.graphCompute(graph) == lk.GGMLStatus.SUCCESS
    lkc.sync()
    lkc.timer_start()
    for _ in range(steps):
        be.graphCompute(graph)
    ms = lkc.timer_stop() / steps
    y = ga.getFloats(out)
    be.forgetGraph(graph)                 # the capture keeps the graph (and through it this allocator's memory) alive
    kv_read = 2 * ctx_len * hidden * 4
    res = {"workload": f"synthetic Llama-2-{'7B' if (hidden, ffn) == (4096, 11008) else f'shaped ({hidden} / {ffn})'} decoder stack, {qtype}, batch 1, " +
                       (f"RoPE + KV cache + decode attention over {ctx_len} positions ({heads} heads), " if ctx_len else "one-position context, ") +
                       f"{layers} layers with distinct weights ({layers * wbytes / 1e9:.2f} GB > L2), {18 if ctx_len else 13} graph nodes per layer",
           "ctx_len": ctx_len, "kv_bytes_read_per_layer": kv_read, "launches_per_token": int(launches),
           "chained_gemv_launches_per_token": int(chains), "of_those_on_the_mma_chain_kernel": int(mma_chains),
           "layers": layers, "ms_per_token": ms, "tok_s": 1e3 / ms, "layer_tok_s": layers * 1e3 / ms,
           "us_per_layer": ms * 1e3 / layers, "weight_bytes_per_layer": wbytes,
           "weight_GBps": layers * wbytes / (ms * 1e-3) / 1e9, "finite_output": bool(np.all(np.isfinite(y)))}
    res["hbm_GBps"] = layers * (wbytes + kv_read) / (ms * 1e-3) / 1e9      # weights + KV rows read per token
    if hbm_peak_gbs:
        res["hbm_frac"] = res["hbm_GBps"] / hbm_peak_gbs
        res["roofline_tok_s"] = hbm_peak_gbs * 1e9 / (layers * (wbytes + kv_read))
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--qtype", default="Q4_0")
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--opt", action="append")
    ap.add_argument("--ctx", type=int, default=0, help="context length: > 0 runs the layer with RoPE, KV cache and decode attention")
    ap.add_argument("--hidden", type=int, default=4096)
    ap.add_argument("--ffn", type=int, default=11008)
    ap.add_argument("--heads", type=int, default=32)
    ap.add_argument("--drop", nargs="*", default=[], help="op names to leave out (timing experiments), e.g. RMS_NORM ADD")
    args = ap.parse_args()
    import llama_kotlin_b200 as lk
    be = lk.GGMLCudaBackend(0)
    for kv in args.opt or []:
        k, v = kv.split("=")
        if k == "nofuse":
            be.fuseOps = not int(v)
        elif k == "max_chain":          # links per chained launch (1 = every GEMV its own launch)
            be.maxChain = max(int(v), 1)
            be.chainOps = None if int(v) == 0 else int(v) > 1      # 0 = the planner's own choice
        else:
            be.lk.set_option(k, int(v))
    peak = 0.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 0))
    except Exception:
        pass
    res = run(lk, be, args.qtype, args.layers, args.steps, hbm_peak_gbs=peak, drop_ops=tuple(args.drop), ctx_len=args.ctx,
              hidden=args.hidden, ffn=args.ffn, heads=args.heads)
    if args.drop:
        res["dropped_ops"] = args.drop
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
