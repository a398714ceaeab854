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
        ffn: int = 11008, hbm_peak_gbs: float = 0.0, drop_ops=()) -> dict:
    from llama_kotlin_b200 import GGMLTensor, GGMLType
    from llama_kotlin_b200.llama_layer import alloc_layer, build_decoder_graph, weight_shapes
    gt = getattr(GGMLType, qtype)
    bb = BLOCK_BYTES[qtype]
    shapes = weight_shapes(hidden, ffn)
    wbytes = sum(M * K // 32 * bb for M, K in shapes.values())
    act_bytes = 13 * ffn * 4 * 2
    ga = lk.GGMLGraphAllocator(be, size=layers * (wbytes + 8 * 256 + 2 * hidden * 4 + act_bytes) + (1 << 20))
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
    graph, out = build_decoder_graph(ga, stack, x, hidden, ffn)
    if drop_ops:     # timing experiments only (the numbers are then meaningless): how much of the step do these ops cost?
        graph.nodes = [n for n in graph.nodes if n.op.name not in drop_ops]
    lkc = be.lk
    for _ in range(warmup):
        assert be.graphCompute(graph) == lk.GGMLStatus.SUCCESS
    lkc.sync()
    lkc.timer_start()
    for _ in range(steps):
        be.graphCompute(graph)
    ms = lkc.timer_stop() / steps
    y = ga.getFloats(out)
    res = {"workload": f"synthetic Llama-2-7B decoder stack, {qtype}, batch 1, one-position context, {layers} layers with "
                       f"distinct weights ({layers * wbytes / 1e9:.2f} GB > L2), 13 graph nodes per layer",
           "layers": layers, "ms_per_token": ms, "tok_s": 1e3 / ms, "layer_tok_s": layers * 1e3 / ms,
           "us_per_layer": ms * 1e3 / layers, "weight_bytes_per_layer": wbytes,
           "weight_GBps": layers * wbytes / (ms * 1e-3) / 1e9, "finite_output": bool(np.all(np.isfinite(y)))}
    if hbm_peak_gbs:
        res["hbm_frac"] = res["weight_GBps"] / hbm_peak_gbs
        res["roofline_tok_s"] = hbm_peak_gbs * 1e9 / (layers * wbytes)
    return res


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--qtype", default="Q4_0")
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--opt", action="append")
    ap.add_argument("--drop", nargs="*", default=[], help="op names to leave out (timing experiments), e.g. RMS_NORM ADD")
    args = ap.parse_args()
    import llama_kotlin_b200 as lk
    be = lk.GGMLCudaBackend(0)
    for kv in args.opt or []:
        k, v = kv.split("=")
        if k == "nofuse":
            be.fuseOps = not int(v)
        else:
            be.lk.set_option(k, int(v))
    peak = 0.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 0))
    except Exception:
        pass
    res = run(lk, be, args.qtype, args.layers, args.steps, hbm_peak_gbs=peak, drop_ops=tuple(args.drop))
    if args.drop:
        res["dropped_ops"] = args.drop
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
