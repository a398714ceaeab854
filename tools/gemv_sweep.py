#!/usr/bin/env python
"""Per-shape timing of the decode GEMV (K1) under tuning options (lk_set_option).

Every launch of a replay reads its own weight copy; the copies of one shape add up to
--rotate-mb (> L2), so each launch streams cold HBM.  Prints one JSON line per
(shape, option set).  Development tool: bench.py is the contract benchmark.

  python tools/gemv_sweep.py --qtype Q4_0 --opts "" "gemv_warps=16" "gemv_stage_bytes=32768,gemv_max_stages=4"
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import BLOCK_BYTES, algo_bytes, synth_blocks  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--qtype", default="Q4_0")
    ap.add_argument("--shapes", nargs="*", default=["4096x4096", "11008x4096", "4096x11008"])
    ap.add_argument("--rotate-mb", type=int, default=512)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--opts", nargs="*", default=[""])
    ap.add_argument("--no-graph", action="store_true")
    ap.add_argument("--min-launches", type=int, default=0,
                    help="launches per replay: the copies are reused round-robin (with a small --rotate-mb the weights stay in L2)")
    args = ap.parse_args()

    import llama_kotlin_b200 as lk
    from llama_kotlin_b200 import GGMLCGraph, GGMLOp, GGMLTensor, GGMLType

    gt = getattr(GGMLType, args.qtype)
    bb = BLOCK_BYTES[args.qtype]
    be = lk.GGMLCudaBackend(0)
    be.useCudaGraphs = not args.no_graph
    lkc = be.lk
    rng = np.random.default_rng(1)
    defaults = {k: 0 for k in ("gemv_warps", "gemv_stage_bytes", "gemv_max_stages", "gemv_smem_budget")}
    defaults.update({"gemv_ctas_per_sm": 1, "gemv_pdl": 1, "gemv_overlap": 1})
    keep = []
    for shp in args.shapes:
        M, K = (int(v) for v in shp.split("x"))
        wb = M * K // 32 * bb
        ncopies = max(2, (args.rotate_mb << 20) // wb)
        nlaunch = max(ncopies, args.min_launches)
        ga = lk.GGMLGraphAllocator(be, size=ncopies * (wb + 256) + (K + (M + 64) * nlaunch) * 4 + (1 << 20))
        proto = synth_blocks(rng, M, K, args.qtype)
        x = rng.normal(0, 1, K).astype(np.float32)
        b = ga.allocateTensor(GGMLTensor(GGMLType.F32, [1, K]))
        ga.setBytes(b, x)
        nodes, first, mats = [], None, []
        for _ in range(ncopies):
            a = ga.allocateTensor(GGMLTensor(gt, [K, M]))
            if first is None:
                ga.setBytes(a, proto)
                first = a
            else:
                ga.deviceBuffer.copyTensor(first, a)
            mats.append(a)
        for i in range(nlaunch):
            d = ga.allocateTensor(GGMLTensor(GGMLType.F32, [1, M], op=GGMLOp.MUL_MAT, src=[mats[i % ncopies], b] + [None] * 8))
            nodes.append(d)
        for optset in args.opts:
            cur = dict(defaults)
            for kv in filter(None, optset.split(",")):
                k, v = kv.split("=")
                cur[k] = int(v)
            for k, v in cur.items():
                lkc.set_option(k, v)
            graph = GGMLCGraph(nodes=nodes, allocator=ga)   # fresh object => fresh capture with these options
            keep.append(graph)
            ok = True
            for _ in range(3):
                ok = ok and be.graphCompute(graph) == lk.GGMLStatus.SUCCESS
            lkc.sync()
            lkc.timer_start()
            for _ in range(args.reps):
                be.graphCompute(graph)
            ms = lkc.timer_stop()
            us = ms * 1e3 / (args.reps * nlaunch)
            got = ga.getFloats(nodes[-1])
            print(json.dumps({"shape": shp, "qtype": args.qtype, "opts": optset or "default", "us_per_launch": round(us, 3),
                              "GBps": round(algo_bytes(M, K, bb) / us / 1e3, 1), "copies": ncopies, "launches": nlaunch, "ok": ok,
                              "checksum": float(np.sum(got.astype(np.float64)))}), flush=True)
        ga.deviceBuffer.free()
    be.free()


if __name__ == "__main__":
    main()
