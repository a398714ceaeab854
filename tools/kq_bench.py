#!/usr/bin/env python
"""Decode GEMV of the reference's K-quant / BitNet 1.58 formats (lk_kquants.cu): algorithmic GB/s per
type and shape.  Every launch of a replay reads its own weight copy (copies add up to --rotate-mb > L2), CUDA events on
the launch stream.  Algorithmic bytes = M K / blockElems * blockBytes + 4 K + 4 M.  One JSON line per (type, shape).

  python tools/kq_bench.py > gpurun_out/kq_bench.jsonl
"""
from __future__ import annotations

import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--types", nargs="*", default=["Q8_K", "Q4_K", "Q2_K", "BITNET_1_58"])
    ap.add_argument("--shapes", nargs="*", default=["4096x4096", "11008x4096", "4096x11008"])
    ap.add_argument("--rotate-mb", type=int, default=512)
    ap.add_argument("--reps", type=int, default=20)
    ap.add_argument("--opt", action="append", help="lk_set_option name=value")
    args = ap.parse_args()

    import llama_kotlin_b200 as lk
    from llama_kotlin_b200 import GGMLCGraph, GGMLOp, GGMLTensor, GGMLType

    peak = None
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"]
    except Exception:
        pass
    be = lk.GGMLCudaBackend(0)
    lkc = be.lk
    for kv in args.opt or []:
        k, v = kv.split("=")
        lkc.set_option(k, int(v))
    rng = np.random.default_rng(1)
    for tname in args.types:
        gt = getattr(GGMLType, tname)
        for shp in args.shapes:
            M, K = (int(v) for v in shp.split("x"))
            K = K // gt.blockElements * gt.blockElements          # 11008 -> 10752 for the 256-element formats
            wb = M * K // gt.blockElements * gt.byteSize
            ncopies = max(2, (args.rotate_mb << 20) // wb)
            ga = lk.GGMLGraphAllocator(be, size=ncopies * (wb + 256) + (K + M * ncopies) * 4 + (1 << 20))
            proto = rng.integers(0, 243, wb, dtype=np.uint8)      # any bytes: timing only (BitNet digits stay valid)
            b = ga.allocateTensor(GGMLTensor(GGMLType.F32, [1, K]))
            ga.setBytes(b, rng.normal(0, 1, K).astype(np.float32))
            nodes, first = [], None
            for _ in range(ncopies):
                a = ga.allocateTensor(GGMLTensor(gt, [K, M]))
                if first is None:
                    ga.setBytes(a, proto)
                    first = a
                else:
                    ga.deviceBuffer.copyTensor(first, a)
                nodes.append(ga.allocateTensor(GGMLTensor(GGMLType.F32, [1, M], op=GGMLOp.MUL_MAT, src=[a, b] + [None] * 8)))
            graph = GGMLCGraph(nodes=nodes, allocator=ga)
            ok = True
            for _ in range(3):
                ok = ok and be.graphCompute(graph) == lk.GGMLStatus.SUCCESS
            lkc.sync()
            lkc.timer_start()
            for _ in range(args.reps):
                be.graphCompute(graph)
            us = lkc.timer_stop() * 1e3 / (args.reps * ncopies)
            algo = wb + 4 * K + 4 * M
            gbps = algo / us / 1e3
            kern = f"gemv_kernel<{tname}> (K1, DP4A planes)" if tname != "BITNET_1_58" else "gemv_kernel<BITNET_1_58> (K1, DP4A planes, trit table)"
            print(json.dumps({"kernel": kern, "shape": f"{M}x{K}", "us_per_launch": round(us, 3),
                              "algorithmic_bytes": algo, "GBps": round(gbps, 1),
                              "hbm_frac": round(gbps / peak, 4) if peak else None, "copies": ncopies, "ok": ok}), flush=True)
            ga.deviceBuffer.free()
    be.free()


if __name__ == "__main__":
    main()
