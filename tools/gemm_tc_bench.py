#!/usr/bin/env python
"""Timing of the tcgen05 prefill GEMM (K2) on device-resident tensors, one JSON line per shape.

TFLOP/s = 2*M*K*N / time of (pack_b_kernel + gemm_tc_kernel); the weight copies rotate so that
consecutive launches do not find their weights in L2 (--rotate-mb > 126).  Development tool:
bench.py is the contract benchmark.

  python tools/gemm_tc_bench.py --qtype Q8_0 --shapes 11008x4096x512
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--qtype", default="Q8_0")
    ap.add_argument("--shapes", nargs="*", default=["11008x4096x512"], help="MxKxN")
    ap.add_argument("--rotate-mb", type=int, default=256)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--cluster", type=int, default=0, help="gemm_tc_cluster option (0 = library default)")
    ap.add_argument("--precision", type=int, default=-1, help="gemm_tc_precision option: 0 / 1 / 2 FP16 parts per operand (-1 = library default)")
    ap.add_argument("--generic", action="store_true", help="time the CUDA-core kernel instead (gemm_tc_min_n=0)")
    args = ap.parse_args()

    import llama_kotlin_b200 as lk
    from llama_kotlin_b200 import GGMLCGraph, GGMLOp, GGMLTensor, GGMLType

    gt = getattr(GGMLType, args.qtype)
    bb = BLOCK_BYTES[args.qtype]
    be = lk.GGMLCudaBackend(0)
    lkc = be.lk
    if args.generic:
        lkc.set_option("gemm_tc_min_n", 0)
    if args.cluster:
        lkc.set_option("gemm_tc_cluster", args.cluster)
    if args.precision >= 0:
        lkc.set_option("gemm_tc_precision", args.precision)
    rng = np.random.default_rng(1)
    for shp in args.shapes:
        M, K, N = (int(v) for v in shp.split("x"))
        wb = M * K // 32 * bb
        ncopies = max(2, (args.rotate_mb << 20) // wb)
        ga = lk.GGMLGraphAllocator(be, size=ncopies * (wb + 256) + (K * N + M * N) * 4 + (1 << 20))
        proto = synth_blocks(rng, M, K, args.qtype)
        x = rng.normal(0, 1, K * N).astype(np.float32)
        b = ga.allocateTensor(GGMLTensor(GGMLType.F32, [N, K]))
        ga.setBytes(b, x)
        dst = ga.allocateTensor(GGMLTensor(GGMLType.F32, [N, M]))
        nodes, first = [], None
        for _ in range(ncopies):
            a = ga.allocateTensor(GGMLTensor(gt, [K, M]))
            if first is None:
                ga.setBytes(a, proto)
                first = a
            else:
                ga.deviceBuffer.copyTensor(first, a)
            nodes.append(GGMLTensor(GGMLType.F32, [N, M], op=GGMLOp.MUL_MAT, src=[a, b] + [None] * 8, bufferId=0,
                                    dataOffset=dst.dataOffset, buffer=ga.deviceBuffer))
        graph = GGMLCGraph(nodes=nodes, allocator=ga)
        before = lkc.get_option("gemm_tc_launches")     # counted when the sequence is captured, not per replay
        for _ in range(3):
            assert be.graphCompute(graph) == lk.GGMLStatus.SUCCESS
        lkc.sync()
        lkc.timer_start()
        for _ in range(args.reps):
            be.graphCompute(graph)
        ms = lkc.timer_stop() / (args.reps * ncopies)
        print(json.dumps({"qtype": args.qtype, "M": M, "K": K, "N": N, "us": round(ms * 1e3, 2),
                          "TFLOPs": round(2.0 * M * K * N / (ms * 1e-3) / 1e12, 1),
                          "weight_GBps": round(wb / (ms * 1e-3) / 1e9, 1), "copies": ncopies,
                          "path": "generic" if args.generic else "tcgen05", "cluster": lkc.get_option("gemm_tc_cluster"),
                          "precision": lkc.get_option("gemm_tc_precision"),
                          "tc_launches_seen": lkc.get_option("gemm_tc_launches") - before}), flush=True)


if __name__ == "__main__":
    main()
