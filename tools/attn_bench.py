#!/usr/bin/env python
"""Times the decode-attention kernel (lk_compute_attention_decode) alone: 32 distinct KV caches (one per layer, > L2 at
long contexts) visited in turn inside one captured sequence, Llama-2-7B heads (32 x 128), F32 cache.
Algorithmic bytes per launch = 2 * ctx * 4096 * 4 (every cached K and V row once) + 2 * 4096 * 4.

  python tools/attn_bench.py --ctx 128 512 2048 8192
"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--ctx", type=int, nargs="*", default=[128, 512, 2048, 8192])
    ap.add_argument("--layers", type=int, default=32)
    ap.add_argument("--steps", type=int, default=20)
    args = ap.parse_args()
    import llama_kotlin_b200 as lk
    from llama_kotlin_b200 import GGMLCGraph, GGMLOp, GGMLTensor, GGMLType
    H, D = 32, 128
    peak = 0.0
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbs", 0))
    except Exception:
        pass
    for ctx in args.ctx:
        be = lk.GGMLCudaBackend(0)
        ga = lk.GGMLGraphAllocator(be, size=args.layers * (2 * ctx * H * D * 4 + 4 * H * D * 4) + (4 << 20))
        rng = np.random.default_rng(1)
        pos = ga.allocateTensor(GGMLTensor(GGMLType.I32, [4]))
        ga.setBytes(pos, np.array([ctx - 1, 0, 0, 0], np.int32))
        fill = rng.normal(0, 1, ctx * H * D).astype(np.float32)
        nodes = []
        import struct
        sbits = struct.unpack("<i", struct.pack("<f", 1.0 / float(np.sqrt(np.float32(D)))))[0]
        for _ in range(args.layers):
            kc = ga.allocateTensor(GGMLTensor(GGMLType.F32, [D, H, ctx]))
            vc = ga.allocateTensor(GGMLTensor(GGMLType.F32, [D, H, ctx]))
            ga.setBytes(kc, fill)
            ga.setBytes(vc, fill)
            q = ga.allocateTensor(GGMLTensor(GGMLType.F32, [H * D]))
            ga.setBytes(q, rng.normal(0, 1, H * D).astype(np.float32))
            o = ga.allocateTensor(GGMLTensor(GGMLType.F32, [H * D], op=GGMLOp.FLASH_ATTN, src=[q, kc, vc, pos] + [None] * 6))
            o.opParams[0] = sbits
            nodes.append(o)
        graph = GGMLCGraph(nodes=nodes, allocator=ga)
        for _ in range(3):
            assert be.graphCompute(graph) == lk.GGMLStatus.SUCCESS
        be.lk.sync()
        be.lk.timer_start()
        for _ in range(args.steps):
            be.graphCompute(graph)
        us = be.lk.timer_stop() * 1e3 / args.steps / args.layers
        nbytes = 2 * ctx * H * D * 4 + 2 * H * D * 4
        res = {"kernel": "attn_decode_kernel<128>", "ctx": ctx, "heads": H, "head_dim": D, "us_per_launch": us,
               "algorithmic_bytes": nbytes, "GBps": nbytes / us / 1e3, "hbm_frac": nbytes / us / 1e3 / peak if peak else None,
               "distinct_caches": args.layers, "cache_bytes_total": args.layers * 2 * ctx * H * D * 4}
        print(json.dumps(res), flush=True)
        be.free()


if __name__ == "__main__":
    main()
