import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import llama_kotlin_b200 as lk
from llama_kotlin_b200 import GGMLCGraph, GGMLOp, GGMLTensor, GGMLType
import layer_bench
import bench
mode = sys.argv[1]
be = lk.GGMLCudaBackend(0)
lkc = be.lk
def go(tag, ctx, chain=None):
    be.chainOps = chain
    r = layer_bench.run(lk, be, "Q4_0", layers=32, steps=30, warmup=3, ctx_len=ctx)
    print(mode, tag, ctx, chain, round(r["us_per_layer"], 2), r["chained_gemv_launches_per_token"], flush=True)
keep = []
if mode in ("headline", "both"):
    layers, SH = 8, bench.SHAPES
    sumK, sumM = sum(K for _, K in SH), sum(M for M, _ in SH)
    ga = lk.GGMLGraphAllocator(be, size=layers * sum(M * K // 32 * 18 + 256 for M, K in SH) + layers * (sumK + sumM) * 4 + (1 << 20))
    rng = np.random.default_rng(1)
    acts = ga.allocateTensor(GGMLTensor(GGMLType.F32, [layers * sumK]))
    dst = ga.allocateTensor(GGMLTensor(GGMLType.F32, [layers * sumM]))
    ga.setBytes(acts, rng.normal(0, 1, layers * sumK).astype(np.float32))
    nodes, koff, moff = [], 0, 0
    for l in range(layers):
        for M, K in SH:
            a = ga.allocateTensor(GGMLTensor(GGMLType.Q4_0, [K, M]))
            ga.setBytes(a, bench.synth_blocks(rng, M, K, "Q4_0"))
            b = GGMLTensor(GGMLType.F32, [1, K], bufferId=0, dataOffset=acts.dataOffset + koff * 4, buffer=ga.deviceBuffer)
            d = GGMLTensor(GGMLType.F32, [1, M], op=GGMLOp.MUL_MAT, src=[a, b] + [None] * 8, bufferId=0,
                           dataOffset=dst.dataOffset + moff * 4, buffer=ga.deviceBuffer)
            koff += K; moff += M
            nodes.append(d)
    g = GGMLCGraph(nodes=nodes, allocator=ga)
    for _ in range(50):
        be.graphCompute(g)
    lkc.sync()
    keep += [ga, g]
if mode in ("pinned", "both"):
    x = bench.pinned_array(lkc, 8 << 20)
    keep.append(x)
go("then", 2048)
go("then again", 2048)
