import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tools"))
import llama_kotlin_b200 as lk
import layer_bench
be = lk.GGMLCudaBackend(0)
for ctx in (0, 128, 512, 2048, 8192):
    out = []
    for chain in (True, False):
        be.chainOps = chain
        r = layer_bench.run(lk, be, "Q4_0", layers=32, steps=30, warmup=3, ctx_len=ctx)
        out.append((round(r["us_per_layer"], 2), round(r["tok_s"], 1)))
    print("ctx", ctx, "chain", out[0], "separate", out[1], flush=True)
