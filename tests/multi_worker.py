"""Worker of tests/test_gpu_multi.py: one process per GPU (torchrun).  Builds the sharded FFN of config 5 at a small size
in every exchange mode, and the same FFN unsharded on rank 0, and writes what it found as one JSON line on rank 0.
SURVEY §8(e): "parity of gathered dst vs single-GPU dst (bit-exact) and vs oracle"."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    rank, world, local_rank = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    json_fd = os.dup(1)
    os.dup2(2, 1)
    import torch
    import torch.distributed as dist

    import llama_kotlin_b200 as lk
    from llama_kotlin_b200 import GGMLTensor, GGMLType
    from llama_kotlin_b200.ffn_sharded import build_sharded_ffn, ffn_buffer_bytes
    from oracle import oracle as O
    O.build()
    torch.cuda.set_device(local_rank)
    dist.init_process_group(backend="cpu:gloo,cuda:nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local_rank))
    hidden, ffn, layers = int(sys.argv[1]), int(sys.argv[2]), 2
    be = lk.GGMLCudaBackend(local_rank)
    be.checkFaults = True

    def exchange(obj):
        out = [None] * world
        dist.all_gather_object(out, obj)
        return out
    group = lk.ShardGroup(be, rank, world, exchange)
    group.init_nccl()        # outside any capture: the communicator cannot be created while a sequence is being recorded
    # full weights, same on every rank (seeded); every rank cuts out its rows
    full = {}
    for li in range(layers):
        for name, (M, K) in (("gate", (ffn, hidden)), ("up", (ffn, hidden)), ("down", (hidden, ffn))):
            w = np.random.default_rng(9000 + 10 * li + len(name)).normal(0, 0.02, M * K).astype(np.float32)
            full[(li, name)] = O.quantize(O.Q4_0, w)
    x0 = np.random.default_rng(9100).normal(0, 1, hidden).astype(np.float32)
    rb = {"gate": hidden // 32 * 18, "up": hidden // 32 * 18, "down": ffn // 32 * 18}

    def run(mode, w, r):
        ga = lk.GGMLGraphAllocator(be, size=ffn_buffer_bytes(w, hidden, ffn, layers, 18) + (1 << 20))
        peers = None
        if w > 1:
            peers = group.map_dst_buffer(ga.deviceBuffer)
            group.barrier()
            be.lk.sync()

        def weights(li, name, rows, K):
            M = ffn if name != "down" else hidden
            row0, n = lk.shard_rows(M, w, r)
            assert n == rows
            return full[(li, name)][row0 * rb[name]:(row0 + rows) * rb[name]]
        S = build_sharded_ffn(ga, group if w > 1 else None, peers, r, w, hidden, ffn, layers, GGMLType.Q4_0, weights, mode)
        ga.setBytes(S.x, x0)
        outs = []
        for rep in range(3):                     # replays: the device-side epochs of the exchange slots must keep working
            assert be.graphCompute(S.graph) == lk.GGMLStatus.SUCCESS, mode
            be.lk.sync()
            outs.append({k: ga.getFloats(S.layers[0][k]).copy() for k in ("g_full", "u_full", "h_full", "y_full")} |
                        {"y_last": ga.getFloats(S.layers[-1]["y_full"]).copy(), "out": ga.getFloats(S.out).copy()})
        be.releaseGraphs()
        return outs

    res = {"world": world, "hidden": hidden, "ffn": ffn}
    bits = lambda a: a.view(np.uint32)   # noqa: E731
    f0, fn_ = lk.shard_rows(ffn, world, rank)
    # g and u are written locally only (this rank's rows; the rest of those vectors is never touched): compare that slice.
    # h, y and everything after them are exchanged: every rank holds all of them.
    FULL = ("h_full", "y_full", "y_last", "out")

    def view(o, k):
        return o[k] if k in FULL else o[k][f0:f0 + fn_]
    got = {}
    for mode in ("sync", "barrier", "nccl"):
        got[mode] = run(mode, world, rank)
        res[mode + "_replays_identical"] = all(np.array_equal(bits(view(got[mode][0], k)), bits(view(got[mode][i], k)))
                                               for i in (1, 2) for k in got[mode][0])
    # every rank holds the same gathered vectors
    digest = {m: [float(np.sum(got[m][0][k].astype(np.float64))) for k in FULL] for m in got}
    alld = exchange(digest)
    res["all_ranks_identical"] = all(d == alld[0] for d in alld)
    # fused == NCCL on this rank, bit for bit, for everything layer 0 writes; the later layer within the 1/rms rounding
    f = got["sync"][0]
    L0 = ("g_full", "u_full", "h_full", "y_full")
    for m in ("barrier", "nccl"):
        same = all(np.array_equal(bits(view(f, k)), bits(view(got[m][0], k))) for k in L0)
        res[f"sync_eq_{m}_layer0"] = all(exchange(bool(same)))
        res[f"sync_vs_{m}_out_err"] = float(np.max(np.abs(f["out"] - got[m][0]["out"])) / max(np.max(np.abs(f["out"])), 1e-30))
    timed_out = exchange(bool(group.timed_out()))
    res["timed_out"] = any(timed_out)
    single = None
    if rank == 0:
        single = run("nccl", 1, 0)[0]            # the unsharded ops on one GPU: all rows of g and u
        res["sync_eq_single_layer0"] = all(np.array_equal(bits(view(f, k)), bits(single[k] if k in FULL else single[k][f0:f0 + fn_])) for k in L0)
        res["sync_vs_single_out_err"] = float(np.max(np.abs(f["out"] - single["out"])) / max(np.max(np.abs(single["out"])), 1e-30))
        # and against the oracle (layer 0)
        g_ref = O.matmul(O.Q4_0, full[(0, "gate")], hidden, ffn, x0, 1, nthreads=4)
        u_ref = O.matmul(O.Q4_0, full[(0, "up")], hidden, ffn, x0, 1, nthreads=4)
        h_ref = (O.silu(g_ref) * u_ref).astype(np.float32)
        y_ref = O.matmul(O.Q4_0, full[(0, "down")], ffn, hidden, h_ref, 1, nthreads=4)
        rel = lambda a, b: float(np.max(np.abs(a - b)) / max(np.max(np.abs(b)), 1e-30))   # noqa: E731
        res["oracle_err"] = max(rel(f["g_full"][f0:f0 + fn_], g_ref[f0:f0 + fn_]), rel(f["u_full"][f0:f0 + fn_], u_ref[f0:f0 + fn_]),
                                rel(f["h_full"], h_ref), rel(f["y_full"], y_ref))
    # computeMatMulSharded alone: gathered dst == single-GPU dst, bit for bit
    M, K = ffn, hidden
    ga = lk.GGMLGraphAllocator(be, size=M * K // 32 * 18 + (K + 2 * M) * 4 + (1 << 20))
    peers = group.map_dst_buffer(ga.deviceBuffer)
    row0, rows = lk.shard_rows(M, world, rank)
    a_sh = ga.allocateTensor(GGMLTensor(GGMLType.Q4_0, [K, rows]))
    ga.setBytes(a_sh, full[(0, "gate")][row0 * rb["gate"]:(row0 + rows) * rb["gate"]])
    xb = ga.allocateTensor(GGMLTensor(GGMLType.F32, [1, K]))
    ga.setBytes(xb, x0)
    dfull = ga.allocateTensor(GGMLTensor(GGMLType.F32, [1, M]))
    ga.setBytes(dfull, np.zeros(M, np.float32))
    dist.barrier()
    group.computeMatMulSharded(ga, a_sh, xb, dfull, row0, M, peers)
    group.barrier()
    be.lk.sync()
    dist.barrier()
    gathered = ga.getFloats(dfull)
    # the gathered dst: the same on every rank, and bit for bit the single-GPU dst
    gd = exchange(float(np.sum(gathered.astype(np.float64))))
    ok = all(v == gd[0] for v in gd)
    if rank == 0:
        ok = ok and bool(np.array_equal(bits(gathered), bits(single["g_full"])))
    res["gathered_gemv_eq_single_gpu"] = ok
    dist.barrier()
    be.lk.sync()
    group.close()
    be.free()
    if rank == 0:
        os.write(json_fd, (json.dumps(res) + "\n").encode())
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
