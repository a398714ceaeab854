"""Host-side logic of the row-sharded matmul (llama.kotlin_b200/sharded.py) on CPU: two processes
over torch.distributed's gloo backend play the two ranks.  Each rank takes its zero-copy byte slice
of the packed weight, computes its rows with the ORACLE (standing in for the GPU kernel, which the
-m gpu tests cover), all-gathers the slices the way bench.py exchanges handles, and rank 0 checks the
gathered vector against the unsharded oracle result bit for bit (no reduction is involved, so the
sharded and unsharded outputs must be identical)."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_main(rank: int, world: int, port: int, out_dir: str):
    import torch.distributed as dist
    sys.path.insert(0, ROOT)
    import llama_kotlin_b200 as lk
    from oracle import oracle as O
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        O.build()
        for qt, bb, M, K in ((2, 18, 96, 256), (3, 20, 50, 128), (6, 34, 33, 64)):
            rng = np.random.default_rng(123)           # same data on every rank
            w = rng.normal(0, 0.02, M * K).astype(np.float32)
            x = rng.normal(0, 1, K).astype(np.float32)
            blocks = O.quantize(qt, w)
            row0, rows = lk.shard_rows(M, world, rank)
            b0, nb = lk.shard_byte_range(M, K, bb, world, rank)
            assert b0 == row0 * (K // 32) * bb and nb == rows * (K // 32) * bb
            mine = O.matmul(qt, blocks[b0:b0 + nb], K, rows, x, 1) if rows else np.zeros(0, np.float32)
            gathered = [None] * world
            dist.all_gather_object(gathered, (row0, rows, mine.tobytes()))
            # the plan tiles [0, M) without gaps or overlap, in rank order
            pos = 0
            full = np.empty(M, np.float32)
            for r0, n, raw in gathered:
                assert r0 == pos
                full[r0:r0 + n] = np.frombuffer(raw, np.float32)
                pos += n
            assert pos == M
            ref = O.matmul(qt, blocks, K, M, x, 1)
            assert np.array_equal(full.view(np.uint32), ref.view(np.uint32)), (qt, rank)
        # exchange() contract used by ShardGroup: list indexed by rank
        got = [None] * world
        dist.all_gather_object(got, ("handle-of-rank", rank))
        assert [g[1] for g in got] == list(range(world))
        open(os.path.join(out_dir, f"ok{rank}"), "w").write("ok")
    finally:
        dist.destroy_process_group()


def test_sharded_matmul_two_ranks_gloo(tmp_path):
    import torch.multiprocessing as mp
    world = 2
    mp.spawn(_rank_main, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    assert all(os.path.exists(tmp_path / f"ok{r}") for r in range(world))


@pytest.mark.parametrize("M,world,unit", [(4096, 1, 1), (4096, 8, 1), (11008, 8, 1), (28672, 8, 1), (8192, 4, 1),
                                          (100, 3, 4), (7, 8, 1)])
def test_shard_plan_tiles_rows(M, world, unit):
    import llama_kotlin_b200 as lk
    pos = 0
    sizes = []
    for r in range(world):
        r0, n = lk.shard_rows(M, world, r, unit)
        assert r0 == pos and n % unit == 0
        pos += n
        sizes.append(n)
    assert pos == M
    assert max(sizes) - min(sizes) <= unit


def test_shard_plan_rejects_bad_arguments():
    import llama_kotlin_b200 as lk
    with pytest.raises(lk.IllegalArgumentException):
        lk.shard_rows(4096, 4, 4)
    with pytest.raises(lk.IllegalArgumentException):
        lk.shard_rows(10, 2, 0, unit=4)
