"""K1c / K1m — a run of dependent decode GEMVs as ONE persistent launch (lk_compute_matmul_chain): the four launches of
LlamaDecoderLayer.forward (model/LlamaModel.kt:253-285) back to back on one CTA per SM.  K1c (lk_gemv.cu
gemv_chain_kernel, DP4A plane sums, a grid-wide barrier between phases) is checked against the same links issued one by
one bit for bit where no RMSNorm prologue is involved.  K1m (lk_gemv_mma.cu, Q4_0 with K % 256 == 0: plane sums on the
int8 MMA, 16 K-slices per row, activations handed over by mailbox instead of a barrier) sums in a different order, so it
is checked to F32 rounding against the separate launches and with both halves of the parity statement against the
oracle."""
import numpy as np
import pytest

import llama_kotlin_b200 as lk
from conftest import same_f32_bits, synth_acts, synth_weights
from llama_kotlin_b200 import GGMLTensor, GGMLType
from llama_kotlin_b200.ggml_compute_ops import computeMatMulChain
from test_gpu_parity import MATMUL_REL_TOL, Q4_0, Q8_0, QT, assert_matmul_parity, rel_err

CHAIN_ORDER_TOL = 2e-6      # K1m against the separate launches: same arithmetic, different F32 summation order

pytestmark = pytest.mark.gpu
BB = {"Q4_0": 18, "Q8_0": 34, "Q4_1": 20}


def _dev(ga):
    def dev(t, shape, data=None):
        tt = ga.allocateTensor(GGMLTensor(t, shape))
        if data is not None:
            ga.setBytes(tt, data)
        return tt
    return dev


@pytest.mark.parametrize("mma", [1, 0])
@pytest.mark.parametrize("tname", ["Q4_0", "Q8_0", "Q4_1"])
@pytest.mark.parametrize("hidden,ffn", [(128, 320), (512, 1408), (512, 1536), (1024, 2816), (4096, 11008)])
def test_chain_equals_the_links_one_by_one(O, tname, hidden, ffn, mma):
    """plain -> residual -> gate/up + SwiGLU -> down + residual: four links.  K1c: every tensor any link writes is
    bit-identical to what the same four lk_compute_matmul_fused calls store.  K1m (mma = 1, where it takes the chain):
    equal to F32 rounding, and the last link's product meets both halves of the parity statement on its own input."""
    if mma == 0 and (hidden % 256 or ffn % 256):
        pytest.skip("K1c takes this chain in the mma = 1 run already")
    gt, ot = QT[tname]
    be = lk.GGMLCudaBackend(0)
    be.lk.set_option("gemv_chain_mma", mma)
    wbytes = (2 * hidden * hidden + 3 * hidden * ffn) // 32 * BB[tname]
    ga = lk.GGMLGraphAllocator(be, size=wbytes + 40 * (hidden + ffn) * 4 + (1 << 20))
    dev = _dev(ga)
    W1 = dev(gt, [hidden, hidden], O.quantize(ot, synth_weights(7000, hidden, hidden)))
    W2 = dev(gt, [hidden, hidden], O.quantize(ot, synth_weights(7001, hidden, hidden)))
    Wg = dev(gt, [hidden, ffn], O.quantize(ot, synth_weights(7002, ffn, hidden)))
    Wu = dev(gt, [hidden, ffn], O.quantize(ot, synth_weights(7003, ffn, hidden)))
    Wd = dev(gt, [ffn, hidden], O.quantize(ot, synth_weights(7004, hidden, ffn)))
    x0, r0 = synth_acts(7005, hidden), synth_acts(7006, hidden)
    x = dev(GGMLType.F32, [1, hidden], x0)
    res = dev(GGMLType.F32, [hidden], r0)

    def outputs():
        z = lambda n, flat=False: dev(GGMLType.F32, [n] if flat else [1, n], np.zeros(n, np.float32))   # noqa: E731
        return dict(y1=z(hidden), y2=z(hidden), s2=z(hidden, True), g=z(ffn), u=z(ffn), act=z(ffn, True), h=z(ffn, True),
                    y3=z(hidden), s3=z(hidden, True))

    def links(o):
        col = lambda t: GGMLTensor(GGMLType.F32, [1, t.numElements()], bufferId=t.bufferId, dataOffset=t.dataOffset, buffer=t.buffer)   # noqa: E731
        s2b, hb = col(o["s2"]), col(o["h"])
        return [(W1, x, o["y1"], None),
                (W2, o["y1"], o["y2"], dict(residual=res, sum_dst=o["s2"])),
                (Wg, s2b, o["g"], dict(extra=[(Wu, o["u"])], silu_dst=o["act"], mul_dst=o["h"])),
                (Wd, hb, o["y3"], dict(residual=o["s2"], sum_dst=o["s3"]))]
    one, ch = outputs(), outputs()
    for a, b, d, kw in links(one):
        if kw is None:
            lk.computeMatMul(ga, None, a, b, d)
        else:
            assert lk.computeMatMulFused(ga, None, a, b, d, **kw)
    n0, m0 = be.lk.get_option("gemv_chain_launches"), be.lk.get_option("gemv_chain_mma_launches")
    assert computeMatMulChain(ga, None, links(ch))
    be.lk.sync()
    assert be.lk.get_option("gemv_chain_launches") == n0 + 1
    assert be.lk.get_option("gemv_chain_timeouts") == 0
    took_mma = be.lk.get_option("gemv_chain_mma_launches") - m0
    if not mma or hidden % 256 or ffn % 256:
        assert took_mma == 0
    elif tname != "Q8_0" or ffn <= 4096:          # (Q8_0 rows of K = 11008 would need five column chunks: K1c is the faster chain there)
        assert took_mma == 1
    for k in one:
        if took_mma:
            assert rel_err(ga.getFloats(ch[k]), ga.getFloats(one[k])) <= CHAIN_ORDER_TOL, k
        else:
            assert same_f32_bits(ga.getFloats(ch[k]), ga.getFloats(one[k])), k
    if took_mma:    # the down projection on the activations the chain itself produced
        wd = O.quantize(ot, synth_weights(7004, hidden, ffn))
        got = ga.getFloats(ch["y3"])
        assert_matmul_parity(O, ot, wd, ffn, hidden, ga.getFloats(ch["h"]), 1, got, tag="K1m down")
    # and the first link against the oracle
    assert rel_err(ga.getFloats(ch["y1"]), O.matmul(ot, O.quantize(ot, synth_weights(7000, hidden, hidden)), hidden, hidden, x0, 1)) <= MATMUL_REL_TOL
    be.free()


def test_chain_with_rms_prologues_and_many_launches(O):
    """Two links with the RMSNorm prologue, launched 50 times back to back (the barrier words are reused by every
    launch): same answer every time, within the 1/rms rounding of the separate launches."""
    hidden, ffn = 512, 1408
    be = lk.GGMLCudaBackend(0)
    ga = lk.GGMLGraphAllocator(be, size=(hidden * hidden + 2 * hidden * ffn) // 32 * 18 + 64 * (hidden + ffn) * 4 + (1 << 20))
    dev = _dev(ga)
    Wq = dev(GGMLType.Q4_0, [hidden, hidden], O.quantize(Q4_0, synth_weights(7100, hidden, hidden)))
    Wg = dev(GGMLType.Q4_0, [hidden, ffn], O.quantize(Q4_0, synth_weights(7101, ffn, hidden)))
    Wu = dev(GGMLType.Q4_0, [hidden, ffn], O.quantize(Q4_0, synth_weights(7102, ffn, hidden)))
    x = dev(GGMLType.F32, [hidden], synth_acts(7103, hidden) * 2)
    w1 = dev(GGMLType.F32, [hidden], (1 + 0.1 * synth_acts(7104, hidden)).astype(np.float32))
    w2 = dev(GGMLType.F32, [hidden], (1 + 0.1 * synth_acts(7105, hidden)).astype(np.float32))

    def run(chain):
        z = lambda n, flat=False: dev(GGMLType.F32, [n] if flat else [1, n], np.zeros(n, np.float32))   # noqa: E731
        n1, q, s, n2, g, u, h = z(hidden), z(hidden), z(hidden, True), z(hidden), z(ffn), z(ffn), z(ffn, True)
        ls = [(Wq, n1, q, dict(norm_x=x, norm_w=w1, eps=1e-6, residual=x, sum_dst=s)),
              (Wg, n2, g, dict(norm_x=s, norm_w=w2, eps=1e-6, extra=[(Wu, u)], mul_dst=h))]
        if chain:
            assert computeMatMulChain(ga, None, ls)
        else:
            for a, b, d, kw in ls:
                assert lk.computeMatMulFused(ga, None, a, b, d, **kw)
        be.lk.sync()
        return [ga.getFloats(t) for t in (n1, q, s, n2, g, u, h)]
    ref = run(False)
    first = run(True)
    for a, b in zip(first, ref):
        assert rel_err(a, b) <= 2e-6
    for _ in range(50):
        again = run(True)
        for a, b in zip(again, first):
            assert same_f32_bits(a, b)
    assert be.lk.get_option("gemv_chain_timeouts") == 0
    be.free()


def test_chain_argument_errors_and_declines(O):
    be = lk.GGMLCudaBackend(0)
    ga = lk.GGMLGraphAllocator(be, size=1 << 22)
    dev = _dev(ga)
    K = M = 128
    a4 = dev(GGMLType.Q4_0, [K, M], O.quantize(Q4_0, synth_weights(1, M, K)))
    a8 = dev(GGMLType.Q8_0, [K, M], O.quantize(Q8_0, synth_weights(2, M, K)))
    x = dev(GGMLType.F32, [1, K], synth_acts(3, K))
    y = dev(GGMLType.F32, [1, M], np.zeros(M, np.float32))
    z = dev(GGMLType.F32, [1, M], np.zeros(M, np.float32))
    with pytest.raises(lk.IllegalArgumentException):
        computeMatMulChain(ga, None, [])
    assert computeMatMulChain(ga, None, [(a4, x, y, None)] * 9) is False            # more links than one launch takes
    assert computeMatMulChain(ga, None, [(a4, x, y, None), (a8, y, z, None)]) is False   # one weight type per chain
    bad = dev(GGMLType.F32, [1, M + 1], np.zeros(M + 1, np.float32))
    with pytest.raises(lk.IllegalArgumentException):                                 # each link is validated like computeMatMul
        computeMatMulChain(ga, None, [(a4, x, y, None), (a4, y, bad, None)])
    be.lk.set_option("gemv_chain", 0)
    assert computeMatMulChain(ga, None, [(a4, x, y, None), (a4, y, z, None)]) is False
    be.lk.set_option("gemv_chain", 1)
    assert computeMatMulChain(ga, None, [(a4, x, y, None), (a4, y, z, None)])
    be.lk.sync()
    ref = O.matmul(Q4_0, O.quantize(Q4_0, synth_weights(1, M, K)), K, M, synth_acts(3, K), 1)
    assert rel_err(ga.getFloats(y), ref) <= MATMUL_REL_TOL
    be.free()


@pytest.mark.parametrize("hidden,ffn", [(256, 704), (4096, 11008)])
def test_layer_graph_planner_chains_the_gemv_launches(O, hidden, ffn):
    """graphCompute with the chain pass on and off: 8 GEMV launches of two layers become one chained launch, every
    tensor the graph names agrees within the RMSNorm rounding, replays repeat bit for bit."""
    from llama_kotlin_b200.llama_layer import alloc_layer, build_decoder_graph, weight_shapes
    outs, launches = [], []
    for chain in (True, False):
        be = lk.GGMLCudaBackend(0)
        be.chainOps = chain
        shapes = weight_shapes(hidden, ffn)
        ga = lk.GGMLGraphAllocator(be, size=2 * (sum(M * K // 32 * 18 for M, K in shapes.values()) + 64 * ffn * 4) + (1 << 20))
        rng = np.random.default_rng(78)
        x = ga.allocateTensor(GGMLTensor(GGMLType.F32, [hidden]))
        ga.setBytes(x, rng.normal(0, 1, hidden).astype(np.float32))
        stack = []
        for li in range(2):
            L = alloc_layer(ga, GGMLType.Q4_0, hidden, ffn)
            for i, (n, (M, K)) in enumerate(shapes.items()):
                ga.setBytes(L.weights[n], O.quantize(Q4_0, synth_weights(650 + 10 * li + i, M, K)))
            ga.setBytes(L.attn_norm, (1 + 0.1 * rng.normal(0, 1, hidden)).astype(np.float32))
            ga.setBytes(L.ffn_norm, (1 + 0.1 * rng.normal(0, 1, hidden)).astype(np.float32))
            stack.append(L)
        graph, out = build_decoder_graph(ga, stack, x, hidden, ffn)
        before = be.lk.launch_count()
        assert be.graphCompute(graph) == lk.GGMLStatus.SUCCESS
        launches.append(be.lk.launch_count() - before)
        first = [ga.getFloats(n) for n in graph.nodes]
        for _ in range(5):
            assert be.graphCompute(graph) == lk.GGMLStatus.SUCCESS
        again = [ga.getFloats(n) for n in graph.nodes]
        for a, b in zip(first, again):
            assert same_f32_bits(a, b)
        if chain:
            assert be.lk.get_option("gemv_chain_launches") == 1
            assert be.lk.get_option("gemv_chain_timeouts") == 0
        outs.append(first)
        be.free()
    assert launches[0] < launches[1]
    for c, u in zip(*outs):
        assert rel_err(c, u) <= 2e-6


@pytest.mark.parametrize("warps", [4, 12, 16])
@pytest.mark.parametrize("hidden,ffn", [(4096, 11008), (512, 1408)])
def test_chain_with_the_gemv_warps_option(O, warps, hidden, ffn):
    """gemv_warps above the chain kernel's 12 consumer warps (16 is what tools/gemv_sweep.py sets) used to give
    a phase with no working warp and a producer that spins for ever.  The target is clamped now; a shape that still
    leaves no warp group is declined.  Either way the call returns and the results are the separate launches'."""
    be = lk.GGMLCudaBackend(0)
    be.lk.set_option("gemv_chain_mma", 0)        # gemv_warps is K1c's option
    ga = lk.GGMLGraphAllocator(be, size=(hidden * hidden + hidden * ffn) // 32 * 18 + 16 * (hidden + ffn) * 4 + (1 << 20))
    dev = _dev(ga)
    W1 = dev(GGMLType.Q4_0, [hidden, hidden], O.quantize(Q4_0, synth_weights(7200, hidden, hidden)))
    W2 = dev(GGMLType.Q4_0, [hidden, ffn], O.quantize(Q4_0, synth_weights(7201, ffn, hidden)))
    x = dev(GGMLType.F32, [1, hidden], synth_acts(7202, hidden))
    z = lambda n: dev(GGMLType.F32, [1, n], np.zeros(n, np.float32))   # noqa: E731
    y1, y2, c1, c2 = z(hidden), z(ffn), z(hidden), z(ffn)
    lk.computeMatMul(ga, None, W1, x, y1)
    lk.computeMatMul(ga, None, W2, y1, y2)
    be.lk.set_option("gemv_warps", warps)
    try:
        took = computeMatMulChain(ga, None, [(W1, x, c1, None), (W2, c1, c2, None)])
        be.lk.sync()
    finally:
        be.lk.set_option("gemv_warps", 0)
    if took:
        assert same_f32_bits(ga.getFloats(c1), ga.getFloats(y1))
        assert same_f32_bits(ga.getFloats(c2), ga.getFloats(y2))
    assert be.lk.get_option("gemv_chain_timeouts") == 0
    be.free()


def _two_links(O, be, ga, hidden, ffn, x0, seed=7300, nib=0):
    dev = _dev(ga)
    w1 = O.quantize(Q4_0, synth_weights(seed, ffn, hidden))
    w2 = O.quantize(Q4_0, synth_weights(seed + 1, hidden, ffn))
    W1 = dev(GGMLType.Q4_0, [hidden, ffn], O.repack_nibbles(Q4_0, w1, True) if nib else w1)
    W2 = dev(GGMLType.Q4_0, [ffn, hidden], O.repack_nibbles(Q4_0, w2, True) if nib else w2)
    x = dev(GGMLType.F32, [1, hidden], x0)
    y1 = dev(GGMLType.F32, [1, ffn], np.zeros(ffn, np.float32))
    y2 = dev(GGMLType.F32, [1, hidden], np.zeros(hidden, np.float32))
    return (w1, w2), (W1, W2), x, y1, y2


def test_mchain_takes_what_it_can_and_declines_the_rest(O):
    """K1m orders its CTAs by the mailbox alone, so it declines chains in which a tensor the launch writes could be read
    (or written again) by another CTA: two phases sharing an output, activations that are an older phase's output.  Those
    go to K1c; either way the results are the oracle's."""
    hidden, ffn = 512, 1536
    be = lk.GGMLCudaBackend(0)
    ga = lk.GGMLGraphAllocator(be, size=8 << 20)
    x0 = synth_acts(7310, hidden)
    (w1, w2), (W1, W2), x, y1, y2 = _two_links(O, be, ga, hidden, ffn, x0)
    m0 = be.lk.get_option("gemv_chain_mma_launches")
    assert computeMatMulChain(ga, None, [(W1, x, y1, None), (W2, y1, y2, None)])
    be.lk.sync()
    assert be.lk.get_option("gemv_chain_mma_launches") == m0 + 1
    r1 = O.matmul(Q4_0, w1, hidden, ffn, x0, 1)
    assert_matmul_parity(O, Q4_0, w1, hidden, ffn, x0, 1, ga.getFloats(y1), ref=r1, tag="link 0")
    assert_matmul_parity(O, Q4_0, w2, ffn, hidden, ga.getFloats(y1), 1, ga.getFloats(y2), tag="link 1")
    # (a) the third link writes the first link's output tensor again: not for K1m
    dev = _dev(ga)
    y3 = dev(GGMLType.F32, [1, hidden], np.zeros(hidden, np.float32))
    first = ga.getFloats(y2).copy()
    assert computeMatMulChain(ga, None, [(W1, x, y1, None), (W2, y1, y2, None), (W1, y2, y1, None)])
    be.lk.sync()
    assert be.lk.get_option("gemv_chain_mma_launches") == m0 + 1          # K1c took it
    assert rel_err(ga.getFloats(y2), first) <= CHAIN_ORDER_TOL
    assert_matmul_parity(O, Q4_0, w1, hidden, ffn, ga.getFloats(y2), 1, ga.getFloats(y1), tag="link 2 via K1c")
    # (b) the third link multiplies with the FIRST link's output (two phases back): not for K1m either
    y4 = dev(GGMLType.F32, [1, hidden], np.zeros(hidden, np.float32))
    assert computeMatMulChain(ga, None, [(W1, x, y1, None), (W2, y1, y2, None), (W2, y1, y4, None)])
    be.lk.sync()
    assert be.lk.get_option("gemv_chain_mma_launches") == m0 + 1
    assert rel_err(ga.getFloats(y4), first) <= CHAIN_ORDER_TOL
    assert be.lk.get_option("gemv_chain_timeouts") == 0
    del y3
    be.free()


def test_mchain_many_launches_graph_replays_and_kernel_switches(O):
    """The mailbox epochs come from the context's monotonic counter, which K1c's barrier advances too: K1m launches,
    K1c launches in between (gemv_chain_mma toggled) and replays of one captured K1m launch must all give the same answer."""
    hidden, ffn = 1024, 2816
    be = lk.GGMLCudaBackend(0)
    ga = lk.GGMLGraphAllocator(be, size=16 << 20)
    x0 = synth_acts(7320, hidden)
    (w1, w2), (W1, W2), x, y1, y2 = _two_links(O, be, ga, hidden, ffn, x0)
    ls = [(W1, x, y1, None), (W2, y1, y2, None)]
    assert computeMatMulChain(ga, None, ls)
    be.lk.sync()
    ref1, ref2 = ga.getFloats(y1).copy(), ga.getFloats(y2).copy()
    for i in range(40):
        be.lk.set_option("gemv_chain_mma", 0 if i % 5 == 4 else 1)
        ga.setBytes(y1, np.zeros(ffn, np.float32)); ga.setBytes(y2, np.zeros(hidden, np.float32))
        assert computeMatMulChain(ga, None, ls)
        be.lk.sync()
        if i % 5 == 4:
            assert rel_err(ga.getFloats(y2), ref2) <= CHAIN_ORDER_TOL
        else:
            assert same_f32_bits(ga.getFloats(y1), ref1) and same_f32_bits(ga.getFloats(y2), ref2), i
    be.lk.set_option("gemv_chain_mma", 1)
    assert be.lk.get_option("gemv_chain_timeouts") == 0
    be.free()


def test_mchain_ggml_nibble_order_and_nonfinite_activations(O):
    hidden, ffn = 512, 1536
    be = lk.GGMLCudaBackend(0)
    be.lk.set_option("nibble_order", 1)
    ga = lk.GGMLGraphAllocator(be, size=8 << 20)
    x0 = synth_acts(7330, hidden)
    (w1, w2), (W1, W2), x, y1, y2 = _two_links(O, be, ga, hidden, ffn, x0, nib=1)
    m0 = be.lk.get_option("gemv_chain_mma_launches")
    assert computeMatMulChain(ga, None, [(W1, x, y1, None), (W2, y1, y2, None)])
    be.lk.sync()
    assert be.lk.get_option("gemv_chain_mma_launches") == m0 + 1
    assert_matmul_parity(O, Q4_0, w1, hidden, ffn, x0, 1, ga.getFloats(y1), tag="ggml order, link 0")
    assert_matmul_parity(O, Q4_0, w2, ffn, hidden, ga.getFloats(y1), 1, ga.getFloats(y2), tag="ggml order, link 1")
    # a NaN and an inf among the activations: every output that touches them is non-finite, like the reference's
    xb = x0.copy(); xb[5] = np.nan; xb[300] = np.inf
    ga.setBytes(x, xb)
    assert computeMatMulChain(ga, None, [(W1, x, y1, None), (W2, y1, y2, None)])
    be.lk.sync()
    assert not np.any(np.isfinite(ga.getFloats(y1))) and not np.any(np.isfinite(ga.getFloats(y2)))
    assert be.lk.get_option("gemv_chain_timeouts") == 0
    be.free()


def test_mchain_links_with_activations_from_outside_the_launch(O):
    """Not every link of a chain reads its predecessor: a link whose activations come from outside the launch is loaded
    directly (no mailbox), a later link may still take the first link's residual-summed output by mailbox.  Three links:
    y1 = W1 x; y2 = W2 z (z from outside); y3 = W3 y2 + y1 as residual (an earlier link's dst with the same row count)."""
    hidden, ffn = 512, 1536
    be = lk.GGMLCudaBackend(0)
    ga = lk.GGMLGraphAllocator(be, size=8 << 20)
    dev = _dev(ga)
    w1 = O.quantize(Q4_0, synth_weights(7400, hidden, hidden))
    w2 = O.quantize(Q4_0, synth_weights(7401, ffn, hidden))
    w3 = O.quantize(Q4_0, synth_weights(7402, hidden, ffn))
    W1, W2, W3 = dev(GGMLType.Q4_0, [hidden, hidden], w1), dev(GGMLType.Q4_0, [hidden, ffn], w2), dev(GGMLType.Q4_0, [ffn, hidden], w3)
    x0, z0 = synth_acts(7403, hidden), synth_acts(7404, hidden)
    x, z = dev(GGMLType.F32, [1, hidden], x0), dev(GGMLType.F32, [1, hidden], z0)
    y1, y2, y3 = (dev(GGMLType.F32, [1, n], np.zeros(n, np.float32)) for n in (hidden, ffn, hidden))
    s3 = dev(GGMLType.F32, [hidden], np.zeros(hidden, np.float32))
    y1_flat = GGMLTensor(GGMLType.F32, [hidden], bufferId=y1.bufferId, dataOffset=y1.dataOffset, buffer=y1.buffer)
    m0 = be.lk.get_option("gemv_chain_mma_launches")
    assert computeMatMulChain(ga, None, [(W1, x, y1, None), (W2, z, y2, None), (W3, y2, y3, dict(residual=y1_flat, sum_dst=s3))])
    be.lk.sync()
    assert be.lk.get_option("gemv_chain_mma_launches") == m0 + 1 and be.lk.get_option("gemv_chain_timeouts") == 0
    r1 = O.matmul(Q4_0, w1, hidden, hidden, x0, 1)
    assert_matmul_parity(O, Q4_0, w1, hidden, hidden, x0, 1, ga.getFloats(y1), ref=r1, tag="link 0")
    assert_matmul_parity(O, Q4_0, w2, hidden, ffn, z0, 1, ga.getFloats(y2), tag="link 1 (outside activations)")
    assert_matmul_parity(O, Q4_0, w3, ffn, hidden, ga.getFloats(y2), 1, ga.getFloats(y3), tag="link 2 (mailbox)")
    assert same_f32_bits(ga.getFloats(s3), (ga.getFloats(y1) + ga.getFloats(y3)).astype(np.float32))
    be.free()


@pytest.mark.parametrize("M", [64, 148, 149, 600])
def test_mchain_fewer_rows_than_ctas_and_ragged_shares(O, M):
    """A link whose matrix has fewer rows than the grid has CTAs (some CTAs own nothing in that phase and still take part
    in the hand-off as consumers of the next one), exactly one row per CTA, and ragged shares."""
    K = 512
    be = lk.GGMLCudaBackend(0)
    ga = lk.GGMLGraphAllocator(be, size=8 << 20)
    dev = _dev(ga)
    w1 = O.quantize(Q4_0, synth_weights(7500 + M, K, K))
    w2 = O.quantize(Q4_0, synth_weights(7600 + M, M, K))
    W1, W2 = dev(GGMLType.Q4_0, [K, K], w1), dev(GGMLType.Q4_0, [K, M], w2)
    x0 = synth_acts(7700 + M, K)
    x = dev(GGMLType.F32, [1, K], x0)
    y1, y2 = dev(GGMLType.F32, [1, K], np.zeros(K, np.float32)), dev(GGMLType.F32, [1, M], np.zeros(M, np.float32))
    m0 = be.lk.get_option("gemv_chain_mma_launches")
    assert computeMatMulChain(ga, None, [(W1, x, y1, None), (W2, y1, y2, None)])
    be.lk.sync()
    assert be.lk.get_option("gemv_chain_mma_launches") == m0 + 1 and be.lk.get_option("gemv_chain_timeouts") == 0
    assert_matmul_parity(O, Q4_0, w1, K, K, x0, 1, ga.getFloats(y1), tag="link 0")
    assert_matmul_parity(O, Q4_0, w2, K, M, ga.getFloats(y1), 1, ga.getFloats(y2), tag=f"link 1, M = {M}")
    be.free()


@pytest.mark.parametrize("tname", ["Q4_0", "Q4_1"])
def test_mchain_llama_13b_shapes(O, tname):
    """hidden 5120, FFN 13824 (Llama-2-13B): K = 13824 is beyond K1c's 12288, K1m stages such rows as column chunks.  Gate/up
    + SwiGLU -> down + residual against the separate launches and, for the down projection, against the oracle."""
    gt, ot = QT[tname]
    hidden, ffn = 5120, 13824
    be = lk.GGMLCudaBackend(0)
    ga = lk.GGMLGraphAllocator(be, size=3 * hidden * ffn // 32 * BB[tname] + 64 * (hidden + ffn) * 4 + (1 << 20))
    dev = _dev(ga)
    wd = O.quantize(ot, synth_weights(7802, hidden, ffn))
    Wg = dev(gt, [hidden, ffn], O.quantize(ot, synth_weights(7800, ffn, hidden)))
    Wu = dev(gt, [hidden, ffn], O.quantize(ot, synth_weights(7801, ffn, hidden)))
    Wd = dev(gt, [ffn, hidden], wd)
    x = dev(GGMLType.F32, [1, hidden], synth_acts(7803, hidden))
    res = dev(GGMLType.F32, [hidden], synth_acts(7804, hidden))

    def run(chain):
        z = lambda n, flat=False: dev(GGMLType.F32, [n] if flat else [1, n], np.zeros(n, np.float32))   # noqa: E731
        g, u, h, y, s = z(ffn), z(ffn), z(ffn, True), z(hidden), z(hidden, True)
        hb = GGMLTensor(GGMLType.F32, [1, ffn], bufferId=h.bufferId, dataOffset=h.dataOffset, buffer=h.buffer)
        ls = [(Wg, x, g, dict(extra=[(Wu, u)], mul_dst=h)), (Wd, hb, y, dict(residual=res, sum_dst=s))]
        if chain:
            m0 = be.lk.get_option("gemv_chain_mma_launches")
            assert computeMatMulChain(ga, None, ls)
            be.lk.sync()
            assert be.lk.get_option("gemv_chain_mma_launches") == m0 + 1 and be.lk.get_option("gemv_chain_timeouts") == 0
        else:
            for a, b, d, kw in ls:
                assert lk.computeMatMulFused(ga, None, a, b, d, **kw)
            be.lk.sync()
        return [ga.getFloats(t) for t in (g, u, h, y, s)]
    one, ch = run(False), run(True)
    for a, b in zip(ch, one):
        assert rel_err(a, b) <= CHAIN_ORDER_TOL
    assert_matmul_parity(O, ot, wd, ffn, hidden, ch[2], 1, ch[3], tag="13B down projection on K1m")
    be.free()
