"""N4 — the attention core around the decode matmuls: ROPE, SCALE, KV-cache append, decode attention and
the 18-node decoder layer, CUDA path (through the C ABI) against the CPU oracle's restatement of
model/LlamaAttention.kt / model/KVCache.kt.  RoPE / SCALE / append: bit-exact.  Attention: <= 1e-5 of max|out|
(chunked softmax vs the reference's sequential F32 sums; the tolerance the repo uses for softmax)."""
import numpy as np
import pytest

import llama_kotlin_b200 as lk
from conftest import same_f32_bits, synth_weights
from llama_kotlin_b200 import GGMLTensor, GGMLType

pytestmark = pytest.mark.gpu

Q4_0, Q8_0 = 2, 6
ATTN_TOL = 1e-5


def rel_err(got, ref):
    ref = np.asarray(ref, np.float64)
    got = np.asarray(got, np.float64)
    return float(np.max(np.abs(got - ref))) / max(float(np.max(np.abs(ref))), 1e-30)


@pytest.fixture(scope="module")
def be():
    b = lk.GGMLCudaBackend(0)
    yield b
    b.free()


@pytest.mark.parametrize("H,D,ntok,pos", [(2, 64, 1, 0), (32, 128, 1, 5), (4, 64, 3, 17), (32, 128, 7, 2040), (8, 16, 2, 1)])
def test_rope_bit_exact(O, H, D, ntok, pos):
    """applyRoPE (model/LlamaAttention.kt:29-72) on host-resident tensors (the op-level drop-in call)."""
    ga = lk.GGMLGraphAllocator(size=8 << 20)
    rng = np.random.default_rng(300 + H + pos)
    x0 = rng.normal(0, 1, H * D * ntok).astype(np.float32)
    maxpos = 2048
    x = ga.allocateTensor(GGMLTensor(GGMLType.F32, [H * D, ntok]))
    tab = ga.allocateTensor(GGMLTensor(GGMLType.F32, [2, D // 2, maxpos]))
    p = ga.allocateTensor(GGMLTensor(GGMLType.I32, [4]))
    d = ga.allocateTensor(GGMLTensor(GGMLType.F32, [H * D, ntok]))
    ga.setBytes(x, x0)
    ga.setBytes(tab, lk.ropeTable(D, 10000.0, maxpos))
    ga.setBytes(p, np.array([pos, 0, 0, 0], np.int32))
    lk.computeRope(ga, None, x, tab, p, D, d)
    assert same_f32_bits(ga.getFloats(d), O.rope(x0, D, H, pos))


def test_rope_against_reference_ggml_rope(O):
    """... and against the reference tree's compiled ggml_rope (mode 0, pairs (2i, 2i+1)); ggml iterates theta in F32,
    so this is a 1e-5 check of the convention (pairing, frequency, sign), the oracle being the bit-exact bar."""
    from oracle import ref as R
    if not R.available():
        pytest.skip("oracle/_ref not shipped")
    H, D, pos = 8, 128, 77
    x0 = np.random.default_rng(9).normal(0, 1, H * D).astype(np.float32)
    ga = lk.GGMLGraphAllocator(size=4 << 20)
    x = ga.allocateTensor(GGMLTensor(GGMLType.F32, [H * D]))
    tab = ga.allocateTensor(GGMLTensor(GGMLType.F32, [2, D // 2, 128]))
    p = ga.allocateTensor(GGMLTensor(GGMLType.I32, [4]))
    d = ga.allocateTensor(GGMLTensor(GGMLType.F32, [H * D]))
    ga.setBytes(x, x0)
    ga.setBytes(tab, lk.ropeTable(D, 10000.0, 128))
    ga.setBytes(p, np.array([pos, 0, 0, 0], np.int32))
    lk.computeRope(ga, None, x, tab, p, D, d)
    assert rel_err(ga.getFloats(d), R.unary(R.OP_ROPE, x0, D, float(pos), 10000.0)) <= 2e-5


def test_rope_argument_errors():
    ga = lk.GGMLGraphAllocator(size=1 << 20)
    x = ga.allocateTensor(GGMLTensor(GGMLType.F32, [96]))
    d = ga.allocateTensor(GGMLTensor(GGMLType.F32, [96]))
    tab = ga.allocateTensor(GGMLTensor(GGMLType.F32, [2, 32, 4]))
    with pytest.raises(lk.IllegalArgumentException):     # 96 is not a multiple of headDim 64
        lk.computeRope(ga, None, x, tab, None, 64, d)
    bad = ga.allocateTensor(GGMLTensor(GGMLType.F32, [2, 16, 4]))
    x2 = ga.allocateTensor(GGMLTensor(GGMLType.F32, [128]))
    d2 = ga.allocateTensor(GGMLTensor(GGMLType.F32, [128]))
    with pytest.raises(lk.IllegalArgumentException):     # table built for another headDim
        lk.computeRope(ga, None, x2, bad, None, 64, d2)


def test_scale_bit_exact(O):
    ga = lk.GGMLGraphAllocator(size=1 << 20)
    x0 = np.random.default_rng(3).normal(0, 3, 4099).astype(np.float32)
    x = ga.allocateTensor(GGMLTensor(GGMLType.F32, [4099]))
    d = ga.allocateTensor(GGMLTensor(GGMLType.F32, [4099]))
    ga.setBytes(x, x0)
    f = 1.0 / float(np.sqrt(np.float32(128)))
    lk.computeScale(ga, None, x, f, d)
    assert same_f32_bits(ga.getFloats(d), O.scale(x0, f))


def _attention_case(O, be, H, D, maxseq, lengths, seed):
    """Fills the cache row by row through KVCache.updateKey / updateValue and checks the attention of the newest query at
    the requested lengths."""
    ga = lk.GGMLGraphAllocator(be, size=2 * maxseq * H * D * 4 + (4 << 20))
    cache = lk.KVCache(maxseq, H, D)
    cache.initialize(ga)
    rng = np.random.default_rng(seed)
    n = max(lengths)
    ks = rng.normal(0, 1, (n, H * D)).astype(np.float32)
    vs = rng.normal(0, 1, (n, H * D)).astype(np.float32)
    qs = rng.normal(0, 1, (n, H * D)).astype(np.float32) * 2.0
    newk = ga.allocateTensor(GGMLTensor(GGMLType.F32, [D, H, 1]))
    newv = ga.allocateTensor(GGMLTensor(GGMLType.F32, [D, H, 1]))
    q = ga.allocateTensor(GGMLTensor(GGMLType.F32, [H * D]))
    out = ga.allocateTensor(GGMLTensor(GGMLType.F32, [H * D]))
    scale = 1.0 / float(np.sqrt(np.float32(D)))
    worst = 0.0
    for s in range(n):
        ga.setBytes(newk, ks[s])
        ga.setBytes(newv, vs[s])
        cache.updateKey(ga, newk)
        cache.updateValue(ga, newv)
        assert cache.getCurrentLength() == s + 1
        if s + 1 in lengths:
            ga.setBytes(q, qs[s])
            cache.setPosition(s)
            lk.computeAttentionDecode(ga, None, q, cache.keys, cache.values, cache.position, scale, out)
            got = ga.getFloats(out)
            ref = O.attention_decode(qs[s], ks[:s + 1], vs[:s + 1], s + 1, D, H)
            worst = max(worst, rel_err(got, ref))
            # run-to-run bit identity (no atomics on data)
            lk.computeAttentionDecode(ga, None, q, cache.keys, cache.values, cache.position, scale, out)
            assert same_f32_bits(ga.getFloats(out), got)
    # the cache holds exactly the appended rows (KVCache.updateKey copy loop)
    kc = ga.getFloats(cache.keys).reshape(maxseq, H * D)
    assert same_f32_bits(kc[:n], ks)
    return worst


@pytest.mark.parametrize("H,D,maxseq,lengths", [
    (2, 64, 64, [1, 2, 33, 64]),                 # one split
    (4, 64, 300, [63, 64, 65, 128, 129, 300]),   # chunk boundaries, ragged last split
    (32, 128, 2048, [1, 500, 2048]),             # Llama-2-7B heads, full context: 32 splits per head
])
def test_attention_decode_matches_oracle(O, be, H, D, maxseq, lengths):
    assert _attention_case(O, be, H, D, maxseq, set(lengths), 40 + H) <= ATTN_TOL


def test_attention_long_context_larger_chunks(O, be):
    """maxSeq above 256 * 64 positions doubles the chunk (splits stay <= 256)."""
    H, D, maxseq, L = 2, 64, 20000, 700
    ga = lk.GGMLGraphAllocator(be, size=2 * maxseq * H * D * 4 + (4 << 20))
    rng = np.random.default_rng(8)
    K = rng.normal(0, 1, (L, H * D)).astype(np.float32)
    V = rng.normal(0, 1, (L, H * D)).astype(np.float32)
    q0 = rng.normal(0, 1, H * D).astype(np.float32)
    kc = ga.allocateTensor(GGMLTensor(GGMLType.F32, [D, H, maxseq]))
    vc = ga.allocateTensor(GGMLTensor(GGMLType.F32, [D, H, maxseq]))
    p = ga.allocateTensor(GGMLTensor(GGMLType.I32, [4]))
    q = ga.allocateTensor(GGMLTensor(GGMLType.F32, [H * D]))
    out = ga.allocateTensor(GGMLTensor(GGMLType.F32, [H * D]))
    rows = ga.allocateTensor(GGMLTensor(GGMLType.F32, [D, H, L]))
    ga.setBytes(p, np.zeros(4, np.int32))
    ga.setBytes(rows, K)
    lk.computeKvAppend(ga, None, rows, kc, p)        # multi-row append (prefill-style)
    ga.setBytes(rows, V)
    lk.computeKvAppend(ga, None, rows, vc, p)
    ga.setBytes(q, q0)
    ga.setBytes(p, np.array([L - 1, 0, 0, 0], np.int32))
    lk.computeAttentionDecode(ga, None, q, kc, vc, p, 0.125, out)
    assert rel_err(ga.getFloats(out), O.attention_decode(q0, K, V, L, D, H)) <= ATTN_TOL


def test_kv_cache_host_logic_and_errors(be):
    ga = lk.GGMLGraphAllocator(be, size=1 << 20)
    c = lk.KVCache(2, 2, 64)
    with pytest.raises(lk.IllegalStateException):
        c.updateKey(ga, GGMLTensor(GGMLType.F32, [64, 2, 1]))          # "KVCache not initialized"
    c.initialize(ga)
    row = ga.allocateTensor(GGMLTensor(GGMLType.F32, [64, 2, 1]))
    with pytest.raises(lk.IllegalArgumentException):                   # "updateKey must be called before updateValue"
        c.updateValue(ga, row)
    c.updateKey(ga, row)
    c.updateKey(ga, row)
    with pytest.raises(lk.IllegalArgumentException):                   # "Sequence length would exceed maximum"
        c.updateKey(ga, row)
    c.clear()
    assert c.getCurrentLength() == 0
    other = ga.allocateTensor(GGMLTensor(GGMLType.F32, [32, 2, 1]))
    with pytest.raises(lk.IllegalArgumentException):                   # [headDim, numHeads] mismatch
        lk.computeKvAppend(ga, None, other, c.keys, c.position)


@pytest.mark.parametrize("tname,hidden,ffn,H", [("Q4_0", 256, 704, 4), ("Q8_0", 128, 320, 2), ("Q4_0", 4096, 11008, 32)])
def test_decoder_layer_with_attention_decodes_tokens(O, tname, hidden, ffn, H):
    """Two stacked 18-node layers (ROPE, KV append, FLASH_ATTN between the q/k/v and o projections) captured ONCE and
    replayed for consecutive tokens — only the device position scalar and the input change — against the same ops
    composed on the oracle with its own KV cache."""
    from llama_kotlin_b200.llama_layer import alloc_attention, alloc_layer, build_decoder_graph, oracle_layer, weight_shapes
    gt, ot = {"Q4_0": (GGMLType.Q4_0, Q4_0), "Q8_0": (GGMLType.Q8_0, Q8_0)}[tname]
    D, maxseq, ntok = hidden // H, 96, (5 if hidden < 4096 else 3)
    be = lk.GGMLCudaBackend(0)
    shapes = weight_shapes(hidden, ffn)
    bb = {"Q4_0": 18, "Q8_0": 34}[tname]
    wbytes = sum(M * K // 32 * bb for M, K in shapes.values())
    ga = lk.GGMLGraphAllocator(be, size=2 * (wbytes + 4096 + 60 * ffn * 4 + 2 * maxseq * hidden * 4) + (2 << 20))
    rng = np.random.default_rng(4101)
    x = ga.allocateTensor(GGMLTensor(GGMLType.F32, [hidden]))
    stack, host = [], []
    for li in range(2):
        L = alloc_layer(ga, gt, hidden, ffn)
        blocks = {n: O.quantize(ot, synth_weights(4101 + 10 * li + i, M, K)) for i, (n, (M, K)) in enumerate(shapes.items())}
        gains = [(1.0 + 0.1 * rng.normal(0, 1, hidden)).astype(np.float32) for _ in range(2)]
        for n in shapes:
            ga.setBytes(L.weights[n], blocks[n])
        ga.setBytes(L.attn_norm, gains[0])
        ga.setBytes(L.ffn_norm, gains[1])
        stack.append(L)
        host.append((blocks, gains, {"kc": np.zeros(maxseq * hidden, np.float32), "vc": np.zeros(maxseq * hidden, np.float32),
                                     "H": H, "D": D}))
    attn = alloc_attention(ga, H, D, maxseq, 2)
    graph, out = build_decoder_graph(ga, stack, x, hidden, ffn, attn=attn)
    assert graph.nNodes == 36
    captured_before = None
    for t in range(ntok):
        xt = rng.normal(0, 1, hidden).astype(np.float32)
        ga.setBytes(x, xt)
        ga.setBytes(attn[0].position, np.array([t, 0, 0, 0], np.int32))
        assert be.graphCompute(graph) == lk.GGMLStatus.SUCCESS
        if captured_before is None:
            captured_before = dict(be._captured)
        else:
            assert be._captured == captured_before          # replayed, not re-captured
        ref = xt
        for blocks, gains, st in host:
            st["pos"] = t
            ref = oracle_layer(O, ot, blocks, gains[0], gains[1], ref, hidden, ffn, attn=st)
        assert rel_err(ga.getFloats(out), ref) <= 1e-3, t
    # the device caches hold the oracle's rows
    for la, (_, _, st) in zip(attn, host):
        assert rel_err(ga.getFloats(la.keys)[:ntok * hidden], st["kc"][:ntok * hidden]) <= 1e-3
    # with attention steps in the graph the planner's own choice is the chained GEMV launch: [Q/K/V], attention,
    # [O, gate/up, down, next layer's Q/K/V] as one launch, attention, [O, gate/up, down]
    assert be.chainOps is None and be.lk.get_option("gemv_chain_launches") == 2
    assert be.lk.get_option("gemv_chain_timeouts") == 0
    be.free()


# ---- fused decode step: ROPE(q) + ROPE(k) + append k / v + attention in one launch -------------------------------
@pytest.mark.parametrize("H,D,maxseq,positions", [
    (2, 64, 64, [0, 1, 63]),
    (4, 64, 300, [0, 63, 64, 65, 128, 299]),      # chunk boundaries: the owner of the new row changes
    (32, 128, 2048, [0, 127, 128, 1000, 2047]),   # Llama-2-7B heads
])
def test_attention_step_bit_identical_to_the_five_ops(be, H, D, maxseq, positions):
    """lk_compute_attention_step == computeRope x 2 + computeKvAppend x 2 + computeAttentionDecode, bit for bit
    (LlamaAttention.forward's core, model/LlamaAttention.kt:87-123), on caches that already hold earlier rows."""
    n = H * D
    ga = lk.GGMLGraphAllocator(be, size=4 * maxseq * n * 4 + maxseq * D * 4 + (4 << 20))
    rng = np.random.default_rng(H + D)
    table = ga.allocateTensor(GGMLTensor(GGMLType.F32, [2, D // 2, maxseq]))
    ga.setBytes(table, lk.ropeTable(D, 10000.0, maxseq))
    pos = ga.allocateTensor(GGMLTensor(GGMLType.I32, [4]))
    caches = [ga.allocateTensor(GGMLTensor(GGMLType.F32, [D, H, maxseq])) for _ in range(4)]   # k, v (5 ops); k, v (fused)
    base_k = rng.normal(0, 1, maxseq * n).astype(np.float32)
    base_v = rng.normal(0, 1, maxseq * n).astype(np.float32)
    vec = lambda: ga.allocateTensor(GGMLTensor(GGMLType.F32, [n]))   # noqa: E731
    q, k, v, qr1, kr1, o1, qr2, kr2, o2 = (vec() for _ in range(9))
    row = lambda t: GGMLTensor(GGMLType.F32, [D, H, 1], bufferId=t.bufferId, dataOffset=t.dataOffset, buffer=t.buffer)   # noqa: E731
    scale = 1.0 / float(np.sqrt(np.float32(D)))
    for p in positions:
        for c, b in zip(caches, (base_k, base_v, base_k, base_v)):
            ga.setBytes(c, b)
        ga.setBytes(pos, np.array([p, 0, 0, 0], np.int32))
        for t in (q, k, v):
            ga.setBytes(t, rng.normal(0, 1, n).astype(np.float32))
        lk.computeRope(ga, None, q, table, pos, D, qr1)
        lk.computeRope(ga, None, k, table, pos, D, kr1)
        lk.computeKvAppend(ga, None, row(kr1), caches[0], pos)
        lk.computeKvAppend(ga, None, row(v), caches[1], pos)
        lk.computeAttentionDecode(ga, None, qr1, caches[0], caches[1], pos, scale, o1)
        lk.computeAttentionStep(ga, None, q, k, v, table, pos, D, caches[2], caches[3], scale, o2, qRopeDst=qr2, kRopeDst=kr2)
        assert same_f32_bits(ga.getFloats(qr2), ga.getFloats(qr1)), p
        assert same_f32_bits(ga.getFloats(kr2), ga.getFloats(kr1)), p
        assert same_f32_bits(ga.getFloats(caches[2]), ga.getFloats(caches[0])), p
        assert same_f32_bits(ga.getFloats(caches[3]), ga.getFloats(caches[1])), p
        assert same_f32_bits(ga.getFloats(o2), ga.getFloats(o1)), p
        # optional stores may be left out
        ga.setBytes(o2, np.zeros(n, np.float32))
        lk.computeAttentionStep(ga, None, q, k, v, table, pos, D, caches[2], caches[3], scale, o2)
        assert same_f32_bits(ga.getFloats(o2), ga.getFloats(o1)), p


def test_attention_step_argument_errors(be):
    ga = lk.GGMLGraphAllocator(be, size=8 << 20)
    H, D, S = 2, 64, 16
    n = H * D
    table = ga.allocateTensor(GGMLTensor(GGMLType.F32, [2, D // 2, S]))
    pos = ga.allocateTensor(GGMLTensor(GGMLType.I32, [4]))
    kc = ga.allocateTensor(GGMLTensor(GGMLType.F32, [D, H, S]))
    vc = ga.allocateTensor(GGMLTensor(GGMLType.F32, [D, H, S]))
    q, k, v, o = (ga.allocateTensor(GGMLTensor(GGMLType.F32, [n])) for _ in range(4))
    short = ga.allocateTensor(GGMLTensor(GGMLType.F32, [n - 2]))
    with pytest.raises(lk.IllegalArgumentException):
        lk.computeAttentionStep(ga, None, q, short, v, table, pos, D, kc, vc, 0.125, o)
    with pytest.raises(lk.IllegalArgumentException):
        lk.computeAttentionStep(ga, None, q, k, v, table, pos, 32, kc, vc, 0.125, o)      # head_dim != cache ne[0]
    fpos = ga.allocateTensor(GGMLTensor(GGMLType.F32, [4]))
    with pytest.raises(lk.IllegalArgumentException):
        lk.computeAttentionStep(ga, None, q, k, v, table, fpos, D, kc, vc, 0.125, o)      # position must be I32


def test_planner_fuses_the_attention_core(be):
    """the backend planner turns the layer's ROPE, ROPE, CPY, CPY, FLASH_ATTN run into one launch"""
    from llama_kotlin_b200 import llama_layer as LL
    hidden, ffn, H = 256, 704, 4
    ga = lk.GGMLGraphAllocator(be, size=64 << 20)
    layers = [LL.alloc_layer(ga, GGMLType.Q4_0, hidden, ffn)]
    attn = LL.alloc_attention(ga, H, hidden // H, 64, 1)
    x = ga.allocateTensor(GGMLTensor(GGMLType.F32, [hidden]))
    graph, _ = LL.build_decoder_graph(ga, layers, x, hidden, ffn, attn=attn)
    be._plan(graph)
    assert len(be._steps) == 1
    (first, idx), = be._steps.items()
    assert [graph.nodes[i].op for i in idx] == [lk.GGMLOp.ROPE, lk.GGMLOp.ROPE, lk.GGMLOp.CPY, lk.GGMLOp.CPY, lk.GGMLOp.FLASH_ATTN]
