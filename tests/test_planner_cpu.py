"""Host logic of GGMLCudaBackend.graphCompute without a GPU: the fusion / chaining plan over the decoder-layer graph
(LlamaDecoderLayer.forward's op order, model/LlamaModel.kt:253-285; LlamaAttention.forward's core,
model/LlamaAttention.kt:87-123) and the replay key.  The library calls are replaced by a recorder, so what is checked
is WHICH calls the planner makes, in which order and with which tensors — the calls themselves are the `-m gpu` tests'
business (tests/test_gpu_parity.py, test_gpu_attention.py, test_gpu_chain.py)."""

import llama_kotlin_b200.ggml_cuda_backend as backend_mod
from llama_kotlin_b200 import GGMLCGraph, GGMLOp, GGMLTensor, GGMLType
from llama_kotlin_b200.ggml_alloc import GGMLGraphAllocator
from llama_kotlin_b200.ggml_cuda_backend import GGMLCudaBackend
from llama_kotlin_b200.llama_layer import alloc_attention, alloc_layer, build_decoder_graph

HIDDEN, FFN, HEADS = 128, 320, 2


class Recorder:
    """Stands in for llama_kotlin_b200.ggml_compute_ops inside the backend module."""

    def __init__(self, chain_ok=True, fused_ok=True):
        self.calls = []
        self.chain_ok, self.fused_ok = chain_ok, fused_ok

    def computeMatMulFused(self, ga, ctx, a, b, dst, **kw):
        self.calls.append(("fused", a, b, dst, kw))
        return self.fused_ok

    def computeMatMulChain(self, ga, ctx, links):
        self.calls.append(("chain", list(links)))
        return self.chain_ok

    def computeAttentionStep(self, *args, **kw):
        self.calls.append(("attn_step",))

    def __getattr__(self, name):              # every other compute* call: remember its name
        if not name.startswith("compute"):
            raise AttributeError(name)
        return lambda *a, **k: self.calls.append((name,))


def make_backend(chain=None):
    be = object.__new__(GGMLCudaBackend)      # no context: the planner itself never touches the library
    be.fuseOps, be.chainOps, be.maxChain = True, chain, 8
    return be


def make_graph(nlayers=2, attention=False):
    ga = GGMLGraphAllocator(None, size=64 << 20)
    x = ga.allocateTensor(GGMLTensor(GGMLType.F32, [HIDDEN]))
    stack = [alloc_layer(ga, GGMLType.Q4_0, HIDDEN, FFN) for _ in range(nlayers)]
    attn = alloc_attention(ga, HEADS, HIDDEN // HEADS, 16, nlayers) if attention else None
    graph, out = build_decoder_graph(ga, stack, x, HIDDEN, FFN, attn=attn)
    return ga, graph


def run(be, ga, graph, rec, monkeypatch):
    monkeypatch.setattr(backend_mod, "ops", rec)
    be._runNodes(ga, graph, fuse=True)
    return rec.calls


def test_plan_of_the_layer_graph():
    ga, graph = make_graph(2)
    assert graph.nNodes == 26
    be = make_backend()
    rms_of, add_of, skip = be._plan(graph)
    ops_of = lambda idx: [graph.nodes[i].op for i in idx]                      # noqa: E731
    # per layer: both RMS_NORMs fold into the matmuls that read them, both ADDs into the matmul that feeds them,
    # Q/K/V and gate/up are stacked, SILU and MUL go into the gate/up epilogue
    assert ops_of(set(rms_of.values())) == [GGMLOp.RMS_NORM] * 4
    assert ops_of(add_of.values()) == [GGMLOp.ADD] * 4
    assert sorted(len(m) for m in be._groups.values()) == [2, 2, 3, 3]
    assert len(be._pairs) == 2 and all(ops_of(p) == [GGMLOp.SILU, GGMLOp.MUL] for p in be._pairs.values())
    launched = [i for i in range(graph.nNodes) if i not in skip]
    assert len(launched) == 8 and all(i in be._gemv for i in launched)         # 26 nodes -> 8 launches
    assert be._steps == {}


def test_one_position_context_keeps_four_launches_per_layer(monkeypatch):
    ga, graph = make_graph(2)
    calls = run(make_backend(None), ga, graph, Recorder(), monkeypatch)
    assert [c[0] for c in calls] == ["fused"] * 8                              # no attention step: nothing is chained
    kws = [c[4] for c in calls]
    assert [sorted(k) for k in kws[:4]] == [["eps", "extra", "norm_w", "norm_x", "store_b"], ["residual", "sum_dst"],
                                            ["eps", "extra", "mul_dst", "norm_w", "norm_x", "silu_dst", "store_b"],
                                            ["residual", "sum_dst"]]
    assert len(kws[0]["extra"]) == 2 and len(kws[2]["extra"]) == 1


def test_explicit_chain_takes_all_dependent_launches(monkeypatch):
    ga, graph = make_graph(3)
    be = make_backend(True)
    calls = run(be, ga, graph, Recorder(), monkeypatch)
    assert [c[0] for c in calls] == ["chain", "chain"]                         # 12 launches: 8 + 4 links
    assert [len(c[1]) for c in calls] == [8, 4]
    # every link reads something the link before it wrote (that is what makes it a chain)
    links = calls[0][1]
    for (a0, b0, d0, k0), (a1, b1, d1, k1) in zip(links, links[1:]):
        written = [d0] + [d for _, d in (k0 or {}).get("extra", [])] + [t for t in ((k0 or {}).get("sum_dst"), (k0 or {}).get("mul_dst")) if t is not None]
        read = [(k1 or {}).get("norm_x", b1), (k1 or {}).get("residual")]
        assert any(r is not None and be._span(r) == be._span(w) for r in read for w in written)
    be.maxChain = 3
    calls = run(be, ga, graph, Recorder(), monkeypatch)
    assert [len(c[1]) for c in calls] == [3, 3, 3, 3]


def test_attention_graph_chains_between_the_attention_steps(monkeypatch):
    ga, graph = make_graph(2, attention=True)
    assert graph.nNodes == 36
    be = make_backend(None)                                                    # the planner's own choice
    calls = run(be, ga, graph, Recorder(), monkeypatch)
    assert [c[0] for c in calls] == ["fused", "attn_step", "chain", "attn_step", "chain"]
    assert [len(c[1]) for c in calls if c[0] == "chain"] == [4, 3]             # [O, gate/up, down, next Q/K/V], [O, gate/up, down]
    assert len(be._steps) == 2
    be.chainOps = False
    calls = run(be, ga, graph, Recorder(), monkeypatch)
    assert [c[0] for c in calls] == ["fused", "attn_step"] + ["fused"] * 4 + ["attn_step"] + ["fused"] * 3


def test_declined_chain_falls_back_to_the_same_single_launches(monkeypatch):
    """LK_ENOTIMPL from lk_compute_matmul_chain: the links go out one by one, with exactly the tensors and fusion
    keywords the unchained plan uses (including which launch stores the shared RMS_NORM result)."""
    ga, graph = make_graph(2, attention=True)
    plain = [c for c in run(make_backend(False), ga, graph, Recorder(), monkeypatch) if c[0] == "fused"]
    declined = [c for c in run(make_backend(True), ga, graph, Recorder(chain_ok=False), monkeypatch) if c[0] == "fused"]
    assert len(plain) == len(declined) == 8
    for p, d in zip(plain, declined):
        assert p[1] is d[1] and p[2] is d[2] and p[3] is d[3]
        assert sorted(p[4]) == sorted(d[4]) and p[4].get("store_b") == d[4].get("store_b")


def test_independent_gemvs_are_never_chained(monkeypatch):
    """The bench step's shape: GEMVs with their own activations and outputs keep the overlapped single launches."""
    ga = GGMLGraphAllocator(None, size=8 << 20)
    nodes = []
    for _ in range(6):
        a = ga.allocateTensor(GGMLTensor(GGMLType.Q4_0, [128, 64]))
        b = ga.allocateTensor(GGMLTensor(GGMLType.F32, [1, 128]))
        d = ga.allocateTensor(GGMLTensor(GGMLType.F32, [1, 64]))
        d.op, d.src = GGMLOp.MUL_MAT, [a, b] + [None] * 8
        nodes.append(d)
    graph = GGMLCGraph(nodes=nodes, allocator=ga)
    calls = run(make_backend(True), ga, graph, Recorder(), monkeypatch)
    assert [c[0] for c in calls] == ["computeMatMul"] * 6


def test_fast_key_follows_node_identity_and_offsets():
    ga, graph = make_graph(1)
    k0 = GGMLCudaBackend._fastKey(graph)
    assert GGMLCudaBackend._fastKey(graph) == k0
    assert GGMLCudaBackend._fastKey(GGMLCGraph(nodes=list(graph.nodes), allocator=ga)) == k0      # same nodes: same key
    graph.nodes[3].dataOffset += 64
    assert GGMLCudaBackend._fastKey(graph) != k0
    graph.nodes[3].dataOffset -= 64
    dropped = GGMLCGraph(nodes=graph.nodes[:-1], allocator=ga)
    assert GGMLCudaBackend._fastKey(dropped) != k0
    n = graph.nodes[5]
    clone = GGMLTensor(n.type, list(n.ne), op=n.op, src=list(n.src), bufferId=n.bufferId, dataOffset=n.dataOffset, buffer=n.buffer)
    swapped = GGMLCGraph(nodes=graph.nodes[:5] + [clone] + graph.nodes[6:], allocator=ga)
    assert GGMLCudaBackend._fastKey(swapped) != k0                              # another object in the same place
    assert GGMLCudaBackend._signature(swapped) == GGMLCudaBackend._signature(graph)   # ... that the full signature equates
