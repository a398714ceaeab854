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
    be.useLibraryPlanner = False              # these tests watch the Python planner's calls; the C planner is compared below
    return be


def make_graph(nlayers=2, attention=False, hidden=HIDDEN, ffn=FFN, qtype=GGMLType.Q4_0):
    ga = GGMLGraphAllocator(None, size=64 << 20)
    x = ga.allocateTensor(GGMLTensor(GGMLType.F32, [hidden]))
    stack = [alloc_layer(ga, qtype, hidden, ffn) for _ in range(nlayers)]
    attn = alloc_attention(ga, HEADS, hidden // HEADS, 16, nlayers) if attention else None
    graph, out = build_decoder_graph(ga, stack, x, hidden, ffn, attn=attn)
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


def test_one_position_context_chains_what_the_mma_chain_kernel_takes(monkeypatch):
    """The planner's own choice: Q4_0 GEMVs with K % 256 == 0 (K1m, lk_gemv_mma.cu, takes them) go out as one chained launch
    even without an attention step (so do the other two block types); one K that is not a multiple of 256 keeps the separate launches."""
    ga, graph = make_graph(2, hidden=256, ffn=768)
    calls = run(make_backend(None), ga, graph, Recorder(), monkeypatch)
    assert [c[0] for c in calls] == ["chain"] and len(calls[0][1]) == 8
    ga, graph = make_graph(2, hidden=256, ffn=704)                  # the down projection's K = 704
    assert [c[0] for c in run(make_backend(None), ga, graph, Recorder(), monkeypatch)] == ["fused"] * 8
    ga, graph = make_graph(2, hidden=256, ffn=768, qtype=GGMLType.Q8_0)          # every type has a chained kernel
    assert [c[0] for c in run(make_backend(None), ga, graph, Recorder(), monkeypatch)] == ["chain"]


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


# ---- the planner behind the C ABI (lk_graph_plan, csrc/lk_planner.cu): same decisions as the Python planner ------------
import pytest

from llama_kotlin_b200 import _binding as B


def _python_launches(be, ga, graph, monkeypatch):
    """What the Python planner issues, as comparable records keyed by node index."""
    idx = {id(n): i for i, n in enumerate(graph.nodes)}
    calls = run(be, ga, graph, Recorder(), monkeypatch)
    out = []

    def fused_rec(a, b, dst, kw, chain):
        kw = kw or {}
        r = dict(kind="fused" if kw else "node", node=idx[id(dst)], chain=chain,
                 rms=None, add=idx[id(kw["sum_dst"])] if "sum_dst" in kw else None,
                 extra=[idx[id(d)] for _, d in kw.get("extra", [])],
                 silu=idx[id(kw["silu_dst"])] if kw.get("silu_dst") is not None else None,
                 mul=idx[id(kw["mul_dst"])] if kw.get("mul_dst") is not None else None,
                 store_b=bool(kw.get("store_b", True)) if "norm_x" in kw else None)
        if "norm_x" in kw:      # the RMS_NORM node whose dst is this launch's b
            r["rms"] = next(i for i, n in enumerate(graph.nodes) if n.op == GGMLOp.RMS_NORM and n.dataOffset == b.dataOffset)
        return r
    cid = 0
    for c in calls:
        if c[0] == "fused":
            out.append(fused_rec(c[1], c[2], c[3], c[4], -1))
        elif c[0] == "chain":
            for (a, b, dst, kw) in c[1]:
                out.append(fused_rec(a, b, dst, kw, cid))
            cid += 1
        elif c[0] == "attn_step":
            out.append(dict(kind="attn"))
        else:
            out.append(dict(kind="node", name=c[0]))
    return out


def _library_launches(ga, graph, chain_mode):
    plan = GGMLCudaBackend.libraryPlan(ga, graph, chain_mode)
    out = []
    for L in plan:
        if L["kind"] == B.LK_PLAN_ATTN_STEP:
            out.append(dict(kind="attn"))
        elif L["kind"] == B.LK_PLAN_NODE and L["chain"] < 0 and graph.nodes[L["node"]].op != GGMLOp.MUL_MAT:
            out.append(dict(kind="node", name=None))
        else:
            nn = lambda v: None if v < 0 else v   # noqa: E731
            out.append(dict(kind="fused" if L["kind"] == B.LK_PLAN_FUSED else "node", node=L["node"], chain=L["chain"], rms=nn(L["rms"]),
                            add=nn(L["add"]), extra=L["extra"], silu=nn(L["silu"]), mul=nn(L["mul"]),
                            store_b=bool(L["store_b"]) if L["rms"] >= 0 else None))
    return out


@pytest.mark.parametrize("dims", [(HIDDEN, FFN), (256, 768)])
@pytest.mark.parametrize("attention", [False, True])
@pytest.mark.parametrize("chain", [None, True, False])
def test_library_planner_makes_the_python_planner_s_launches(monkeypatch, attention, chain, dims):
    """lk_graph_plan through ctypes on the decoder-layer graph (with and without the attention core, chaining on / off /
    automatic): launch for launch what GGMLCudaBackend._plan / _runNodes issue — fused neighbours, stacked groups, the
    SwiGLU pair, which launch stores the RMS_NORM's dst, which launches form one chain, the attention steps."""
    ga, graph = make_graph(2, attention=attention, hidden=dims[0], ffn=dims[1])
    py = _python_launches(make_backend(chain), ga, graph, monkeypatch)
    lib = _library_launches(ga, graph, -1 if chain is None else int(chain))
    assert len(py) == len(lib)
    for a, b in zip(py, lib):
        assert a["kind"] == b["kind"], (a, b)
        if a["kind"] == "fused" or (a["kind"] == "node" and "node" in a and "node" in b):
            for k in ("node", "chain", "rms", "add", "extra", "silu", "mul", "store_b"):
                assert a[k] == b[k], (k, a, b)


def test_library_planner_plain_graph_and_argument_errors():
    ga = GGMLGraphAllocator(None, size=1 << 20)
    a = ga.allocateTensor(GGMLTensor(GGMLType.F32, [8]))
    b = ga.allocateTensor(GGMLTensor(GGMLType.F32, [8]))
    d = GGMLTensor(GGMLType.F32, [8], op=GGMLOp.ADD, src=[a, b] + [None] * 8, bufferId=0, dataOffset=ga.allocateTensor(GGMLTensor(GGMLType.F32, [8])).dataOffset)
    g = GGMLCGraph(nodes=[d], allocator=ga)
    plan = GGMLCudaBackend.libraryPlan(ga, g)
    assert len(plan) == 1 and plan[0]["kind"] == B.LK_PLAN_NODE and plan[0]["node"] == 0 and plan[0]["chain"] == -1
    n = B.C.c_int32()
    assert B.lib().lk_graph_plan(None, 3, -1, None, 0, B.C.byref(n)) == B.LK_EINVAL_ARG
    assert B.lib().lk_graph_plan(None, 0, -1, None, 0, B.C.byref(n)) == B.LK_OK and n.value == 0
