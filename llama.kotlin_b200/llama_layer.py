"""Decode-step graph of one Llama decoder layer, in the op order of the reference's model code
(BASELINE config 4).

`LlamaDecoderLayer.forward` (model/LlamaModel.kt:253-285) supplies the order and the shapes:
RMSNorm -> q/k/v linear (MUL_MAT) -> attention -> o_proj -> residual ADD -> RMSNorm ->
`LlamaMLP.forward` (:159-176: gate/up MUL_MAT, SILU, MUL, down MUL_MAT) -> residual ADD.
The reference only allocates these nodes; nothing executes them, so the numbers a
layer produces are defined by composing the op-level oracles in the same order (`oracle_layer`).

Attention core, two forms:
  * `attn=None` (BASELINE config 4 as first built): the context holds one position, softmax([s]) = 1 and the
    attention output is v itself; q and k are still computed and the graph wires v into o_proj.  13 nodes.
  * `attn=LayerAttention(...)`: `LlamaAttention.computeAttention` (model/LlamaAttention.kt:77-124) with a KV cache
    (model/KVCache.kt): ROPE(q), ROPE(k), CPY k -> key cache, CPY v -> value cache, FLASH_ATTN (per head
    softmax(q.K^T / sqrt(d)).V over positions 0..pos) -> o_proj.  18 nodes.  The position is a device I32 scalar
    shared by all layers, so one captured graph serves every token.

All nodes are written into tensors of one `GGMLGraphAllocator` buffer.  No arithmetic happens in this module.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass
from typing import Dict, List

import numpy as np

from .ggml_types import GGMLCGraph, GGMLOp, GGMLTensor, GGMLType

LLAMA2_7B = dict(hidden=4096, ffn=11008)      # model/LlamaModel.kt LlamaConfig defaults: hidden 4096, intermediate 11008
RMS_EPS = 1e-6                                 # LlamaConfig.rmsNormEps
WEIGHT_NAMES = ("wq", "wk", "wv", "wo", "w_gate", "w_up", "w_down")


def weight_shapes(hidden: int, ffn: int) -> Dict[str, tuple]:
    """name -> (M rows, K) of the matmul weight (ne = [K, M])."""
    return {"wq": (hidden, hidden), "wk": (hidden, hidden), "wv": (hidden, hidden), "wo": (hidden, hidden),
            "w_gate": (ffn, hidden), "w_up": (ffn, hidden), "w_down": (hidden, ffn)}


@dataclass
class DecoderLayer:
    """Device tensors of one layer (weights + the norm gains)."""
    weights: Dict[str, GGMLTensor]
    attn_norm: GGMLTensor
    ffn_norm: GGMLTensor


@dataclass
class LayerAttention:
    """Per-layer attention state: the KV cache tensors (ne = [headDim, numHeads, maxSeq]) plus what all layers share
    (position scalar, RoPE table)."""
    keys: GGMLTensor
    values: GGMLTensor
    position: GGMLTensor     # I32, element 0 = index of the token being decoded
    table: GGMLTensor        # F32 [2, headDim/2, maxSeq]: (cos, sin) per position, ops.ropeTable
    numHeads: int
    headDim: int


def alloc_attention(ga, numHeads: int, headDim: int, maxSeq: int, nlayers: int, ropeTheta: float = 10000.0):
    """KV caches for `nlayers` layers, one shared position scalar and RoPE table (uploaded).  Caches start zeroed
    (`lk_buf_clear` semantics are the allocator's; rows beyond the current position are never read)."""
    from .ggml_compute_ops import ropeTable
    pos = ga.allocateTensor(GGMLTensor(GGMLType.I32, [4]))
    table = ga.allocateTensor(GGMLTensor(GGMLType.F32, [2, headDim // 2, maxSeq]))
    ga.setBytes(table, ropeTable(headDim, ropeTheta, maxSeq))
    ga.setBytes(pos, np.zeros(4, dtype=np.int32))
    out = []
    for _ in range(nlayers):
        k = ga.allocateTensor(GGMLTensor(GGMLType.F32, [headDim, numHeads, maxSeq]))
        v = ga.allocateTensor(GGMLTensor(GGMLType.F32, [headDim, numHeads, maxSeq]))
        out.append(LayerAttention(k, v, pos, table, numHeads, headDim))
    return out


def _eps_bits(eps: float) -> int:
    return struct.unpack("<i", struct.pack("<f", eps))[0]


def alloc_layer(ga, qtype: GGMLType, hidden: int, ffn: int) -> DecoderLayer:
    ws = {n: ga.allocateTensor(GGMLTensor(qtype, [K, M])) for n, (M, K) in weight_shapes(hidden, ffn).items()}
    return DecoderLayer(ws, ga.allocateTensor(GGMLTensor(GGMLType.F32, [hidden])),
                        ga.allocateTensor(GGMLTensor(GGMLType.F32, [hidden])))


def append_layer_nodes(ga, layer: DecoderLayer, x: GGMLTensor, nodes: List[GGMLTensor], hidden: int, ffn: int,
                       eps: float = RMS_EPS, attn: "LayerAttention | None" = None) -> GGMLTensor:
    """Appends the 13 (or, with `attn`, 18) nodes of one layer to `nodes`; returns the layer's output tensor."""
    def node(op, ne, src, params=None):
        t = ga.allocateTensor(GGMLTensor(GGMLType.F32, ne, op=op, src=list(src) + [None] * (10 - len(src))))
        if params is not None:
            t.opParams[0] = params
        nodes.append(t)
        return t

    def col(t):     # the same bytes as the matmul's b operand: ne = [N = 1, K] (core/GGMLComputeOps.kt:1436-1440)
        return GGMLTensor(GGMLType.F32, [1, t.numElements()], bufferId=t.bufferId, dataOffset=t.dataOffset, buffer=t.buffer)

    def flat(t):    # ... and a matmul result [1, M] as the flat vector the elementwise ops and RMSNorm rows take
        return GGMLTensor(GGMLType.F32, [t.numElements()], bufferId=t.bufferId, dataOffset=t.dataOffset, buffer=t.buffer)

    w = layer.weights
    h1 = node(GGMLOp.RMS_NORM, [hidden], [x, layer.attn_norm], _eps_bits(eps))
    q = node(GGMLOp.MUL_MAT, [1, hidden], [w["wq"], col(h1)])   # noqa: F841  (degenerate context: unused)
    k = node(GGMLOp.MUL_MAT, [1, hidden], [w["wk"], col(h1)])   # noqa: F841  (what the KV cache would append)
    v = node(GGMLOp.MUL_MAT, [1, hidden], [w["wv"], col(h1)])
    if attn is None:
        o = node(GGMLOp.MUL_MAT, [1, hidden], [w["wo"], v])     # one-position context: attention output = v
    else:
        H, D = attn.numHeads, attn.headDim
        qr = node(GGMLOp.ROPE, [hidden], [flat(q), attn.position, attn.table], D)
        kr = node(GGMLOp.ROPE, [hidden], [flat(k), attn.position, attn.table], D)

        def append(new, cache):     # the node is a view of the cache it writes (KVCache.updateKey / updateValue)
            t = GGMLTensor(GGMLType.F32, list(cache.ne), op=GGMLOp.CPY, bufferId=cache.bufferId, dataOffset=cache.dataOffset,
                           buffer=cache.buffer,
                           src=[GGMLTensor(GGMLType.F32, [D, H, 1], bufferId=new.bufferId, dataOffset=new.dataOffset, buffer=new.buffer),
                                cache, attn.position] + [None] * 7)
            nodes.append(t)
            return t
        kc = append(kr, attn.keys)
        vc = append(v, attn.values)
        at = node(GGMLOp.FLASH_ATTN, [hidden], [qr, kc, vc, attn.position], _eps_bits(1.0 / float(np.sqrt(np.float32(D)))))
        o = node(GGMLOp.MUL_MAT, [1, hidden], [w["wo"], col(at)])
    x1 = node(GGMLOp.ADD, [hidden], [x, flat(o)])
    h2 = node(GGMLOp.RMS_NORM, [hidden], [x1, layer.ffn_norm], _eps_bits(eps))
    g = node(GGMLOp.MUL_MAT, [1, ffn], [w["w_gate"], col(h2)])
    u = node(GGMLOp.MUL_MAT, [1, ffn], [w["w_up"], col(h2)])
    a = node(GGMLOp.SILU, [ffn], [flat(g)])
    m = node(GGMLOp.MUL, [ffn], [a, flat(u)])
    d = node(GGMLOp.MUL_MAT, [1, hidden], [w["w_down"], col(m)])
    return node(GGMLOp.ADD, [hidden], [x1, flat(d)])


def build_decoder_graph(ga, layers: List[DecoderLayer], x: GGMLTensor, hidden: int, ffn: int, attn=None) -> tuple:
    """GGMLCGraph of the stacked layers (each feeding the next) and the final output tensor.  attn: None or one
    LayerAttention per layer (alloc_attention)."""
    nodes: List[GGMLTensor] = []
    cur = x
    for i, layer in enumerate(layers):
        cur = append_layer_nodes(ga, layer, cur, nodes, hidden, ffn, attn=attn[i] if attn is not None else None)
    return GGMLCGraph(nodes=nodes, allocator=ga), cur


def oracle_layer(O, otype: int, blocks: Dict[str, np.ndarray], attn_norm, ffn_norm, x, hidden: int, ffn: int,
                 eps: float = RMS_EPS, nthreads: int = 8, attn: dict | None = None) -> np.ndarray:
    """The same op sequence on the CPU oracle (test infrastructure: tests / bench checker only).
    attn = {"kc", "vc": float32 arrays [maxSeq * hidden] (updated in place), "pos", "H", "D", "theta"}."""
    h1 = O.rms_norm(x, attn_norm, eps, hidden)
    v = O.matmul(otype, blocks["wv"], hidden, hidden, h1, 1, nthreads=nthreads)
    if attn is not None:
        q = O.matmul(otype, blocks["wq"], hidden, hidden, h1, 1, nthreads=nthreads)
        k = O.matmul(otype, blocks["wk"], hidden, hidden, h1, 1, nthreads=nthreads)
        pos, H, D = attn["pos"], attn["H"], attn["D"]
        qr, kr = O.rope(q, D, H, pos, attn.get("theta", 10000.0)), O.rope(k, D, H, pos, attn.get("theta", 10000.0))
        attn["kc"][pos * hidden:(pos + 1) * hidden] = kr
        attn["vc"][pos * hidden:(pos + 1) * hidden] = v
        v = O.attention_decode(qr, attn["kc"], attn["vc"], pos + 1, D, H)
    o = O.matmul(otype, blocks["wo"], hidden, hidden, v, 1, nthreads=nthreads)
    x1 = O.add(x, o)
    h2 = O.rms_norm(x1, ffn_norm, eps, hidden)
    g = O.matmul(otype, blocks["w_gate"], hidden, ffn, h2, 1, nthreads=nthreads)
    u = O.matmul(otype, blocks["w_up"], hidden, ffn, h2, 1, nthreads=nthreads)
    m = O.mul(O.silu(g), u)
    d = O.matmul(otype, blocks["w_down"], ffn, hidden, m, 1, nthreads=nthreads)
    return O.add(x1, d)
