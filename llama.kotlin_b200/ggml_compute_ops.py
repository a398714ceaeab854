"""Destination-based compute ops with the reference's names and argument meaning
(core/GGMLComputeOps.kt): computeMatMul :1435, computeAdd :735, computeMul :811,
dequantizeTensor :918, plus the RMS_NORM / SILU / SOFT_MAX ops the reference leaves
unimplemented (semantics from its tests).  Every op is a C-ABI call into
libllamakotlin_cuda.so — there is no host arithmetic and no fallback in this file."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _binding as B
from .ggml_types import GGMLContext, GGMLTensor, GGMLType

_default_ctx = None


class LkContext:
    """Owns one lk_ctx (device + stream)."""

    def __init__(self, device: int = 0):
        h = C.c_void_p()
        B.check(B.lib().lk_ctx_create(device, C.byref(h)))
        self.handle = h
        self.device = device

    def sync(self) -> None:
        B.check(B.lib().lk_sync(self.handle), self.handle)

    def set_option(self, name: str, value: int) -> None:
        B.check(B.lib().lk_set_option(self.handle, name.encode(), int(value)), self.handle)

    def get_option(self, name: str) -> int:
        return int(B.lib().lk_get_option(self.handle, name.encode()))

    def launch_count(self) -> int:
        return int(B.lib().lk_ctx_launch_count(self.handle))

    def sm_count(self) -> int:
        return int(B.lib().lk_ctx_sm_count(self.handle))

    def timer_start(self) -> None:
        B.check(B.lib().lk_timer_start(self.handle), self.handle)

    def timer_stop(self) -> float:
        ms = C.c_float()
        B.check(B.lib().lk_timer_stop(self.handle, C.byref(ms)), self.handle)
        return float(ms.value)

    def close(self) -> None:
        if self.handle:
            B.lib().lk_ctx_destroy(self.handle)
            self.handle = None


def default_context(device: int = 0) -> LkContext:
    global _default_ctx
    if _default_ctx is None or _default_ctx.handle is None:
        _default_ctx = LkContext(device)
    return _default_ctx


def _ctx_of(graphAllocator) -> LkContext:
    be = getattr(graphAllocator, "backend", None)
    if be is not None and getattr(be, "lk", None) is not None:
        return be.lk
    return default_context()


def describe(graphAllocator, t: GGMLTensor) -> B.lk_tensor:
    """GGMLTensor -> lk_tensor: (buffer, dataOffset) pair, never tensor.data."""
    d = B.lk_tensor()
    d.type = int(t.type)
    for i in range(4):
        d.ne[i] = int(t.ne[i])
        d.nb[i] = 0 if t.type.isBlockType else int(t.nb[i])
    d.data_offset = int(t.dataOffset)
    if t.buffer is not None:                      # device-resident (GGMLCudaBuffer)
        d.dev_buf = t.buffer.handle
        d.buffer_size = t.buffer.getSize()
        return d
    bufs = graphAllocator.buffers
    if t.bufferId < 0 or t.bufferId >= len(bufs) or bufs[t.bufferId] is None:
        # core/GGMLTypes.kt:552-553
        raise B.IllegalStateException(f"Tensor buffer not found for bufferId {t.bufferId} for tensor '{t.name}'.")
    arr = bufs[t.bufferId]
    d.host_base = arr.ctypes.data
    d.buffer_size = arr.size
    return d


def computeMatMul(graphAllocator, context: GGMLContext, a: GGMLTensor, b: GGMLTensor, dst: GGMLTensor) -> None:
    lk = _ctx_of(graphAllocator)
    B.check(B.lib().lk_compute_matmul(lk.handle, describe(graphAllocator, a), describe(graphAllocator, b),
                                      describe(graphAllocator, dst)), lk.handle)


def _fusion_struct(graphAllocator, norm_x=None, norm_w=None, eps: float = 0.0, residual=None, sum_dst=None,
                   store_b: bool = True, extra=(), silu_dst=None, mul_dst=None):
    """lk_matmul_fusion for one launch; returns (struct, keep-alive list of the tensor descriptors it points to)."""
    keep = [describe(graphAllocator, t) if t is not None else None for t in (norm_x, norm_w, residual, sum_dst)]
    f = B.lk_matmul_fusion()
    f.norm_x = C.pointer(keep[0]) if keep[0] is not None else None
    f.norm_w = C.pointer(keep[1]) if keep[1] is not None else None
    f.eps = float(eps)
    f.residual = C.pointer(keep[2]) if keep[2] is not None else None
    f.sum_dst = C.pointer(keep[3]) if keep[3] is not None else None
    f.flags = 0 if store_b else 1        # LK_FUSE_SKIP_B_STORE
    # stacked matmuls on the same activations: extra = [(a1, dst1), ...]; SwiGLU epilogue with silu_dst / mul_dst
    f.n_extra = len(extra)
    for i, (ae, de) in enumerate(extra):
        keep.append(describe(graphAllocator, ae))
        f.a_extra[i] = C.pointer(keep[-1])
        keep.append(describe(graphAllocator, de))
        f.dst_extra[i] = C.pointer(keep[-1])
    for name, t in (("silu_dst", silu_dst), ("mul_dst", mul_dst)):
        if t is not None:
            keep.append(describe(graphAllocator, t))
            setattr(f, name, C.pointer(keep[-1]))
    return f, keep


def computeMatMulFused(graphAllocator, context, a, b, dst, **fusion) -> bool:
    """computeMatMul with the RMS_NORM that produces `b` and / or the ADD that consumes `dst` folded into the GEMV
    launch (lk_compute_matmul_fused); keywords: norm_x, norm_w, eps, residual, sum_dst, store_b, extra, silu_dst,
    mul_dst.  Returns False when the library declines the shape (LK_ENOTIMPL): the caller then issues the unfused
    ops."""
    f, _keep = _fusion_struct(graphAllocator, **fusion)
    ctx = _ctx_of(graphAllocator)
    st = B.lib().lk_compute_matmul_fused(ctx.handle, describe(graphAllocator, a), describe(graphAllocator, b),
                                         describe(graphAllocator, dst), C.byref(f))
    if st == B.LK_ENOTIMPL:
        return False
    B.check(st, ctx.handle)
    return True


def computeMatMulChain(graphAllocator, context, links) -> bool:
    """A run of dependent decode matmuls as ONE persistent launch (lk_compute_matmul_chain): `links` is a list of
    (a, b, dst, fusion keywords or None), each exactly one computeMatMulFused / computeMatMul (N == 1) call; they run in
    order.  Returns False when the library declines (LK_ENOTIMPL): the caller then issues the links one by one."""
    arr = (B.lk_matmul_link * len(links))()
    keep = []
    for i, (a, b, dst, fusion) in enumerate(links):
        d = [describe(graphAllocator, a), describe(graphAllocator, b), describe(graphAllocator, dst)]
        keep.append(d)
        arr[i].a, arr[i].b, arr[i].dst = C.pointer(d[0]), C.pointer(d[1]), C.pointer(d[2])
        if fusion:
            f, k = _fusion_struct(graphAllocator, **fusion)
            keep.append((f, k))
            arr[i].fusion = C.pointer(f)
    ctx = _ctx_of(graphAllocator)
    st = B.lib().lk_compute_matmul_chain(ctx.handle, arr, len(links))
    if st == B.LK_ENOTIMPL:
        return False
    B.check(st, ctx.handle)
    return True


def computeAdd(graphAllocator, context, a, b, dst) -> None:
    lk = _ctx_of(graphAllocator)
    B.check(B.lib().lk_compute_add(lk.handle, describe(graphAllocator, a), describe(graphAllocator, b),
                                   describe(graphAllocator, dst)), lk.handle)


def computeMul(graphAllocator, context, a, b, dst) -> None:
    lk = _ctx_of(graphAllocator)
    B.check(B.lib().lk_compute_mul(lk.handle, describe(graphAllocator, a), describe(graphAllocator, b),
                                   describe(graphAllocator, dst)), lk.handle)


def computeSilu(graphAllocator, context, a, dst) -> None:
    lk = _ctx_of(graphAllocator)
    B.check(B.lib().lk_compute_silu(lk.handle, describe(graphAllocator, a), describe(graphAllocator, dst)), lk.handle)


def computeSiluMul(graphAllocator, context, gate, up, dst) -> None:
    lk = _ctx_of(graphAllocator)
    B.check(B.lib().lk_compute_silu_mul(lk.handle, describe(graphAllocator, gate), describe(graphAllocator, up),
                                        describe(graphAllocator, dst)), lk.handle)


def computeRmsNorm(graphAllocator, context, a, weight, eps: float, dst) -> None:
    lk = _ctx_of(graphAllocator)
    w = describe(graphAllocator, weight) if weight is not None else None
    B.check(B.lib().lk_compute_rms_norm(lk.handle, describe(graphAllocator, a), w, float(eps),
                                        describe(graphAllocator, dst)), lk.handle)


def computeSoftMax(graphAllocator, context, a, dst) -> None:
    lk = _ctx_of(graphAllocator)
    B.check(B.lib().lk_compute_softmax(lk.handle, describe(graphAllocator, a), describe(graphAllocator, dst)), lk.handle)


# ---- attention core: LlamaAttention / KVCache, model/LlamaAttention.kt, model/KVCache.kt -------------
def computeScale(graphAllocator, context, a, factor: float, dst) -> None:
    """LlamaAttention.scale (model/LlamaAttention.kt:166-195): value * factor."""
    lk = _ctx_of(graphAllocator)
    B.check(B.lib().lk_compute_scale(lk.handle, describe(graphAllocator, a), float(factor), describe(graphAllocator, dst)), lk.handle)


def ropeTable(headDim: int, ropeTheta: float, maxPositions: int) -> np.ndarray:
    """(cos, sin) of applyRoPE (model/LlamaAttention.kt:53-58) for positions [0, maxPositions): F32, ne = [2, headDim/2,
    maxPositions], filled by the library in the reference's F64 operations."""
    out = np.empty(maxPositions * (headDim // 2) * 2, dtype=np.float32)
    st = B.lib().lk_rope_table(int(headDim), float(ropeTheta), int(maxPositions), out.ctypes.data)
    if st != B.LK_OK:
        raise B.IllegalArgumentException(f"lk_rope_table({headDim}, {ropeTheta}, {maxPositions}) -> {st}")
    return out


def computeRope(graphAllocator, context, a, table, position, headDim: int, dst) -> None:
    """LlamaAttention.applyRoPE (model/LlamaAttention.kt:29-72); `position` is an I32 tensor (or None = 0)."""
    lk = _ctx_of(graphAllocator)
    pos = describe(graphAllocator, position) if position is not None else None
    B.check(B.lib().lk_compute_rope(lk.handle, describe(graphAllocator, a), describe(graphAllocator, table), pos,
                                    int(headDim), describe(graphAllocator, dst)), lk.handle)


def computeKvAppend(graphAllocator, context, src, cache, position) -> None:
    """The copy loops of KVCache.updateKey / updateValue (model/KVCache.kt:47-55, :88-96)."""
    lk = _ctx_of(graphAllocator)
    pos = describe(graphAllocator, position) if position is not None else None
    B.check(B.lib().lk_compute_kv_append(lk.handle, describe(graphAllocator, src), describe(graphAllocator, cache), pos), lk.handle)


def computeAttentionDecode(graphAllocator, context, q, kcache, vcache, position, scale: float, dst) -> None:
    """LlamaAttention.computeAttention for one query position (model/LlamaAttention.kt:104-123), heads per the KVCache
    layout [headDim, numHeads, maxSeq]; length = position + 1."""
    lk = _ctx_of(graphAllocator)
    B.check(B.lib().lk_compute_attention_decode(lk.handle, describe(graphAllocator, q), describe(graphAllocator, kcache),
                                                describe(graphAllocator, vcache), describe(graphAllocator, position),
                                                float(scale), describe(graphAllocator, dst)), lk.handle)


def computeAttentionStep(graphAllocator, context, q, k, v, table, position, headDim: int, kcache, vcache, scale: float,
                         dst, qRopeDst=None, kRopeDst=None) -> None:
    """One decode step of LlamaAttention.forward's core in one launch (model/LlamaAttention.kt:87-99, :104-123):
    applyRoPE(q), applyRoPE(k), KVCache.updateKey / updateValue at `position`, computeAttention.  Bit-identical to
    computeRope x 2 + computeKvAppend x 2 + computeAttentionDecode."""
    lk = _ctx_of(graphAllocator)
    d = lambda t: describe(graphAllocator, t) if t is not None else None   # noqa: E731
    B.check(B.lib().lk_compute_attention_step(lk.handle, d(q), d(k), d(v), d(table), d(position), int(headDim), d(kcache),
                                              d(vcache), float(scale), d(qRopeDst), d(kRopeDst), d(dst)), lk.handle)


class KVCache:
    """model/KVCache.kt:9-115 on device memory: keys / values ne = [headDim, numHeads, maxSequenceLength, 1] F32 in the
    allocator's device buffer, `currentLength` kept by the host (the `require`s of :44 and :85 live here), the
    position handed to the kernels through a one-element I32 device tensor so that a captured layer graph can be
    replayed for every token.  Unlike the reference, update* return nothing: the cache itself is what the attention
    kernel reads (the reference re-copies the whole cache into a fresh tensor per token, :57-71)."""

    def __init__(self, maxSequenceLength: int, numHeads: int, headDim: int):
        self.maxSequenceLength, self.numHeads, self.headDim = maxSequenceLength, numHeads, headDim
        self.keys = self.values = self.position = None
        self.currentLength = 0

    def initialize(self, graphAllocator) -> None:
        ne = [self.headDim, self.numHeads, self.maxSequenceLength, 1]
        self.keys = graphAllocator.allocateTensor(GGMLTensor(GGMLType.F32, ne))
        self.values = graphAllocator.allocateTensor(GGMLTensor(GGMLType.F32, ne))
        self.position = graphAllocator.allocateTensor(GGMLTensor(GGMLType.I32, [4]))   # 16 bytes; element 0 is read
        self.currentLength = 0
        self._ga = graphAllocator
        self.setPosition(0)

    def setPosition(self, pos: int) -> None:
        self._ga.setBytes(self.position, np.array([pos, 0, 0, 0], dtype=np.int32))

    def updateKey(self, graphAllocator, newKey: GGMLTensor) -> None:
        if self.keys is None:
            raise B.IllegalStateException("KVCache not initialized")
        n = max(int(newKey.ne[2]), 1)
        if self.currentLength + n > self.maxSequenceLength:
            raise B.IllegalArgumentException("Sequence length would exceed maximum")
        self.setPosition(self.currentLength)
        computeKvAppend(graphAllocator, None, newKey, self.keys, self.position)
        self.currentLength += n

    def updateValue(self, graphAllocator, newValue: GGMLTensor) -> None:
        if self.values is None:
            raise B.IllegalStateException("KVCache not initialized")
        n = max(int(newValue.ne[2]), 1)
        if self.currentLength < n:
            raise B.IllegalArgumentException("updateKey must be called before updateValue in a step")
        self.setPosition(self.currentLength - n)
        computeKvAppend(graphAllocator, None, newValue, self.values, self.position)

    def clear(self) -> None:
        self.currentLength = 0

    def getCurrentLength(self) -> int:
        return self.currentLength


def dequantizeTensor(graphAllocator, tensor: GGMLTensor) -> GGMLTensor:
    """core/GGMLComputeOps.kt:918 — returns an F32 tensor whose `data` holds the floats
    (like the reference, the result lives in result.data, not in the allocator: SURVEY F8)."""
    lk = _ctx_of(graphAllocator)
    n = tensor.numElements()
    out = np.empty(n, dtype=np.float32)
    res = GGMLTensor(type=GGMLType.F32, ne=list(tensor.ne))
    d = B.lk_tensor()
    d.type = int(GGMLType.F32)
    for i in range(4):
        d.ne[i] = int(tensor.ne[i])
    d.host_base = out.ctypes.data
    d.buffer_size = out.nbytes
    B.check(B.lib().lk_compute_dequantize(lk.handle, describe(graphAllocator, tensor), d), lk.handle)
    res.data = out
    return res


def computeQuantize(graphAllocator, context, src: GGMLTensor, dst: GGMLTensor) -> None:
    """Destination-based quantizeTensor (lk_compute_quantize): src F32 -> dst Q8_0 / Q4_0 / Q4_1 in the allocator's buffer,
    bit-exact with core/GGMLComputeOps.kt:1058-1087 / :1162-1203."""
    lk = _ctx_of(graphAllocator)
    B.check(B.lib().lk_compute_quantize(lk.handle, describe(graphAllocator, src), describe(graphAllocator, dst)), lk.handle)


def quantizeTensor(graphAllocator, tensor: GGMLTensor, targetType: GGMLType) -> GGMLTensor:
    """core/GGMLComputeOps.kt:1040 — returns a tensor of `targetType` whose `data` holds the packed blocks (like the
    reference, the result lives in result.data with bufferId = -1, SURVEY F8: copy it into an allocator buffer before
    handing it to computeMatMul)."""
    lk = _ctx_of(graphAllocator)
    targetType = GGMLType(targetType)
    n = tensor.numElements()
    out = np.empty(n // targetType.blockElements * targetType.byteSize, dtype=np.uint8)
    res = GGMLTensor(type=targetType, ne=list(tensor.ne))
    d = B.lk_tensor()
    d.type = int(targetType)
    for i in range(4):
        d.ne[i] = int(tensor.ne[i])
    d.host_base = out.ctypes.data
    d.buffer_size = max(out.nbytes, 1)
    B.check(B.lib().lk_compute_quantize(lk.handle, describe(graphAllocator, tensor), d), lk.handle)
    res.data = out
    return res
