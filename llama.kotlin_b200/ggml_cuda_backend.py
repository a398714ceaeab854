"""GGMLCudaBackend — the graph-level plug-in seam of the reference
(interface GGMLBackend core/GGMLBackend.kt:90-157, GGMLBackendBuffer :45-75,
GGMLBackendBufferType :24-39) implemented over the C ABI.  Tensors placed in a
GGMLCudaBuffer stay in HBM; graphCompute walks graph.nodes like
GGMLComputeOps.computeGraph (core/GGMLComputeOps.kt:2515-2550) and, for a graph it
has seen before, replays a captured op sequence (one CUDA-graph launch)."""
from __future__ import annotations

import ctypes as C
import struct

import numpy as np

from . import _binding as B
from . import ggml_compute_ops as ops
from .ggml_types import GGMLCGraph, GGMLOp, GGMLStatus, GGMLTensor


class GGMLCudaBufferType:
    def __init__(self, backend: "GGMLCudaBackend"):
        self.backend = backend

    def getName(self) -> str:
        return "CUDA"

    def allocBuffer(self, size: int) -> "GGMLCudaBuffer":
        return GGMLCudaBuffer(self.backend, size)

    def getAlignment(self) -> int:
        return 256

    def getMaxSize(self) -> int:
        return 180 * (1 << 30)

    def isHost(self) -> bool:
        return False


class GGMLCudaBuffer:
    """GGMLBackendBuffer on device memory (lk_buf)."""

    def __init__(self, backend: "GGMLCudaBackend", size: int, handle=None):
        self.backend = backend
        self.lk = backend.lk
        if handle is None:
            h = C.c_void_p()
            B.check(B.lib().lk_buf_alloc(self.lk.handle, int(size), C.byref(h)), self.lk.handle)
            handle = h
        self.handle = handle
        self._size = int(size)

    def getType(self) -> GGMLCudaBufferType:
        return self.backend.getDefaultBufferType()

    def getName(self) -> str:
        return "CUDA"

    def getBase(self):
        return B.lib().lk_buf_device_ptr(self.handle)

    def getSize(self) -> int:
        return self._size

    def initTensor(self, tensor: GGMLTensor) -> None:
        tensor.buffer = self

    def setTensor(self, tensor: GGMLTensor, data, offset: int, size: int) -> None:
        raw = np.ascontiguousarray(data).view(np.uint8).reshape(-1)
        B.check(B.lib().lk_buf_upload(self.handle, int(tensor.dataOffset) + offset, raw.ctypes.data, int(size)),
                self.lk.handle)

    def getTensor(self, tensor: GGMLTensor, data: np.ndarray, offset: int, size: int) -> None:
        B.check(B.lib().lk_buf_download(self.handle, int(tensor.dataOffset) + offset, data.ctypes.data, int(size)),
                self.lk.handle)

    def copyTensor(self, src: GGMLTensor, dst: GGMLTensor) -> bool:
        if not isinstance(src.buffer, GGMLCudaBuffer):
            return False
        B.check(B.lib().lk_buf_copy(self.handle, int(dst.dataOffset), src.buffer.handle, int(src.dataOffset),
                                    src.byteSize()), self.lk.handle)
        return True

    def copyFrom(self, other: "GGMLCudaBuffer", n: int) -> None:
        B.check(B.lib().lk_buf_copy(self.handle, 0, other.handle, 0, int(n)), self.lk.handle)

    def clear(self, value: int) -> None:
        B.check(B.lib().lk_buf_clear(self.handle, value & 0xFF), self.lk.handle)

    def free(self) -> None:
        if self.handle:
            B.lib().lk_buf_free(self.handle)
            self.handle = None


class GGMLCudaBackend:
    """GGMLBackend 'CUDA' (registry name as in GGMLBackendRegistry.register, core/GGMLBackend.kt:162-181)."""

    _SUPPORTED = {GGMLOp.MUL_MAT, GGMLOp.ADD, GGMLOp.MUL, GGMLOp.RMS_NORM, GGMLOp.SILU, GGMLOp.SOFT_MAX,
                  GGMLOp.NONE}

    def __init__(self, device: int = 0):
        self.lk = ops.LkContext(device)
        self._bt = GGMLCudaBufferType(self)
        self._captured = {}      # id(graph) -> (lk_graph handle, node signature)
        self.useCudaGraphs = True

    # -- GGMLBackend ---------------------------------------------------------------
    def getGuid(self) -> str:
        return "llamakotlin-cuda-sm100a"

    def getName(self) -> str:
        return "CUDA"

    def getDefaultBufferType(self) -> GGMLCudaBufferType:
        return self._bt

    def allocBuffer(self, size: int) -> GGMLCudaBuffer:
        return self._bt.allocBuffer(size)

    def supportsOp(self, tensor: GGMLTensor) -> bool:
        return tensor.op in self._SUPPORTED

    def supportsBufferType(self, bufferType) -> bool:
        return isinstance(bufferType, GGMLCudaBufferType)

    def offloadOp(self, tensor: GGMLTensor) -> bool:
        return tensor.op == GGMLOp.MUL_MAT

    def synchronize(self) -> None:
        self.lk.sync()

    def releaseGraphs(self) -> None:
        """Destroys the captured CUDA graphs (a graph holding NCCL kernels must go before its communicator)."""
        self.lk.sync()
        for h, _ in self._captured.values():
            B.lib().lk_graph_destroy(h)
        self._captured.clear()

    def free(self) -> None:
        self.releaseGraphs()
        self.lk.close()

    # -- graph execution -------------------------------------------------------------
    def _computeNode(self, ga, node: GGMLTensor) -> None:
        op = node.op
        ex = node.extra if isinstance(node.extra, dict) else None
        if ex is not None and "shard_barrier" in ex:      # exchange point of a row-sharded matmul (sharded.py)
            ex["shard_barrier"].barrier()
            return
        if ex is not None and "shard_allgather" in ex:    # baseline exchange: in-place ncclAllGather of per-rank slices
            grp, count = ex["shard_allgather"]
            grp.allgather(node.buffer, node.dataOffset, count)
            return
        if ex is not None and "shard_push" in ex:         # the same exchange as peer stores (follow with shard_barrier)
            grp, count, peers = ex["shard_push"]
            grp.push(node.buffer, node.dataOffset, count, peers)
            return
        if op == GGMLOp.NONE:
            return
        if op == GGMLOp.MUL_MAT and ex is not None and "shard" in ex:
            grp, row0, m_total, peers = ex["shard"]
            grp.computeMatMulSharded(ga, node.src[0], node.src[1], node, row0, m_total, peers)
        elif op == GGMLOp.MUL_MAT:       # computeMulMat, core/GGMLComputeOps.kt:2647-2652
            ops.computeMatMul(ga, None, node.src[0], node.src[1], node)
        elif op == GGMLOp.ADD:
            ops.computeAdd(ga, None, node.src[0], node.src[1], node)
        elif op == GGMLOp.MUL:
            ops.computeMul(ga, None, node.src[0], node.src[1], node)
        elif op == GGMLOp.SILU:
            ops.computeSilu(ga, None, node.src[0], node)
        elif op == GGMLOp.RMS_NORM:    # rmsNormOp: src[1] = weight, opParams[0] = eps bits
            eps = struct.unpack("<f", struct.pack("<i", int(node.opParams[0])))[0]
            ops.computeRmsNorm(ga, None, node.src[0], node.src[1], eps, node)
        elif op == GGMLOp.SOFT_MAX:
            ops.computeSoftMax(ga, None, node.src[0], node)
        else:                          # core/GGMLComputeOps.kt:2548
            raise NotImplementedError(f"Operation {op} not implemented")

    @staticmethod
    def _signature(graph: GGMLCGraph):
        sig = []
        for n in graph.nodes:
            sig.append((n.op, int(n.type), tuple(n.ne), n.dataOffset, id(n.buffer), id(n.extra),
                        tuple((id(s.buffer), s.dataOffset, int(s.type), tuple(s.ne)) for s in n.src if s is not None)))
        return tuple(sig)

    def _allDevice(self, graph: GGMLCGraph) -> bool:
        for n in graph.nodes:
            if n.op == GGMLOp.NONE:
                continue
            for t in [n] + [s for s in n.src if s is not None]:
                if not isinstance(t.buffer, GGMLCudaBuffer):
                    return False
        return True

    def graphCompute(self, graph: GGMLCGraph) -> GGMLStatus:
        """core/GGMLBackend.kt:145; exceptions become FAILED like GGMLCpuBackend.kt:168-175."""
        ga = graph.allocator
        try:
            if self.useCudaGraphs and self._allDevice(graph):
                sig = self._signature(graph)
                ent = self._captured.get(id(graph))
                if ent is None or ent[1] != sig:
                    if ent is not None:
                        B.lib().lk_graph_destroy(ent[0])
                    B.check(B.lib().lk_capture_begin(self.lk.handle), self.lk.handle)
                    try:
                        for node in graph.nodes:
                            self._computeNode(ga, node)
                    finally:
                        h = C.c_void_p()
                        st = B.lib().lk_capture_end(self.lk.handle, C.byref(h))
                    B.check(st, self.lk.handle)
                    ent = (h, sig)
                    self._captured[id(graph)] = ent
                B.check(B.lib().lk_graph_launch(self.lk.handle, ent[0]), self.lk.handle)
            else:
                for node in graph.nodes:
                    self._computeNode(ga, node)
            return GGMLStatus.SUCCESS
        except (B.IllegalArgumentException, B.IndexOutOfBoundsException, B.IllegalStateException,
                NotImplementedError) as e:
            print(f"CUDA backend graph computation failed: {e}")
            return GGMLStatus.FAILED
