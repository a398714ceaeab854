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
from .ggml_compute_ops import describe as _describe
from .ggml_types import GGMLCGraph, GGMLOp, GGMLStatus, GGMLTensor, GGMLType


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

    def setTensorAsync(self, tensor: GGMLTensor, data: np.ndarray, offset: int, size: int) -> None:
        """setTensor without the wait (lk_buf_upload_async): the copy runs beside the compute stream, every op enqueued
        after this call waits for it; `data` (pinned for the overlap to exist) must stay untouched until the backend's
        next synchronize().  It does not wait for ops already enqueued (see the header)."""
        B.check(B.lib().lk_buf_upload_async(self.handle, int(tensor.dataOffset) + offset, data.ctypes.data, int(size)),
                self.lk.handle)

    def getTensorAsync(self, tensor: GGMLTensor, data: np.ndarray, offset: int, size: int) -> None:
        """getTensor without the wait (lk_buf_download_async): starts when everything enqueued so far has finished;
        `data` is valid after the backend's next synchronize()."""
        B.check(B.lib().lk_buf_download_async(self.handle, int(tensor.dataOffset) + offset, data.ctypes.data, int(size)),
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
                  GGMLOp.SCALE, GGMLOp.ROPE, GGMLOp.CPY, GGMLOp.FLASH_ATTN, GGMLOp.NONE}

    def __init__(self, device: int = 0):
        self.lk = ops.LkContext(device)
        self.fuseOps = True          # fold RMS_NORM / ADD neighbours into the decode GEMV inside captured sequences
        # runs of DEPENDENT decode GEMVs as one persistent launch (lk_compute_matmul_chain): True / False, or None = where
        # it was measured to pay: graphs with an attention step between the runs (+19 % tok/s at 2048 cached positions;
        # with a one-position context four separate launches are 2 % faster than K1c, profiles/r01t_gemv_chain_ab.txt) and
        # graphs whose GEMVs the MMA chain K1m takes (Q4_0, K % 256 == 0: 8 % faster than separate launches at ctx 0)
        self.chainOps = None
        self.maxChain = 8            # links per launch (the library's limit)
        self._bt = GGMLCudaBufferType(self)
        self._captured = {}      # id(graph) -> (lk_graph handle, node signature, fast key)
        self._ioKeys = {}        # id(hostIO list) -> (the list, its key): see graphCompute
        self._graphRefs = {}     # id(graph) -> graph
        self.useCudaGraphs = True
        # graphCompute is stream-ordered (asynchronous) like every device-resident op.  A device-side fault of the step —
        # the chained GEMV's grid barrier or a cross-GPU flag barrier giving up after its bounded wait — is reported by
        # the next call that synchronises (synchronize(), getTensor).  With checkFaults = True graphCompute synchronises
        # itself and returns GGMLStatus.FAILED for the step that faulted (tests, benches, debugging).
        self.checkFaults = False
        # graphs without backend-private nodes (sharded matmuls, exchanges) are planned and issued by the library itself
        # (lk_graph_compute: the same algorithm as _plan / _runNodes below, behind the C ABI so that the Kotlin binding gets
        # the fused and chained launches without a port); False keeps the Python planner for every graph
        self.useLibraryPlanner = True

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
        for ent in self._captured.values():
            B.lib().lk_graph_destroy(ent[0])
        self._captured.clear()
        self._graphRefs.clear()

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
        if ex is not None and "shard_swiglu" in ex:       # sharded gate/up + SwiGLU (+ RMSNorm prologue) as one launch
            d = ex["shard_swiglu"]
            d["group"].computeSwigluSharded(ga, d["gate"], d["up"], d["x"], d["g_full"], d["u_full"], node, d["row0"],
                                            d["M_total"], d.get("peers"), silu_full=d.get("silu_full"),
                                            norm_w=d.get("norm_w"), eps=d.get("eps", 0.0), normed_dst=d.get("normed_dst"),
                                            sync=d.get("sync"))
            return
        if op == GGMLOp.NONE:
            return
        if op == GGMLOp.MUL_MAT and ex is not None and "shard" in ex:
            grp, row0, m_total, peers = ex["shard"][:4]
            grp.computeMatMulSharded(ga, node.src[0], node.src[1], node, row0, m_total, peers,
                                     sync=ex["shard"][4] if len(ex["shard"]) > 4 else None)
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
        elif op == GGMLOp.SCALE:       # scaleOp: opParams[0] = factor bits (model/GGMLIntegration.kt:118-139)
            ops.computeScale(ga, None, node.src[0], struct.unpack("<f", struct.pack("<i", int(node.opParams[0])))[0], node)
        elif op == GGMLOp.ROPE:        # src = x, position (I32), (cos, sin) table; opParams[0] = headDim
            ops.computeRope(ga, None, node.src[0], node.src[2], node.src[1], int(node.opParams[0]), node)
        elif op == GGMLOp.CPY:         # KVCache append: src = new rows, cache, position; the node is a view of the cache
            ops.computeKvAppend(ga, None, node.src[0], node.src[1], node.src[2])
        elif op == GGMLOp.FLASH_ATTN:  # src = q, key cache, value cache, position; opParams[0] = scale bits
            ops.computeAttentionDecode(ga, None, node.src[0], node.src[1], node.src[2], node.src[3],
                                       struct.unpack("<f", struct.pack("<i", int(node.opParams[0])))[0], node)
        else:                          # core/GGMLComputeOps.kt:2548
            raise NotImplementedError(f"Operation {op} not implemented")

    # -- fusion of a matmul with its graph neighbours -----------------------------------
    @staticmethod
    def _key(t: GGMLTensor):
        return (id(t.buffer), int(t.dataOffset), int(t.numElements()))

    @staticmethod
    def _span(t: GGMLTensor):
        return (id(t.buffer), int(t.dataOffset), int(t.dataOffset) + int(t.byteSize()))

    def _plan(self, graph: GGMLCGraph):
        """Which RMS_NORM nodes fold into the matmuls that consume them and which ADD nodes fold into the matmul
        that feeds them.  Returns (rms_of, add_of, skip): matmul index -> RMS_NORM index / ADD index, and the set
        of node indices that are not launched on their own.  Every fused-away node's dst is still written by the
        fused launch, so the graph's intermediates stay valid whatever reads them later."""
        nodes = graph.nodes
        n = len(nodes)

        def is_gemv(t):
            return (t.op == GGMLOp.MUL_MAT and not isinstance(t.extra, dict) and t.src[0] is not None and
                    t.src[0].type in (GGMLType.Q4_0, GGMLType.Q4_1, GGMLType.Q8_0) and t.src[1] is not None and
                    int(t.src[1].ne[0]) == 1 and
                    t.type == GGMLType.F32)

        def overlaps(a, b):
            return a[0] == b[0] and a[1] < b[2] and b[1] < a[2]

        consumers = {}
        for j, t in enumerate(nodes):
            for sidx, sr in enumerate(t.src):
                if sr is not None:
                    consumers.setdefault(self._key(sr), []).append((j, sidx))
        rms_of, add_of, skip = {}, {}, set()
        for i, t in enumerate(nodes):
            if t.op == GGMLOp.RMS_NORM and t.type == GGMLType.F32 and int(t.numElements()) == int(t.ne[0]):
                cons = [c for c in consumers.get(self._key(t), []) if c[0] > i]
                # a later writer of the same storage before the last consumer would change what they read
                if cons and all(is_gemv(nodes[j]) and sidx == 1 for j, sidx in cons):
                    last = max(j for j, _ in cons)
                    clean = all(not overlaps(self._span(nodes[k]), self._span(t)) and
                                not overlaps(self._span(nodes[k]), self._span(t.src[0])) and
                                (t.src[1] is None or not overlaps(self._span(nodes[k]), self._span(t.src[1])))
                                for k in range(i + 1, last + 1) if nodes[k].op != GGMLOp.NONE)
                    if clean:
                        for j, _ in cons:
                            rms_of[j] = i
                        skip.add(i)
        for i, t in enumerate(nodes):
            if not is_gemv(t):
                continue
            cons = [c for c in consumers.get(self._key(t), []) if c[0] > i]
            if len(cons) != 1:
                continue
            j, sidx = cons[0]
            a = nodes[j]
            if a.op != GGMLOp.ADD or a.type != GGMLType.F32 or j in skip:
                continue
            res = a.src[1 - sidx]
            if res is None or int(res.numElements()) != int(t.numElements()):
                continue
            # the ADD moves up to the matmul's position: nothing in between may write its operand or touch its dst
            if any(overlaps(self._span(nodes[k]), self._span(res)) or overlaps(self._span(nodes[k]), self._span(a))
                   for k in range(i, j) if nodes[k].op != GGMLOp.NONE):
                continue
            if any(s2 is not None and overlaps(self._span(s2), self._span(a)) for k in range(i, j) for s2 in nodes[k].src):
                continue
            add_of[i] = j
            skip.add(j)
        # stacked matmuls: consecutive GEMVs (nothing but fused-away nodes between them) with the same activations
        # and weights of one type and shape run as one launch at the first one's position; a gate/up pair whose
        # only readers are SILU(gate) and MUL(silu, up) also takes those two nodes into its epilogue
        group_of, pair_of = {}, {}
        i = 0
        while i < n:
            t = nodes[i]
            if i in skip or not is_gemv(t) or i in add_of:
                i += 1
                continue
            members, k = [i], i + 1
            while k < n and len(members) < 3:
                if k in skip:
                    k += 1
                    continue
                u = nodes[k]
                if (is_gemv(u) and k not in add_of and self._key(u.src[1]) == self._key(t.src[1]) and
                        u.src[0].type == t.src[0].type and tuple(u.src[0].ne) == tuple(t.src[0].ne) and
                        rms_of.get(k) == rms_of.get(i) and
                        not any(overlaps(self._span(u.src[1]), self._span(nodes[m])) for m in members)):
                    members.append(k)
                    k += 1
                else:
                    break
            if len(members) > 1:
                group_of[i] = members
                for m in members[1:]:
                    skip.add(m)
                if len(members) == 2:
                    g, u2 = members
                    cg = [c for c in consumers.get(self._key(nodes[g]), []) if c[0] > g]
                    cu = [c for c in consumers.get(self._key(nodes[u2]), []) if c[0] > u2]
                    if len(cg) == 1 and len(cu) == 1 and nodes[cg[0][0]].op == GGMLOp.SILU and nodes[cu[0][0]].op == GGMLOp.MUL:
                        si, mi = cg[0][0], cu[0][0]
                        cs = [c for c in consumers.get(self._key(nodes[si]), []) if c[0] > si]
                        if (si == u2 + 1 and mi == si + 1 and len(cs) == 1 and cs[0][0] == mi and
                                nodes[mi].type == GGMLType.F32 and int(nodes[mi].numElements()) == int(nodes[g].numElements())):
                            pair_of[i] = (si, mi)
                            skip.update((si, mi))
            i = members[-1] + 1
        self._groups, self._pairs = group_of, pair_of
        self._gemv = {i for i, t in enumerate(nodes) if is_gemv(t)}
        self._steps = self._planAttentionSteps(nodes, skip)
        return rms_of, add_of, skip

    def _planAttentionSteps(self, nodes, skip):
        """ROPE(q), ROPE(k), CPY(k' -> key cache), CPY(v -> value cache), FLASH_ATTN — the core of one decode step of
        LlamaAttention.forward (model/LlamaAttention.kt:87-123) — as ONE launch (lk_compute_attention_step) at the first
        node's position.  Only the exact five-node run with one token, one position tensor and one table; anything
        else runs node by node.  Returns {index of ROPE(q): (ROPE(q), ROPE(k), CPY k, CPY v, FLASH_ATTN indices)}."""
        steps = {}
        n = len(nodes)
        for i in range(n - 4):
            idx = tuple(range(i, i + 5))
            if any(k in skip for k in idx):
                continue
            rq, rk, ck, cv, fa = (nodes[k] for k in idx)
            if (rq.op, rk.op, ck.op, cv.op, fa.op) != (GGMLOp.ROPE, GGMLOp.ROPE, GGMLOp.CPY, GGMLOp.CPY, GGMLOp.FLASH_ATTN):
                continue
            if any(isinstance(t.extra, dict) for t in (rq, rk, ck, cv, fa)):
                continue
            pos, table, hd = rq.src[1], rq.src[2], int(rq.opParams[0])
            if pos is None or table is None or rk.src[1] is None or rk.src[2] is None:
                continue
            same = lambda a, b: a is not None and b is not None and self._span(a) == self._span(b)   # noqa: E731
            if not (same(rk.src[1], pos) and same(rk.src[2], table) and int(rk.opParams[0]) == hd and
                    same(ck.src[2], pos) and same(cv.src[2], pos) and same(fa.src[3], pos)):
                continue
            kc, vc = ck.src[1], cv.src[1]
            if kc is None or vc is None or int(kc.ne[0]) != hd or tuple(kc.ne) != tuple(vc.ne):
                continue
            rowlen = int(kc.ne[0]) * int(kc.ne[1])
            if not (int(rq.numElements()) == rowlen and int(rk.numElements()) == rowlen and
                    int(ck.src[0].numElements()) == rowlen and int(cv.src[0].numElements()) == rowlen and
                    int(fa.numElements()) == rowlen):
                continue                                   # one token only
            if not (same(ck.src[0], rk) and same(fa.src[0], rq) and same(fa.src[1], kc) and same(fa.src[2], vc)):
                continue
            # the five dsts and the raw inputs must not alias each other
            spans = [self._span(t) for t in (rq, rk, fa, kc, vc, rq.src[0], rk.src[0], cv.src[0])]
            if any(a[0] == b[0] and a[1] < b[2] and b[1] < a[2] for x, a in enumerate(spans) for b in spans[x + 1:]):
                continue
            steps[i] = idx
            skip.update(idx[1:])
        return steps

    # -- the graph as the C ABI takes it -------------------------------------------------------------------
    @staticmethod
    def nodeArray(ga, graph: GGMLCGraph):
        """lk_graph_node[] of a graph, or None if it holds a node only this Python backend understands (`extra` dicts of
        sharded.py) or an op outside the ABI's list."""
        nodes = graph.nodes
        arr = (B.lk_graph_node * max(len(nodes), 1))()
        for i, nd in enumerate(nodes):
            if isinstance(nd.extra, dict) or nd.op.name not in B.LK_OP:
                return None
            srcs = [s_ for s_ in nd.src[:4]]
            while srcs and srcs[-1] is None:
                srcs.pop()
            if any(s_ is None for s_ in srcs):
                return None
            arr[i].op = B.LK_OP[nd.op.name]
            arr[i].n_src = len(srcs)
            arr[i].dst = _describe(ga, nd)
            for k, s_ in enumerate(srcs):
                arr[i].src[k] = _describe(ga, s_)
            for k in range(4):
                arr[i].op_params[k] = C.c_int32(int(nd.opParams[k]) & 0xFFFFFFFF).value
        return arr

    @staticmethod
    def libraryPlan(ga, graph: GGMLCGraph, chain_mode: int = -1):
        """lk_graph_plan of the graph as a list of dicts (tests; tools): the launches the library would make."""
        arr = GGMLCudaBackend.nodeArray(ga, graph)
        if arr is None:
            return None
        n = C.c_int32()
        out = (B.lk_plan_launch * (len(graph.nodes) + 1))()
        B.check(B.lib().lk_graph_plan(arr, len(graph.nodes), chain_mode, out, len(graph.nodes) + 1, C.byref(n)))
        return [dict(kind=L.kind, node=L.node, rms=L.rms, add=L.add, extra=list(L.extra[:L.n_extra]), silu=L.silu, mul=L.mul,
                     store_b=L.store_b, chain=L.chain, attn=list(L.attn) if L.kind == B.LK_PLAN_ATTN_STEP else None) for L in out[:n.value]]

    @staticmethod
    def _ioArray(hostIO):
        """hostIO: [("in" | "out", tensor, host ndarray), ...] -> lk_host_io[] (the region is the array's bytes at the
        tensor's offset in its device buffer)."""
        arr = (B.lk_host_io * len(hostIO))()
        for k, (direction, t, host) in enumerate(hostIO):
            if direction not in ("in", "out") or not isinstance(t.buffer, GGMLCudaBuffer):
                raise B.IllegalArgumentException("hostIO entries are ('in' | 'out', device-resident tensor, host array)")
            arr[k].direction = B.LK_IO_UPLOAD if direction == "in" else B.LK_IO_DOWNLOAD
            arr[k].buf = t.buffer.handle
            arr[k].offset = int(t.dataOffset)
            arr[k].bytes = int(host.nbytes)
            arr[k].host = host.ctypes.data
        return arr

    def _runNodes(self, ga, graph: GGMLCGraph, fuse: bool, hostIO=None) -> None:
        nodes = graph.nodes
        if self.useLibraryPlanner and hasattr(B.lib(), "lk_graph_compute"):     # (an older build under LK_LIBRARY may lack it)
            arr = self.nodeArray(ga, graph)
            if arr is not None:
                flags = 0 if fuse else B.LK_GRAPH_NO_FUSION
                if fuse and self.chainOps is not None:
                    flags |= B.LK_GRAPH_CHAIN_ALWAYS if self.chainOps else B.LK_GRAPH_CHAIN_NEVER
                if hostIO:
                    io = self._ioArray(hostIO)
                    B.check(B.lib().lk_graph_compute_io(self.lk.handle, arr, len(nodes), flags, io, len(hostIO)), self.lk.handle)
                else:
                    B.check(B.lib().lk_graph_compute(self.lk.handle, arr, len(nodes), flags), self.lk.handle)
                return
        if hostIO:      # the Python planner does not place copies between launches: all uploads first, all downloads last
            for direction, t, host in hostIO:
                if direction == "in":
                    t.buffer.setTensorAsync(t, host, 0, host.nbytes)
            try:
                self._runNodes(ga, graph, fuse, None)
            finally:
                for direction, t, host in hostIO:
                    if direction == "out":
                        t.buffer.getTensorAsync(t, host, 0, host.nbytes)
            return
        if not fuse:
            for node in nodes:
                self._computeNode(ga, node)
            return
        rms_of, add_of, skip = self._plan(graph)
        stored = set()                       # RMS_NORM nodes whose dst has been written
        pending = []                         # decode GEMVs (with their fusions) waiting to go out as one chained launch

        def flush():
            items, pending[:] = list(pending), []
            for c in range(0, len(items), self.maxChain):
                self._runChain(ga, nodes, items[c:c + self.maxChain], rms_of, add_of, stored)

        # chainOps None: graphs with an attention step, and graphs whose decode GEMVs all have K % 256 == 0 (lk_planner.cu has the same rule)
        k1m = bool(self._gemv) and all(int(nodes[i].src[0].ne[0]) % 256 == 0 for i in self._gemv)
        chain = (bool(self._steps) or k1m) if self.chainOps is None else bool(self.chainOps)
        for i, node in enumerate(nodes):
            if i in skip or (pending and node.op == GGMLOp.NONE and not isinstance(node.extra, dict)):
                continue
            if chain and i in self._gemv:
                # only DEPENDENT launches are chained (a link reads what the link before it wrote); independent GEMVs
                # keep the pairwise-overlapped launches, which stream faster than any serial order
                if pending and not self._dependsOn(nodes, i, pending[-1], rms_of, add_of):
                    flush()
                pending.append(i)
                continue
            flush()
            self._runPlanned(ga, nodes, i, rms_of, add_of, stored)
        flush()

    def _linkSpans(self, nodes, i, rms_of, add_of):
        """(reads, writes) of the launch the planner makes at GEMV node i, weights left out."""
        node = nodes[i]
        reads, writes = [], [self._span(node)]
        if i in rms_of:
            r = nodes[rms_of[i]]
            reads.append(self._span(r.src[0]))
            writes.append(self._span(r))
        else:
            reads.append(self._span(node.src[1]))
        if i in self._groups:
            writes += [self._span(nodes[m]) for m in self._groups[i][1:]]
            if i in self._pairs:
                writes += [self._span(nodes[k]) for k in self._pairs[i]]
        elif i in add_of:
            a = nodes[add_of[i]]
            reads.append(self._span(a.src[1] if self._key(a.src[0]) == self._key(node) else a.src[0]))
            writes.append(self._span(a))
        return reads, writes

    def _dependsOn(self, nodes, i, j, rms_of, add_of) -> bool:
        reads, _ = self._linkSpans(nodes, i, rms_of, add_of)
        _, writes = self._linkSpans(nodes, j, rms_of, add_of)
        return any(r[0] == w[0] and r[1] < w[2] and w[1] < r[2] for r in reads for w in writes)

    def _linkArgs(self, nodes, i, rms_of, add_of, stored):
        """(a, b, dst, fusion keywords) of the launch the planner makes at GEMV node i: one link of a chain."""
        node, kw = nodes[i], {}
        if i in self._groups:
            kw["extra"] = [(nodes[m].src[0], nodes[m]) for m in self._groups[i][1:]]
            if i in self._pairs:
                si, mi = self._pairs[i]
                kw.update(silu_dst=nodes[si], mul_dst=nodes[mi])
        elif i in add_of:
            a = nodes[add_of[i]]
            kw.update(residual=a.src[1] if self._key(a.src[0]) == self._key(node) else a.src[0], sum_dst=a)
        if i in rms_of:
            r = nodes[rms_of[i]]
            eps = struct.unpack("<f", struct.pack("<i", int(r.opParams[0])))[0]
            kw.update(norm_x=r.src[0], norm_w=r.src[1], eps=eps, store_b=rms_of[i] not in stored)
            stored.add(rms_of[i])
        return (node.src[0], node.src[1], node, kw or None)

    def _runChain(self, ga, nodes, items, rms_of, add_of, stored) -> None:
        """Consecutive GEMV launches of the plan as one lk_compute_matmul_chain; one by one when the library declines."""
        if len(items) >= 2:
            saved = set(stored)
            links = [self._linkArgs(nodes, i, rms_of, add_of, stored) for i in items]
            if ops.computeMatMulChain(ga, None, links):
                return
            stored.clear()
            stored.update(saved)
        for i in items:
            self._runPlanned(ga, nodes, i, rms_of, add_of, stored)

    def _runPlanned(self, ga, nodes, i, rms_of, add_of, stored) -> None:
        """Node i as the plan has it: attention step, stacked group, fused single launch or the plain op."""
        node = nodes[i]
        if i in self._steps:
            rq, rk, ck, cv, fa = (nodes[k] for k in self._steps[i])
            scale = struct.unpack("<f", struct.pack("<i", int(fa.opParams[0])))[0]
            try:
                ops.computeAttentionStep(ga, None, rq.src[0], rk.src[0], cv.src[0], rq.src[2], rq.src[1], int(rq.opParams[0]),
                                         ck.src[1], cv.src[1], scale, fa, qRopeDst=rq, kRopeDst=rk)
            except NotImplementedError:      # the library declined this shape: the five ops, in graph order
                for k in self._steps[i]:
                    self._computeNode(ga, nodes[k])
            return
        if i in self._groups:
            if self._runGroup(ga, nodes, i, rms_of, stored):
                return
            # declined: every member (and the SILU / MUL pair) on its own, in graph order
            for m in self._groups[i]:
                self._runOne(ga, nodes, m, rms_of, add_of, stored)
            if i in self._pairs:
                for k in self._pairs[i]:
                    self._computeNode(ga, nodes[k])
            return
        self._runOne(ga, nodes, i, rms_of, add_of, stored)

    def _runGroup(self, ga, nodes, i, rms_of, stored) -> bool:
        members = self._groups[i]
        node = nodes[i]
        kw = dict(extra=[(nodes[m].src[0], nodes[m]) for m in members[1:]])
        if i in rms_of:
            r = nodes[rms_of[i]]
            eps = struct.unpack("<f", struct.pack("<i", int(r.opParams[0])))[0]
            kw.update(norm_x=r.src[0], norm_w=r.src[1], eps=eps, store_b=rms_of[i] not in stored)
        if i in self._pairs:
            si, mi = self._pairs[i]
            kw.update(silu_dst=nodes[si], mul_dst=nodes[mi])
        if not ops.computeMatMulFused(ga, None, node.src[0], node.src[1], node, **kw):
            return False
        if i in rms_of:
            stored.add(rms_of[i])
        return True

    def _runOne(self, ga, nodes, i, rms_of, add_of, stored) -> None:
        node = nodes[i]
        if True:
            if i not in rms_of and i not in add_of:
                self._computeNode(ga, node)
                return
            r = nodes[rms_of[i]] if i in rms_of else None
            a = nodes[add_of[i]] if i in add_of else None
            kw = {}
            if r is not None:
                eps = struct.unpack("<f", struct.pack("<i", int(r.opParams[0])))[0]
                kw.update(norm_x=r.src[0], norm_w=r.src[1], eps=eps, store_b=rms_of[i] not in stored)
            if a is not None:
                res = a.src[1] if self._key(a.src[0]) == self._key(node) else a.src[0]
                kw.update(residual=res, sum_dst=a)
            if ops.computeMatMulFused(ga, None, node.src[0], node.src[1], node, **kw):
                if r is not None:
                    stored.add(rms_of[i])
                return
            # the library declined this shape: the unfused ops, in graph order
            if r is not None and rms_of[i] not in stored:
                self._computeNode(ga, r)
                stored.add(rms_of[i])
            self._computeNode(ga, node)
            if a is not None:
                self._computeNode(ga, a)

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

    @staticmethod
    def _fastKey(graph: GGMLCGraph):
        """What a replay checks before it trusts its capture: the same node objects at the same offsets.  The full
        signature walks every node and source in Python — 1.7 ms for a 32-layer decode graph, as long as the GPU
        step itself — so it is taken only when this key (or the graph object) has changed.  A caller that edits a
        captured node's sources or shape in place calls forgetGraph() / releaseGraphs()."""
        nodes = graph.nodes
        return (len(nodes), hash(tuple(map(id, nodes))), sum(n.dataOffset for n in nodes))

    def forgetGraph(self, graph: GGMLCGraph) -> None:
        """Drops the capture of one graph (and the reference that keeps the graph object alive for it)."""
        ent = self._captured.pop(id(graph), None)
        self._graphRefs.pop(id(graph), None)
        if ent is not None:
            self.lk.sync()
            B.lib().lk_graph_destroy(ent[0])

    def graphCompute(self, graph: GGMLCGraph, hostIO=None) -> GGMLStatus:
        """core/GGMLBackend.kt:145; exceptions become FAILED like GGMLCpuBackend.kt:168-175.

        hostIO (optional): [("in" | "out", tensor, host ndarray), ...] — host copies that belong to the step.  Each "in"
        array is uploaded to its tensor right before the first launch that reads it, each "out" tensor downloaded right
        after the last launch that writes it, on copy streams beside the compute stream (lk_graph_compute_io), so the
        reference's setTensor -> graphCompute -> getTensor sequence costs one chunk of each copy at the ends of the step
        instead of both in full.  Split the arrays by consumer (one per layer) for the overlap to exist; use pinned
        arrays; the "out" arrays are valid after synchronize().  The host addresses are part of a captured sequence; pass the
        same list object step after step (its key is computed once per object — build a new list to change it)."""
        ga = graph.allocator
        try:
            iokey = None
            if hostIO:      # the same list object step after step: its key (ctypes address look-ups) is computed once
                memo = self._ioKeys.get(id(hostIO))
                if memo is None or memo[0] is not hostIO or len(hostIO) != len(memo[1]):
                    memo = (hostIO, tuple((d, id(t.buffer), int(t.dataOffset), int(h.nbytes), int(h.ctypes.data)) for d, t, h in hostIO))
                    if len(self._ioKeys) > 64:
                        self._ioKeys.clear()
                    self._ioKeys[id(hostIO)] = memo
                iokey = memo[1]
            ent = self._captured.get(id(graph)) if self.useCudaGraphs else None
            if ent is not None and self._graphRefs.get(id(graph)) is graph and ent[2] == (self._fastKey(graph), iokey):
                B.check(B.lib().lk_graph_launch(self.lk.handle, ent[0]), self.lk.handle)      # replay: one launch
            elif self.useCudaGraphs and self._allDevice(graph):
                sig = (self._signature(graph), iokey)
                if ent is None or ent[1] != sig:
                    if ent is not None:
                        B.lib().lk_graph_destroy(ent[0])
                    B.check(B.lib().lk_capture_begin(self.lk.handle), self.lk.handle)
                    try:
                        self._runNodes(ga, graph, fuse=self.fuseOps, hostIO=hostIO)
                    finally:
                        h = C.c_void_p()
                        st = B.lib().lk_capture_end(self.lk.handle, C.byref(h))
                    B.check(st, self.lk.handle)
                    ent = (h, sig, (self._fastKey(graph), iokey))
                else:
                    ent = (ent[0], sig, (self._fastKey(graph), iokey))
                self._captured[id(graph)] = ent
                self._graphRefs[id(graph)] = graph       # the id stays this graph's for as long as the capture exists
                B.check(B.lib().lk_graph_launch(self.lk.handle, ent[0]), self.lk.handle)
            else:
                self._runNodes(ga, graph, fuse=False, hostIO=hostIO)
            if self.checkFaults:
                self.lk.sync()                   # raises IllegalStateException if a bounded device-side wait gave up
            return GGMLStatus.SUCCESS
        except (B.IllegalArgumentException, B.IndexOutOfBoundsException, B.IllegalStateException,
                NotImplementedError) as e:
            print(f"CUDA backend graph computation failed: {e}")
            return GGMLStatus.FAILED
