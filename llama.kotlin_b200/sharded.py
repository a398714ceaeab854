"""Weight-row sharding of the decode matmul over the GPUs of one box.

The reference has no multi-device path; the layout rules come from how
computeMatMul addresses the weight (core/GGMLComputeOps.kt:1520-1523, core/GGMLTypes.kt:549):
output rows are independent dot products and weight rows are contiguous byte ranges, so rank r
owns rows [row0, row0 + rows) of `a` as a zero-copy slice of the on-disk layout, `b` is
replicated, and rank r's outputs land at dst[row0 : row0 + rows] on EVERY rank:

  * fused path (default): the GEMV epilogue stores each output into every peer's dst through
    NVLink peer mappings (CUDA IPC); the exchange point is either folded into the launches
    themselves (`sync=(wait_slot, signal_slot)`: the producer's last CTA arrives on every rank's
    counter, the consumer waits in its prologue, after its weight prefetch has been issued) or
    a separate flag-barrier launch (`barrier()`);
  * baseline: `allgather()` = in-place ncclAllGather of the per-rank slices.

Reuse rule for a vector that peers write into (dst of a sharded launch): it may be rewritten only
by a launch that comes after an exchange point LATER than the one its readers waited on — every
reader has then provably finished.  [SwiGLU -> exchange -> down -> exchange -> next SwiGLU] satisfies
it with one buffer per vector; a loop with a SINGLE exchange point that reuses one dst does not
(a fast rank's step i+1 stores can overtake a slow rank's step-i reads): alternate two buffers by
step parity or add a second exchange after the consumer.

One process per GPU.  `exchange(obj) -> [obj_rank0, ...]` is any all-gather of small picklable
objects between the ranks (torch.distributed.all_gather_object on a gloo group in bench.py and
the tests); this module does not import torch.
"""
from __future__ import annotations

import ctypes as C
from typing import Callable, List, Sequence, Tuple

from . import _binding as B
from .ggml_compute_ops import describe
from .ggml_types import GGMLTensor

FLAG_BYTES = B.LK_FLAG_BYTES   # first 64 bytes = the 16 words of barrier() (one epoch slot per rank, word 14 = time-out mark,
                               # word 15 = epoch counter); then the arrival / ticket counters of the exchange slots


def shard_rows(M: int, world: int, rank: int, unit: int = 1) -> Tuple[int, int]:
    """Rows [row0, row0 + rows) of an M-row weight that rank `rank` of `world` owns: contiguous,
    balanced to within one `unit` (unit = rows whose packed bytes are a multiple of 16, so every
    shard starts on a TMA-aligned byte offset; 1 for all BASELINE shapes)."""
    if world < 1 or not 0 <= rank < world:
        raise B.IllegalArgumentException(f"bad rank {rank} of {world}")
    if unit < 1 or M % unit:
        raise B.IllegalArgumentException(f"M={M} is not a multiple of the shard unit {unit}")
    units = M // unit
    base, rem = divmod(units, world)
    u0 = rank * base + min(rank, rem)
    n = base + (1 if rank < rem else 0)
    return u0 * unit, n * unit


def shard_byte_range(M: int, K: int, block_bytes: int, world: int, rank: int, unit: int = 1) -> Tuple[int, int]:
    """Byte range of rank's rows inside the packed weight (row i = bytes [i*K/32*B, (i+1)*K/32*B))."""
    row0, rows = shard_rows(M, world, rank, unit)
    rb = K // 32 * block_bytes
    return row0 * rb, rows * rb


class ShardGroup:
    """The ranks that share one sharded matmul: owns the peer mappings of the symmetric dst buffer
    and of the barrier flags."""

    def __init__(self, backend, rank: int, world: int, exchange: Callable[[object], List[object]]):
        if world > 8:
            raise B.IllegalArgumentException("at most 8 ranks (one NVSwitch box)")
        self.backend, self.lk = backend, backend.lk
        self.rank, self.world, self.exchange = rank, world, exchange
        self._imported = []
        self.flags = backend.allocBuffer(FLAG_BYTES)
        self.flags.clear(0)
        self.lk.sync()
        self.peer_flags = self._map_peers(self.flags)
        self._nccl = False
        self._barrier_ready = False

    # -- peer mappings ---------------------------------------------------------------------
    def _map_peers(self, buf) -> List[object]:
        """lk_buf handles of `buf` on every rank (own entry = buf's own handle)."""
        h = (C.c_uint8 * B.LK_IPC_HANDLE_BYTES)()
        B.check(B.lib().lk_buf_ipc_export(buf.handle, h), self.lk.handle)
        handles = self.exchange((bytes(h), buf.getSize()))
        out = []
        for r, (raw, size) in enumerate(handles):
            if r == self.rank:
                out.append(buf.handle)
                continue
            arr = (C.c_uint8 * B.LK_IPC_HANDLE_BYTES).from_buffer_copy(raw)
            ph = C.c_void_p()
            B.check(B.lib().lk_buf_ipc_import(self.lk.handle, arr, int(size), C.byref(ph)), self.lk.handle)
            self._imported.append(ph)
            out.append(ph)
        return out

    def map_dst_buffer(self, buf) -> List[object]:
        """Peer handles of a dst buffer that every rank allocated with the same size and layout."""
        return self._map_peers(buf)

    # -- ops ----------------------------------------------------------------------------------
    def _sync(self, sync):
        """lk_shard_sync for sync = (wait_slot or None, signal_slot or None); returns (pointer or None, keep-alive)."""
        if sync is None:
            return None, None
        wait_slot, signal_slot = sync
        arr = (C.c_void_p * self.world)(*[p if isinstance(p, C.c_void_p) else C.c_void_p(p) for p in self.peer_flags])
        st = B.lk_shard_sync(arr, self.world, self.rank, -1 if wait_slot is None else int(wait_slot),
                             -1 if signal_slot is None else int(signal_slot))
        return C.pointer(st), (arr, st)

    def computeMatMulSharded(self, graphAllocator, a_shard: GGMLTensor, b: GGMLTensor, dst_full: GGMLTensor,
                             row0: int, M_total: int, peer_dst: Sequence[object] = None, sync=None) -> None:
        """dst_full[row0 : row0 + a_shard.ne[1]] = a_shard x b on this rank and, when peer_dst is
        given (map_dst_buffer of dst_full's buffer), in every peer's copy of dst_full.
        sync = (wait_slot, signal_slot): the exchange points folded into this launch (lk_shard_sync): wait for every
        rank's arrival on wait_slot before reading b; arrive on every rank's signal_slot after the peer stores."""
        n = len(peer_dst) if peer_dst else 0
        arr = (C.c_void_p * max(n, 1))(*[p if isinstance(p, C.c_void_p) else C.c_void_p(p) for p in (peer_dst or [])])
        sp, _keep = self._sync(sync)
        B.check(B.lib().lk_compute_matmul_sharded_sync(self.lk.handle, describe(graphAllocator, a_shard),
                                                       describe(graphAllocator, b), describe(graphAllocator, dst_full),
                                                       int(row0), int(M_total), arr if n else None, n, sp), self.lk.handle)

    def computeSwigluSharded(self, graphAllocator, gate_shard: GGMLTensor, up_shard: GGMLTensor, x: GGMLTensor,
                             g_full: GGMLTensor, u_full: GGMLTensor, h_full: GGMLTensor, row0: int, M_total: int,
                             peer_h: Sequence[object] = None, silu_full: GGMLTensor = None, norm_w: GGMLTensor = None,
                             eps: float = 0.0, normed_dst: GGMLTensor = None, sync=None) -> None:
        """The sharded SwiGLU half of LlamaMLP.forward (model/LlamaModel.kt:159-176) as one launch: this rank's rows of
        g = gate x', u = up x', h = silu(g) * u (x' = RMSNorm(x, norm_w, eps) when norm_w is given), h also stored into
        every peer's copy of h_full (peer_h = map_dst_buffer of h_full's buffer); follow with barrier(), or fold the
        exchange points into the launches with sync = (wait_slot, signal_slot) as in computeMatMulSharded."""
        n = len(peer_h) if peer_h else 0
        arr = (C.c_void_p * max(n, 1))(*[p if isinstance(p, C.c_void_p) else C.c_void_p(p) for p in (peer_h or [])])
        d = lambda t: describe(graphAllocator, t) if t is not None else None   # noqa: E731
        sp, _keep = self._sync(sync)
        B.check(B.lib().lk_compute_swiglu_sharded_sync(self.lk.handle, d(gate_shard), d(up_shard), d(x), d(norm_w), float(eps),
                                                       d(normed_dst), d(g_full), d(u_full), d(silu_full), d(h_full), int(row0),
                                                       int(M_total), arr if n else None, n, sp), self.lk.handle)

    def barrier(self) -> None:
        """Every rank's preceding peer stores are visible everywhere after this (stream-ordered)."""
        arr = (C.c_void_p * self.world)(*[p if isinstance(p, C.c_void_p) else C.c_void_p(p) for p in self.peer_flags])
        B.check(B.lib().lk_peer_barrier(self.lk.handle, self.flags.handle, arr, self.world, self.rank, 0), self.lk.handle)
        self._barrier_ready = True

    def push(self, buf, offset: int, count: int, peer_bufs: Sequence[object]) -> None:
        """All-gather as peer stores: this rank's `count` floats at `offset` of `buf` go to the same offset of
        every peer's buffer (map_dst_buffer of `buf`); follow with barrier()."""
        arr = (C.c_void_p * len(peer_bufs))(*[p if isinstance(p, C.c_void_p) else C.c_void_p(p) for p in peer_bufs])
        B.check(B.lib().lk_peer_push_f32(self.lk.handle, buf.handle, int(offset), int(count), arr, len(peer_bufs)),
                self.lk.handle)

    def timed_out(self) -> bool:
        """True if a bounded cross-GPU wait gave up (barrier() word 14, exchange slots' word at byte 256).  The library also
        reports this once through the status of the next synchronising call (IllegalStateException)."""
        import numpy as np
        w = np.zeros(FLAG_BYTES // 4, np.uint32)
        try:
            B.check(B.lib().lk_buf_download(self.flags.handle, 0, w.ctypes.data, FLAG_BYTES), self.lk.handle)
        except B.IllegalStateException:
            return True
        return bool(w[14]) or bool(w[64])

    # -- NCCL baseline ----------------------------------------------------------------------------
    def init_nccl(self) -> None:
        uid = (C.c_uint8 * B.LK_COMM_ID_BYTES)()
        if self.rank == 0:
            B.check(B.lib().lk_comm_unique_id(uid))
        ids = self.exchange(bytes(uid))
        arr = (C.c_uint8 * B.LK_COMM_ID_BYTES).from_buffer_copy(ids[0])
        B.check(B.lib().lk_comm_init(self.lk.handle, self.world, self.rank, arr), self.lk.handle)
        self._nccl = True

    def allgather(self, buf, offset: int, count_per_rank: int) -> None:
        """In-place ncclAllGather: rank r's `count_per_rank` floats sit at offset + r*count*4."""
        if not self._nccl:
            self.init_nccl()
        B.check(B.lib().lk_allgather_f32(self.lk.handle, buf.handle, int(offset), int(count_per_rank)), self.lk.handle)

    def close(self) -> None:
        self.lk.sync()
        for ph in self._imported:
            B.lib().lk_buf_free(ph)
        self._imported = []
        if self._nccl:
            B.lib().lk_comm_destroy(self.lk.handle)
            self._nccl = False
        self.flags.free()
