"""Minimal mirror of GGMLGraphAllocator (core/GGMLAlloc.kt:266-719): it only owns the
(buffer, dataOffset) contract the hot path sees — `buffers[bufferId]` holds the bytes,
tensors are suballocated at 16-byte aligned offsets (:122,:149).  Host buffers are numpy
uint8 arrays (the ByteArray analogue); for a device backend buffers[i] is None
(:300-304) and the tensor's `buffer` field holds the GGMLCudaBuffer."""
from __future__ import annotations

import numpy as np

from .ggml_types import GGMLTensor, calculateContiguousStrides

ALIGNMENT = 16


def _align(x: int, a: int = ALIGNMENT) -> int:
    return (x + a - 1) // a * a


class GGMLGraphAllocator:
    def __init__(self, backend=None, size: int = 1 << 20):
        self.backend = backend
        self.context = None
        self._offset = 0
        if backend is None:
            self.buffers = [np.zeros(size, dtype=np.uint8)]
            self.deviceBuffer = None
        else:
            self.buffers = [None]
            self.deviceBuffer = backend.allocBuffer(size)

    def reset(self) -> None:
        self._offset = 0

    def reserve(self, size: int) -> None:
        """ensureBufferCapacity (:361-394) — grows buffer 0; contents are kept."""
        if self.backend is None:
            if self.buffers[0].size < size:
                nb = np.zeros(size, dtype=np.uint8)
                nb[: self.buffers[0].size] = self.buffers[0]
                self.buffers[0] = nb
        elif self.deviceBuffer.getSize() < size:
            old = self.deviceBuffer
            self.deviceBuffer = self.backend.allocBuffer(size)
            self.deviceBuffer.copyFrom(old, old.getSize())
            old.free()

    def allocateTensor(self, tensor: GGMLTensor) -> GGMLTensor:
        """allocateTensor (:486,:508): place the tensor at the next 16-byte aligned offset."""
        nbytes = tensor.byteSize()
        off = _align(self._offset)
        self.reserve(off + nbytes)
        self._offset = off + nbytes
        tensor.bufferId = 0
        tensor.dataOffset = off
        if not any(tensor.nb):
            tensor.nb = calculateContiguousStrides(tensor.ne, tensor.type)
        tensor.buffer = self.deviceBuffer if self.backend is not None else None
        return tensor

    # -- host-side byte access used by tests (ByteArray get/set analogue) -----------
    def setBytes(self, tensor: GGMLTensor, data) -> None:
        raw = np.ascontiguousarray(data).view(np.uint8).reshape(-1)
        if tensor.buffer is not None:
            tensor.buffer.setTensor(tensor, raw, 0, raw.size)
        else:
            self.buffers[tensor.bufferId][tensor.dataOffset: tensor.dataOffset + raw.size] = raw

    def getBytes(self, tensor: GGMLTensor) -> np.ndarray:
        n = tensor.byteSize()
        if tensor.buffer is not None:
            out = np.empty(n, dtype=np.uint8)
            tensor.buffer.getTensor(tensor, out, 0, n)
            return out
        return self.buffers[tensor.bufferId][tensor.dataOffset: tensor.dataOffset + n].copy()

    def getFloats(self, tensor: GGMLTensor) -> np.ndarray:
        return self.getBytes(tensor).view(np.float32)
