"""Host-side mirror of the reference's tensor descriptors for the hot path.

Mirrors (names, field meaning) core/GGMLTypes.kt: GGMLType :99-153, GGMLTensor :251-267,
GGMLOp (subset), GGMLCGraph, GGMLContext; strides per core/GGMLOps.kt:3-28.
Only descriptors live here — no arithmetic (that is CUDA, behind the C ABI).
"""
from __future__ import annotations

import enum
from typing import Any, List, Optional

GGML_MAX_DIMS = 4
QK4_0 = QK4_1 = QK8_0 = 32


class GGMLType(enum.IntEnum):
    """Ordinals as in core/GGMLTypes.kt:145-153; byteSize = bytes per block for Q types (:99-133)."""
    F32 = 0
    F16 = 1
    Q4_0 = 2
    Q4_1 = 3
    Q8_0 = 6
    Q2_K = 8            # N4: the reference's own K-quant formats (:117-122) and BitNet 1.58 (:127)
    Q4_K = 10
    Q8_K = 13
    I32 = 17
    BITNET_1_58 = 40    # no fromValue id in the reference; this ABI's own (include/llamakotlin_cuda.h)

    @property
    def byteSize(self) -> int:
        return {0: 4, 1: 2, 2: 18, 3: 20, 6: 34, 8: 84, 10: 144, 13: 292, 17: 4, 40: 10}[int(self)]

    @property
    def blockElements(self) -> int:
        """getNumBlocks' elementsPerBlock, core/GGMLTypes.kt:507-520 (1 for plain types)"""
        return {2: 32, 3: 32, 6: 32, 8: 256, 10: 256, 13: 256, 40: 32}.get(int(self), 1)

    @property
    def isBlockType(self) -> bool:
        return self.blockElements > 1


class GGMLOp(enum.Enum):
    NONE = "NONE"
    ADD = "ADD"
    MUL = "MUL"
    MUL_MAT = "MUL_MAT"
    RMS_NORM = "RMS_NORM"
    SILU = "SILU"
    SOFT_MAX = "SOFT_MAX"
    SCALE = "SCALE"
    ROPE = "ROPE"
    CPY = "CPY"                # here: KVCache append (src[0] = new rows, src[1] = cache, src[2] = position)
    FLASH_ATTN = "FLASH_ATTN"  # decode attention over the KV cache (src = q, kcache, vcache, position)
    DUP = "DUP"


class GGMLStatus(enum.Enum):
    """core/GGMLBackend.kt GGMLStatus"""
    SUCCESS = 0
    FAILED = 1
    ALLOC_FAILED = 2
    ABORTED = 3


class GGMLContext:
    """Unused by the compute ops, kept for signature parity (core/GGMLComputeOps.kt:1435)."""


def calculateContiguousStrides(ne, type_: GGMLType, rank: int = GGML_MAX_DIMS) -> List[int]:
    """core/GGMLOps.kt:3-28 (block types: nb[0] = block bytes, as the reference writes it)."""
    nb = [0] * GGML_MAX_DIMS
    nb[0] = type_.byteSize
    for d in range(1, GGML_MAX_DIMS):
        dim = ne[d - 1] if d - 1 < len(ne) else 1
        nb[d] = nb[d - 1] * (dim if dim > 0 else 1)
    return nb


class GGMLTensor:
    """core/GGMLTypes.kt:251-267 — the fields the hot path reads."""

    def __init__(self, type: GGMLType = GGMLType.F32, ne=None, nb=None, op: GGMLOp = GGMLOp.NONE,
                 opParams=None, src=None, bufferId: int = -1, dataOffset: int = 0, name: str = "",
                 buffer: Any = None, data: Any = None, extra: Any = None):
        self.type = GGMLType(type)
        self.ne = list(ne) + [1] * (GGML_MAX_DIMS - len(ne)) if ne is not None else [0] * GGML_MAX_DIMS
        self.nb = list(nb) if nb is not None else [0] * GGML_MAX_DIMS
        self.op = op
        self.opParams = list(opParams) if opParams is not None else [0] * 16
        self.src: List[Optional["GGMLTensor"]] = list(src) if src is not None else [None] * 10
        self.bufferId = bufferId
        self.dataOffset = dataOffset
        self.name = name
        self.buffer = buffer      # GGMLBackendBuffer for device-resident tensors
        self.data = data
        self.extra = extra        # backend-private annotations (the reference's GGMLTensor.extra slot)

    def rank(self) -> int:
        if all(v <= 1 for v in self.ne):
            return 1 if any(v > 0 for v in self.ne) else 0
        return max(i for i, v in enumerate(self.ne) if v > 1) + 1

    def numElements(self) -> int:
        r = self.rank()
        if r == 0:
            return 1 if all(v <= 1 for v in self.ne) and any(v > 0 for v in self.ne) else 0
        count = 1
        for i in range(max(r, 1)):
            if self.ne[i] == 0 and r > 1:
                return 0
            if self.ne[i] > 0:
                count *= self.ne[i]
        return count

    def byteSize(self) -> int:
        """calculateTensorByteSize, core/GGMLTypes.kt:1136-1190"""
        n = self.numElements()
        return n // self.type.blockElements * self.type.byteSize


class GGMLCGraph:
    """core/GGMLTypes.kt GGMLCGraph: nodes in execution order."""

    def __init__(self, nodes=None, leafs=None, allocator=None):
        self.nodes: List[GGMLTensor] = list(nodes) if nodes else []
        self.leafs: List[GGMLTensor] = list(leafs) if leafs else []
        self.allocator = allocator

    @property
    def nNodes(self) -> int:
        return len(self.nodes)
