"""GGUF file -> device tensors: the step directly before the hot path.

Host-side mirror of the reference's parser, same names and error behaviour:
  GGUFParser.parse        gguf/GGUFParser.kt:19-56   (magic, version, counts, KV pairs, tensor infos, alignment, data offset)
  readKeyValue / readArray  :58-126                  (the 13 GGUFType value kinds; nested arrays rejected)
  GGUFContext               gguf/GGUFContext.kt:6-106 (getStringValue / getIntValue / ..., findTensor, getTensorData)
  GGUFConstants             gguf/GGUFTypes.kt:72-75   (MAGIC "GGUF", DEFAULT_ALIGNMENT 32)
and of the part the reference stops short of — ModelLoader.loadTensorData (gguf/ModelLoader.kt:78-100) only copies F32
payloads and prints a warning for every quantized type.  Here `uploadTensor` streams the payload bytes of F32 / F16 /
Q4_0 / Q4_1 / Q8_0 tensors straight from the (memory-mapped) file into the allocator's device buffer; no repacking:
the on-disk block layout IS the layout the kernels read.

Two reference defects are not mirrored (recorded in DESIGN.md):
  * type ids: a GGUF file stores ggml_type ids (Q8_0 = 8); the Kotlin enum's `fromValue` maps 6 -> Q8_0 and 8 -> Q5_1
    (core/GGMLTypes.kt:145-170), so it misreads real files.  `type_ids="ggml"` (default) reads real files;
    `type_ids="kotlin"` reads files written with the Kotlin ordinals.
  * tensor byte size: GGUFContext.calculateTensorSize multiplies the ELEMENT count by the block size
    (gguf/GGUFContext.kt:101-104); here it is n/32 * blockBytes (calculateTensorByteSize, core/GGMLTypes.kt:1136-1190).
Real GGUF files hold Q4 blocks in the canonical ggml nibble order: set the context option `nibble_order` = 1 (N1)
before computing with tensors uploaded from them (`ModelLoader.configure` does).

No arithmetic happens in this module; it never touches tensor values."""
from __future__ import annotations

import mmap
import struct
from dataclasses import dataclass
from typing import Any, Dict, List, Optional

import numpy as np

from . import _binding as B
from .ggml_types import GGMLTensor, GGMLType

MAGIC = "GGUF"                 # GGUFConstants.MAGIC
DEFAULT_ALIGNMENT = 32         # GGUFConstants.DEFAULT_ALIGNMENT

# GGUFType, gguf/GGUFTypes.kt:6-20
UINT8, INT8, UINT16, INT16, UINT32, INT32, FLOAT32, BOOL, STRING, ARRAY, UINT64, INT64, FLOAT64 = range(13)
_SCALAR = {UINT8: "<B", INT8: "<b", UINT16: "<H", INT16: "<h", UINT32: "<I", INT32: "<i", FLOAT32: "<f",
           UINT64: "<Q", INT64: "<q", FLOAT64: "<d"}
# ggml_type id (what files store) -> the reference's GGMLType ordinal used everywhere else in this package
_GGML_IDS = {0: GGMLType.F32, 1: GGMLType.F16, 2: GGMLType.Q4_0, 3: GGMLType.Q4_1, 8: GGMLType.Q8_0}
_KOTLIN_IDS = {0: GGMLType.F32, 1: GGMLType.F16, 2: GGMLType.Q4_0, 3: GGMLType.Q4_1, 6: GGMLType.Q8_0}


@dataclass
class GGUFKeyValue:            # gguf/GGUFTypes.kt:33-37
    key: str
    type: int
    value: Any


@dataclass
class GGUFTensorInfo:          # gguf/GGUFTypes.kt:42-47
    name: str
    dimensions: List[int]
    type: Optional[GGMLType]   # None: a type this package has no kernels for (the info is kept, upload refuses)
    offset: int
    typeId: int = 0            # the id as stored in the file

    def numElements(self) -> int:
        n = 1
        for d in self.dimensions:
            n *= d
        return n

    def byteSize(self) -> int:
        if self.type is None:
            raise NotImplementedError(f"tensor '{self.name}': type id {self.typeId} is not supported on this device path")
        n = self.numElements()
        return n // self.type.blockElements * self.type.byteSize


class GGUFContext:             # gguf/GGUFContext.kt:6-13
    def __init__(self, version, metadata, tensors, dataOffset, alignment, rawData):
        self.version: int = version
        self.metadata: Dict[str, GGUFKeyValue] = metadata
        self.tensors: List[GGUFTensorInfo] = tensors
        self.dataOffset: int = dataOffset
        self.alignment: int = alignment
        self.rawData = rawData      # bytes-like (bytes, mmap or numpy uint8)

    def getMetadataValue(self, key: str):
        kv = self.metadata.get(key)
        return kv.value if kv is not None else None

    def getStringValue(self, key: str) -> Optional[str]:
        v = self.getMetadataValue(key)
        return v if isinstance(v, str) else None

    def getIntValue(self, key: str) -> Optional[int]:
        v = self.getMetadataValue(key)
        return int(v) if isinstance(v, int) and not isinstance(v, bool) else None

    getLongValue = getIntValue

    def getFloatValue(self, key: str) -> Optional[float]:
        v = self.getMetadataValue(key)
        return float(v) if isinstance(v, float) else None

    def getBooleanValue(self, key: str) -> Optional[bool]:
        v = self.getMetadataValue(key)
        return v if isinstance(v, bool) else None

    def findTensor(self, name: str) -> Optional[GGUFTensorInfo]:
        for t in self.tensors:
            if t.name == name:
                return t
        return None

    def getTensorData(self, info: GGUFTensorInfo) -> np.ndarray:
        """gguf/GGUFContext.kt:85-96 — a zero-copy uint8 view of the payload (the reference copies)."""
        size = info.byteSize()
        start = self.dataOffset + info.offset
        end = start + size
        if end > len(self.rawData):
            raise B.IndexOutOfBoundsException(f"Tensor data extends beyond file: {end} > {len(self.rawData)}")
        return np.frombuffer(self.rawData, dtype=np.uint8, count=size, offset=start)

    def getArchitecture(self) -> Optional[str]:
        return self.getStringValue("general.architecture")

    def getModelName(self) -> Optional[str]:
        return self.getStringValue("general.name")


class GGUFParser:
    """gguf/GGUFParser.kt:13-201."""

    def __init__(self, data, type_ids: str = "ggml"):
        self.data = data
        self.position = 0
        self._ids = _GGML_IDS if type_ids == "ggml" else _KOTLIN_IDS

    # -- primitive readers (:128-196); running off the end is IndexOutOfBoundsException as in the reference
    def _take(self, n: int) -> bytes:
        if n < 0 or self.position + n > len(self.data):
            raise B.IndexOutOfBoundsException("End of data reached")
        b = bytes(self.data[self.position:self.position + n])
        self.position += n
        return b

    def _scalar(self, t: int):
        fmt = _SCALAR[t]
        return struct.unpack(fmt, self._take(struct.calcsize(fmt)))[0]

    def _string(self, length: int) -> str:
        if length < 0 or self.position + length > len(self.data):
            raise B.IndexOutOfBoundsException(f"String length {length} exceeds available data")
        return self._take(length).decode("utf-8", "replace")

    def _value(self, t: int, in_array: bool = False):
        if t == BOOL:
            return self._scalar(UINT8) != 0
        if t == STRING:
            return self._string(self._scalar(UINT64))
        if t == ARRAY:
            if in_array:
                raise B.IllegalArgumentException("Nested arrays not supported")       # :123
            at = self._scalar(UINT32)
            if at not in range(13):
                raise B.IllegalArgumentException(f"Unknown GGUF type: {at}")
            n = self._scalar(UINT64)
            return [self._value(at, True) for _ in range(n)]
        if t in _SCALAR:
            return self._scalar(t)
        raise B.IllegalArgumentException(f"Unknown GGUF type: {t}")                    # gguf/GGUFTypes.kt:24-26

    def parse(self) -> GGUFContext:
        magic = self._take(4).decode("latin-1")
        if magic != MAGIC:
            raise B.IllegalArgumentException(f"Invalid GGUF magic: {magic}")           # :23-25
        version = self._scalar(UINT32)
        tensorCount = self._scalar(UINT64)
        kvCount = self._scalar(UINT64)
        metadata: Dict[str, GGUFKeyValue] = {}
        for _ in range(kvCount):
            key = self._string(self._scalar(UINT64))
            t = self._scalar(UINT32)
            metadata[key] = GGUFKeyValue(key, t, self._value(t))
        tensors: List[GGUFTensorInfo] = []
        for _ in range(tensorCount):
            name = self._string(self._scalar(UINT64))
            nd = self._scalar(UINT32)
            dims = [self._scalar(UINT64) for _ in range(nd)]
            tid = self._scalar(UINT32)
            off = self._scalar(UINT64)
            tensors.append(GGUFTensorInfo(name, dims, self._ids.get(tid), off, tid))
        a = metadata.get("general.alignment")
        alignment = int(a.value) if a is not None and isinstance(a.value, int) else DEFAULT_ALIGNMENT   # :47
        dataOffset = (self.position + alignment - 1) // alignment * alignment                            # :199-201
        return GGUFContext(version, metadata, tensors, dataOffset, alignment, self.data)


class ModelLoader:
    """gguf/ModelLoader.kt: GGUF tensors -> GGMLTensor in a GGMLGraphAllocator — here a DEVICE allocator, and with the
    quantized payloads the reference leaves unloaded (:94-97)."""

    def __init__(self, ggufContext: GGUFContext):
        self.ggufContext = ggufContext

    @classmethod
    def fromFile(cls, path: str, type_ids: str = "ggml") -> "ModelLoader":
        f = open(path, "rb")
        mm = mmap.mmap(f.fileno(), 0, access=mmap.ACCESS_READ)
        return cls(GGUFParser(mm, type_ids).parse())

    def getTensorNames(self) -> List[str]:
        return [t.name for t in self.ggufContext.tensors]

    def configure(self, backend) -> None:
        """Real GGUF files hold Q4 blocks in the canonical ggml nibble order."""
        backend.lk.set_option("nibble_order", 1)

    def deviceBytes(self, names=None) -> int:
        """Bytes the named tensors need in the allocator (16-byte aligned placements, core/GGMLAlloc.kt:122)."""
        return sum((t.byteSize() + 15) // 16 * 16 for t in self.ggufContext.tensors if names is None or t.name in names)

    def uploadTensor(self, graphAllocator, name: str) -> GGMLTensor:
        """loadTensor + loadTensorData for any supported type: descriptor from the tensor info (ne = dimensions, GGUF
        order: ne[0] fastest), payload bytes file -> HBM as they are."""
        info = self.ggufContext.findTensor(name)
        if info is None:
            raise B.IllegalArgumentException(f"Tensor not found: {name}")
        if len(info.dimensions) < 1 or len(info.dimensions) > 4:
            raise B.IllegalArgumentException(f"Unsupported tensor dimension count: {len(info.dimensions)}")   # :64
        if info.type is None:
            raise NotImplementedError(f"tensor '{name}': GGUF type id {info.typeId} has no kernels on this device path")
        if info.type.isBlockType and info.dimensions[0] % 32:
            raise B.IllegalArgumentException(f"tensor '{name}': ne[0] = {info.dimensions[0]} is not a multiple of the block size 32")
        payload = self.ggufContext.getTensorData(info)
        t = graphAllocator.allocateTensor(GGMLTensor(info.type, list(info.dimensions), name=name))
        graphAllocator.setBytes(t, payload)
        return t

    def uploadAll(self, graphAllocator, names=None) -> Dict[str, GGMLTensor]:
        return {t.name: self.uploadTensor(graphAllocator, t.name) for t in self.ggufContext.tensors
                if (names is None or t.name in names) and t.type is not None}
