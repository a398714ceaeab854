"""GGUF reader (llama.kotlin_b200/gguf.py) against (1) the byte stream of the reference's own
TestGGUFGenerator with the assertions of its GGUFTest.kt, (2) a file written by the reference tree's own C writer
(gguf_write_to_file through oracle/_ref; fixture + generating script under tests/golden).  CPU only, no device."""
import hashlib
import json
import os
import struct

import numpy as np
import pytest

from llama_kotlin_b200 import IllegalArgumentException, IndexOutOfBoundsException
from llama_kotlin_b200.ggml_types import GGMLType
from llama_kotlin_b200.gguf import GGUFParser, ModelLoader

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _kotlin_file() -> bytes:
    with open(os.path.join(G, "reference_testfile.gguf"), "rb") as f:
        return f.read()


def test_reference_test_file_header_metadata_tensors():
    """nativeTest gguf/GGUFTest.kt:14-65 on TestGGUFGenerator.generateTestFile()."""
    c = GGUFParser(_kotlin_file()).parse()
    assert c.version == 3 and len(c.tensors) == 2 and len(c.metadata) == 3              # :21-23
    assert c.getStringValue("general.architecture") == "test"                            # :35
    assert c.getStringValue("general.name") == "test-model"                              # :36
    assert c.getLongValue("general.alignment") == 32                                     # :37
    t0, t1 = c.findTensor("weight.0"), c.findTensor("weight.1")
    assert t0.name == "weight.0" and t0.type == GGMLType.F32 and t0.dimensions == [2, 2] and t0.offset == 0    # :50-56
    assert t1.type == GGMLType.F32 and t1.dimensions == [3, 3] and t1.offset == 16                              # :59-64
    assert c.findTensor("nope") is None
    assert c.getArchitecture() == "test" and c.getModelName() == "test-model"
    assert c.dataOffset % 32 == 0


def test_reference_test_file_tensor_data():
    """GGUFTest.kt:68-86: 16 bytes, 1, 2, 3, 4; and the identity-like second tensor."""
    c = GGUFParser(_kotlin_file()).parse()
    d = c.getTensorData(c.findTensor("weight.0"))
    assert d.size == 16 and list(d.view(np.float32)) == [1.0, 2.0, 3.0, 4.0]
    assert list(c.getTensorData(c.findTensor("weight.1")).view(np.float32)) == [1, 0, 0, 0, 1, 0, 0, 0, 1]


def test_invalid_inputs_raise_the_reference_exceptions():
    good = _kotlin_file()
    with pytest.raises(IllegalArgumentException, match="Invalid GGUF magic"):            # GGUFParser.kt:23-25
        GGUFParser(b"GGML" + good[4:]).parse()
    with pytest.raises(IndexOutOfBoundsException):                                       # truncated KV section
        GGUFParser(good[:40]).parse()
    c = GGUFParser(good[:-8]).parse()                                                    # payload cut short
    with pytest.raises(IndexOutOfBoundsException, match="beyond file"):                  # GGUFContext.kt:90-92
        c.getTensorData(c.findTensor("weight.1"))

    def s(x):
        b = x.encode()
        return struct.pack("<Q", len(b)) + b
    nested = b"GGUF" + struct.pack("<IQQ", 3, 0, 1) + s("k") + struct.pack("<I", 9) + struct.pack("<IQ", 9, 1) + struct.pack("<IQ", 4, 0)
    with pytest.raises(IllegalArgumentException, match="Nested arrays"):                 # GGUFParser.kt:123
        GGUFParser(nested).parse()
    unknown = b"GGUF" + struct.pack("<IQQ", 3, 0, 1) + s("k") + struct.pack("<I", 99)
    with pytest.raises(IllegalArgumentException, match="Unknown GGUF type"):             # GGUFTypes.kt:24-26
        GGUFParser(unknown).parse()


def test_default_alignment_when_key_absent():
    def s(x):
        b = x.encode()
        return struct.pack("<Q", len(b)) + b
    buf = b"GGUF" + struct.pack("<IQQ", 3, 1, 0) + s("w") + struct.pack("<IQIQ", 1, 4, 0, 0)
    c = GGUFParser(buf + b"\0" * 64).parse()
    assert c.alignment == 32 and c.dataOffset == (len(buf) + 31) // 32 * 32              # GGUFConstants.DEFAULT_ALIGNMENT


def test_file_written_by_the_reference_c_writer(O):
    """Real ggml type ids (Q8_0 = 8), ggml's padding, 5 KV value kinds, a string array."""
    ml = ModelLoader.fromFile(os.path.join(G, "ggml_written.gguf"))
    c = ml.ggufContext
    want = json.load(open(os.path.join(G, "ggml_written.json")))
    assert c.version == 3 and c.alignment == want["alignment"] and c.dataOffset % c.alignment == 0
    assert c.getStringValue("general.architecture") == "llama"
    assert c.getIntValue("llama.block_count") == 1 and c.getIntValue("llama.embedding_length") == 128
    assert c.getFloatValue("llama.rope.freq_base") == 10000.0
    assert c.getBooleanValue("lk.flag") is True
    assert c.getMetadataValue("tokenizer.ggml.tokens") == ["<s>", "</s>", "hello"]
    assert ml.getTensorNames() == list(want["tensors"])
    for name, w in want["tensors"].items():
        info = c.findTensor(name)
        assert int(info.type) == w["type"] and info.dimensions == w["ne"] and info.offset % c.alignment == 0
        data = c.getTensorData(info)
        assert data.nbytes == w["nbytes"] and hashlib.sha256(data.tobytes()).hexdigest() == w["sha256"]
        # the payload decodes (ggml nibble order) to the seeded weights it was quantized from
        src = np.random.default_rng(w["seed"]).normal(0, 0.02, w["ne"][0] * w["ne"][1]).astype(np.float32)
        deq = data.view(np.float32) if w["type"] == 0 else O.dequantize_ggml(w["type"], data, src.size)
        tol = {0: 0.0, 2: 0.008, 3: 0.004, 6: 0.0005}[w["type"]]
        assert np.max(np.abs(deq - src)) <= tol
    with pytest.raises(IllegalArgumentException, match="Tensor not found"):
        ml.uploadTensor(None, "missing")


def test_kotlin_type_ids_mode():
    """A file written with the Kotlin enum's ordinals (Q8_0 = 6) is read with type_ids='kotlin'; the same id in a real
    file is ggml's Q5_0, which this device path has no kernels for (info kept, type None)."""
    def s(x):
        b = x.encode()
        return struct.pack("<Q", len(b)) + b
    buf = b"GGUF" + struct.pack("<IQQ", 3, 1, 0) + s("w") + struct.pack("<IQIQ", 1, 32, 6, 0)
    buf += b"\0" * ((32 - len(buf) % 32) % 32) + bytes(34)
    assert GGUFParser(buf, type_ids="kotlin").parse().tensors[0].type == GGMLType.Q8_0
    t = GGUFParser(buf, type_ids="ggml").parse().tensors[0]
    assert t.type is None and t.typeId == 6
    with pytest.raises(NotImplementedError):
        t.byteSize()
