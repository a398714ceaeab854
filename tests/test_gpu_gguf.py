"""N2 end to end on the device: a GGUF file written by the reference tree's own C writer (tests/golden/ggml_written.gguf)
-> GGUFParser -> payload bytes streamed into the allocator's device buffer -> the hot path with nibble_order = 1."""
import os

import numpy as np
import pytest

import llama_kotlin_b200 as lk
from conftest import same_f32_bits, synth_acts
from llama_kotlin_b200 import GGMLTensor, GGMLType
from llama_kotlin_b200.gguf import ModelLoader

pytestmark = pytest.mark.gpu
G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ggml_written.gguf")


def test_gguf_upload_and_compute(O):
    be = lk.GGMLCudaBackend(0)
    ml = ModelLoader.fromFile(G)
    ml.configure(be)                                   # real GGUF => canonical ggml nibble order
    assert be.lk.get_option("nibble_order") == 1
    ga = lk.GGMLGraphAllocator(be, size=ml.deviceBytes() + (1 << 20))
    ts = ml.uploadAll(ga)
    assert set(ts) == set(ml.getTensorNames())
    c = ml.ggufContext
    for name, t in ts.items():
        info = c.findTensor(name)
        payload = c.getTensorData(info)
        assert t.buffer is ga.deviceBuffer and list(t.ne[:2]) == info.dimensions
        assert np.array_equal(ga.getBytes(t), payload)                      # bytes in HBM == bytes in the file
        if not t.type.isBlockType:
            continue
        K, M = info.dimensions
        ot = int(t.type)
        # K0 on the uploaded blocks: bit-exact with dequantize_row_q* semantics
        assert same_f32_bits(lk.dequantizeTensor(ga, t).data, O.dequantize_ggml(ot, payload, K * M))
        # K1 on the uploaded blocks against the F32 oracle on the same matrix (repacked to the oracle's order)
        x0 = synth_acts(90 + ot, K)
        x = ga.allocateTensor(GGMLTensor(GGMLType.F32, [1, K]))
        d = ga.allocateTensor(GGMLTensor(GGMLType.F32, [1, M]))
        ga.setBytes(x, x0)
        lk.computeMatMul(ga, None, t, x, d)
        ref = O.matmul(ot, O.repack_nibbles(ot, payload, to_ggml=False), K, M, x0, 1)
        assert np.max(np.abs(ga.getFloats(d) - ref)) <= 1e-3 * np.max(np.abs(ref))
    be.free()
