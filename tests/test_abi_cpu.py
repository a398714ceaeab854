"""CPU-side checks of the drop-in boundary: the C-ABI library loads, exports every symbol
include/llamakotlin_cuda.h declares, the pure (no-device) helpers answer like the reference,
and the product path fails loudly without a GPU (no CPU fallback)."""
import ctypes as C
import os

import numpy as np
import pytest

import llama_kotlin_b200 as lk
from llama_kotlin_b200 import _binding as B
from llama_kotlin_b200 import GGMLTensor, GGMLType, calculateContiguousStrides


def _has_gpu() -> bool:
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def test_library_exports_every_declared_symbol():
    L = B.lib()
    names = B.declared_symbols()
    assert len(names) >= 40
    for n in names:
        assert hasattr(L, n), f"{n} declared in the header but not exported"
    assert L.lk_abi_version() == 3


def test_binding_covers_every_declared_symbol():
    # every header function has ctypes argtypes set (the binding is complete)
    L = B.lib()
    for n in B.declared_symbols():
        assert getattr(L, n).argtypes is not None, n


def test_block_bytes_match_reference_byteSize():
    # core/GGMLTypes.kt:99-133
    L = B.lib()
    for t in GGMLType:
        assert L.lk_type_block_bytes(int(t)) == t.byteSize
    assert L.lk_type_block_bytes(99) == 0


def test_tensor_nbytes_matches_calculateTensorByteSize():
    L = B.lib()
    for t, ne in ((GGMLType.Q4_0, [4096, 4096]), (GGMLType.Q4_1, [4096, 11008]), (GGMLType.Q8_0, [64, 3]),
                  (GGMLType.F32, [7, 5, 3]), (GGMLType.F16, [8])):
        ten = GGMLTensor(t, ne)
        d = B.lk_tensor()
        d.type = int(t)
        for i in range(4):
            d.ne[i] = ten.ne[i]
        assert L.lk_tensor_nbytes(C.byref(d)) == ten.byteSize()
    # figures quoted in SURVEY.md §8(d)
    assert GGMLTensor(GGMLType.Q4_0, [4096, 4096]).byteSize() == 9437184
    assert GGMLTensor(GGMLType.Q8_0, [4096, 11008]).byteSize() == 47906816


def test_contiguous_strides_follow_reference():
    # core/GGMLOps.kt:3-28: nb[0] = type size (block bytes for Q types), nb[d] = nb[d-1]*ne[d-1]
    assert calculateContiguousStrides([3, 2, 1, 1], GGMLType.F32) == [4, 12, 24, 24]
    assert calculateContiguousStrides([64, 4, 1, 1], GGMLType.Q8_0) == [34, 34 * 64, 34 * 64 * 4, 34 * 64 * 4]


def test_allocator_offsets_are_16_byte_aligned():
    ga = lk.GGMLGraphAllocator(size=1 << 16)
    a = ga.allocateTensor(GGMLTensor(GGMLType.Q4_0, [64, 3]))     # 108 bytes
    b = ga.allocateTensor(GGMLTensor(GGMLType.F32, [3, 64]))
    assert a.dataOffset == 0 and b.dataOffset == 112 and b.dataOffset % 16 == 0
    ga.setBytes(b, np.arange(192, dtype=np.float32))
    assert np.array_equal(ga.getFloats(b), np.arange(192, dtype=np.float32))


def test_tensor_descriptor_uses_buffer_and_offset_not_data():
    # the boundary takes (buffer, dataOffset), never tensor.data
    from llama_kotlin_b200.ggml_compute_ops import describe
    ga = lk.GGMLGraphAllocator(size=1 << 12)
    t = ga.allocateTensor(GGMLTensor(GGMLType.F32, [4, 4]))
    d = describe(ga, t)
    assert d.host_base == ga.buffers[0].ctypes.data and d.data_offset == t.dataOffset and not d.dev_buf
    orphan = GGMLTensor(GGMLType.F32, [4, 4], data=np.zeros(16, np.float32))   # bufferId = -1
    with pytest.raises(lk.IllegalStateException):
        describe(ga, orphan)


@pytest.mark.skipif(_has_gpu(), reason="checks the no-GPU failure mode")
def test_no_cpu_fallback_ctx_create_fails_without_gpu():
    h = C.c_void_p()
    st = B.lib().lk_ctx_create(0, C.byref(h))
    assert st == B.LK_ECUDA and not h.value
    assert b"no CPU fallback" in B.lib().lk_last_error(None)
    ga = lk.GGMLGraphAllocator(size=1 << 12)
    a = ga.allocateTensor(GGMLTensor(GGMLType.F32, [2, 2]))
    with pytest.raises(lk.IllegalStateException):
        lk.computeAdd(ga, None, a, a, a)


def test_null_arguments_are_rejected_not_crashed():
    L = B.lib()
    assert L.lk_sync(None) == B.LK_EINVAL_ARG
    assert L.lk_compute_matmul(None, None, None, None) == B.LK_EINVAL_ARG
    assert L.lk_buf_upload(None, 0, None, 0) == B.LK_EINVAL_ARG
    assert L.lk_capture_begin(None) == B.LK_EINVAL_ARG
    assert L.lk_ctx_destroy(None) == B.LK_OK
    assert L.lk_get_option(None, b"x") == -1


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "llama.kotlin_b200")
    for dirpath, _, files in os.walk(pkg):
        for fn in files:
            if fn.endswith((".py", ".cu", ".cuh", ".h")):
                with open(os.path.join(dirpath, fn)) as f:
                    src = f.read()
                assert "lk_oracle" not in src and "from oracle" not in src and "import oracle" not in src, fn
