"""lk_compute_attention_decode_noalloc — the entry point that allocates nothing (scratch memory in a file-scope static
device array, include/llamakotlin_cuda.h; no counterpart in the reference).
Checked through the C ABI against the existing lk_compute_attention_decode on the same device-resident tensors, on two
calls in a row (the second finds the first call's contents in the shared scratch array), within the attention tests'
tolerance (1e-5 of max |out|; the kernel is the same, so the results are expected to be bit-identical).

This file is named to be collected LAST: a failure here must not hide other tests.
It must pass both alone in a fresh process and after the rest of the suite."""
import ctypes as C

import numpy as np
import pytest

import llama_kotlin_b200 as lk
from llama_kotlin_b200 import GGMLTensor, GGMLType
from llama_kotlin_b200 import _binding as B
from llama_kotlin_b200.ggml_compute_ops import describe

pytestmark = pytest.mark.gpu
ATTN_TOL = 1e-5          # tests/test_gpu_attention.py


def rel_err(got, ref):
    den = max(float(np.max(np.abs(ref))), 1e-30)
    return float(np.max(np.abs(np.asarray(got, np.float64) - np.asarray(ref, np.float64)))) / den


def _noalloc(be, ga, q, kc, vc, pos, scale, out):
    return B.lib().lk_compute_attention_decode_noalloc(be.lk.handle, describe(ga, q), describe(ga, kc), describe(ga, vc),
                                                       describe(ga, pos), C.c_float(scale), describe(ga, out))


@pytest.mark.parametrize("H,D,maxseq,lengths", [(32, 128, 2048, (1500, 77)), (8, 64, 300, (300, 33))])
def test_noalloc_attention_matches_the_existing_entry_point(H, D, maxseq, lengths):
    be = lk.GGMLCudaBackend(0)
    ga = lk.GGMLGraphAllocator(be, size=2 * maxseq * H * D * 4 + (4 << 20))
    rng = np.random.default_rng(1234 + H)
    kc = ga.allocateTensor(GGMLTensor(GGMLType.F32, [D, H, maxseq]))
    vc = ga.allocateTensor(GGMLTensor(GGMLType.F32, [D, H, maxseq]))
    ga.setBytes(kc, rng.normal(0, 1, maxseq * H * D).astype(np.float32))
    ga.setBytes(vc, rng.normal(0, 1, maxseq * H * D).astype(np.float32))
    q = ga.allocateTensor(GGMLTensor(GGMLType.F32, [H * D]))
    pos = ga.allocateTensor(GGMLTensor(GGMLType.I32, [4]))
    ref_out = ga.allocateTensor(GGMLTensor(GGMLType.F32, [H * D]))
    out = ga.allocateTensor(GGMLTensor(GGMLType.F32, [H * D]))
    scale = 1.0 / float(np.sqrt(np.float32(D)))
    assert B.LK_OK == 0
    for length in lengths:                      # two calls in a row, different query and context length
        ga.setBytes(q, (rng.normal(0, 1, H * D) * 2.0).astype(np.float32))
        ga.setBytes(pos, np.array([length - 1, 0, 0, 0], np.int32))
        ga.setBytes(out, np.full(H * D, np.nan, np.float32))
        lk.computeAttentionDecode(ga, None, q, kc, vc, pos, scale, ref_out)
        want = ga.getFloats(ref_out).copy()
        st = _noalloc(be, ga, q, kc, vc, pos, scale, out)
        assert st == B.LK_OK, B.lib().lk_last_error(be.lk.handle)
        be.lk.sync()
        got = ga.getFloats(out)
        assert np.all(np.isfinite(got)) and float(np.max(np.abs(want))) > 0
        assert rel_err(got, want) <= ATTN_TOL, (length, rel_err(got, want))
    be.free()


def test_noalloc_attention_rejects_what_it_cannot_serve():
    be = lk.GGMLCudaBackend(0)
    ga = lk.GGMLGraphAllocator(be, size=8 << 20)
    D, H, maxseq = 128, 8, 64
    kc = ga.allocateTensor(GGMLTensor(GGMLType.F32, [D, H, maxseq]))
    vc = ga.allocateTensor(GGMLTensor(GGMLType.F32, [D, H, maxseq]))
    q = ga.allocateTensor(GGMLTensor(GGMLType.F32, [H * D]))
    out = ga.allocateTensor(GGMLTensor(GGMLType.F32, [H * D]))
    pos = ga.allocateTensor(GGMLTensor(GGMLType.I32, [4]))
    # a host-resident tensor would need the staging arena: refused, nothing launched
    host = np.zeros(H * D, np.float32)
    hq = B.lk_tensor()
    hq.type = int(GGMLType.F32)
    for i, v in enumerate([H * D, 1, 1, 1]):
        hq.ne[i] = v
    hq.host_base = host.ctypes.data
    hq.buffer_size = host.nbytes
    st = B.lib().lk_compute_attention_decode_noalloc(be.lk.handle, hq, describe(ga, kc), describe(ga, vc), describe(ga, pos),
                                                     C.c_float(0.1), describe(ga, out))
    assert st == B.LK_EINVAL_ARG and b"device-resident" in B.lib().lk_last_error(be.lk.handle)
    # a shape whose scratch need exceeds the static array (numHeads * 4 bytes of tickets alone pass 512 KiB): clear error
    Hbig = 140000
    big = B.lk_tensor()
    big.type = int(GGMLType.F32)
    for i, v in enumerate([64, Hbig, 1, 1]):
        big.ne[i] = v
    big.dev_buf = ga.deviceBuffer.handle
    big.buffer_size = ga.deviceBuffer.getSize()
    vec = B.lk_tensor()
    vec.type = int(GGMLType.F32)
    for i, v in enumerate([64 * Hbig, 1, 1, 1]):
        vec.ne[i] = v
    vec.dev_buf = ga.deviceBuffer.handle
    vec.buffer_size = ga.deviceBuffer.getSize()
    ga2 = lk.GGMLGraphAllocator(be, size=64 * Hbig * 4 * 4 + (1 << 20))
    for t in (big, vec):
        t.dev_buf = ga2.deviceBuffer.handle
        t.buffer_size = ga2.deviceBuffer.getSize()
    p2 = ga2.allocateTensor(GGMLTensor(GGMLType.I32, [4]))
    st = B.lib().lk_compute_attention_decode_noalloc(be.lk.handle, vec, big, big, describe(ga2, p2), C.c_float(0.1), vec)
    msg = B.lib().lk_last_error(be.lk.handle)
    assert st == B.LK_EINVAL_ARG and b"scratch" in msg and b"524288" in msg, msg
    be.free()
