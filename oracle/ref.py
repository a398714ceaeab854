"""ctypes binding for oracle/_ref/libggml_ref.so — the reference tree's OWN ggml C sources
(/root/reference/ggml/src/*.c), compiled where they lie by `make -C oracle ref` together with
oracle/ref_harness.c.

TEST INFRASTRUCTURE ONLY (same rule as oracle.py).  The library is built in the authoring container
(where /root/reference is mounted) and travels to the GPU box as a prebuilt file; nothing here reads
/root/reference at run time.  `available()` is False where the file is missing: callers skip.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "_ref", "libggml_ref.so")
_REF_TREE = "/root/reference/ggml/src"

Q4_0, Q4_1, Q8_0 = 2, 3, 6
BLOCK_BYTES = {Q4_0: 18, Q4_1: 20, Q8_0: 34}
OP_RMS_NORM, OP_SILU, OP_SOFT_MAX, OP_ROPE = 0, 1, 2, 3


def build(force: bool = False) -> str | None:
    """Compile the reference's ggml C sources + our harness (only where the reference tree is mounted)."""
    if not os.path.isdir(_REF_TREE):
        return _LIB_PATH if os.path.exists(_LIB_PATH) else None
    src = os.path.join(_HERE, "ref_harness.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-B", "ref"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


def available() -> bool:
    return os.path.exists(_LIB_PATH)


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        L = C.CDLL(_LIB_PATH)
        vp, i64, f32, i32 = C.c_void_p, C.c_int64, C.c_float, C.c_int
        L.ref_init.restype = None
        L.ref_fp16_to_fp32.restype = f32
        L.ref_fp16_to_fp32.argtypes = [C.c_uint16]
        L.ref_fp32_to_fp16.restype = C.c_uint16
        L.ref_fp32_to_fp16.argtypes = [f32]
        L.ref_dequantize.argtypes = [i32, vp, i64, vp]
        L.ref_quantize.argtypes = [i32, vp, i64, vp]
        L.ref_unary.argtypes = [i32, vp, i64, i64, f32, f32, vp, i32]
        L.ref_mul_mat.argtypes = [i32, vp, i64, i64, vp, i64, vp, i32, i32, C.POINTER(C.c_double)]
        L.ref_gguf_write.argtypes = [C.c_char_p, i32, i32, C.POINTER(C.c_char_p), C.POINTER(i32), C.POINTER(i64),
                                     C.POINTER(i64), C.POINTER(vp)]
        L.ref_init()
        _lib = L
    return _lib


def _p(a: np.ndarray) -> C.c_void_p:
    return C.c_void_p(a.ctypes.data)


def fp16_to_fp32(h: int) -> float:
    return float(lib().ref_fp16_to_fp32(int(h) & 0xFFFF))


def fp32_to_fp16(f: float) -> int:
    return int(lib().ref_fp32_to_fp16(float(np.float32(f))))


def dequantize(type_: int, blocks: np.ndarray, n: int) -> np.ndarray:
    """dequantize_row_q4_0 / q4_1 / q8_0 (ggml nibble order)."""
    blocks = np.ascontiguousarray(blocks, dtype=np.uint8)
    out = np.empty(n, dtype=np.float32)
    assert lib().ref_dequantize(type_, _p(blocks), n, _p(out)) == 0
    return out


def quantize(type_: int, x) -> np.ndarray:
    """quantize_row_q*_ref (ggml's quantizers; Q4_0 differs from the Kotlin one)."""
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1)
    out = np.zeros(x.size // 32 * BLOCK_BYTES[type_], dtype=np.uint8)
    assert lib().ref_quantize(type_, _p(x), x.size, _p(out)) == 0
    return out


def unary(op: int, x, ne0: int, p0: float = 0.0, p1: float = 0.0, nthreads: int = 1) -> np.ndarray:
    x = np.ascontiguousarray(x, dtype=np.float32).reshape(-1)
    out = np.empty_like(x)
    assert lib().ref_unary(op, _p(x), ne0, x.size // ne0, p0, p1, _p(out), nthreads) == 0
    return out


def mul_mat(a_type: int, a: np.ndarray, K: int, M: int, b, N: int, nthreads: int = 1, reps: int = 1):
    """ggml_mul_mat(a[K,M], b[K,N]) -> (dst[N][M] row-major i.e. M fastest, seconds per compute)."""
    a = np.ascontiguousarray(a)
    b = np.ascontiguousarray(b, dtype=np.float32).reshape(-1)
    assert b.size == K * N
    dst = np.zeros(M * N, dtype=np.float32)
    sec = C.c_double(0.0)
    rc = lib().ref_mul_mat(a_type, _p(a), K, M, _p(b), N, _p(dst), nthreads, reps, C.byref(sec))
    if rc != 0:
        raise RuntimeError(f"ref_mul_mat rc={rc}")
    return dst, sec.value


def gguf_write(path: str, tensors, alignment: int = 32) -> None:
    """Writes a GGUF file with the reference tree's own writer (gguf_write_to_file in ggml.c).
    tensors: list of (name, type id (0 F32, 2 Q4_0, 3 Q4_1, 6 Q8_0), ne0, ne1, payload uint8/float32 array)."""
    n = len(tensors)
    keep = [np.ascontiguousarray(t[4]) for t in tensors]
    names = (C.c_char_p * n)(*[t[0].encode() for t in tensors])
    types = (C.c_int * n)(*[int(t[1]) for t in tensors])
    ne0 = (C.c_int64 * n)(*[int(t[2]) for t in tensors])
    ne1 = (C.c_int64 * n)(*[int(t[3]) for t in tensors])
    ptrs = (C.c_void_p * n)(*[k.ctypes.data for k in keep])
    rc = lib().ref_gguf_write(path.encode(), alignment, n, names, types, ne0, ne1, ptrs)
    if rc != 0:
        raise RuntimeError(f"ref_gguf_write rc={rc}")
