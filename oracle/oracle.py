"""ctypes binding for the C oracle (oracle/liblk_oracle.so).

TEST INFRASTRUCTURE ONLY.  Importable from tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs — never from the product
package (llama.kotlin_b200/), which has no CPU fallback.

Parity status: pinned by the reference's golden vectors (tests/golden,
tests/test_oracle_golden.py); see lk_oracle.h.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "liblk_oracle.so")

F32, F16, Q4_0, Q4_1, Q8_0 = 0, 1, 2, 3, 6
Q2_K, Q4_K, Q8_K, BITNET_1_58 = 8, 10, 13, 40
BLOCK_BYTES = {Q4_0: 18, Q4_1: 20, Q8_0: 34}
KQ_BLOCK_BYTES = {Q2_K: 84, Q4_K: 144, Q8_K: 292, BITNET_1_58: 10}
KQ_BLOCK_ELEMS = {Q2_K: 256, Q4_K: 256, Q8_K: 256, BITNET_1_58: 32}
OK, EINVAL_DIM, EINVAL_TYPE, EBOUNDS, ENOTIMPL = 0, 1, 2, 3, 4


def build(force: bool = False) -> str:
    """Compile the oracle with gcc (oracle/Makefile).  Building the checker is
    not using it; __graft_entry__.build() calls this."""
    srcs = [os.path.join(_HERE, f) for f in ("lk_oracle.c", "lk_oracle_kq.c", "lk_oracle.h", "Makefile")]
    stale = not os.path.exists(_LIB_PATH) or any(os.path.getmtime(_LIB_PATH) < os.path.getmtime(f) for f in srcs)
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "-B", "liblk_oracle.so"],
                              stdout=subprocess.DEVNULL)
    return _LIB_PATH


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(_LIB_PATH):
            build()
        L = C.CDLL(_LIB_PATH)
        vp, i64, f32, i32 = C.c_void_p, C.c_int64, C.c_float, C.c_int
        L.lko_half_to_float.restype = f32
        L.lko_half_to_float.argtypes = [C.c_uint16]
        L.lko_float_to_half.restype = C.c_uint16
        L.lko_float_to_half.argtypes = [f32]
        L.lko_block_bytes.argtypes = [i32]
        L.lko_quantize.argtypes = [i32, vp, i64, vp]
        L.lko_dequantize.argtypes = [i32, vp, i64, vp]
        L.lko_repack_nibbles.argtypes = [i32, vp, i64, vp, i32]
        L.lko_dequantize_ggml.argtypes = [i32, vp, i64, vp]
        L.lko_dot.restype = f32
        L.lko_dot.argtypes = [i32, vp, i64, i64, vp, i64, i64]
        L.lko_matmul.argtypes = [i32, vp, i64, i64, vp, i64, vp]
        L.lko_matmul_mt.argtypes = [i32, vp, i64, i64, vp, i64, vp, i32]
        L.lko_matmul_rows.argtypes = [i32, vp, i64, i64, vp, i64, vp, i64, i64, i32]
        L.lko_rms_norm.restype = None
        L.lko_rms_norm.argtypes = [vp, vp, f32, i64, i64, vp]
        for name in ("lko_silu",):
            getattr(L, name).restype = None
            getattr(L, name).argtypes = [vp, i64, vp]
        for name in ("lko_silu_mul", "lko_add", "lko_mul"):
            getattr(L, name).restype = None
            getattr(L, name).argtypes = [vp, vp, i64, vp]
        L.lko_softmax.restype = None
        L.lko_softmax.argtypes = [vp, i64, i64, vp]
        L.lko_rope.restype = None
        L.lko_rope.argtypes = [vp, i32, i32, i32, i32, f32, vp]
        L.lko_scale.restype = None
        L.lko_scale.argtypes = [vp, i64, f32, vp]
        L.lko_attention_decode.restype = None
        L.lko_attention_decode.argtypes = [vp, vp, vp, i64, i32, i32, vp]
        L.lko_kq_block_bytes.argtypes = [i32]
        L.lko_kq_block_elems.argtypes = [i32]
        L.lko_kq_quantize.argtypes = [i32, vp, i64, vp]
        L.lko_kq_dequantize.argtypes = [i32, vp, i64, vp]
        L.lko_kq_dot.restype = f32
        L.lko_kq_dot.argtypes = [i32, vp, i64, i64, vp, i64, i64]
        L.lko_kq_matmul.argtypes = [i32, vp, i64, i64, vp, i64, vp]
        _lib = L
    return _lib


def _p(a: np.ndarray) -> C.c_void_p:
    return C.c_void_p(a.ctypes.data)


def _f32(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float32)


def half_to_float(h: int) -> float:
    return float(lib().lko_half_to_float(int(h) & 0xFFFF))


def float_to_half(f: float) -> int:
    return int(lib().lko_float_to_half(float(np.float32(f))))


def quantize(type_: int, x) -> np.ndarray:
    x = _f32(x).reshape(-1)
    out = np.zeros(x.size // 32 * BLOCK_BYTES[type_], dtype=np.uint8)
    rc = lib().lko_quantize(type_, _p(x), x.size, _p(out))
    if rc != OK:
        raise ValueError(f"lko_quantize rc={rc}")
    return out


def dequantize(type_: int, blocks: np.ndarray, n: int) -> np.ndarray:
    blocks = np.ascontiguousarray(blocks, dtype=np.uint8)
    out = np.empty(n, dtype=np.float32)
    rc = lib().lko_dequantize(type_, _p(blocks), n, _p(out))
    if rc != OK:
        raise ValueError(f"lko_dequantize rc={rc}")
    return out


def repack_nibbles(type_: int, blocks: np.ndarray, to_ggml: bool) -> np.ndarray:
    """Kotlin accessor order <-> canonical ggml / GGUF order of the Q4 nibbles."""
    blocks = np.ascontiguousarray(blocks, dtype=np.uint8).reshape(-1)
    out = np.empty_like(blocks)
    rc = lib().lko_repack_nibbles(type_, _p(blocks), blocks.size // BLOCK_BYTES[type_], _p(out), int(bool(to_ggml)))
    if rc != OK:
        raise ValueError(f"lko_repack_nibbles rc={rc}")
    return out


def dequantize_ggml(type_: int, blocks: np.ndarray, n: int) -> np.ndarray:
    blocks = np.ascontiguousarray(blocks, dtype=np.uint8)
    out = np.empty(n, dtype=np.float32)
    rc = lib().lko_dequantize_ggml(type_, _p(blocks), n, _p(out))
    if rc != OK:
        raise ValueError(f"lko_dequantize_ggml rc={rc}")
    return out


def matmul(a_type: int, a: np.ndarray, K: int, M: int, b, N: int, nthreads: int = 1,
           rows: tuple[int, int] | None = None) -> np.ndarray:
    """dst (flat, ne=[N,M], element (j,i) at j + i*N).  b is flat ne=[N,K]."""
    a = np.ascontiguousarray(a)
    b = _f32(b).reshape(-1)
    assert b.size == N * K
    dst = np.zeros(N * M, dtype=np.float32)
    r0, r1 = rows if rows is not None else (0, M)
    rc = lib().lko_matmul_rows(a_type, _p(a), K, M, _p(b), N, _p(dst), r0, r1, nthreads)
    if rc != OK:
        raise ValueError(f"lko_matmul rc={rc}")
    return dst


def dot(a_type: int, a: np.ndarray, K: int, i: int, b, N: int, j: int) -> float:
    a = np.ascontiguousarray(a)
    b = _f32(b).reshape(-1)
    return float(lib().lko_dot(a_type, _p(a), K, i, _p(b), N, j))


def rms_norm(x, weight, eps: float, ne0: int) -> np.ndarray:
    x = _f32(x).reshape(-1)
    out = np.empty_like(x)
    w = _f32(weight).reshape(-1) if weight is not None else None
    lib().lko_rms_norm(_p(x), _p(w) if w is not None else None, float(eps), ne0,
                       x.size // max(ne0, 1), _p(out))
    return out


def silu(x) -> np.ndarray:
    x = _f32(x).reshape(-1)
    out = np.empty_like(x)
    lib().lko_silu(_p(x), x.size, _p(out))
    return out


def _binary(name, a, b):
    a = _f32(a).reshape(-1)
    b = _f32(b).reshape(-1)
    assert a.size == b.size
    out = np.empty_like(a)
    getattr(lib(), name)(_p(a), _p(b), a.size, _p(out))
    return out


def silu_mul(gate, up) -> np.ndarray:
    return _binary("lko_silu_mul", gate, up)


def add(a, b) -> np.ndarray:
    return _binary("lko_add", a, b)


def mul(a, b) -> np.ndarray:
    return _binary("lko_mul", a, b)


def softmax(x, ne0: int) -> np.ndarray:
    x = _f32(x).reshape(-1)
    out = np.empty_like(x)
    lib().lko_softmax(_p(x), ne0, x.size // max(ne0, 1), _p(out))
    return out


def rope(x, head_dim: int, nheads: int, position: int, theta: float = 10000.0) -> np.ndarray:
    x = _f32(x).reshape(-1)
    out = np.empty_like(x)
    lib().lko_rope(_p(x), head_dim, nheads, x.size // (head_dim * nheads), position, float(theta), _p(out))
    return out


def scale(x, factor: float) -> np.ndarray:
    x = _f32(x).reshape(-1)
    out = np.empty_like(x)
    lib().lko_scale(_p(x), x.size, float(factor), _p(out))
    return out


def attention_decode(q, kc, vc, length: int, head_dim: int, nheads: int) -> np.ndarray:
    q, kc, vc = _f32(q).reshape(-1), _f32(kc).reshape(-1), _f32(vc).reshape(-1)
    out = np.empty(head_dim * nheads, dtype=np.float32)
    lib().lko_attention_decode(_p(q), _p(kc), _p(vc), length, head_dim, nheads, _p(out))
    return out


# ---- N4: the reference's K-quant / BitNet 1.58 formats (oracle/lk_oracle_kq.c) ----
def kq_quantize(type_: int, x) -> np.ndarray:
    x = _f32(x).reshape(-1)
    out = np.zeros(x.size // KQ_BLOCK_ELEMS[type_] * KQ_BLOCK_BYTES[type_], dtype=np.uint8)
    rc = lib().lko_kq_quantize(type_, _p(x), x.size, _p(out))
    if rc != OK:
        raise ValueError(f"lko_kq_quantize rc={rc}")
    return out


def kq_dequantize(type_: int, blocks: np.ndarray, n: int) -> np.ndarray:
    blocks = np.ascontiguousarray(blocks, dtype=np.uint8)
    out = np.empty(n, dtype=np.float32)
    rc = lib().lko_kq_dequantize(type_, _p(blocks), n, _p(out))
    if rc != OK:
        raise ValueError(f"lko_kq_dequantize rc={rc}")
    return out


def kq_matmul(type_: int, a: np.ndarray, K: int, M: int, b, N: int) -> np.ndarray:
    a = np.ascontiguousarray(a, dtype=np.uint8)
    b = _f32(b).reshape(-1)
    assert b.size == N * K
    dst = np.zeros(N * M, dtype=np.float32)
    rc = lib().lko_kq_matmul(type_, _p(a), K, M, _p(b), N, _p(dst))
    if rc != OK:
        raise ValueError(f"lko_kq_matmul rc={rc}")
    return dst
