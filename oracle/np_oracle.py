"""Second, independent restatement of the reference path in numpy / pure Python.

TEST INFRASTRUCTURE ONLY.  Exists to cross-check oracle/lk_oracle.c (a typo in
one restatement is unlikely to be repeated in the other).  Slow; small cases.

Reference paths relative to
/root/reference/src/nativeMain/kotlin/ai/solace/llamakotlin/.
"""
from __future__ import annotations

import struct

import numpy as np

F32, Q4_0, Q4_1, Q8_0 = 0, 2, 3, 6
BLOCK_BYTES = {Q4_0: 18, Q4_1: 20, Q8_0: 34}


def _i32(x: int) -> int:
    x &= 0xFFFFFFFF
    return x - (1 << 32) if x & 0x80000000 else x


def _shl(x: int, s: int) -> int:           # Kotlin Int shl: count masked to 5 bits, wraps
    return _i32(x << (s & 31))


def _ushr(x: int, s: int) -> int:          # Kotlin Int ushr
    return _i32((x & 0xFFFFFFFF) >> (s & 31))


def half_to_float(h: int) -> float:
    """core/NumericConversions.kt:9-54 (IEEE-correct, so numpy is a valid restatement
    for non-NaN; NaNs get the quiet bit set at :44)."""
    h &= 0xFFFF
    exp, mant = (h >> 10) & 0x1F, h & 0x3FF
    if exp == 0x1F and mant != 0:
        bits = ((h >> 15) << 31) | 0x7F800000 | (mant << 13) | (1 << 22)
        return struct.unpack("<f", struct.pack("<I", bits))[0]
    return float(np.array([h], dtype=np.uint16).view(np.float16)[0])


def float_to_half(f: float) -> int:
    """core/NumericConversions.kt:61-124, line by line, with Kotlin Int semantics."""
    f32bits = _i32(struct.unpack("<I", struct.pack("<f", float(np.float32(f))))[0])
    fSign = _ushr(f32bits, 16) & 0x8000
    absF = f32bits & 0x7FFFFFFF
    if absF > 0x47FFEFFF:
        return (fSign | 0x7C00 | (0x0200 if (absF & 0x007FFFFF) != 0 else 0)) & 0xFFFF
    if absF < 0x38800000:
        fMant = (absF & 0x007FFFFF) | 0x00800000
        shift = 127 - _ushr(absF, 23)
        hMant = _ushr(fMant, shift) if shift < 24 else 0
        roundBits = fMant & _i32(_shl(1, shift) - 1)
        halfway = _shl(1, shift - 1)
        if roundBits > halfway or (roundBits == halfway and (hMant & 1) != 0):
            t = hMant + 1
            if t == 0x0400:
                return (fSign | (1 << 10)) & 0xFFFF
            return (fSign | t) & 0xFFFF
        return (fSign | hMant) & 0xFFFF
    hExp = (_ushr(absF, 23) - 112) << 10
    hMant = _ushr(absF & 0x007FFFFF, 13)
    if (absF & 0x00001000) != 0:
        if (absF & 0x00000FFF) != 0 or (hMant & 1) != 0:
            hMant += 1
            if hMant == 0x0400:
                return (fSign | (hExp + (1 << 10))) & 0xFFFF
    return (fSign | hExp | hMant) & 0xFFFF


def _round_clamp(v: np.ndarray, lo: int, hi: int) -> np.ndarray:
    r = np.rint(v.astype(np.float32))          # ties-to-even == kotlin.math.round
    r = np.where(np.isnan(r), 0.0, r)
    return np.clip(r, lo, hi).astype(np.int32)


def quantize(type_: int, x) -> np.ndarray:
    """core/GGMLComputeOps.kt:1058-1087, :1162-1203 — vectorised per block; scales go
    through the scalar float_to_half above."""
    x = np.asarray(x, dtype=np.float32).reshape(-1, 32)
    nb = x.shape[0]
    out = np.zeros((nb, BLOCK_BYTES[type_]), dtype=np.uint8)
    one = np.float32(1.0)
    if type_ == Q8_0:
        amax = np.abs(x).max(axis=1)
        scale = np.where(amax == 0, one, amax / np.float32(127.0)).astype(np.float32)
        inv = (one / scale).astype(np.float32)
        q = _round_clamp(x * inv[:, None], -128, 127)
        out[:, 2:] = q.astype(np.int8).view(np.uint8)
        hs = [float_to_half(s) for s in scale]
    elif type_ == Q4_0:
        amax = np.abs(x).max(axis=1)
        scale = np.where(amax == 0, one, amax / np.float32(8.0)).astype(np.float32)
        inv = (one / scale).astype(np.float32)
        q = _round_clamp((x * inv[:, None]).astype(np.float32) + np.float32(8.0), 0, 15)
        out[:, 2:] = (q[:, 0::2] | (q[:, 1::2] << 4)).astype(np.uint8)
        hs = [float_to_half(s) for s in scale]
    elif type_ == Q4_1:
        mn, mx = x.min(axis=1), x.max(axis=1)
        d = ((mx - mn).astype(np.float32) / np.float32(15.0)).astype(np.float32)
        d = np.where(d == 0, one, d).astype(np.float32)
        inv = (one / d).astype(np.float32)
        q = _round_clamp(((x - mn[:, None]).astype(np.float32) * inv[:, None]).astype(np.float32), 0, 15)
        out[:, 4:] = (q[:, 0::2] | (q[:, 1::2] << 4)).astype(np.uint8)
        hs = [float_to_half(s) for s in d]
        hm = np.array([float_to_half(s) for s in mn], dtype=np.uint16)
        out[:, 2] = hm & 0xFF
        out[:, 3] = hm >> 8
    else:
        raise NotImplementedError(type_)
    hs = np.array(hs, dtype=np.uint16)
    out[:, 0] = hs & 0xFF
    out[:, 1] = hs >> 8
    return out.reshape(-1)


def _halves(lo: np.ndarray, hi: np.ndarray) -> np.ndarray:
    bits = lo.astype(np.uint16) | (hi.astype(np.uint16) << 8)
    return bits.view(np.float16).astype(np.float32)


def dequantize(type_: int, blocks, n: int) -> np.ndarray:
    """core/GGMLComputeOps.kt:932-964 with the Kotlin nibble order
    (core/GGMLTypes.kt:647-651): element e -> (qs[e//2] >> 4*(e&1)) & 15."""
    b = np.asarray(blocks, dtype=np.uint8).reshape(-1, BLOCK_BYTES[type_])
    d = _halves(b[:, 0].copy(), b[:, 1].copy())
    if type_ == Q8_0:
        q = b[:, 2:].view(np.int8).astype(np.float32)
        y = d[:, None] * q
    else:
        off = 2 if type_ == Q4_0 else 4
        qs = b[:, off:]
        nib = np.empty((b.shape[0], 32), dtype=np.float32)
        nib[:, 0::2] = (qs & 0x0F).astype(np.float32)
        nib[:, 1::2] = (qs >> 4).astype(np.float32)
        if type_ == Q4_0:
            y = d[:, None] * (nib - np.float32(8.0))
        else:
            m = _halves(b[:, 2].copy(), b[:, 3].copy())
            y = (d[:, None] * nib).astype(np.float32) + m[:, None]
    return y.astype(np.float32).reshape(-1)[:n]


def matmul(a_type: int, a, K: int, M: int, b, N: int) -> np.ndarray:
    """core/GGMLComputeOps.kt:1435-1565: dst[j + i*N] = sum_k deq(a)[i,k] * b[j + k*N],
    F32, sequential in k (vectorised over (i, j), looped over k)."""
    if a_type == F32:
        w = np.asarray(a, dtype=np.float32).reshape(M, K)
    else:
        w = dequantize(a_type, a, M * K).reshape(M, K)
    bm = np.asarray(b, dtype=np.float32).reshape(K, N)
    acc = np.zeros((M, N), dtype=np.float32)
    for k in range(K):
        acc = (acc + (w[:, k:k + 1] * bm[k:k + 1, :]).astype(np.float32)).astype(np.float32)
    return acc.reshape(-1)


def matmul_f64(a_type: int, a, K: int, M: int, b, N: int) -> np.ndarray:
    """Exact-ish (f64) value of the same product; used to report which F32 summation
    order is closer to the truth, never as the parity target."""
    w = (np.asarray(a, dtype=np.float32).reshape(M, K) if a_type == F32
         else dequantize(a_type, a, M * K).reshape(M, K)).astype(np.float64)
    return (w @ np.asarray(b, dtype=np.float32).reshape(K, N).astype(np.float64)).reshape(-1)
