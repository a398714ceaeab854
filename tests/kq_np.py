"""Second, independent restatement (vectorised numpy, F32 operations in the reference's order) of the reference's
K-quant / BitNet 1.58 dequantisation — core/GGMLComputeOps.kt:2125-2155 (Q2_K), :2207-2250 (Q4_K), :2356-2368 (Q8_K),
:1021-1032 + core/GGMLTypes.kt:953-1000 (BitNet 1.58).  Test infrastructure: checks oracle/lk_oracle_kq.c bit for bit."""
import numpy as np

np.seterr(all="ignore")      # random bytes hold inf / NaN scales on purpose
f32 = np.float32


def _h2f(b2: np.ndarray) -> np.ndarray:           # [nb, 2] uint8 little-endian half -> f32 (exact)
    return np.ascontiguousarray(b2).view("<f2").astype(np.float32).reshape(-1)


def dequant_q2_k(raw: np.ndarray) -> np.ndarray:
    blk = np.ascontiguousarray(raw, np.uint8).reshape(-1, 84)
    d, dmin = _h2f(blk[:, 80:82])[:, None], _h2f(blk[:, 82:84])[:, None]
    sm = blk[:, :16]
    scale = ((sm & 15).astype(f32) / f32(15.0)) * d                 # [nb, 16]
    mn = (sm >> 4).astype(f32) * d + dmin
    qs = blk[:, 16:80]                                              # byte b: elements 4b..4b+3, 2 bits each
    q = ((qs[:, :, None] >> (np.arange(4, dtype=np.uint8) * 2)) & 3).reshape(-1, 256).astype(f32)
    return ((q / f32(3.0)) * np.repeat(scale, 16, axis=1) + np.repeat(mn, 16, axis=1)).reshape(-1).astype(f32)


def dequant_q4_k(raw: np.ndarray, partial: bool = False) -> np.ndarray:
    blk = np.ascontiguousarray(raw, np.uint8).reshape(-1, 144)
    d, dmin = _h2f(blk[:, 0:2])[:, None], _h2f(blk[:, 2:4])[:, None]
    sc = blk[:, 4:16].astype(np.int32)
    sub = np.arange(8)
    sb = sc[:, sub]
    hi = np.where(sub[None, :] <= 5, sc[:, np.minimum(sub * 2 + 1, 11)] & 15, 0)
    qmin = (sb >> 6) | (hi << 2)
    scale = ((sb & 63).astype(f32) / f32(63.0)) * d
    mn = (qmin.astype(f32) / f32(63.0)) * d + dmin
    qs = blk[:, 16:144]
    q = np.stack([qs & 15, qs >> 4], axis=2).reshape(-1, 256).astype(f32)   # byte j: elements 2j (lo), 2j+1 (hi)
    off = np.broadcast_to(dmin, (blk.shape[0], 256)) if partial else np.repeat(mn, 32, axis=1)
    return ((q / f32(15.0)) * np.repeat(scale, 32, axis=1) + off).reshape(-1).astype(f32)


def dequant_q8_k(raw: np.ndarray) -> np.ndarray:
    blk = np.ascontiguousarray(raw, np.uint8).reshape(-1, 292)
    d = np.ascontiguousarray(blk[:, 0:4]).view("<f4").reshape(-1, 1)
    return (blk[:, 4:260].view(np.int8).astype(f32) * d).reshape(-1).astype(f32)


def dequant_bitnet(raw: np.ndarray) -> np.ndarray:
    blk = np.ascontiguousarray(raw, np.uint8).reshape(-1, 10)
    d = _h2f(blk[:, 0:2])[:, None]
    by = blk[:, 2:9].astype(np.int32)                               # 7 bytes x 5 digits, least significant first
    pw = np.array([1, 3, 9, 27, 81])
    dig = (by[:, :, None] // pw) % np.array([3, 3, 3, 3, 256])      # the leading digit is not reduced (3 = invalid)
    dig = dig.reshape(-1, 35)[:, :32]
    out = d * (dig - 1).astype(f32)
    return np.where(dig > 2, f32(np.nan), out).reshape(-1).astype(f32)


DEQUANT = {8: dequant_q2_k, 10: dequant_q4_k, 13: dequant_q8_k, 40: dequant_bitnet}
