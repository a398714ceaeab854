"""N4: the oracle's restatement of the reference's K-quant / BitNet 1.58 formats (oracle/lk_oracle_kq.c)
against (1) a second, independent numpy restatement (tests/kq_np.py), bit for bit, (2) hand-computed block KATs and
(3) what the reference's own tests assert for these formats (nativeTest core/GGMLKQuantAccuracyTest.kt,
GGMLBitNet158Test.kt).  Note (recorded in DESIGN.md): a literal restatement of the reference's Q2_K and Q4_K quantizer +
dequantizer pair does NOT meet the MSE thresholds of testQ2_KAccuracy / testQ4_KAccuracy (:150-158, :190-197) — the
dequantizers scale by q/3 resp. q/15 what the quantizers already divided by 3 resp. 15, and the Q4_K scale bytes of
different sub-blocks overlap (:1907-1912) — so those thresholds cannot pin anything; Q8_K's and BitNet's do."""
import numpy as np
import pytest

import kq_np

Q2_K, Q4_K, Q8_K, BITNET = 8, 10, 13, 40
TYPES = [Q2_K, Q4_K, Q8_K, BITNET]
BB = {Q2_K: 84, Q4_K: 144, Q8_K: 292, BITNET: 10}
BE = {Q2_K: 256, Q4_K: 256, Q8_K: 256, BITNET: 32}


def same_bits(a, b):
    a, b = np.asarray(a, np.float32), np.asarray(b, np.float32)
    na, nb = np.isnan(a), np.isnan(b)
    return np.array_equal(na, nb) and np.array_equal(a.view(np.uint32)[~na], b.view(np.uint32)[~nb])


def random_blocks(t, nblocks, seed, valid_bitnet=True):
    """Random bytes with sane scales (random normal halves); BitNet digit bytes < 243 unless asked otherwise."""
    rng = np.random.default_rng(seed)
    raw = rng.integers(0, 256, nblocks * BB[t], dtype=np.uint8).reshape(nblocks, BB[t])
    h = rng.normal(0, 0.05, (nblocks, 2)).astype(np.float16).view(np.uint8).reshape(nblocks, 4)
    if t == Q2_K:
        raw[:, 80:84] = h
    elif t == Q4_K:
        raw[:, 0:4] = h
    elif t == Q8_K:
        raw[:, 0:4] = rng.normal(0, 0.01, nblocks).astype("<f4").view(np.uint8).reshape(nblocks, 4)
    else:
        raw[:, 0:2] = h[:, :2]
        if valid_bitnet:
            raw[:, 2:9] = rng.integers(0, 243, (nblocks, 7), dtype=np.uint8)
    return raw.reshape(-1)


@pytest.mark.parametrize("t", TYPES)
def test_dequantize_two_restatements_agree(O, t):
    raw = random_blocks(t, 64, seed=3 + t)
    assert same_bits(O.kq_dequantize(t, raw, 64 * BE[t]), kq_np.DEQUANT[t](raw))


@pytest.mark.parametrize("t", [Q2_K, Q4_K, Q8_K])
def test_dequantize_any_bytes_agree(O, t):
    """every byte pattern, incl. subnormal / inf / NaN half scales"""
    raw = np.random.default_rng(9).integers(0, 256, 128 * BB[t], dtype=np.uint8)
    assert same_bits(O.kq_dequantize(t, raw, 128 * 256), kq_np.DEQUANT[t](raw))


def test_block_kats(O):
    """hand-computed: one block per format"""
    # Q8_K: d = 0.5, q[i] = i - 128 -> 0.5 * (i - 128)
    blk = np.zeros(292, np.uint8)
    blk[0:4] = np.array([0.5], "<f4").view(np.uint8)
    blk[4:260] = (np.arange(256) - 128).astype(np.int8).view(np.uint8)
    assert np.array_equal(O.kq_dequantize(Q8_K, blk, 256), 0.5 * (np.arange(256, dtype=np.float32) - 128))
    # BitNet: d = 2, byte 0 = 0 + 1*3 + 2*9 + 2*27 + 0*81 = 75 -> digits 0,1,2,2,0 -> -2, 0, 2, 2, -2; other bytes 121 = 1+3+9+27+81 -> 0
    blk = np.zeros(10, np.uint8)
    blk[0:2] = np.array([2.0], np.float16).view(np.uint8)
    blk[2] = 75
    blk[3:9] = 121
    got = O.kq_dequantize(BITNET, blk, 32)
    assert np.array_equal(got[:5], [-2, 0, 2, 2, -2]) and not np.any(got[5:])
    # Q2_K: d = 3, dmin = -1; sub-block 0: scale nibble 15, min nibble 2 -> scale = 3, min = 2*3 - 1 = 5; byte 0b11100100 = q 0,1,2,3
    blk = np.zeros(84, np.uint8)
    blk[80:82] = np.array([3.0], np.float16).view(np.uint8)
    blk[82:84] = np.array([-1.0], np.float16).view(np.uint8)
    blk[0] = 15 | (2 << 4)
    blk[16] = 0b11100100
    got = O.kq_dequantize(Q2_K, blk, 256)
    assert np.array_equal(got[:4], np.float32([5, 6, 7, 8]))          # q/3 * 3 + 5
    assert np.array_equal(got[16:20], np.float32([-1, -1, -1, -1]))   # sub-block 1: scale 0, min 0*3 - 1
    # Q4_K: d = 63, dmin = 1; sub 0: scale byte 15 | (1 << 6), scales[1] = 2 -> qscale 15, qmin = 1 | (2 << 2) = 9 -> scale = 15, min = 9 + 1 = 10
    blk = np.zeros(144, np.uint8)
    blk[0:2] = np.array([63.0], np.float16).view(np.uint8)
    blk[2:4] = np.array([1.0], np.float16).view(np.uint8)
    blk[4] = 15 | (1 << 6)
    blk[5] = 2
    blk[16] = 0xF0 | 3
    got = O.kq_dequantize(Q4_K, blk, 256)
    assert np.array_equal(got[:2], np.float32([np.float32(3) / np.float32(15) * np.float32(15) + 10, 25]))
    # sub-block 1 reads scales[1] = 2 as its scale byte (qscale 2, qmin_lo 0) and scales[3] = 0 as its min-high
    assert np.array_equal(got[32:34], np.float32([1, 1]))


def test_bitnet_invalid_digit(O):
    """a byte >= 243 decodes to a leading digit 3: IllegalStateException in the reference (GGMLTypes.kt:993-998)"""
    blk = np.zeros(10, np.uint8)
    blk[0:2] = np.array([1.0], np.float16).view(np.uint8)
    blk[2] = 250
    with pytest.raises(ValueError):
        O.kq_dequantize(BITNET, blk, 32)


def _q8k_data():
    i = np.arange(256)
    return np.where(i < 64, i / np.float32(64) * np.float32(10), np.where(i < 128, -(i - 64) / np.float32(64) * np.float32(10), np.where(
        i < 192, np.where(i % 2 == 0, 50.0, -50.0), (i - 192) * np.float32(0.1) - 1))).astype(np.float32)


def test_q8_k_accuracy_thresholds(O):
    """GGMLKQuantAccuracyTest.kt:201-237: MSE < 0.02, MAD < 0.1 on its data"""
    x = _q8k_data()
    d = O.kq_dequantize(Q8_K, O.kq_quantize(Q8_K, x), 256).astype(np.float64)
    assert np.mean((d - x) ** 2) < 0.02 and np.mean(np.abs(d - x)) < 0.1


def test_q8_k_bsums(O):
    blk = O.kq_quantize(Q8_K, _q8k_data())
    q = blk[4:260].view(np.int8).astype(np.int32).reshape(16, 16).sum(axis=1)
    assert np.array_equal(blk[260:292].view("<i2"), q.astype(np.int16))


def test_bitnet_ternary_values(O):
    """GGMLBitNet158Test.kt:232-258: [-1.5, -0.1, 0, 0.1, 1.5, -2, 0, 2] -> scale 2, normalised values in {-1, 0, 1}"""
    x = np.zeros(32, np.float32)
    x[:8] = [-1.5, -0.1, 0.0, 0.1, 1.5, -2.0, 0.0, 2.0]
    blk = O.kq_quantize(BITNET, x)
    assert O.half_to_float(int(blk[0]) | (int(blk[1]) << 8)) == 2.0
    d = O.kq_dequantize(BITNET, blk, 32)
    assert np.array_equal(d[:8] / 2.0, [-1, 0, 0, 0, 1, -1, 0, 1])


@pytest.mark.parametrize("t,M,K,N", [(Q2_K, 4, 256, 3), (Q4_K, 2, 256, 2), (Q8_K, 3, 512, 2), (BITNET, 2, 64, 3)])
def test_matmul_equals_dequantized_matmul(O, t, M, K, N):
    """GGMLKQuantAccuracyTest.kt:240-338 / GGMLBitNet158Test.kt:261-330 assert finite, non-zero results; the stronger
    property checked here: the direct dot product equals the sequential F32 dot of the dequantised row, bit for bit
    (same operations in the same order whenever K is a multiple of the block size)."""
    rng = np.random.default_rng(t)
    w = rng.normal(0, 1, M * K).astype(np.float32)
    raw = O.kq_quantize(t, w)
    b = rng.normal(0, 1, K * N).astype(np.float32)
    got = O.kq_matmul(t, raw, K, M, b, N)
    assert np.all(np.isfinite(got)) and np.any(got != 0)
    deq = O.kq_dequantize(t, raw, M * K).reshape(M, K)
    ref = np.zeros(M * N, np.float32)
    for i in range(M):
        for j in range(N):
            s = np.float32(0)
            for k in range(K):
                s = np.float32(s + np.float32(deq[i, k] * b[j + k * N]))
            ref[j + i * N] = s
    assert same_bits(got, ref)


@pytest.mark.parametrize("t", [Q2_K, Q4_K, Q8_K])
def test_irregular_k_block_addressing(O, t):
    """K = 384 (not a multiple of 256), M = 2: chunk [0, 256) of row i reads block (i K)/256 from its element 0, the
    trailing 128 elements are addressed by flat index; Q4_K's trailing chunk adds dmin instead of the sub-block min
    (core/GGMLComputeOps.kt:170-176, :308-336)."""
    M, K = 2, 384
    raw = random_blocks(t, 3, seed=21)
    x = np.random.default_rng(5).normal(0, 1, K).astype(np.float32)
    got = O.kq_matmul(t, raw, K, M, x, 1)
    full = kq_np.DEQUANT[t](raw).reshape(3, 256)
    part = kq_np.dequant_q4_k(raw, partial=True).reshape(3, 256) if t == Q4_K else full
    ref = []
    for i in range(M):
        w = np.empty(K, np.float32)
        w[:256] = full[(i * K) // 256]
        flat = i * K + np.arange(256, K)
        w[256:] = part[flat // 256, flat % 256]
        s = np.float32(0)
        for k in range(K):
            s = np.float32(s + np.float32(w[k] * x[k]))
        ref.append(s)
    assert same_bits(got, np.float32(ref))


def test_matmul_rejects_partial_blocks(O):
    raw = np.zeros(292, np.uint8)
    with pytest.raises(ValueError):
        O.kq_matmul(Q8_K, raw, 100, 1, np.zeros(100, np.float32), 1)


# ---- pinning against the reference's own tests and hand-worked blocks ---------------------------------
import json
import os


def _kq_golden():
    with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "kquant_reference_tests.json")) as f:
        return json.load(f)


def _ref_test_input(name):
    """The FloatArray initialisers of nativeTest core/GGMLKQuantAccuracyTest.kt, evaluated in F32 like Kotlin does."""
    f32 = np.float32
    if name == "q2k":       # :122-130
        out = []
        for i in range(512):
            r = i % 256
            if r < 64:
                out.append(f32(r) / f32(64) * f32(2.0) - f32(1.0))
            elif r < 128:
                out.append(f32(0.5) if r % 2 == 0 else f32(-0.5))
            elif r < 192:
                out.append(f32(r - 128) / f32(64) * f32(0.25))
            else:
                out.append((f32(r - 192) / f32(64)) * f32(3.0) - f32(1.5))
        return np.array(out, np.float32)
    if name == "q4k":       # :165-170
        return np.array([f32(i % 256) / f32(256) * f32(4.0) - f32(2.0) if i // 256 == 0 else (f32(1.0) if (i % 256) % 16 < 8 else f32(-1.0))
                         for i in range(512)], np.float32)
    return _q8k_data()      # :204-211


def test_hand_worked_q2_k_block(O):
    """A block derived BY HAND from core/GGMLComputeOps.kt:1741-1801 / :2125-2156 (tests/golden/make_kquant_golden.py holds
    the derivation line by line): x = 0, 1, 2, 3 repeated.  Bytes: 16 scale bytes 0x0F, 64 quant bytes 0xE4, d = 1.0, dmin = 0.
    The dequantizer returns q / 3, not q: the reference's pair is not an inverse, MSE = 14 / 9 on exactly representable data."""
    g = _kq_golden()["hand_q2_k"]
    x = np.array(g["x"], np.float32)
    blk = O.kq_quantize(Q2_K, x)
    assert bytes(blk) == bytes.fromhex(g["expect_blocks_hex"])
    deq = O.kq_dequantize(Q2_K, blk, 256)
    want = np.tile(np.float32([0, 1, 2, 3]) / np.float32(3) * np.float32(1) + np.float32(0), 64)
    assert same_bits(deq, want)
    assert abs(float(np.mean((deq.astype(np.float64) - x) ** 2)) - g["expect_mse"]) < 1e-6
    assert same_bits(kq_np.DEQUANT[Q2_K](np.frombuffer(bytes.fromhex(g["expect_blocks_hex"]), np.uint8)), want)


def test_hand_worked_q4_k_block(O):
    """x = 0..15 repeated, derived by hand from :1867-1931 / :2207-2250: d = half(15 / 255) = 0x2B88, all eight scale bytes 0x3F
    (round(scale / d * 63) = 1071 saturates at 63), bytes 5 / 7 / 9 / 11 are BOTH a sub-block's scale and another one's high min
    bits, quants 0x10 0x32 ... 0xFE.  Dequantized: (q / 15) * d + (60 / 63) * d for sub-blocks 0..3, (q / 15) * d for 4..7 — inputs
    0..15 come back as 0.056..0.115: the reference's Q4_K round trip loses the sub-block scale."""
    g = _kq_golden()["hand_q4_k"]
    x = np.array(g["x"], np.float32)
    blk = O.kq_quantize(Q4_K, x)
    assert bytes(blk) == bytes.fromhex(g["expect_blocks_hex"])
    deq = O.kq_dequantize(Q4_K, blk, 256).reshape(8, 32)
    d = np.float32(O.half_to_float(g["d_half_bits"]))
    q = np.float32(np.arange(32) % 16)
    for sb in range(8):
        scale = np.float32(np.float32(63) / np.float32(63)) * d                       # :2231
        mn = np.float32(np.float32(np.float32(g["expect_min_numerator"][sb]) / np.float32(63)) * d) + np.float32(0)   # :2232
        assert same_bits(deq[sb], np.float32(np.float32(q / np.float32(15)) * scale) + mn), sb     # :2243
    assert float(np.mean((deq.reshape(-1).astype(np.float64) - x) ** 2)) > 70.0


def test_reference_accuracy_thresholds_where_the_reference_code_can_meet_them(O):
    """Inputs and thresholds of testQ2_KAccuracy / testQ4_KAccuracy / testQ8_KAccuracy.  Q8_K's hold.  Q2_K's and Q4_K's
    MSE thresholds do not hold for the reference's own arithmetic — the two hand-worked blocks above show why, on inputs
    a reader can follow by hand — so what is pinned for those two is the measured outcome of the literal restatement
    (it must not drift)."""
    g = _kq_golden()["reference_tests"]
    res = {}
    for name, t in (("testQ2_KAccuracy", Q2_K), ("testQ4_KAccuracy", Q4_K), ("testQ8_KAccuracy", Q8_K)):
        x = _ref_test_input(g[name]["formula"])
        assert x.size == g[name]["n"]
        d = O.kq_dequantize(t, O.kq_quantize(t, x), x.size).astype(np.float64)
        res[name] = (float(np.mean((d - x) ** 2)), float(np.mean(np.abs(d - x))))
    assert res["testQ8_KAccuracy"][0] < g["testQ8_KAccuracy"]["mse_lt"] and res["testQ8_KAccuracy"][1] < g["testQ8_KAccuracy"]["mad_lt"]
    # the reference's code on the reference's data: MSE 0.1365 (threshold 0.1), MAD 0.3047 (threshold 0.3): both missed
    assert 0.130 < res["testQ2_KAccuracy"][0] < 0.143 and 0.295 < res["testQ2_KAccuracy"][1] < 0.315
    # Q4_K: MSE 3.6 (threshold 0.05)
    assert 3.0 < res["testQ4_KAccuracy"][0] < 4.2
