"""Pins the oracle (oracle/lk_oracle.c) to every golden vector / known-answer test the
reference's own tests hold for this path and to the independent numpy
restatement.  CPU only."""
import math

import numpy as np
import pytest

from conftest import same_f32_bits

F32, Q4_0, Q4_1, Q8_0 = 0, 2, 3, 6
TYPES = {"Q8_0": Q8_0, "Q4_0": Q4_0, "Q4_1": Q4_1}


def test_f32_matmul_2x3_3x2(O, golden):
    g = golden["transcribed"]["f32_matmul_2x3_3x2"]
    K, M = g["a_ne"]
    N = g["b_ne"][0]
    got = O.matmul(F32, np.array(g["a"], np.float32), K, M, g["b"], N)
    assert np.allclose(got, g["expected"], atol=g["tol"])


def test_f32_matmul_4x4_manual_triple_loop(O, golden):
    g = golden["transcribed"]["f32_matmul_4x4"]
    a, b = np.array(g["a"], np.float32), np.array(g["b"], np.float32)
    got = O.matmul(F32, a, 4, 4, b, 4).reshape(4, 4)
    # TestTensorOps.kt:150-160: expected[i][j] = sum_k a[k + i*4] * b[j + k*4]
    exp = np.array([[sum(a[k + i * 4] * b[j + k * 4] for k in range(4)) for j in range(4)] for i in range(4)])
    assert np.allclose(got, exp, atol=g["tol"])


def test_identity_matmul_plus_c(O, golden):
    eye = np.eye(4, dtype=np.float32).reshape(-1)
    ab = O.matmul(F32, eye, 4, 4, eye, 4)
    res = O.add(ab, np.full(16, 0.1, np.float32)).reshape(4, 4)
    tol = golden["transcribed"]["identity_plus_c"]["tol"]
    assert np.allclose(np.diag(res), 1.1, atol=tol)
    assert np.allclose(res - np.diag(np.diag(res)), 0.1 * (1 - np.eye(4)), atol=tol)


def test_q80_dot_cos_rel_error(O, golden):
    g = golden["transcribed"]["q80_dot_cos"]
    n = g["n"]
    i = np.arange(n, dtype=np.float32)
    v1 = (0.1 + 2.0 * np.cos(i)).astype(np.float32)
    v2 = (0.1 + 2.0 * np.cos(i + np.float32(1.0))).astype(np.float32)
    ref = float(np.float32(np.sum(v1.astype(np.float64) * v2.astype(np.float64))))
    q = O.quantize(Q8_0, v1)
    got = O.dot(Q8_0, q, n, 0, v2, 1, 0)
    assert abs(ref - got) / abs(ref) < g["rel_tol"]


@pytest.mark.parametrize("tname", ["Q8_0", "Q4_0", "Q4_1"])
def test_direct_vs_dequantized_matmul(O, golden, tname):
    g = golden["transcribed"]["direct_vs_dequant"]
    t = TYPES[tname]
    for M, K, N in g["shapes"]:
        idx = np.arange(M * K)
        w = {"Q8_0": (idx % 127) - 63.0, "Q4_0": (idx % 15) - 7.0, "Q4_1": (idx % 20) * 1.0}[tname].astype(np.float32)
        a = O.quantize(t, w)
        b = ((np.arange(K * N) % 10) + 1.0).astype(np.float32)   # GGMLMatMulOptimizationTest.kt:52-55
        direct = O.matmul(t, a, K, M, b, N)
        fallback = O.matmul(F32, O.dequantize(t, a, M * K), K, M, b, N)
        assert np.max(np.abs(direct - fallback)) < g["tol"]


def _metrics(x, y):
    err = x.astype(np.float64) - y.astype(np.float64)
    mse = float(np.mean(err ** 2))
    mad = float(np.mean(np.abs(err)))
    snr = 10 * math.log10(float(np.mean(x.astype(np.float64) ** 2)) / mse) if mse > 0 else float("inf")
    return mse, mad, snr


@pytest.mark.parametrize("tname", ["Q8_0", "Q4_0", "Q4_1"])
def test_roundtrip_thresholds_cos(O, golden, tname):
    th = golden["transcribed"]["roundtrip_thresholds"][tname]
    x = (0.1 + 2.0 * np.cos(np.arange(512, dtype=np.float32))).astype(np.float32)   # n=512 as in :256
    t = TYPES[tname]
    y = O.dequantize(t, O.quantize(t, x), x.size)
    mse, mad, snr = _metrics(x, y)
    assert mad < th["mad"]
    if tname == "Q4_0":
        # Reference defect, recorded not mirrored: the Kotlin Q4_0 quantizer (d = amax/8, q clamped to 15,
        # GGMLComputeOps.kt:1079-1083) loses the +amax element by a full step, so its own
        # (uncompilable, F7) test threshold MSE < 0.01 is missed by a hair (0.0107).  Two independent
        # restatements agree on that value; we assert the restated value instead of the unmet threshold.
        assert 0.0100 < mse < 0.0115
    else:
        assert mse < th["mse"]
    if "snr_db" in th:
        assert snr >= th["snr_db"]


def test_silu_rmsnorm_vectors(O, golden):
    s = golden["transcribed"]["silu"]
    assert np.allclose(O.silu(s["x"]), s["expected"], atol=s["tol"])
    r = golden["transcribed"]["rms_norm"]
    assert np.allclose(O.rms_norm(r["x"], None, r["eps"], 4), r["expected"], atol=r["tol"])


def test_softmax_properties(O):
    # no softmax test exists in the reference: pinned by source + properties
    x = np.random.default_rng(7).normal(0, 3, 5 * 37).astype(np.float32)
    y = O.softmax(x, 37).reshape(5, 37)
    assert np.allclose(y.sum(axis=1), 1.0, atol=1e-5)
    assert np.allclose(O.softmax(x + np.float32(100.0), 37), y.reshape(-1), atol=1e-6)       # shift invariance
    big = O.softmax(np.array([1e30, -1e30, 0], np.float32), 3)
    assert np.allclose(big, [1, 0, 0]) and not np.any(np.isnan(big))


def test_float_to_half_f11_and_all_halves(O, NP, golden):
    for v, h in golden["transcribed"]["float_to_half_F11"]["cases"]:
        assert O.float_to_half(v) == h, (v, hex(O.float_to_half(v)), hex(h))
    for h in range(0, 65536, 7):
        a, b = O.half_to_float(h), NP.half_to_float(h)
        assert a == b or (math.isnan(a) and math.isnan(b))


def test_block_kats(O, golden):
    for name, k in golden["kat"].items():
        x = np.array(k["x"], np.float32)
        blocks = O.quantize(k["type"], x)
        assert blocks.tobytes().hex() == k["blocks_hex"], name
        deq = O.dequantize(k["type"], blocks, x.size)
        exp = np.frombuffer(bytes.fromhex(k["dequant_f32_hex"]), dtype="<f4")
        assert same_f32_bits(deq, exp), name


def test_q80_ramp_hand_computed(O):
    # k-16: amax=16, d=16/127=0.12598425 -> half 0x3008; -16*(127/16) = -127; 15*7.9375 = 119.06 -> 119
    b = O.quantize(Q8_0, np.arange(32, dtype=np.float32) - 16)
    assert b[0] == 0x08 and b[1] == 0x30
    q = b[2:].view(np.int8)
    assert q[0] == -127 and q[16] == 0 and q[31] == 119 and q[17] == 8
    # Q4_0 of the same data: d = 16/8 = 2 (0x4000); q = round(x/2 + 8): x=-16 -> 0, x=15 -> round(15.5)=16 -> clamp 15
    b4 = O.quantize(Q4_0, np.arange(32, dtype=np.float32) - 16)
    assert b4[0] == 0x00 and b4[1] == 0x40
    # element 0 -> low nibble of byte 0, element 1 (x=-15 -> round(0.5)=0 ties-to-even) -> high nibble
    assert (b4[2] & 15) == 0 and (b4[2] >> 4) == 0
    # element 3 (x=-13 -> round(1.5) = 2), element 2 (x=-14 -> 1)
    assert (b4[3] & 15) == 1 and (b4[3] >> 4) == 2
    assert (b4[17] >> 4) == 15


@pytest.mark.parametrize("t", [Q8_0, Q4_0, Q4_1])
def test_c_oracle_equals_numpy_restatement(O, NP, t):
    rng = np.random.default_rng(11 + t)
    M, K, N = 5, 96, 3
    w = rng.normal(0, 0.02, M * K).astype(np.float32)
    a = O.quantize(t, w)
    assert np.array_equal(a, NP.quantize(t, w))
    assert np.array_equal(O.dequantize(t, a, M * K), NP.dequantize(t, a, M * K))
    b = rng.normal(0, 1, K * N).astype(np.float32)
    ref = NP.matmul(t, a, K, M, b, N)
    assert np.array_equal(O.matmul(t, a, K, M, b, N), ref)
    assert np.array_equal(O.matmul(t, a, K, M, b, N, nthreads=4), ref)     # MT variant is bit-identical


def test_oracle_error_codes(O):
    with pytest.raises(ValueError):
        O.quantize(Q8_0, np.zeros(6, np.float32))   # 6 elements is not a multiple of 32 (GGMLComputeOps.kt:1059)
