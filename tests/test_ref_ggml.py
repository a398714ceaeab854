"""Pins the oracle against the reference tree's OWN compiled C (oracle/_ref/libggml_ref.so = the unmodified
/root/reference/ggml/src/*.c built by `make -C oracle ref`): f16 conversions over all 65536 patterns,
dequantize_row_q4_0/q4_1/q8_0 bit for bit (the canonical ggml / GGUF nibble order),
the Q8_0 / Q4_1 quantizers, and ggml's rms_norm / silu / soft_max / rope / mul_mat executed through
ggml_graph_compute.  CPU only; skipped where the prebuilt library is absent."""
import numpy as np
import pytest

from conftest import same_f32_bits, synth_acts, synth_weights

from oracle import ref as R

pytestmark = pytest.mark.skipif(R.build() is None, reason="oracle/_ref not built (reference tree not mounted)")

Q4_0, Q4_1, Q8_0 = 2, 3, 6
TYPES = [Q4_0, Q4_1, Q8_0]


def test_half_to_float_all_patterns(O):
    # core/NumericConversions.kt:9-54 vs ggml_fp16_to_fp32, every bit pattern
    for h in range(65536):
        a, b = O.half_to_float(h), R.fp16_to_fp32(h)
        assert a == b or (a != a and b != b), hex(h)


def test_float_to_half_normal_range(O):
    # floatToHalf agrees with ggml_fp32_to_fp16 on normal halves (the Kotlin port's half-denormal off-by-one,
    # SURVEY F11, is a documented divergence and is excluded: |x| >= 2^-14)
    rng = np.random.default_rng(7)
    xs = np.concatenate([rng.normal(0, 1, 2000), rng.normal(0, 1e-3, 2000), rng.uniform(-6e4, 6e4, 2000),
                         [0.0, 1.0, -1.0, 65504.0, 6.1035156e-05]]).astype(np.float32)
    xs = xs[(np.abs(xs) >= 6.1035156e-05) | (xs == 0)]
    for x in xs:
        assert O.float_to_half(float(x)) == R.fp32_to_fp16(float(x)), float(x)


@pytest.mark.parametrize("t", TYPES)
def test_dequantize_ggml_order_bit_exact(O, t):
    rng = np.random.default_rng(100 + t)
    bb = O.BLOCK_BYTES[t]
    nblk = 4096
    raw = rng.integers(0, 256, nblk * bb, dtype=np.uint8)   # every byte pattern incl. subnormal / inf / NaN scales
    want = R.dequantize(t, raw, nblk * 32)
    assert same_f32_bits(O.dequantize_ggml(t, raw, nblk * 32), want)
    # the Kotlin-order oracle on the repacked bytes is the same function
    assert same_f32_bits(O.dequantize(t, O.repack_nibbles(t, raw, to_ggml=False), nblk * 32), want)


@pytest.mark.parametrize("t", [Q4_0, Q4_1])
def test_repack_is_a_bijection(O, t):
    raw = np.random.default_rng(5).integers(0, 256, 1000 * O.BLOCK_BYTES[t], dtype=np.uint8)
    g = O.repack_nibbles(t, raw, to_ggml=True)
    assert not np.array_equal(g, raw)
    assert np.array_equal(O.repack_nibbles(t, g, to_ggml=False), raw)
    # hand KAT: Kotlin byte j = (e[2j], e[2j+1]); ggml byte j = (e[j], e[j+16])
    e = np.arange(32) % 16
    blk = np.zeros(O.BLOCK_BYTES[t], np.uint8)
    q0 = 2 if t == Q4_0 else 4
    blk[q0:] = e[0::2] | (e[1::2] << 4)
    gg = O.repack_nibbles(t, blk, to_ggml=True)
    assert np.array_equal(gg[q0:], (e[:16] | (e[16:] << 4)).astype(np.uint8))


@pytest.mark.parametrize("t", [Q8_0, Q4_1])
def test_quantizers_match_ggml_ref(O, t):
    # quantizeTensor's Q8_0 (core/GGMLComputeOps.kt:1058-1071) and Q4_1 (:1162-1203) are the ggml formulas;
    # Q4_0 is NOT (the Kotlin one scales by amax/8) and is pinned by source only.
    for seed in (1, 2, 3):
        x = synth_weights(seed, 64, 256)
        mine = O.repack_nibbles(t, O.quantize(t, x), to_ggml=True)
        assert np.array_equal(mine, R.quantize(t, x))


def test_rms_norm_silu_softmax_against_ggml(O):
    x = synth_acts(11, 4096 * 4)
    got, want = O.rms_norm(x, None, 1e-5, 4096), R.unary(R.OP_RMS_NORM, x, 4096, 1e-5)
    assert np.max(np.abs(got - want)) <= 1e-6 * np.max(np.abs(want))
    got, want = O.silu(x), R.unary(R.OP_SILU, x, 4096)
    assert np.max(np.abs(got - want)) <= 1e-6
    got, want = O.softmax(x * 4, 2048), R.unary(R.OP_SOFT_MAX, x * 4, 2048)
    assert np.max(np.abs(got - want)) <= 1e-5 * np.max(want)     # ggml: SIMD expf approximation, f64 row sum
    assert np.allclose(want.reshape(-1, 2048).sum(1), 1.0, atol=1e-5)


@pytest.mark.parametrize("t", TYPES)
def test_matmul_against_ggml_mul_mat(O, t):
    # ggml quantizes the activations to Q8_0 first, so the two agree to ~1e-2 only; this checks layout and
    # nibble order end to end (a wrong order gives O(1) relative error), not rounding
    K, M = 1024, 96
    w, x = synth_weights(21, M, K), synth_acts(21, K)
    qk = O.quantize(t, w)
    want, _ = R.mul_mat(t, O.repack_nibbles(t, qk, to_ggml=True), K, M, x, 1)
    got = O.matmul(t, qk, K, M, x, 1)
    assert np.max(np.abs(got - want)) <= 2e-2 * np.max(np.abs(want))


def test_rope_against_ggml_rope(O):
    # applyRoPE (model/LlamaAttention.kt:29-72): same pairing (2i, 2i+1), frequency base^(-2i/d) and sign as ggml_rope
    # mode 0; ggml iterates theta in F32 (theta *= theta_scale), hence 1e-5, not bit-exact
    for H, D, pos in ((4, 128, 7), (2, 64, 1000), (8, 32, 0)):
        x = synth_acts(31 + pos, H * D)
        got, want = O.rope(x, D, H, pos), R.unary(R.OP_ROPE, x, D, float(pos), 10000.0)
        assert np.max(np.abs(got - want)) <= 2e-5 * np.max(np.abs(want))


def test_attention_oracle_against_f64(O):
    rng = np.random.default_rng(1)
    D, H, L = 64, 3, 77
    q, kc, vc = (rng.standard_normal(n).astype(np.float32) for n in (D * H, L * D * H, L * D * H))
    got = O.attention_decode(q, kc, vc, L, D, H)
    K, V, Q = kc.reshape(L, H, D).astype(np.float64), vc.reshape(L, H, D).astype(np.float64), q.reshape(H, D).astype(np.float64)
    s = np.einsum("hd,lhd->hl", Q, K) / np.sqrt(D)
    w = np.exp(s - s.max(1, keepdims=True))
    w /= w.sum(1, keepdims=True)
    assert np.max(np.abs(np.einsum("hl,lhd->hd", w, V).reshape(-1) - got)) <= 1e-6
