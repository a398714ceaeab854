"""N4: the reference's K-quant / BitNet 1.58 formats on the CUDA path (lk_kquants.cu) through the C ABI,
against the oracle (oracle/lk_oracle_kq.c).  Dequantisation: bit-exact.  Matmul: <= 1e-3 of max|ref| (the main path's
bar: the GPU sums in a different order than the sequential reference)."""
import numpy as np
import pytest

import llama_kotlin_b200 as lk
from conftest import same_f32_bits
from llama_kotlin_b200 import GGMLTensor, GGMLType
from test_oracle_kquants import BB, BE, BITNET, Q2_K, Q4_K, Q8_K, TYPES, random_blocks

pytestmark = pytest.mark.gpu
TOL = 1e-3


def rel_err(got, ref):
    ref, got = np.asarray(ref, np.float64), np.asarray(got, np.float64)
    return float(np.max(np.abs(got - ref))) / max(float(np.max(np.abs(ref))), 1e-30)


@pytest.fixture(scope="module")
def ga():
    return lk.GGMLGraphAllocator(size=96 << 20)


def gpu_matmul(ga, t, raw, K, M, x, N):
    ga.reset()
    a = ga.allocateTensor(GGMLTensor(GGMLType(t), [K, M]))
    b = ga.allocateTensor(GGMLTensor(GGMLType.F32, [N, K]))
    d = ga.allocateTensor(GGMLTensor(GGMLType.F32, [N, M]))
    ga.setBytes(a, raw)
    ga.setBytes(b, x)
    lk.computeMatMul(ga, None, a, b, d)
    return ga.getFloats(d)


@pytest.mark.parametrize("t", TYPES)
@pytest.mark.parametrize("nblocks", [1, 3, 4096])
def test_dequantize_bit_exact(O, ga, t, nblocks):
    raw = random_blocks(t, nblocks, seed=nblocks + t)
    n = nblocks * BE[t]
    ga.reset()
    a = ga.allocateTensor(GGMLTensor(GGMLType(t), [n]))
    ga.setBytes(a, raw)
    assert same_f32_bits(lk.dequantizeTensor(ga, a).data, O.kq_dequantize(t, raw, n))


@pytest.mark.parametrize("t", [Q2_K, Q4_K, Q8_K])
def test_dequantize_any_bytes_bit_exact(O, ga, t):
    """every byte pattern incl. subnormal / inf / NaN scales"""
    raw = np.random.default_rng(17).integers(0, 256, 2048 * BB[t], dtype=np.uint8)
    n = 2048 * 256
    ga.reset()
    a = ga.allocateTensor(GGMLTensor(GGMLType(t), [n]))
    ga.setBytes(a, raw)
    assert same_f32_bits(lk.dequantizeTensor(ga, a).data, O.kq_dequantize(t, raw, n))


@pytest.mark.parametrize("t", TYPES)
def test_dequantize_of_quantized_weights(O, ga, t):
    w = np.random.default_rng(4).normal(0, 0.02, 256 * 64).astype(np.float32)
    raw = O.kq_quantize(t, w)
    ga.reset()
    a = ga.allocateTensor(GGMLTensor(GGMLType(t), [w.size]))
    ga.setBytes(a, raw)
    assert same_f32_bits(lk.dequantizeTensor(ga, a).data, O.kq_dequantize(t, raw, w.size))


@pytest.mark.parametrize("t", TYPES)
@pytest.mark.parametrize("M,K", [(1, 256), (4, 256), (7, 512), (64, 4096), (1000, 2048), (33, 11008 // 256 * 256)])
def test_gemv_fast_path(O, ga, t, M, K):
    """N == 1, rows of whole blocks: kq_gemv_kernel"""
    raw = O.kq_quantize(t, np.random.default_rng(M + K).normal(0, 0.02, M * K).astype(np.float32))
    x = np.random.default_rng(K).normal(0, 1, K).astype(np.float32)
    assert rel_err(gpu_matmul(ga, t, raw, K, M, x, 1), O.kq_matmul(t, raw, K, M, x, 1)) <= TOL


@pytest.mark.parametrize("t", TYPES)
def test_gemv_random_blocks(O, ga, t):
    """random block bytes (every scale / min nibble, every digit) with sane super-block scales"""
    M, K = 96, 1024
    raw = random_blocks(t, M * K // BE[t], seed=31)
    x = np.random.default_rng(8).normal(0, 1, K).astype(np.float32)
    assert rel_err(gpu_matmul(ga, t, raw, K, M, x, 1), O.kq_matmul(t, raw, K, M, x, 1)) <= TOL


def test_bitnet_gemv_k_multiple_of_32(O, ga):
    M, K = 40, 32 * 37
    raw = random_blocks(BITNET, M * K // 32, seed=2)
    x = np.random.default_rng(3).normal(0, 1, K).astype(np.float32)
    assert rel_err(gpu_matmul(ga, BITNET, raw, K, M, x, 1), O.kq_matmul(BITNET, raw, K, M, x, 1)) <= TOL


@pytest.mark.parametrize("t", TYPES)
@pytest.mark.parametrize("M,K,N", [(4, 256, 3), (2, 256, 2), (16, 512, 9), (8, 1024, 32)])
def test_matmul_batched(O, ga, t, M, K, N):
    """GGMLKQuantAccuracyTest.kt:240-338 shapes (4 x 256 x 3, 2 x 256 x 2) and wider ones: generic kernel"""
    raw = O.kq_quantize(t, np.random.default_rng(M * N).normal(0, 1, M * K).astype(np.float32))
    x = np.random.default_rng(N).normal(0, 1, K * N).astype(np.float32)
    got = gpu_matmul(ga, t, raw, K, M, x, N)
    assert np.all(np.isfinite(got)) and np.any(got != 0)
    assert rel_err(got, O.kq_matmul(t, raw, K, M, x, N)) <= TOL


@pytest.mark.parametrize("t", [Q2_K, Q4_K, Q8_K])
@pytest.mark.parametrize("M,K,N", [(2, 384, 1), (4, 320, 2), (2, 128, 1)])
def test_matmul_irregular_k(O, ga, t, M, K, N):
    """K not a multiple of 256 (M K is): the reference's literal block addressing, Q4_K's dmin-only tail
    (core/GGMLComputeOps.kt:170-176, :308-336)"""
    raw = random_blocks(t, M * K // 256, seed=K)
    x = np.random.default_rng(K).normal(0, 1, K * N).astype(np.float32)
    assert rel_err(gpu_matmul(ga, t, raw, K, M, x, N), O.kq_matmul(t, raw, K, M, x, N)) <= TOL


def test_bitnet_irregular_k(O, ga):
    M, K = 4, 40       # rows are not whole blocks: flat block addressing
    raw = random_blocks(BITNET, M * K // 32, seed=6)
    x = np.random.default_rng(7).normal(0, 1, K).astype(np.float32)
    assert rel_err(gpu_matmul(ga, BITNET, raw, K, M, x, 1), O.kq_matmul(BITNET, raw, K, M, x, 1)) <= TOL


def test_bitnet_invalid_byte_is_nan(ga):
    """byte >= 243: the reference throws IllegalStateException (GGMLTypes.kt:993-998); the device path poisons the
    affected outputs with NaN instead of returning numbers"""
    M, K = 4, 64
    raw = random_blocks(BITNET, M * K // 32, seed=1).reshape(-1, 10)
    raw[2, 4] = 250                                     # row 1, first block
    x = np.ones(K, np.float32)
    got = gpu_matmul(ga, BITNET, raw.reshape(-1), K, M, x, 1)
    assert np.isnan(got[1]) and np.all(np.isfinite(np.delete(got, 1)))


# ---- BitNet 1.58 on the streaming DP4A GEMV (K1, lk_gemv.cu block_dot<T_BITNET_1_58>): N == 1, K % 64 == 0 ----------------
def _bitnet_exact(raw, K, M, x):
    """float64 reference from the raw blocks: w = d * (digit - 1), core/GGMLTypes.kt:975-990"""
    blk = raw.reshape(-1, 10)
    d = blk[:, 0:2].copy().view(np.float16).astype(np.float64).reshape(-1)
    dig = np.empty((blk.shape[0], 35), np.int64)
    v = blk[:, 2:9].astype(np.int64)
    for p in range(5):
        dig[:, p::5] = v % 3
        v //= 3
    w = ((dig[:, :32] - 1) * d[:, None]).reshape(M, K)
    return w @ x.astype(np.float64)


@pytest.mark.parametrize("M,K", [(4, 64), (148, 128), (37 * 4, 1024), (4096, 4096), (1000, 11008), (4096, 28672)])
def test_bitnet_streaming_gemv(O, ga, M, K):
    raw = random_blocks(BITNET, M * K // 32, seed=M + K)
    x = np.random.default_rng(K).normal(0, 1, K).astype(np.float32)
    got = gpu_matmul(ga, BITNET, raw, K, M, x, 1)
    want = _bitnet_exact(raw, K, M, x)
    assert rel_err(got, want) <= 1e-5           # exact integer block sums, 24-bit activations: far inside the 1e-3 bar
    if M * K <= 1 << 22:
        assert rel_err(got, O.kq_matmul(BITNET, raw, K, M, x, 1)) <= TOL


def test_bitnet_every_byte_value_in_every_position(ga):
    """each digit byte 0..242 at each of the 7 positions of an even and of an odd block, one non-zero activation per
    decoded element: the product is exactly d * (digit - 1) * x"""
    K = 64
    rows = []
    for pos in range(7):
        for blk in range(2):
            for v in range(243):
                r = np.zeros((2, 10), np.uint8)
                r[:, 0:2] = np.array([1.0], np.float16).view(np.uint8)
                r[:, 2:9] = 121                      # 121 = 11111 base 3: every weight 0
                r[blk, 2 + pos] = v
                rows.append(r.reshape(-1))
    raw = np.concatenate(rows)
    M = len(rows)
    pad = (-M) % 4
    raw = np.concatenate([raw, np.tile(rows[0], pad)])
    M += pad
    x = (np.arange(K) + 1).astype(np.float32)        # distinct integers: the sum identifies the digits
    got = gpu_matmul(ga, BITNET, raw, K, M, x, 1)
    want = _bitnet_exact(raw, K, M, x)
    assert np.array_equal(got.astype(np.float64), want)


@pytest.mark.parametrize("blk", [0, 1, 2, 3])
@pytest.mark.parametrize("pos", range(7))
def test_bitnet_streaming_invalid_byte_is_nan(ga, blk, pos):
    M, K = 8, 128
    raw = random_blocks(BITNET, M * K // 32, seed=5).reshape(M, K // 32, 10)
    raw[:, :, 9] = 255                                # the unused byte may hold anything
    raw[3, blk, 2 + pos] = 243 + (pos * 2 + blk) % 13
    x = np.random.default_rng(1).normal(0, 1, K).astype(np.float32)
    got = gpu_matmul(ga, BITNET, raw.reshape(-1), K, M, x, 1)
    assert np.isnan(got[3]) and np.all(np.isfinite(np.delete(got, 3)))


def test_bitnet_range_fixup(ga):
    """a block whose large activation meets a zero weight: the small elements' products must survive (K1 range fix-up)"""
    M, K = 8, 256
    raw = random_blocks(BITNET, M * K // 32, seed=9).reshape(M, K // 32, 10)
    raw[:, :, 0:2] = np.array([0.5], np.float16).view(np.uint8)
    raw[:, 3, 2] = (raw[:, 3, 2] // 3) * 3 + 1        # digit 0 of byte 0 = 1: weight of element 0 of block 3 is zero
    x = np.random.default_rng(2).normal(0, 1e-9, K).astype(np.float32)
    x[3 * 32] = 3.0e4                                 # 2^45 above its neighbours
    got = gpu_matmul(ga, BITNET, raw.reshape(-1), K, M, x, 1)
    want = _bitnet_exact(raw.reshape(-1), K, M, x)
    assert float(np.max(np.abs(want))) < 1e-6         # the large element contributes nothing
    assert rel_err(got, want) <= 1e-5


@pytest.mark.parametrize("t", TYPES)
def test_errors(ga, t):
    gt = GGMLType(t)
    ga.reset()
    a = ga.allocateTensor(GGMLTensor(gt, [256, 2]))
    b = ga.allocateTensor(GGMLTensor(GGMLType.F32, [1, 128]))
    d = ga.allocateTensor(GGMLTensor(GGMLType.F32, [1, 2]))
    with pytest.raises(lk.IllegalArgumentException):      # K mismatch, core/GGMLComputeOps.kt:1440
        lk.computeMatMul(ga, None, a, b, d)
    b2 = ga.allocateTensor(GGMLTensor(GGMLType.F32, [1, 256]))
    d2 = ga.allocateTensor(GGMLTensor(GGMLType.F32, [2, 2]))
    with pytest.raises(lk.IllegalArgumentException):      # dst shape, :1444-1446
        lk.computeMatMul(ga, None, a, b2, d2)
    d3 = ga.allocateTensor(GGMLTensor(gt, [256 if BE[t] == 256 else 32, 2]))
    d3.ne = [1, 2, 1, 1]
    with pytest.raises(lk.IllegalArgumentException):      # dst type, :1483 / :1494 / :1505
        lk.computeMatMul(ga, None, a, b2, d3)


@pytest.mark.parametrize("t", TYPES)
def test_linearity_full_size(O, ga, t):
    """size-independent property at a Llama-2-7B shape: A(2x) == 2 Ax bit for bit, run-to-run identical"""
    M, K = 4096, 4096
    raw = random_blocks(t, M * K // BE[t], seed=77)
    x = np.random.default_rng(78).normal(0, 1, K).astype(np.float32)
    y1 = gpu_matmul(ga, t, raw, K, M, x, 1)
    y2 = gpu_matmul(ga, t, raw, K, M, 2 * x, 1)
    y3 = gpu_matmul(ga, t, raw, K, M, x, 1)
    assert same_f32_bits(y1, y3) and same_f32_bits(2 * y1, y2)
    rows = [0, 1, 2047, 4095]
    ref = O.kq_matmul(t, raw.reshape(M, -1)[rows].reshape(-1), K, len(rows), x, 1)
    assert rel_err(y1[rows], ref) <= TOL
