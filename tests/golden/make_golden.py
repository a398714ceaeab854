"""Regenerates tests/golden/reference_vectors.json.

Two kinds of entries:
 (1) vectors and tolerances TRANSCRIBED from the reference's own tests (file:line cited
     in each entry; paths relative to /root/reference/src/nativeTest/kotlin/ai/solace/llamakotlin/),
 (2) block-level known-answer tests: inputs from the reference tests' data formulas,
     expected bytes produced by the pure-Python/numpy restatement (oracle/np_oracle.py),
     frozen here so that the C oracle and the CUDA path are checked against fixed bytes.
     The Kotlin reference itself cannot run in this environment (no JVM/Konan), so (2) is
     pinned by source citation, not by executing the reference.
Run:  python tests/golden/make_golden.py
"""
import json
import math
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import np_oracle as NP  # noqa: E402

F32, Q4_0, Q4_1, Q8_0 = 0, 2, 3, 6
out = {"transcribed": {}, "kat": {}}
T = out["transcribed"]

T["f32_matmul_2x3_3x2"] = {
    "cite": "core/GGMLComputeOpsDestinationTest.kt:219-265",
    "a": [1, 2, 3, 4, 5, 6], "a_ne": [3, 2], "b": [7, 8, 9, 10, 11, 12], "b_ne": [2, 3],
    "expected": [58, 64, 139, 154], "tol": 1e-3}
T["f32_matmul_4x4"] = {
    "cite": "nativeMain core/TestTensorOps.kt:122-168 (a[i][j]=4i+j, b=a+1, vs manual triple loop, 1e-4)",
    "a": [float(i) for i in range(16)], "b": [float(i + 1) for i in range(16)], "tol": 1e-4}
T["identity_plus_c"] = {
    "cite": "core/GGMLIntegrationTest.kt:159-196 ((I*I)+0.1 -> diag 1.1, off-diag 0.1, 1e-3)", "tol": 1e-3}
T["q80_dot_cos"] = {
    "cite": "core/GGMLStandardizedQuantizationTest.kt:363-394 (n=128, 0.1+2cos(i), 0.1+2cos(i+1), rel err < 0.02 vs f64)",
    "n": 128, "rel_tol": 0.02}
T["direct_vs_dequant"] = {
    "cite": "core/GGMLMatMulOptimizationTest.kt:233-357 (|direct - dequantize->F32 matmul| < 1e-3; data :60-95)",
    "shapes": [[4, 64, 3], [3, 64, 4], [2, 32, 2]], "tol": 1e-3,
    "data": {"Q8_0": "(idx%127)-63", "Q4_0": "(idx%15)-7", "Q4_1": "idx%20"}}
T["roundtrip_thresholds"] = {
    "cite": "core/GGMLStandardizedQuantizationTest.kt:35-46,205-208,264-266,300-301 (data 0.1+2cos(i), :64-68)",
    "Q8_0": {"mse": 1e-4, "mad": 0.01, "snr_db": 40.0},
    "Q4_0": {"mse": 0.01, "mad": 0.2, "snr_db": 20.0},
    "Q4_1": {"mse": 0.015, "mad": 0.1}}
T["accuracy_thresholds_128"] = {
    "cite": "core/GGMLQuantizationAccuracyTest.kt:202-393 (Q8_0 MSE<0.05 MAD<0.2; Q4_0 MSE<0.02 MAD<0.25; Q4_1 MSE<0.015 MAD<0.1)",
    "Q8_0": {"mse": 0.05, "mad": 0.2}, "Q4_0": {"mse": 0.02, "mad": 0.25}, "Q4_1": {"mse": 0.015, "mad": 0.1}}
T["silu"] = {
    "cite": "core/GGMLComputeOpsTest.kt:491-499,607-625", "x": [-5, -1, 0, 1, 5],
    "expected": [x * (1.0 / (1.0 + math.exp(-x))) for x in [-5.0, -1.0, 0.0, 1.0, 5.0]], "tol": 1e-3}
rms = math.sqrt((1 + 4 + 9 + 16) / 4 + 1e-5)
T["rms_norm"] = {
    "cite": "core/GGMLComputeOpsTest.kt:502-521,650-672", "x": [1, 2, 3, 4], "eps": 1e-5,
    "expected": [v / rms for v in [1.0, 2.0, 3.0, 4.0]], "tol": 1e-4}
T["float_to_half_F11"] = {
    "cite": "nativeMain core/NumericConversions.kt:83 (denormal shift off-by-one; SURVEY F11 values)",
    "cases": [[2.0 ** -15, 0x0100], [6e-5, 0x01F7], [5.96e-8, 0x0000], [1.0, 0x3C00], [65504.0, 0x7BFF],
              [-2.0, 0xC000], [0.0, 0x0000], [65520.0, 0x7C00]]}

# ---- block KATs ------------------------------------------------------------------
K = out["kat"]


def kat(name, type_, x, cite):
    x = np.asarray(x, dtype=np.float32)
    blocks = NP.quantize(type_, x)
    deq = NP.dequantize(type_, blocks, x.size)
    K[name] = {"cite": cite, "type": type_, "x": [float(v) for v in x], "blocks_hex": blocks.tobytes().hex(),
               "dequant_f32_hex": deq.astype("<f4").tobytes().hex()}


idx = np.arange(128)
kat("q80_ramp", Q8_0, np.arange(32) - 16.0, "hand KAT: k-16 -> amax 16, d=16/127 (SURVEY §8c)")
kat("q80_mod127", Q8_0, (idx % 127) - 63.0, "GGMLMatMulOptimizationTest.kt:60-62 data")
kat("q40_mod15", Q4_0, (idx % 15) - 7.0, "GGMLMatMulOptimizationTest.kt:77-79 data")
kat("q41_mod20", Q4_1, (idx % 20).astype(np.float32), "GGMLMatMulOptimizationTest.kt:93-95 data")
cosd = (0.1 + 2.0 * np.cos(idx.astype(np.float32))).astype(np.float32)
for nm, t in (("q80_cos", Q8_0), ("q40_cos", Q4_0), ("q41_cos", Q4_1)):
    kat(nm, t, cosd, "GGMLStandardizedQuantizationTest.kt:64-68 data")
zeros = np.zeros(32, dtype=np.float32)
for nm, t in (("q80_zero", Q8_0), ("q40_zero", Q4_0), ("q41_zero", Q4_1)):
    kat(nm, t, zeros, "all-zero block: scale=1.0 path, GGMLComputeOps.kt:1065,1079,1180")
tiny = (np.arange(32, dtype=np.float32) - 16) * np.float32(1e-6)
for nm, t in (("q80_denorm_scale", Q8_0), ("q40_denorm_scale", Q4_0)):
    kat(nm, t, tiny, "denormal-half scale (F11 bug path), NumericConversions.kt:75-100")
edge = np.array([0.0, -0.0, 1.4e-45, -1.4e-45, 1, -1, 3.4e35, -3.4e35, 10, -10, 100, -100, 1e-3, -1e-3,
                 math.pi, -math.pi, math.e, -math.e] + [0.0] * 14, dtype=np.float32)
for nm, t in (("q80_edge", Q8_0), ("q40_edge", Q4_0), ("q41_edge", Q4_1)):
    kat(nm, t, edge, "GGMLStandardizedQuantizationTest.kt:83-95 edge-case list, padded to one block")

with open(os.path.join(os.path.dirname(os.path.abspath(__file__)), "reference_vectors.json"), "w") as f:
    json.dump(out, f, indent=1)
print("wrote", len(T), "transcribed and", len(K), "KAT entries")
