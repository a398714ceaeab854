import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a real B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def golden():
    with open(os.path.join(ROOT, "tests", "golden", "reference_vectors.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def O():
    """The C oracle (test infrastructure; built on demand with gcc)."""
    from oracle import oracle
    oracle.build()
    return oracle


@pytest.fixture(scope="session")
def NP():
    from oracle import np_oracle
    return np_oracle


def synth_weights(seed: int, M: int, K: int) -> np.ndarray:
    """F32 ~ N(0, 0.02^2), numpy default_rng(seed)."""
    return np.random.default_rng(seed).normal(0.0, 0.02, M * K).astype(np.float32)


def synth_acts(seed: int, K: int, N: int = 1) -> np.ndarray:
    """activations F32 ~ N(0,1), seed + 500."""
    return np.random.default_rng(seed + 500).normal(0.0, 1.0, K * N).astype(np.float32)


def same_f32_bits(a, b) -> bool:
    """Bit-exact F32 comparison; NaNs compare equal to NaNs (NaN sign/payload is not a parity target:
    it depends on x86 operand order, not on the algorithm)."""
    a = np.ascontiguousarray(a, dtype=np.float32).reshape(-1)
    b = np.ascontiguousarray(b, dtype=np.float32).reshape(-1)
    if a.size != b.size:
        return False
    na, nb = np.isnan(a), np.isnan(b)
    if not np.array_equal(na, nb):
        return False
    return bool(np.array_equal(a.view(np.uint32)[~na], b.view(np.uint32)[~nb]))
