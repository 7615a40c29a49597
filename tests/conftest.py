import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


@pytest.fixture(scope="session")
def oracle():
    import oraclelib
    if not os.path.exists(oraclelib.ORACLE_SO):
        subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")])
    return oraclelib.Oracle()


@pytest.fixture(scope="session")
def ctx():
    import simd_dct_b200 as sd
    c = sd.Context(0)
    yield c
    c.close()


def make_image(kind, h, w, seed=0, comps=None):
    """'rand' = uniform u8 (worst case); 'smooth' = SURVEY 8d config-2 synthetic."""
    rng = np.random.default_rng(seed)
    shape = (h, w) if comps is None else (h, w, comps)
    if kind == "rand":
        return rng.integers(0, 256, shape, dtype=np.uint8)
    y, x = np.mgrid[0:h, 0:w]
    base = 128 + 60 * np.sin(.05 * x + seed) + 50 * np.cos(.031 * y)
    if comps is not None:
        base = base[..., None] + np.arange(comps) * 7.0
    return np.clip(base + rng.integers(0, 9, shape), 0, 255).astype(np.uint8)
