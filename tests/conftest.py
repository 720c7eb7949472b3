import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden_dir():
    return os.path.join(ROOT, "tests", "golden")


def assert_pgrad_matches_fp64(got, g, key, rel=1e-4):
    """Parameter gradient `got` of parameter `key` against the fp64 run of the REFERENCE class stored in the fixture
    (``pgrad64_*``: unclamped, tests/golden/make_golden.py), clamped to +-0.1 like the reference's gradient hook
    (voxel_perturber.py:465-475).  The reference's own fp32 run agrees with its fp64 run to 1e-7, so the bound is
    the stated bar: 1e-4 of the tensor's largest entry, with an absolute floor of 1e-6 (a Linear bias in front of a
    BatchNorm has a gradient of exactly zero; fp32 leaves 1e-8 of rounding noise there)."""
    import numpy as np
    ref = np.clip(np.nan_to_num(np.asarray(g[f"pgrad64_{key}"], np.float64), nan=0.0), -0.1, 0.1)
    scale = max(float(np.abs(ref).max()), 1e-6)
    np.testing.assert_allclose(np.asarray(got, np.float64), ref, rtol=rel, atol=max(rel * scale, 1e-6), err_msg=key)
