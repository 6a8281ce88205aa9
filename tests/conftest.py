"""pytest configuration: registers the ``gpu`` marker and puts the repo root on sys.path.

``-m "not gpu"`` runs here on CPU (oracle vs golden vectors, host logic, C-ABI symbol
checks, gloo multi-process tests); ``-m gpu`` runs on a B200 and calls the CUDA path through
the C-ABI library.
"""
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)


@pytest.fixture(scope="session")
def golden():
    return load_golden


def assert_close_rel(new, ref, tol=1e-5, what=""):
    """SURVEY.md section 8(c) tolerance: max|new-ref| <= tol * max|ref| (plus elementwise
    rtol=tol, atol=tol*max|ref|)."""
    new = np.asarray(new, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    assert new.shape == ref.shape, f"{what}: shape {new.shape} != {ref.shape}"
    if ref.size == 0:
        return
    scale = max(float(np.max(np.abs(ref))), 1e-30)
    err = float(np.max(np.abs(new - ref)))
    assert err <= tol * scale, f"{what}: max abs err {err:.3e} > {tol:g} * max|ref| {scale:.3e}"
