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


def unpack_masks(g):
    shape = tuple(int(v) for v in g["masks_shape"])
    n = int(np.prod(shape))
    return np.unpackbits(g["masks_packed"])[:n].reshape(shape).astype(bool)


def assert_close_f32(actual, expected, tol=1e-5, what="", scale=None):
    """|a-e| <= tol * max(|e|, 1[, scale]): the 1e-5 relative bar with an absolute floor.
    `scale` (broadcastable) raises the floor where a result is a difference of large terms, e.g. a box
    corner = centre - width/2: there the error is relative to the box extent, not to the corner value."""
    actual = np.asarray(actual, dtype=np.float64)
    expected = np.asarray(expected, dtype=np.float64)
    assert actual.shape == expected.shape, f"{what}: shape {actual.shape} vs {expected.shape}"
    if actual.size == 0:
        return
    den = np.maximum(np.abs(expected), 1.0)
    if scale is not None:
        den = np.maximum(den, np.broadcast_to(np.asarray(scale, dtype=np.float64), den.shape))
    err = np.abs(actual - expected) / den
    assert np.nanmax(err) <= tol, f"{what}: max scaled error {np.nanmax(err):.3e} > {tol}"


def assert_masks_close(actual, expected, frac=1e-3, what=""):
    """Binarised masks may differ in at most 0.1 % of pixels (BASELINE.json north_star)."""
    assert actual.shape == expected.shape, f"{what}: shape {actual.shape} vs {expected.shape}"
    if actual.size == 0:
        return
    bad = float(np.mean(actual != expected))
    assert bad <= frac, f"{what}: {bad:.5%} of mask pixels differ"


@pytest.fixture(scope="session")
def golden():
    return load_golden
