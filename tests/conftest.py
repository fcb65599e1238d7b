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


def check_pasted(pasted, g, prefix, sub_step, tol=1e-5, what="paste"):
    """Compare a full pasted-mask tensor [m,1,H,W] (numpy) with a golden that stores a sub-sampled copy, per-mask float64
    sums, the exact zero pattern (row / column occupancy + non-zero counts)."""
    sy, sx = sub_step
    key = lambda k: g[(prefix + k) if prefix else k]
    assert_close_f32(pasted[:, :, ::sy, ::sx], key("sub"), tol, what + " sub-sample")
    nz = pasted != 0
    assert np.array_equal(nz.sum(axis=(1, 2, 3)), key("nonzero")), what + ": non-zero counts"
    assert np.array_equal(np.packbits(nz.any(axis=3)), key("rows")), what + ": occupied rows"
    assert np.array_equal(np.packbits(nz.any(axis=2)), key("cols")), what + ": occupied columns"
    sums = pasted.astype(np.float64).sum(axis=(1, 2, 3))
    ref = key("sum" if prefix else "sums")
    assert np.all(np.abs(sums - ref) <= 1e-5 * np.maximum(np.abs(ref), 1.0)), what + ": per-mask sums"


# ------------------------------------------------------------------ guard bands (stand-in for compute-sanitizer on this pool)
BAND = 4096
BAND_BYTE = 0xA5


class BandedAllocator:
    """Drop-in for the `torch` name inside an isb module: `empty()` hands out views into buffers that carry BAND bytes of a
    known pattern before and after the payload, so that a kernel writing outside its output is caught by `check()`."""

    def __init__(self):
        import torch
        self._torch = torch
        self.raws = []

    def __getattr__(self, name):
        return getattr(self._torch, name)

    def empty(self, *shape, dtype=None, device=None, **kw):
        torch = self._torch
        if len(shape) == 1 and isinstance(shape[0], (tuple, list, torch.Size)):
            shape = tuple(shape[0])
        dtype = dtype or torch.float32
        if device is None or torch.device(device).type != "cuda":
            return torch.empty(shape, dtype=dtype, device=device, **kw)
        n = int(np.prod(shape, dtype=np.int64)) * torch.empty((), dtype=dtype).element_size()
        pad = (n + 15) // 16 * 16
        raw = torch.full((BAND + pad + BAND,), BAND_BYTE, dtype=torch.uint8, device=device)
        self.raws.append((raw, pad))
        return raw[BAND:BAND + n].view(dtype).view(shape)

    def check(self):
        assert self.raws, "no banded allocation was made"
        for i, (raw, pad) in enumerate(self.raws):
            assert bool((raw[:BAND] == BAND_BYTE).all()), f"allocation {i}: bytes BEFORE the output were overwritten"
            assert bool((raw[BAND + pad:] == BAND_BYTE).all()), f"allocation {i}: bytes AFTER the output were overwritten"
