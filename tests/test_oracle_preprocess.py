"""The numpy restatement of the wrappers' pre-processing (oracle/preprocess_np.py: Pillow's 8-bit bicubic resize +
utils.py:32-36 normalisation) against outputs of Pillow / the reference's helpers (tests/golden/preprocess.npz,
oracle/gen_golden_preprocess.py).  Integer path: bit-exact; the float32 input: bit-exact too (the division is done in
float64 and rounded once, like the reference)."""
import hashlib

import numpy as np
import pytest

from conftest import load_golden
from oracle import preprocess_np as P
from oracle import synth


def _sha(a):
    return hashlib.sha1(np.ascontiguousarray(a).tobytes()).hexdigest()


def test_small_cases_bit_exact():
    g = load_golden("preprocess")
    for i in range(int(g["num_small"])):
        img, resized, net_in = g[f"s{i}_in"], g[f"s{i}_resized"], g[f"s{i}_input"]
        out = P.resize_u8(img, resized.shape[1], resized.shape[0])
        assert np.array_equal(out, resized), f"case {i}: {img.shape} -> {resized.shape}"
        assert np.array_equal(P.normalize_chw(out), net_in)
        assert np.array_equal(P.image_to_input(img, resized.shape[:2]), net_in)


@pytest.mark.parametrize("tag", ["full544", "full550", "up"])
def test_full_size_bit_exact(tag):
    g = load_golden("preprocess")
    h, w = [int(v) for v in g[f"{tag}_hw"]]
    shape = tuple(int(v) for v in g[f"{tag}_shape"])
    img = synth.picture(int(g[f"{tag}_seed"]), h, w)
    assert _sha(img) == str(g[f"{tag}_in_sha"]), "synthetic picture differs from the one the golden was made from"
    resized = P.resize_u8(img, shape[1], shape[0])
    assert np.array_equal(resized[::9, ::7], g[f"{tag}_resized_sub"])
    assert _sha(resized) == str(g[f"{tag}_resized_sha"])
    net_in = P.normalize_chw(resized)
    assert net_in.dtype == np.float32 and net_in.shape == (1, 3) + shape
    assert np.array_equal(net_in[:, :, ::9, ::7], g[f"{tag}_input_sub"])
    assert _sha(net_in) == str(g[f"{tag}_input_sha"])


def test_coefficients_known_answers():
    """Resample.c precompute_coeffs: identity scale gives the 5-tap window with a single unit coefficient; a 2:1
    reduction has support 4 (9 taps) and every row of fixed-point coefficients sums to 2^22 within rounding."""
    ksize, bounds, kk = P.precompute_coeffs(8, 8)
    assert ksize == 5
    for xx in range(8):
        row = np.zeros(8, np.int64)
        row[bounds[xx, 0]:bounds[xx, 0] + bounds[xx, 1]] = kk[xx, :bounds[xx, 1]]
        assert row[xx] == 1 << P.PRECISION_BITS and np.count_nonzero(row) == 1
    ksize, bounds, kk = P.precompute_coeffs(1080, 544)
    assert ksize == 9
    assert np.all(np.abs(kk.sum(1) - (1 << P.PRECISION_BITS)) <= ksize)
    assert bounds[:, 0].min() == 0 and (bounds[:, 0] + bounds[:, 1]).max() == 1080


def test_against_pillow_when_available():
    Image = pytest.importorskip("PIL.Image")
    rng = np.random.RandomState(3)
    for (h, w), (ow, oh) in [((45, 61), (30, 22)), ((16, 16), (37, 41)), ((211, 97), (97, 211))]:
        img = rng.randint(0, 256, (h, w, 3), dtype=np.uint8)
        ref = np.array(Image.fromarray(img).resize((ow, oh), Image.BICUBIC))
        assert np.array_equal(P.resize_u8(img, ow, oh), ref)
