"""CPU: the RLE oracle (cocoapi maskApi.c restated; parity unpinned -- pycocotools is absent) against hand-derived known
answers and round trips, and the library's host-side string encoder against the oracle."""
import ctypes as C

import numpy as np

from oracle import rle_np as R


def test_known_answers():
    # 3x3, column-major: col0 = [0,1,1], col1 = [1,0,0], col2 = [0,0,1]  ->  0 | 111 | 0000 | 1
    m = np.array([[0, 1, 0], [1, 0, 0], [1, 0, 1]], bool)
    assert R.rle_counts(m).tolist() == [1, 3, 4, 1]
    assert R.rle_counts(np.ones((2, 3), bool)).tolist() == [0, 6]
    assert R.rle_counts(np.zeros((2, 3), bool)).tolist() == [6]
    # string encoding: small values are one char ('0' + x); 4th count is stored as a difference to the 2nd
    assert R.rle_to_string([1, 3, 4, 1]) == bytes([48 + 1, 48 + 3, 48 + 4, 48 + ((1 - 3) & 0x1F)])
    # 32 needs two groups: low 5 bits 0 with the continuation bit, then 1
    assert R.rle_to_string([32]) == bytes([48 + 0x20, 48 + 1])
    # 16 has bit 0x10 set in its only group, so a second (zero) group must follow to keep the sign positive
    assert R.rle_to_string([16]) == bytes([48 + 0x30, 48 + 0])


def test_round_trips():
    g = np.random.default_rng(7)
    for h, w in ((1, 1), (5, 7), (64, 48), (550, 550)):
        blob = g.random((h, w)) > 0.5 if h < 100 else np.zeros((h, w), bool)
        if h >= 100:
            blob[100:400, 50:300] = True
            blob[200:250, 100:150] = False
        cnt = R.rle_counts(blob)
        assert cnt.sum() == h * w
        s = R.rle_to_string(cnt)
        back = R.rle_from_string(s)
        assert np.array_equal(back, cnt)
        assert np.array_equal(R.rle_decode(back, h, w), blob)


def test_library_string_encoder_matches_oracle():
    import __graft_entry__ as g
    g.build()
    from isb import _lib
    lib = _lib.load()
    rng = np.random.default_rng(11)
    for n in (1, 2, 3, 4, 50, 1000):
        cnt = rng.integers(0, 3000, n).astype(np.uint32)
        cnt[rng.integers(0, n)] = 302500
        buf = C.create_string_buffer(7 * n + 8)
        k = lib.isb_rle_to_string(cnt.ctypes.data_as(C.c_void_p), n, buf, len(buf))
        assert k > 0 and buf.raw[:k] == R.rle_to_string(cnt)
    assert lib.isb_rle_to_string(cnt.ctypes.data_as(C.c_void_p), n, buf, 3) == -2
