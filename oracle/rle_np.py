"""CPU restatement of COCO run-length encoding (test infrastructure; see oracle/__init__.py).

The reference encodes every instance mask with `pycocotools.mask.encode(np.asfortranarray(mask.astype(np.uint8)))`
(inst_model/yolact/utils/utils_map.py:33-34).  pycocotools is a third-party dependency that is NOT vendored under
/root/reference and is NOT installed in this image, and the reference pins no version (no requirements file); the
functions below restate the published algorithm of cocoapi 2.0 `common/maskApi.c` (rleEncode, rleDecode, rleToString,
rleFrString).  PARITY UNPINNED for this row: there is no reference output to compare with here; the tests rely on
hand-derived known answers and on encode -> string -> parse -> decode round trips.
"""
from __future__ import annotations

import numpy as np


def rle_counts(mask):
    """maskApi.c rleEncode for one [H,W] mask: column-major scan, lengths of alternating runs, zeros first."""
    flat = np.asarray(mask).astype(np.uint8).reshape(mask.shape[0], mask.shape[1]).flatten(order="F")
    if flat.size == 0:
        return np.zeros(0, np.int64)
    change = np.nonzero(flat[1:] != flat[:-1])[0] + 1
    edges = np.concatenate([[0], change, [flat.size]])
    counts = np.diff(edges)
    if flat[0] != 0:
        counts = np.concatenate([[0], counts])
    return counts.astype(np.int64)


def rle_to_string(counts):
    """maskApi.c rleToString: from the 4th count on the difference to the count two places back is stored; each value is a
    little-endian sequence of 5-bit groups, 0x20 = more follow, + 48."""
    out = bytearray()
    cnts = [int(c) for c in counts]
    for i, x in enumerate(cnts):
        if i > 2:
            x -= cnts[i - 2]
        more = True
        while more:
            c = x & 0x1F
            x >>= 5
            more = (x != -1) if (c & 0x10) else (x != 0)
            if more:
                c |= 0x20
            out.append(c + 48)
    return bytes(out)


def rle_from_string(s):
    """maskApi.c rleFrString."""
    cnts = []
    p = 0
    while p < len(s):
        x, k, more = 0, 0, True
        while more:
            c = s[p] - 48
            x |= (c & 0x1F) << (5 * k)
            more = bool(c & 0x20)
            p += 1
            k += 1
            if not more and (c & 0x10):
                x |= -1 << (5 * k)
        if len(cnts) > 2:
            x += cnts[-2]
        cnts.append(x)
    return np.asarray(cnts, np.int64)


def rle_decode(counts, h, w):
    """maskApi.c rleDecode -> bool [H,W]."""
    flat = np.zeros(h * w, np.uint8)
    pos, v = 0, 0
    for c in counts:
        flat[pos:pos + int(c)] = v
        pos += int(c)
        v ^= 1
    assert pos == h * w
    return flat.reshape((h, w), order="F").astype(bool)


def encode(masks):
    """[H,W,m] -> list of {'size', 'counts'} like pycocotools.mask.encode on a Fortran-ordered [H,W,m] array."""
    masks = np.asarray(masks)
    return [{"size": [int(masks.shape[0]), int(masks.shape[1])], "counts": rle_to_string(rle_counts(masks[:, :, j]))}
            for j in range(masks.shape[2])]
