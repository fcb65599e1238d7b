"""Generate tests/golden/fnms_lists.npz by running the REAL reference's fast_non_max_suppression (utils_bbox.py:50-100,
imported from /root/reference) on the list structures the single-CTA Fast NMS kernel treats differently: many short class
lists, lists longer than top_k, a few long lists, a full 4096-candidate buffer, a handful of candidates.

    PYTHONDONTWRITEBYTECODE=1 python -m oracle.gen_golden_fnms     (build container only)

Inputs are re-drawn from the recorded seeds by the tests (`fnms_case_inputs`) and checked against a stored digest."""
from __future__ import annotations

import numpy as np
import torch

from .gen_golden import _ref, _save, digest

# (n rows, c foreground classes, top_k, max_detections)
CASES = [(40, 80, 200, 100), (60, 30, 20, 50), (1000, 4, 200, 100), (300, 10, 50, 20), (4096, 1, 100, 100), (7, 3, 200, 100)]


def fnms_case_inputs(n, c):
    """box [n,4] (heavily overlapping), score [n,c], coef [n,32]: torch CPU generator, seed 7 n + c."""
    g = torch.Generator().manual_seed(7 * n + c)
    ctr = torch.rand(n, 2, generator=g)
    wh = torch.rand(n, 2, generator=g) * 0.2 + 0.1
    box = torch.cat([ctr - wh / 2, ctr + wh / 2], 1).contiguous()
    score = torch.rand(n, c, generator=g).contiguous()
    coef = torch.tanh(torch.randn(n, 32, generator=g))
    return box, score, coef


def main():
    bu = _ref()
    out = {}
    for i, (n, c, top_k, max_det) in enumerate(CASES):
        box, score, coef = fnms_case_inputs(n, c)
        r = bu.fast_non_max_suppression(box, score, coef, 0.5, top_k, max_det)
        out[f"c{i}_params"] = np.array([n, c, top_k, max_det])
        out[f"c{i}_digest"] = digest(box.numpy(), score.numpy(), coef.numpy())
        out[f"c{i}_box"], out[f"c{i}_score"], out[f"c{i}_cls"], out[f"c{i}_coef"] = [t.numpy() for t in r]
    _save("fnms_lists", **out)


if __name__ == "__main__":
    main()
