"""Seeded synthetic head outputs for the YOLACT / Mask R-CNN post-processing path.

Test/bench infrastructure (see oracle/__init__.py).  The generator follows SURVEY.md
section 8(d): a background-dominated softmax with a long noise tail plus a few "objects",
each a run of neighbouring anchors boosted on one class so that Fast NMS has real work.
Everything is drawn from a CPU `torch.Generator`, so the same seed gives the same tensors
in the build container and on the GPU box (same torch build).
"""
from __future__ import annotations

import numpy as np
import torch

from .yolact_np import get_anchors


def yolact_dims(input_size=550, num_classes=81, num_coef=32):
    """(A, P) for a square YOLACT input: 19248/138 at 550, 18525/136 at 544."""
    anchors = get_anchors((input_size, input_size))
    # ProtoNet upsamples P3 by 2 (reference nets/yolact.py:70-93): P = 2 * size(P3).
    from .yolact_np import get_img_output_length
    p3 = int(get_img_output_length(input_size, input_size)[0][0])
    return anchors.shape[0], 2 * p3


def yolact_heads(seed, num_anchors, num_classes=81, num_coef=32, proto=138, objects=12, run=40,
                 box_std=0.3, proto_hw=None):
    """One image worth of head tensors (torch CPU float32):
    box [A,4], conf [A,C] (softmaxed, column 0 = background), coef [A,K], proto [P,Q,K]."""
    g = torch.Generator().manual_seed(int(seed))
    a, c = num_anchors, num_classes
    box = torch.randn(a, 4, generator=g) * box_std
    logits = torch.randn(a, c, generator=g)
    logits[:, 0] += 7.0
    for _ in range(objects):
        cls = int(torch.randint(1, c, (1,), generator=g))
        r = min(run, a)
        start = int(torch.randint(0, a - r + 1, (1,), generator=g))
        boost = torch.rand(r, generator=g) * 8.0 + 5.0
        logits[start:start + r, cls] += boost
    conf = torch.softmax(logits, dim=1)
    coef = torch.tanh(torch.randn(a, num_coef, generator=g))
    ph, pw = (proto, proto) if proto_hw is None else proto_hw
    prot = torch.relu(torch.randn(ph, pw, num_coef, generator=g))
    return box.contiguous(), conf.contiguous(), coef.contiguous(), prot.contiguous()


def yolact_batch(seed0, batch, num_anchors, **kw):
    """Stack `batch` images (seeds seed0, seed0+1, ...) into [B,...] tensors."""
    per = [yolact_heads(seed0 + i, num_anchors, **kw) for i in range(batch)]
    return tuple(torch.stack([p[k] for p in per]).contiguous() for k in range(4))


def tie_free(conf, confidence, top_k=200, max_det=100):
    """True when no two candidate scores (> confidence) are equal inside a class window or
    across the pool the global top-`max_det` is drawn from.  conf [A,C] numpy/torch."""
    s = np.asarray(conf)[:, 1:]
    pool = []
    for c in range(s.shape[1]):
        v = np.sort(s[:, c][s[:, c] > confidence])[::-1][:top_k]
        if len(np.unique(v)) != len(v):
            return False
        pool.append(v)
    pool = np.concatenate(pool) if pool else np.zeros(0, np.float32)
    return len(np.unique(pool)) == len(pool)


def verseg_boxes(seed, n=5000, num_coef=32):
    """Config-5 stress: n vertebra-like boxes of one class with heavy overlap.
    Returns box [n,4] normalised corners, score [n,1], coef [n,K] (torch CPU float32)."""
    g = torch.Generator().manual_seed(int(seed))
    ctr = torch.rand(n, 2, generator=g)
    w = torch.rand(n, generator=g) * 0.08 + 0.01
    h = torch.rand(n, generator=g) * 0.05 + 0.01
    box = torch.stack([ctr[:, 0] - w / 2, ctr[:, 1] - h / 2, ctr[:, 0] + w / 2, ctr[:, 1] + h / 2], 1)
    score = torch.rand(n, 1, generator=g) * 0.9 + 0.05
    coef = torch.tanh(torch.randn(n, num_coef, generator=g))
    return box.contiguous(), score.contiguous(), coef.contiguous()
