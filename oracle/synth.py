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
                 box_std=0.3, proto_hw=None, return_logits=False):
    """One image worth of head tensors (torch CPU float32):
    box [A,4], conf [A,C] (softmaxed, column 0 = background), coef [A,K], proto [P,Q,K].
    return_logits=True puts the raw class logits in place of conf (same random draws)."""
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
    return box.contiguous(), (logits if return_logits else conf).contiguous(), coef.contiguous(), prot.contiguous()


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


# --------------------------------------------------------------------------- Mask R-CNN side
FPN_STRIDES = (4, 8, 16, 32, 64)


def fpn_shapes(image_hw=(800, 1344), strides=FPN_STRIDES):
    """Feature sizes of '0'..'3','pool' for a padded image (ceil division, as conv/pool strides give)."""
    h, w = image_hw
    out = []
    for s in strides:
        out.append((-(-h // s), -(-w // s)))
    return out


def rpn_anchors(image_hw=(800, 1344), sizes=(32, 64, 128, 256, 512), ratios=(0.5, 1.0, 2.0)):
    """Plausible RPN anchor grid (level-major, cell-major, ratio-minor): [A,4] float32 corners and
    anchors-per-level.  800x1344 gives 201600/50400/12600/3150/819 = 268569 anchors."""
    per_level, out = [], []
    for (fh, fw), stride, size in zip(fpn_shapes(image_hw), FPN_STRIDES, sizes):
        r = torch.tensor(ratios)
        hr = torch.sqrt(r)
        ws, hs = size / hr, size * hr
        base = torch.stack([-ws, -hs, ws, hs], 1) / 2
        ys, xs = torch.meshgrid(torch.arange(fh) * stride, torch.arange(fw) * stride, indexing="ij")
        shift = torch.stack([xs, ys, xs, ys], -1).reshape(-1, 1, 4).float()
        a = (shift + base[None]).reshape(-1, 4)
        out.append(a)
        per_level.append(a.shape[0])
    return torch.cat(out).contiguous(), per_level


def rpn_heads(seed, image_hw=(800, 1344), objects=25):
    """Synthetic RPN outputs for one image: anchors [A,4], deltas [A,4], objectness logits [A],
    anchors_per_level.  Logits are noise around -4 with `objects` clusters of boosted anchors, so the
    per-level top-1000 are spatially clustered and NMS has real work."""
    g = torch.Generator().manual_seed(int(seed))
    anchors, per_level = rpn_anchors(image_hw)
    a = anchors.shape[0]
    deltas = torch.randn(a, 4, generator=g) * torch.tensor([0.1, 0.1, 0.2, 0.2])
    logits = torch.randn(a, generator=g) * 1.5 - 4.0
    ctr = (anchors[:, :2] + anchors[:, 2:]) / 2
    for _ in range(objects):
        c = torch.rand(2, generator=g) * torch.tensor([float(image_hw[1]), float(image_hw[0])])
        rad = float(torch.rand(1, generator=g)) * 120 + 30
        near = ((ctr - c).abs().max(dim=1).values < rad)
        logits[near] += torch.rand(int(near.sum()), generator=g) * 6.0 + 2.0
    return anchors, deltas.contiguous(), logits.contiguous(), per_level


def roi_boxes(seed, n, image_hw=(800, 1333)):
    """n plausible proposal boxes inside the image, log-uniform sizes 8..700 px (all FPN levels)."""
    g = torch.Generator().manual_seed(int(seed))
    h, w = image_hw
    ctr = torch.rand(n, 2, generator=g) * torch.tensor([float(w), float(h)])
    size = torch.exp(torch.rand(n, 2, generator=g) * 4.5 + 2.0)
    b = torch.cat([ctr - size / 2, ctr + size / 2], 1)
    b[:, 0::2] = b[:, 0::2].clamp(0, w)
    b[:, 1::2] = b[:, 1::2].clamp(0, h)
    keep = ((b[:, 2] - b[:, 0]) >= 2) & ((b[:, 3] - b[:, 1]) >= 2)
    b[~keep] = torch.tensor([10.0, 10.0, 60.0, 80.0])
    return b.contiguous()


def fpn_features(seed, channels=256, image_hw=(800, 1344), levels=4):
    """Synthetic FPN maps '0'..'3': list of [1,C,H_l,W_l] float32."""
    g = torch.Generator().manual_seed(int(seed))
    return [torch.randn(1, channels, fh, fw, generator=g) for (fh, fw) in fpn_shapes(image_hw)[:levels]]


def box_head_outputs(seed, n=1000, num_classes=21, image_hw=(800, 1333), hot=60):
    """Synthetic box-head outputs: proposals [n,4], class_logits [n,C], box_regression [n,4C]."""
    g = torch.Generator().manual_seed(int(seed))
    proposals = roi_boxes(seed + 1, n, image_hw)
    # clustered duplicates so per-class NMS suppresses
    src = torch.randint(0, max(n // 8, 1), (n,), generator=g)
    proposals = (proposals[src] + torch.randn(n, 4, generator=g) * 4.0)
    proposals = torch.cat([torch.min(proposals[:, :2], proposals[:, 2:] - 2), proposals[:, 2:]], 1).contiguous()
    logits = torch.randn(n, num_classes, generator=g)
    logits[:, 0] += 3.0
    rows = torch.randint(0, n, (hot * 4,), generator=g)
    cls = torch.randint(1, num_classes, (hot * 4,), generator=g)
    logits[rows, cls] += torch.rand(hot * 4, generator=g) * 6.0 + 2.0
    reg = torch.randn(n, 4 * num_classes, generator=g) * torch.tensor([1.0, 1.0, 0.5, 0.5]).repeat(num_classes)
    return proposals, logits.contiguous(), reg.contiguous()


def picture(seed, h, w):
    """Seeded photo-like uint8 [h, w, 3] picture: linear colour ramps, a few hard-edged rectangles and integer noise.
    Only integer draws from numpy's RandomState and exact integer arithmetic, so the GPU box re-draws the same
    bytes whatever its CPU / libm.  Input of the pre-processing row."""
    rs = np.random.RandomState(int(seed))
    yy, xx = np.mgrid[0:h, 0:w].astype(np.int64)
    img = np.zeros((h, w, 3), np.int64)
    for c in range(3):
        ay, ax, base = rs.randint(-200, 201), rs.randint(-200, 201), rs.randint(60, 196)
        img[..., c] = base + (ay * yy) // max(h, 1) + (ax * xx) // max(w, 1)
    for _ in range(12):
        y0, x0 = rs.randint(0, max(h, 2)), rs.randint(0, max(w, 2))
        y1, x1 = y0 + rs.randint(1, max(h // 3, 2)), x0 + rs.randint(1, max(w // 3, 2))
        img[y0:y1, x0:x1] = rs.randint(0, 256, 3)
    img += rs.randint(-8, 9, img.shape)
    return np.clip(img, 0, 255).astype(np.uint8)


# --------------------------------------------------------------------------- Mask R-CNN chain test
class ChainBoxHead(torch.nn.Module):
    """Stand-in for TwoMLPHead (faster_rcnn_framework.py:203): spatial mean of the 7x7 RoI features in float64, so the
    result does not depend on the device's summation order beyond one final rounding.  [R,256,7,7] -> [R,256]."""

    def forward(self, x):
        return x.double().mean(dim=(2, 3)).float()


class ChainBoxPredictor(torch.nn.Module):
    """Stand-in for FastRCNNPredictor (faster_rcnn_framework.py:226): element-wise maps of selected channels -- no
    matmul, so CPU and GPU agree to the last bit given equal inputs.  -> (class_logits [R,C], box_regression [R,4C])."""

    def __init__(self, num_classes):
        super().__init__()
        self.c = int(num_classes)

    def forward(self, x):
        c = self.c
        logits = x[:, :c] * 14.0
        logits = torch.cat([logits[:, :1] + 2.0, logits[:, 1:]], 1)
        reg = x[:, c:5 * c] * torch.tensor([6.0, 6.0, 3.0, 3.0], device=x.device).repeat(c)
        return logits.contiguous(), reg.contiguous()


class ChainMaskPredictor(torch.nn.Module):
    """Stand-in for MaskRCNNHeads + MaskRCNNPredictor (mask_rcnn.py:62-105): nearest 2x up-sampling of the first C
    channels of the 14x14 RoI features, scaled.  [m,256,14,14] -> [m,C,28,28]."""

    def __init__(self, num_classes):
        super().__init__()
        self.c = int(num_classes)

    def forward(self, x):
        return (x[:, :self.c].repeat_interleave(2, dim=2).repeat_interleave(2, dim=3) * 3.0).contiguous()


def chain_inputs(seed, image_hw=(800, 1344)):
    """Inputs of the chained Mask R-CNN test at configs[3] size: FPN features, RPN anchors / deltas / logits."""
    feats = fpn_features(seed, 256, image_hw)
    anchors, deltas, logits, per_level = rpn_heads(seed + 1, image_hw)
    return feats, anchors, deltas, logits, per_level
