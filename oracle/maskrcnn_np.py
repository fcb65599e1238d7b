"""numpy restatement of the Mask R-CNN post-processing path (oracle; test infrastructure only).

Follows /root/reference/inst_model/Mask_RCNN/net/network_files/{det_utils,boxes,rpn_function,
roi_head,transform}.py.  Two pieces of arithmetic live in a third-party dependency that is NOT under
/root/reference: torchvision's `nms` / `roi_align` C++ kernels and `MultiScaleRoIAlign` glue.  The
reference pins no version (README says "torchvision 0.4", which cannot run this code); the oracle
restates the published algorithm of the installed torchvision 0.26.0 (ops/poolers.py is readable;
the kernels' algorithm is the well-known Detectron one) and is anchored on outputs of the real ops
generated in the build container (tests/golden/maskrcnn_*.npz, oracle/gen_golden.py).

Tie rule as in yolact_np: stable sorts (score descending, then index ascending) -- torchvision's CPU
nms sorts with stable=True, torch.topk has no tie contract.
"""
from __future__ import annotations

import math

import numpy as np

F32 = np.float32
BBOX_XFORM_CLIP = math.log(1000.0 / 16)   # det_utils.py:146


def _desc_stable(x):
    from .yolact_np import _argsort_desc_stable
    return _argsort_desc_stable(np.asarray(x), axis=0)    # NaN first, like torch.sort(descending=True)


# --------------------------------------------------------------------------- M1
def decode_single(rel_codes, boxes, weights=(1.0, 1.0, 1.0, 1.0), clip=BBOX_XFORM_CLIP):
    """det_utils.py:224-267 -- rel_codes [N, 4k], boxes [N, 4] -> [N, 4k] (x1,y1,x2,y2 per k)."""
    rel = np.asarray(rel_codes, dtype=F32)
    b = np.asarray(boxes, dtype=F32)
    widths = b[:, 2] - b[:, 0]
    heights = b[:, 3] - b[:, 1]
    ctr_x = b[:, 0] + F32(0.5) * widths
    ctr_y = b[:, 1] + F32(0.5) * heights
    wx, wy, ww, wh = [F32(w) for w in weights]
    dx = rel[:, 0::4] / wx
    dy = rel[:, 1::4] / wy
    dw = np.minimum(rel[:, 2::4] / ww, F32(clip))
    dh = np.minimum(rel[:, 3::4] / wh, F32(clip))
    pcx = dx * widths[:, None] + ctr_x[:, None]
    pcy = dy * heights[:, None] + ctr_y[:, None]
    pw = np.exp(dw).astype(F32) * widths[:, None]
    ph = np.exp(dh).astype(F32) * heights[:, None]
    half = F32(0.5)
    out = np.stack([pcx - half * pw, pcy - half * ph, pcx + half * pw, pcy + half * ph], axis=2)
    return out.reshape(rel.shape[0], -1).astype(F32)


# --------------------------------------------------------------------------- boxes.py
def clip_boxes_to_image(boxes, size):
    """boxes.py:108-138 -- clamp x to [0, W], y to [0, H]."""
    h, w = size
    b = np.asarray(boxes, dtype=F32).copy()
    b[..., 0::2] = np.clip(b[..., 0::2], F32(0), F32(w))
    b[..., 1::2] = np.clip(b[..., 1::2], F32(0), F32(h))
    return b


def remove_small_boxes(boxes, min_size):
    """boxes.py:86-105 -- indices with w >= min_size and h >= min_size."""
    b = np.asarray(boxes, dtype=F32)
    ws, hs = b[:, 2] - b[:, 0], b[:, 3] - b[:, 1]
    return np.where((ws >= F32(min_size)) & (hs >= F32(min_size)))[0]


def nms(boxes, scores, iou_threshold):
    """boxes.py:7-35 -> torch.ops.torchvision.nms (CPU kernel algorithm, torchvision 0.26).

    Greedy: visit boxes by stable-descending score; a visited, unsuppressed box suppresses every later
    box whose IoU with it is > threshold.  IoU = inter / ((area_i + area_j) - inter) in float32 with
    inter = max(0, xx2-xx1) * max(0, yy2-yy1); the comparison is against the threshold as a double.
    max / min are the C++ std::max(a, b) = (a < b) ? b : a and std::min(a, b) = (b < a) ? b : a with a = the visited
    (higher-ranked) box, which fixes what a NaN coordinate does; `ovr > thr` is false for a NaN overlap, so NaN never
    suppresses here (the opposite of Fast NMS).
    Returns kept indices (int64) in decreasing score order.
    """
    b = np.asarray(boxes, dtype=F32)
    s = np.asarray(scores, dtype=F32)
    n = b.shape[0]
    if n == 0:
        return np.zeros((0,), np.int64)
    x1, y1, x2, y2 = b[:, 0], b[:, 1], b[:, 2], b[:, 3]
    areas = (x2 - x1) * (y2 - y1)
    order = _desc_stable(s)
    suppressed = np.zeros(n, bool)
    keep = []
    thr = np.float64(iou_threshold)
    for pos in range(n):
        i = order[pos]
        if suppressed[i]:
            continue
        keep.append(i)
        rest = order[pos + 1:]
        with np.errstate(invalid="ignore"):
            xx1 = np.where(x1[i] < x1[rest], x1[rest], x1[i])
            yy1 = np.where(y1[i] < y1[rest], y1[rest], y1[i])
            xx2 = np.where(x2[rest] < x2[i], x2[rest], x2[i])
            yy2 = np.where(y2[rest] < y2[i], y2[rest], y2[i])
            dw, dh = xx2 - xx1, yy2 - yy1
            w = np.where(F32(0) < dw, dw, F32(0))
            h = np.where(F32(0) < dh, dh, F32(0))
            inter = w * h
        with np.errstate(divide="ignore", invalid="ignore"):
            ovr = inter / ((areas[i] + areas[rest]) - inter)
        suppressed[rest[ovr.astype(np.float64) > thr]] = True
    return np.asarray(keep, dtype=np.int64)


def batched_nms(boxes, scores, idxs, iou_threshold):
    """boxes.py:38-83 -- offset every box by idx * (max_coordinate + 1) in float32, then nms."""
    b = np.asarray(boxes, dtype=F32)
    if b.size == 0:
        return np.zeros((0,), np.int64)
    max_coordinate = b.max()                                   # propagates NaN like torch's boxes.max()
    with np.errstate(invalid="ignore", over="ignore"):
        offsets = np.asarray(idxs).astype(F32) * (max_coordinate + F32(1))
        shifted = b + offsets[:, None]
    return nms(shifted, scores, iou_threshold)


# --------------------------------------------------------------------------- M2, M3
def get_top_n_idx(objectness, num_anchors_per_level, pre_nms_top_n):
    """rpn_function.py:434-458 -- per-level top-k indices (descending score), concatenated."""
    ob = np.asarray(objectness, dtype=F32)
    out, offset = [], 0
    for n in num_anchors_per_level:
        k = min(pre_nms_top_n, n)
        out.append(_desc_stable(ob[offset:offset + n])[:k] + offset)
        offset += n
    return np.concatenate(out) if out else np.zeros((0,), np.int64)


def sigmoid(x):
    x = np.asarray(x, dtype=F32)
    with np.errstate(over="ignore"):
        return (F32(1) / (F32(1) + np.exp(-x).astype(F32))).astype(F32)


def filter_proposals(proposals, objectness, image_shape, num_anchors_per_level, pre_nms_top_n=1000,
                     post_nms_top_n=1000, nms_thresh=0.7, score_thresh=0.0, min_size=1.0):
    """rpn_function.py:460-529 for one image: proposals [A,4] (decoded), objectness [A] (logits).
    Returns (boxes [k,4], scores [k])."""
    proposals = np.asarray(proposals, dtype=F32)
    objectness = np.asarray(objectness, dtype=F32).reshape(-1)
    levels = np.concatenate([np.full((n,), i, np.int64) for i, n in enumerate(num_anchors_per_level)])
    top = get_top_n_idx(objectness, num_anchors_per_level, pre_nms_top_n)          # :491
    scores = sigmoid(objectness[top])                                              # :502
    lvl = levels[top]
    boxes = clip_boxes_to_image(proposals[top], image_shape)                       # :509
    keep = remove_small_boxes(boxes, min_size)                                     # :512
    boxes, scores, lvl = boxes[keep], scores[keep], lvl[keep]
    keep = np.where(scores >= F32(score_thresh))[0]                                # :517
    boxes, scores, lvl = boxes[keep], scores[keep], lvl[keep]
    keep = batched_nms(boxes, scores, lvl, nms_thresh)[:post_nms_top_n]            # :521-524
    return boxes[keep], scores[keep]


# --------------------------------------------------------------------------- M6
def softmax(x):
    x = np.asarray(x, dtype=F32)
    e = np.exp(x - x.max(axis=-1, keepdims=True)).astype(F32)
    return (e / e.sum(axis=-1, keepdims=True, dtype=F32)).astype(F32)


def postprocess_detections(class_logits, box_regression, proposals, image_shape, weights=(10.0, 10.0, 5.0, 5.0),
                           score_thresh=0.05, nms_thresh=0.5, detections_per_img=100, scores=None):
    """roi_head.py:366-457 for one image.  `scores` may carry precomputed softmax probabilities (so a
    test can feed both sides identical fp32 scores)."""
    logits = np.asarray(class_logits, dtype=F32)
    num_classes = logits.shape[-1]
    pred_boxes = decode_single(box_regression, proposals, weights).reshape(-1, num_classes, 4)   # :398
    pred_scores = softmax(logits) if scores is None else np.asarray(scores, dtype=F32)          # :401
    boxes = clip_boxes_to_image(pred_boxes, image_shape)                                         # :413
    labels = np.broadcast_to(np.arange(num_classes, dtype=np.int64)[None, :], pred_scores.shape)
    boxes, sc, labels = boxes[:, 1:].reshape(-1, 4), pred_scores[:, 1:].reshape(-1), labels[:, 1:].reshape(-1)
    inds = np.where(sc > F32(score_thresh))[0]                                                   # :437
    boxes, sc, labels = boxes[inds], sc[inds], labels[inds]
    keep = remove_small_boxes(boxes, 1.0)                                                        # :442
    boxes, sc, labels = boxes[keep], sc[keep], labels[keep]
    keep = batched_nms(boxes, sc, labels, nms_thresh)[:detections_per_img]                       # :447-451
    return boxes[keep], sc[keep], labels[keep]


# --------------------------------------------------------------------------- M5
def infer_scale(feat_hw, image_hw):
    """poolers.py `_infer_scale`: 2 ** round(log2(feat / img)) from the first (height) axis."""
    approx = float(feat_hw[0]) / float(image_hw[0])
    return 2.0 ** float(np.round(np.log2(np.float32(approx))))


def level_mapper(boxes, k_min, k_max, canonical_scale=224, canonical_level=4, eps=1e-6):
    """poolers.py `LevelMapper.__call__`: floor(lvl0 + log2(sqrt(area)/s0) + eps) clamped, minus k_min."""
    b = np.asarray(boxes, dtype=F32)
    area = (b[:, 2] - b[:, 0]) * (b[:, 3] - b[:, 1])
    s = np.sqrt(area).astype(F32)
    with np.errstate(divide="ignore", invalid="ignore"):
        t = np.floor(F32(canonical_level) + np.log2(s / F32(canonical_scale)).astype(F32) + F32(eps))
    t = np.clip(t, F32(k_min), F32(k_max))
    return t.astype(np.int64) - k_min


def roi_align(feat, rois, spatial_scale, out_h, out_w, sampling_ratio):
    """torchvision.ops.roi_align (aligned=False) forward, CPU kernel algorithm.
    feat [N,C,H,W], rois [K,5] (batch_idx, x1,y1,x2,y2) -> [K,C,out_h,out_w]."""
    feat = np.asarray(feat, dtype=F32)
    rois = np.asarray(rois, dtype=F32)
    _, c, height, width = feat.shape
    k = rois.shape[0]
    out = np.zeros((k, c, out_h, out_w), F32)
    ss = F32(spatial_scale)
    for r in range(k):
        bi = int(rois[r, 0])
        sw, sh, ew, eh = rois[r, 1] * ss, rois[r, 2] * ss, rois[r, 3] * ss, rois[r, 4] * ss
        rw, rh = max(ew - sw, F32(1)), max(eh - sh, F32(1))
        bh, bw = F32(rh) / F32(out_h), F32(rw) / F32(out_w)
        gh = sampling_ratio if sampling_ratio > 0 else int(math.ceil(rh / out_h))
        gw = sampling_ratio if sampling_ratio > 0 else int(math.ceil(rw / out_w))
        count = F32(max(gh * gw, 1))
        ph = np.arange(out_h, dtype=F32)[:, None]
        iy = np.arange(gh, dtype=F32)[None, :]
        ys = (sh + ph * bh + ((iy + F32(0.5)) * bh) / F32(gh)).astype(F32).reshape(-1)     # [out_h*gh]
        pw = np.arange(out_w, dtype=F32)[:, None]
        ix = np.arange(gw, dtype=F32)[None, :]
        xs = (sw + pw * bw + ((ix + F32(0.5)) * bw) / F32(gw)).astype(F32).reshape(-1)     # [out_w*gw]

        def prep(v, size):
            bad = (v < F32(-1.0)) | (v > F32(size))
            v = np.where(v <= 0, F32(0), v).astype(F32)
            lo = v.astype(np.int64)
            top = lo >= size - 1
            hi = np.where(top, size - 1, lo + 1)
            lo = np.where(top, size - 1, lo)
            v = np.where(top, lo.astype(F32), v).astype(F32)
            l = (v - lo.astype(F32)).astype(F32)
            return bad, lo, hi, l, (F32(1) - l).astype(F32)

        ybad, ylo, yhi, ly, hy = prep(ys, height)
        xbad, xlo, xhi, lx, hx = prep(xs, width)
        bad = ybad[:, None] | xbad[None, :]
        w1 = np.where(bad, F32(0), hy[:, None] * hx[None, :]).astype(F32)
        w2 = np.where(bad, F32(0), hy[:, None] * lx[None, :]).astype(F32)
        w3 = np.where(bad, F32(0), ly[:, None] * hx[None, :]).astype(F32)
        w4 = np.where(bad, F32(0), ly[:, None] * lx[None, :]).astype(F32)
        f = feat[bi]
        v1 = f[:, ylo[:, None], xlo[None, :]]
        v2 = f[:, ylo[:, None], xhi[None, :]]
        v3 = f[:, yhi[:, None], xlo[None, :]]
        v4 = f[:, yhi[:, None], xhi[None, :]]
        val = ((w1 * v1 + w2 * v2) + w3 * v3) + w4 * v4                                   # [C, out_h*gh, out_w*gw]
        val = val.reshape(c, out_h, gh, out_w, gw)
        acc = np.zeros((c, out_h, out_w), F32)
        for a in range(gh):                       # accumulate in the kernel's (iy, ix) order
            for b_ in range(gw):
                acc = acc + val[:, :, a, :, b_]
        out[r] = acc / count
    return out


def multiscale_roi_align(feats, boxes, image_shape, out_size, sampling_ratio=2, canonical_scale=224, canonical_level=4):
    """torchvision MultiScaleRoIAlign.forward for ONE image: feats = list of [1,C,H_l,W_l] (finest
    first), boxes [K,4], image_shape (H, W).  Returns [K,C,oh,ow]."""
    oh, ow = (out_size, out_size) if isinstance(out_size, int) else out_size
    scales = [infer_scale(f.shape[-2:], image_shape) for f in feats]
    k_min = int(-np.log2(np.float32(scales[0])))
    k_max = int(-np.log2(np.float32(scales[-1])))
    boxes = np.asarray(boxes, dtype=F32)
    rois = np.concatenate([np.zeros((boxes.shape[0], 1), F32), boxes], axis=1)
    if len(feats) == 1:
        return roi_align(feats[0], rois, scales[0], oh, ow, sampling_ratio)
    levels = level_mapper(boxes, k_min, k_max, canonical_scale, canonical_level)
    out = np.zeros((boxes.shape[0], feats[0].shape[1], oh, ow), F32)
    for lvl, (f, sc) in enumerate(zip(feats, scales)):
        idx = np.where(levels == lvl)[0]
        if len(idx):
            out[idx] = roi_align(f, rois[idx], sc, oh, ow, sampling_ratio)
    return out


# --------------------------------------------------------------------------- M7
def maskrcnn_inference(mask_logits, labels):
    """roi_head.py:59-91 -- sigmoid, then pick channel = label: [m,C,S,S] -> [m,1,S,S]."""
    x = sigmoid(mask_logits)
    idx = np.arange(x.shape[0])
    return x[idx, np.asarray(labels, dtype=np.int64)][:, None]


# --------------------------------------------------------------------------- M8, M9
def resize_boxes(boxes, original_size, new_size):
    """transform.py:462-484 -- per-axis ratio new/orig in float32."""
    rh = F32(new_size[0]) / F32(original_size[0])
    rw = F32(new_size[1]) / F32(original_size[1])
    b = np.asarray(boxes, dtype=F32)
    return np.stack([b[:, 0] * rw, b[:, 1] * rh, b[:, 2] * rw, b[:, 3] * rh], axis=1).astype(F32)


def expand_boxes(boxes, scale):
    """transform.py:145-163."""
    b = np.asarray(boxes, dtype=F32)
    w_half = (b[:, 2] - b[:, 0]) * F32(0.5)
    h_half = (b[:, 3] - b[:, 1]) * F32(0.5)
    x_c = (b[:, 2] + b[:, 0]) * F32(0.5)
    y_c = (b[:, 3] + b[:, 1]) * F32(0.5)
    w_half = w_half * F32(scale)
    h_half = h_half * F32(scale)
    return np.stack([x_c - w_half, y_c - h_half, x_c + w_half, y_c + h_half], axis=1).astype(F32)


def paste_masks_in_image(masks, boxes, img_shape, padding=1):
    """transform.py:212-230 (+ expand_masks :171, paste_mask_in_image :182).
    masks [m,1,S,S] float32 probabilities, boxes [m,4] -> float32 [m,1,H,W]."""
    from .yolact_np import upsample_bilinear
    masks = np.asarray(masks, dtype=F32)
    m, s = masks.shape[0], masks.shape[-1]
    im_h, im_w = int(img_shape[0]), int(img_shape[1])
    scale = float(s + 2 * padding) / s                                           # :178
    padded = np.pad(masks, ((0, 0), (0, 0), (padding, padding), (padding, padding)))
    ib = expand_boxes(boxes, scale).astype(np.int64)                             # :218 trunc toward zero
    out = np.zeros((m, 1, im_h, im_w), F32)
    for i in range(m):
        x0b, y0b, x1b, y1b = [int(v) for v in ib[i]]
        w = max(x1b - x0b + 1, 1)                                                # :187-190
        h = max(y1b - y0b + 1, 1)
        res = upsample_bilinear(padded[i], h, w)[0]                              # :197
        x_0, x_1 = max(x0b, 0), min(x1b + 1, im_w)                               # :202-205
        y_0, y_1 = max(y0b, 0), min(y1b + 1, im_h)
        if x_1 > x_0 and y_1 > y_0:
            out[i, 0, y_0:y_1, x_0:x_1] = res[(y_0 - y0b):(y_1 - y0b), (x_0 - x0b):(x_1 - x0b)]
    return out


# --------------------------------------------------------------------------- next row
def masks_class_image(pasted, labels, scores, confidence, image_shape):
    """inst_model/Mask_RCNN/mask_rcnn.py:174-198 -- the wrapper's class image from the model's outputs:
    pasted [m,1,Hp,Wp] float32 (paste_masks_in_image), labels int64 [m], scores [m].  Detections with
    score <= confidence are dropped (:174-181), the rest are resized to image_shape (bilinear, align_corners=False,
    :186), arg-max over detections (:188), > 0.5 (:189), class image = (label + 1) of the arg-max detection where its
    mask is set (:195-197).  Returns (masks_class int64 [H,W], masks_arg int64 [H,W]) or (None, None)."""
    from .yolact_np import masks_class, upsample_bilinear
    keep = np.asarray(scores, dtype=F32) > F32(confidence)
    if not keep.any():
        return None, None
    maps = np.asarray(pasted, dtype=F32)[keep][:, 0]
    classes = np.asarray(labels).astype(np.int64)[keep]
    up = upsample_bilinear(maps, int(image_shape[0]), int(image_shape[1]))    # [m', H, W]
    hwm = np.ascontiguousarray(np.transpose(up, (1, 2, 0)))                   # :187
    arg = np.argmax(hwm, axis=-1)                                             # first maximum, like torch.argmax
    return masks_class(hwm > F32(0.5), classes, arg), arg


# --------------------------------------------------------------------------- next row
def transform_output_size(h, w, min_size, max_size):
    """transform.py:88-97 + F.interpolate(scale_factor=..., recompute_scale_factor=True): `self_min_size / min_size`
    is python-float / float32-tensor, which torch evaluates as reciprocal(tensor) * float in float32
    (Tensor.__rtruediv__); the scale is the float32 minimum of the two ratios taken to a Python float (.item()), and the
    output size is int(size * scale) per axis in double (torch/nn/functional.py interpolate)."""
    lo, hi = F32(min(h, w)), F32(max(h, w))
    scale = float(min((F32(1) / lo) * F32(min_size), (F32(1) / hi) * F32(max_size)))
    return int(h * scale), int(w * scale)


def rcnn_transform(images, min_size, max_size, image_mean, image_std, size_divisible=32):
    """GeneralizedRCNNTransform.forward for inference (transform.py:436-459): per image normalize (:264-270, float32),
    resize (:284-315 -> _resize_image_and_masks :72-110, bilinear, align_corners=False, sizes recomputed), then
    batch_images (:352-386): zero canvas [B, 3, ceil(maxH / 32) * 32, ceil(maxW / 32) * 32], images at the top-left.
    images: list of float32 [3, H, W].  Returns (batch float32 [B,3,Hp,Wp], [(h_i, w_i)])."""
    import math
    from .yolact_np import upsample_bilinear
    mean = np.asarray(image_mean, dtype=F32)[:, None, None]
    std = np.asarray(image_std, dtype=F32)[:, None, None]
    resized, sizes = [], []
    for img in images:
        x = (np.asarray(img, dtype=F32) - mean) / std
        oh, ow = transform_output_size(x.shape[1], x.shape[2], min_size, max_size)
        resized.append(upsample_bilinear(x, oh, ow))
        sizes.append((oh, ow))
    stride = float(size_divisible)
    hp = int(math.ceil(float(max(s[0] for s in sizes)) / stride) * stride)
    wp = int(math.ceil(float(max(s[1] for s in sizes)) / stride) * stride)
    batch = np.zeros((len(images), 3, hp, wp), F32)
    for i, r in enumerate(resized):
        batch[i, :, : r.shape[1], : r.shape[2]] = r
    return batch, sizes
