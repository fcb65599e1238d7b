"""numpy restatement of the YOLACT Detect + mask-assembly path (oracle; test infrastructure only).

Follows /root/reference/inst_model/yolact/utils/utils_bbox.py (`BBoxUtility`) and
/root/reference/inst_model/yolact/utils/anchors.py.  Every op is done in float32 with one
rounding per elementary operation, in the reference's order, so integer / index results
(keep sets, class ids, argmax) are bit-comparable and float results agree to a few ulp
(`exp`, `sigmoid` and the K=32 dot product are the only places where library rounding may
differ from torch's CPU kernels).

Tie rule: torch's CPU `sort(descending=True)` is unstable, so the reference defines no tie
order.  This oracle (and the CUDA path) use the order the reference would produce with
`stable=True` sorts: score descending, then class ascending, then row index ascending.
"""
from __future__ import annotations

import math

import numpy as np

F32 = np.float32


# --------------------------------------------------------------------------- anchors
def get_img_output_length(height, width):
    """anchors.py:25-37 -- sizes of the five FPN levels P3..P7 (seven stride-2 stages)."""
    kernel = [7, 3, 3, 3, 3, 3, 3]
    pad = [3, 1, 1, 1, 1, 1, 1]
    hs, ws = [], []
    for k, p in zip(kernel, pad):
        height = (height + 2 * p - k) // 2 + 1
        width = (width + 2 * p - k) // 2 + 1
        hs.append(height)
        ws.append(width)
    return np.array(hs[-5:]), np.array(ws[-5:])


def get_anchors(input_shape=(550, 550), anchors_size=(24, 48, 96, 192, 384)):
    """anchors.py:7-48 -- priors (cx, cy, w, h) normalised, float64 [A, 4].

    Level order P3->P7, cells row-major, aspect ratios [1, 1/2, 2] with
    w = s*sqrt(ar)/W, h = s/sqrt(ar)/H.
    """
    fh, fw = get_img_output_length(input_shape[0], input_shape[1])
    out = []
    for lvl in range(len(fh)):
        ch, cw, scale = int(fh[lvl]), int(fw[lvl]), anchors_size[lvl]
        for j in range(ch):
            for i in range(cw):
                x = (i + 0.5) / cw
                y = (j + 0.5) / ch
                for ar in (1, 1 / 2, 2):
                    r = math.sqrt(ar)
                    out.append((x, y, scale * r / input_shape[1], scale / r / input_shape[0]))
    return np.asarray(out, dtype=np.float64).reshape(-1, 4)


# --------------------------------------------------------------------------- Y2
def decode_boxes(pred_box, anchors, variances=(0.1, 0.2)):
    """utils_bbox.py:11-25 -- centre/size decode then corner form, all float32."""
    p = np.asarray(pred_box, dtype=F32)
    a = np.asarray(anchors, dtype=F32)
    v0, v1 = F32(variances[0]), F32(variances[1])
    cxy = a[:, :2] + (p[:, :2] * v0) * a[:, 2:]
    wh = a[:, 2:] * np.exp(p[:, 2:] * v1).astype(F32)
    x1y1 = cxy - wh / F32(2)
    x2y2 = wh + x1y1
    return np.concatenate([x1y1, x2y2], axis=1).astype(F32)


# --------------------------------------------------------------------------- Y5
def jaccard(box_a, box_b):
    """utils_bbox.py:27-48 -- pairwise IoU, inter / ((area_a + area_b) - inter), no eps.

    Accepts [A,4],[B,4] or batched [n,A,4],[n,B,4].
    """
    a = np.asarray(box_a, dtype=F32)
    b = np.asarray(box_b, dtype=F32)
    squeeze = a.ndim == 2
    if squeeze:
        a, b = a[None], b[None]
    max_xy = np.minimum(a[:, :, None, 2:], b[:, None, :, 2:])
    min_xy = np.maximum(a[:, :, None, :2], b[:, None, :, :2])
    d = np.maximum(max_xy - min_xy, F32(0))
    inter = d[..., 0] * d[..., 1]
    area_a = ((a[:, :, 2] - a[:, :, 0]) * (a[:, :, 3] - a[:, :, 1]))[:, :, None]
    area_b = ((b[:, :, 2] - b[:, :, 0]) * (b[:, :, 3] - b[:, :, 1]))[:, None, :]
    union = (area_a + area_b) - inter
    with np.errstate(divide="ignore", invalid="ignore"):
        out = (inter / union).astype(F32)
    return out[0] if squeeze else out


def _argsort_desc_stable(x, axis=-1):
    """Descending sort order with ties kept in ascending index order; NaN ranks before everything else, +inf included,
    as in torch.sort(descending=True)."""
    x = np.asarray(x)
    nan = np.isnan(x)
    if not nan.any():
        return np.argsort(-x, axis=axis, kind="stable")
    xm, nm = np.moveaxis(x, axis, -1), np.moveaxis(nan, axis, -1)
    out = np.empty(xm.shape, np.int64)
    for i in np.ndindex(xm.shape[:-1]):
        out[i] = np.lexsort((np.where(nm[i], 0, -xm[i]), ~nm[i]))     # primary key: NaN first; then value descending
    return np.moveaxis(out, -1, axis)


# --------------------------------------------------------------------------- Y4-Y6
def fast_non_max_suppression(box_thre, class_thre, mask_thre, nms_iou=0.5, top_k=200, max_detections=100):
    """utils_bbox.py:50-100 -- per-class top-k, upper-triangular IoU column max, global top-N.

    box_thre [n,4], class_thre [n,C] (C foreground classes), mask_thre [n,K].
    Returns box [m,4] f32, score [m] f32, class_id [m] int64, coef [m,K] f32.
    """
    box_thre = np.asarray(box_thre, dtype=F32)
    mask_thre = np.asarray(mask_thre, dtype=F32)
    scores = np.ascontiguousarray(np.asarray(class_thre, dtype=F32).T)      # :55  [C, n]
    order = _argsort_desc_stable(scores, axis=1)                            # :61
    idx = order[:, :top_k]                                                  # :63
    scores = np.take_along_axis(scores, idx, axis=1)                        # :64
    num_classes, num_dets = idx.shape
    boxes = box_thre[idx.reshape(-1)].reshape(num_classes, num_dets, 4)     # :72
    coefs = mask_thre[idx.reshape(-1)].reshape(num_classes, num_dets, -1)   # :73
    iou = jaccard(boxes, boxes)                                             # :75
    iu = np.triu_indices(num_dets, k=1)
    upper = np.zeros_like(iou)
    upper[:, iu[0], iu[1]] = iou[:, iu[0], iu[1]]                           # :80 triu_(1)
    # torch.max propagates NaN; np.max does too.
    iou_max = upper.max(axis=1) if num_dets > 0 else np.zeros((num_classes, 0), F32)   # :81
    with np.errstate(invalid="ignore"):
        keep = iou_max <= F32(nms_iou)                                      # :86
    class_ids = np.broadcast_to(np.arange(num_classes, dtype=np.int64)[:, None], keep.shape)
    box_nms, score_nms, cls_nms, coef_nms = boxes[keep], scores[keep], class_ids[keep], coefs[keep]  # :89-92
    order2 = _argsort_desc_stable(score_nms, axis=0)[:max_detections]       # :94-95
    return box_nms[order2], score_nms[order2], cls_nms[order2], coef_nms[order2]


# --------------------------------------------------------------------------- Y15
def traditional_non_max_suppression(box_thre, class_thre, mask_thre, pred_class_max, nms_iou, max_detections):
    """utils_bbox.py:102-140 -- class = argmax of the row, per-class torchvision nms on the row maxima, concatenation in
    class order, stable descending sort by score, first max_detections."""
    from .maskrcnn_np import nms
    box_thre = np.asarray(box_thre, dtype=F32)
    class_thre = np.asarray(class_thre, dtype=F32)
    mask_thre = np.asarray(mask_thre, dtype=F32)
    pred_class_max = np.asarray(pred_class_max, dtype=F32)
    labels = np.argmax(class_thre, axis=-1)                                 # :104
    bs, ss, cs, ms = [], [], [], []
    for c in range(class_thre.shape[1]):                                    # :107
        sel = np.where(labels == c)[0]
        idx = nms(box_thre[sel], pred_class_max[sel], nms_iou)              # :123
        bs.append(box_thre[sel][idx]); ss.append(pred_class_max[sel][idx])
        cs.append(np.full(len(idx), c, np.int64)); ms.append(mask_thre[sel][idx])
    b, s2, c2, m2 = np.concatenate(bs), np.concatenate(ss), np.concatenate(cs), np.concatenate(ms)
    order = _argsort_desc_stable(s2, axis=0)[:max_detections]               # :136
    return b[order], s2[order], c2[order], m2[order]


# --------------------------------------------------------------------------- Y8
def yolact_correct_boxes(boxes, image_shape):
    """utils_bbox.py:142-152 -- scale to pixels, order corners, clamp to the image.

    Note :149 takes the max against the *already overwritten* x1y1 (kept as is).
    """
    h, w = int(image_shape[0]), int(image_shape[1])
    size = np.array([w, h], dtype=F32)
    b = np.asarray(boxes, dtype=F32) * np.concatenate([size, size])
    b = b.copy()
    b[:, :2] = np.minimum(b[:, :2], b[:, 2:])
    b[:, 2:] = np.maximum(b[:, :2], b[:, 2:])
    b[:, :2] = np.maximum(b[:, :2], F32(0))
    b[:, 2:] = np.minimum(b[:, 2:], size[None, :])
    return b


# --------------------------------------------------------------------------- Y9
def sigmoid(x):
    """torch.sigmoid in float32: 1 / (1 + exp(-x))."""
    x = np.asarray(x, dtype=F32)
    with np.errstate(over="ignore"):
        return (F32(1) / (F32(1) + np.exp(-x).astype(F32))).astype(F32)


def mask_logits(proto, coef):
    """utils_bbox.py:223 (inner matmul) -- proto [P,Q,K] @ coef[m,K]^T -> [P,Q,m] float32."""
    return np.matmul(np.asarray(proto, dtype=F32), np.asarray(coef, dtype=F32).T).astype(F32)


# --------------------------------------------------------------------------- Y10
def _linear_index_weights(in_size, out_size):
    """aten `area_pixel_compute_source_index` + `guard_index_and_lambda`, align_corners=False.

    (torch 2.11 ATen/native/UpSample.h; reached from utils_bbox.py:228.)  All in float32:
    scale = in/out; src = scale*(dst+0.5)-0.5 clamped at 0; i0 = min(floor(src), in-1);
    lambda1 = clamp(src - i0, 0, 1); lambda0 = 1 - lambda1; i1 = i0 + (i0 < in-1).
    The product and the subtraction of `src` are ONE fused multiply-add in torch's binaries (the vectorised CPU kernel
    and the CUDA kernel are both compiled with contraction on): with separate roundings the index is off by an ulp of
    ~in_size, i.e. the weights by up to 1e-4 on a 1000-pixel axis (measured: 1.5e-4 max output error at 1080 -> 749,
    5e-7 with the fused form).  The fused operation is emulated exactly in float64 (the product of two float32 is
    exact there) with one rounding to float32.
    """
    scale = F32(in_size) / F32(out_size)
    dst = np.arange(out_size, dtype=F32)
    src = (np.float64(scale) * (dst + F32(0.5)).astype(np.float64) - 0.5).astype(F32)
    src = np.where(src < 0, F32(0), src).astype(F32)
    i0 = np.minimum(np.floor(src).astype(np.int64), in_size - 1)
    lam1 = np.minimum(np.maximum(src - i0.astype(F32), F32(0)), F32(1)).astype(F32)
    lam0 = (F32(1) - lam1).astype(F32)
    i1 = i0 + (i0 < in_size - 1)
    return i0, i1, lam0, lam1


def upsample_bilinear(maps, out_h, out_w):
    """utils_bbox.py:227-229 -- F.interpolate(mode='bilinear', align_corners=False).

    maps [m, P, Q] -> [m, out_h, out_w];
    out = h0*(w0*p00 + w1*p01) + h1*(w0*p10 + w1*p11) (aten cpu_upsample_generic nesting).
    """
    maps = np.asarray(maps, dtype=F32)
    y0, y1, hy0, hy1 = _linear_index_weights(maps.shape[1], out_h)
    x0, x1, wx0, wx1 = _linear_index_weights(maps.shape[2], out_w)
    # horizontal pass on the small map first: row[py, x] = w0*p[py, x0] + w1*p[py, x1]; the vertical
    # pass then combines two such rows -- the same values in the same order as the nested form.
    rows = maps[:, :, x0] * wx0 + maps[:, :, x1] * wx1
    return (rows[:, y0, :] * hy0[None, :, None] + rows[:, y1, :] * hy1[None, :, None]).astype(F32)


# --------------------------------------------------------------------------- Y11
def crop(masks, boxes):
    """utils_bbox.py:154-168 -- zero everything outside each detection's pixel box.

    masks [H,W,m] float32, boxes [m,4] pixel corners; tests col>=x1, col<x2, row>=y1, row<y2
    in float32.
    """
    masks = np.asarray(masks, dtype=F32)
    boxes = np.asarray(boxes, dtype=F32)
    h, w, _ = masks.shape
    xs = np.arange(w, dtype=F32)[None, :, None]
    ys = np.arange(h, dtype=F32)[:, None, None]
    inside = (xs >= boxes[:, 0]) & (xs < boxes[:, 2]) & (ys >= boxes[:, 1]) & (ys < boxes[:, 3])
    return masks * inside.astype(F32)


# --------------------------------------------------------------------------- Y9-Y13
def assemble_masks(proto, coef, boxes_px, image_shape):
    """utils_bbox.py:223-241 -- matmul, sigmoid, upsample, crop, argmax, >0.5.

    Returns (masks_arg int64 [H,W], masks bool [H,W,m]).
    """
    h, w = int(image_shape[0]), int(image_shape[1])
    s = sigmoid(mask_logits(proto, coef))                        # [P,Q,m]     :223
    s = np.ascontiguousarray(np.transpose(s, (2, 0, 1)))         # [m,P,Q]     :227
    up = upsample_bilinear(s, h, w)                              # [m,H,W]     :228
    up = np.ascontiguousarray(np.transpose(up, (1, 2, 0)))       # [H,W,m]     :229
    up = crop(up, boxes_px)                                      #             :230
    masks_arg = np.argmax(up, axis=-1).astype(np.int64)          # first max   :236
    return masks_arg, up > F32(0.5)                              #             :241


# --------------------------------------------------------------------------- Y14
def decode_nms(outputs, anchors, confidence, nms_iou, image_shape, traditional_nms=False, max_detections=100):
    """utils_bbox.py:170-243 for one image (Fast NMS by default, per-class greedy NMS with traditional_nms=True).

    outputs = (box [1,A,4] | [A,4], conf [.,A,C] softmaxed incl. background column 0,
    coef [.,A,K], proto [.,P,Q,K]).  Returns the reference's 5-tuple or five Nones.
    Note `decode_nms` never forwards `max_detections`/`top_k` to Fast NMS (:207), so the
    function defaults 200 / 100 apply; the parameter here is ignored in the same way.
    """
    pred_box, pred_class, pred_coef, pred_proto = [np.squeeze(np.asarray(o, dtype=F32)) for o in outputs]
    if pred_box.ndim == 1:                       # A == 1 collapses under squeeze(); keep 2-D
        pred_box, pred_class, pred_coef = pred_box[None], pred_class[None], pred_coef[None]
    boxes = decode_boxes(pred_box, anchors)                                  # :186
    fg = pred_class[:, 1:]                                                   # :193
    keep = fg.max(axis=1) > F32(confidence)                                  # :194-195
    box_thre, class_thre, coef_thre = boxes[keep], fg[keep], pred_coef[keep]  # :200-202
    if class_thre.shape[0] == 0:                                             # :203
        return None, None, None, None, None
    if traditional_nms:                                                      # :213
        box_thre, score, class_ids, coef_thre = traditional_non_max_suppression(
            box_thre, class_thre, coef_thre, fg.max(axis=1)[keep], nms_iou, max_detections)
    else:
        box_thre, score, class_ids, coef_thre = fast_non_max_suppression(box_thre, class_thre, coef_thre, nms_iou)  # :207
        keep2 = score > F32(confidence)                                      # :208
        box_thre, score, class_ids, coef_thre = box_thre[keep2], score[keep2], class_ids[keep2], coef_thre[keep2]
    box_px = yolact_correct_boxes(box_thre, image_shape)                     # :216
    masks_arg, masks = assemble_masks(pred_proto, coef_thre, box_px, image_shape)
    return box_px, score, class_ids, masks_arg, masks


# --------------------------------------------------------------------------- next row
def masks_class(masks_sigmoid, class_ids, masks_arg):
    """inst_model/yolact/yolact.py:167-169 -- the wrapper's per-pixel class image built from decode_nms' outputs:
    every pixel takes (class id + 1) of its arg-max detection if that detection's binary mask is set there, else 0.
    The wrapper multiplies bool [H,W,m] by (class_ids + 1) and gathers along m with masks_arg; gathering first and
    multiplying afterwards is the same thing."""
    masks_sigmoid = np.asarray(masks_sigmoid)
    class_ids = np.asarray(class_ids).astype(np.int64)
    masks_arg = np.asarray(masks_arg).astype(np.int64)
    picked = np.take_along_axis(masks_sigmoid, masks_arg[:, :, None], axis=2)[:, :, 0]
    return np.where(picked, class_ids[masks_arg] + 1, 0).astype(np.int64)
