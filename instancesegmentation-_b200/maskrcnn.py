"""Drop-ins for the Mask R-CNN post-processing functions of the reference
(inst_model/Mask_RCNN/net/network_files/...); the work is done by libisb.so on the tensors' CUDA device and
current stream.

Module-level functions with the reference's names, argument order and return types (rebound by isb.patch):
    nms, batched_nms                  boxes.py:7, :38
    MultiScaleRoIAlign                torchvision module constructed at faster_rcnn_framework.py:326, mask_rcnn.py:150
    maskrcnn_inference                roi_head.py:59
    resize_boxes, paste_masks_in_image  transform.py:462, :212

METHOD drop-ins: same `(self, ...)` signatures as the reference's methods, reading the same attributes of `self`;
isb.patch binds them onto the reference's classes, so a patched reference model reaches the fused kernels
without any edit of the reference tree:
    BoxCoder.decode_single            det_utils.py:224   -> boxcoder_decode_single    (isb_box_decode)
    RegionProposalNetwork.filter_proposals  rpn_function.py:460 -> rpn_filter_proposals (isb_rpn_filter per image)
    RoIHeads.postprocess_detections   roi_head.py:366    -> roiheads_postprocess_detections (isb_box_postprocess)
    GeneralizedRCNNTransform.postprocess  transform.py:388 -> transform_postprocess   (isb_resize_boxes + isb_paste_masks)

Functional forms with explicit parameters (one image, used by the stage tests and benches):
    decode_single, filter_proposals, postprocess_detections
"""
from __future__ import annotations

import ctypes as C
import math

import numpy as np
import torch
import torchvision

from . import _lib
from ._lib import RoiAlignCfg, get_handle
from .yolact import _cuda_f32, _ptr, _stream

BBOX_XFORM_CLIP = math.log(1000.0 / 16)  # det_utils.py:146


def _call(dev, fn):
    h = get_handle(dev.index)
    with _lib.on_device(dev):
        h.check(fn(h))


# ---------------------------------------------------------------------------- boxes.py
def nms(boxes, scores, iou_threshold):
    """boxes.py:7-35: int64 indices of kept boxes, in decreasing order of score."""
    boxes = _cuda_f32(boxes, "boxes")
    scores = _cuda_f32(scores, "scores").to(boxes.device)
    dev, n = boxes.device, boxes.shape[0]
    keep = torch.empty((max(n, 1),), dtype=torch.int64, device=dev)
    count = torch.zeros((1,), dtype=torch.int32, device=dev)
    _call(dev, lambda h: h.lib.isb_nms(h.ptr, _ptr(boxes), _ptr(scores), n, float(iou_threshold), _ptr(keep), _ptr(count),
                                       _stream(dev)))
    return keep[: int(count.item())]


def nms_static(boxes, scores, iou_threshold, keep, count):
    """`nms` for callers that may not allocate (no counterpart in the reference):
    float32 CUDA `boxes` [n,4] / `scores` [n], n <= ISB_NMS_STATIC_MAX, and PREALLOCATED outputs `keep`
    (int64 [>= n]) and `count` (int32 [1]) on the same device.  Enqueues `isb_nms_static` on the current stream: no handle,
    no torch or CUDA allocation, nothing synchronised -- the kept indices are keep[:count].  Every call on a GPU shares one
    fixed scratch array, so calls that overlap on the GPU race."""
    dev, n = boxes.device, boxes.shape[0]
    for t, dt, name in ((boxes, torch.float32, "boxes"), (scores, torch.float32, "scores"), (keep, torch.int64, "keep"),
                        (count, torch.int32, "count")):
        if not (t.is_cuda and t.device == dev and t.dtype == dt and t.is_contiguous()):
            raise ValueError(f"nms_static: {name} must be a contiguous {dt} tensor on {dev} (nothing is converted or copied)")
    if keep.numel() < n or count.numel() < 1:
        raise ValueError("nms_static: keep needs n elements and count one")
    err = C.create_string_buffer(256)
    with _lib.on_device(dev):
        rc = _lib.load().isb_nms_static(_ptr(boxes), _ptr(scores), n, float(iou_threshold), _ptr(keep), _ptr(count), _stream(dev),
                                        err, len(err))
    if rc != 0:
        raise _lib.IsbError(rc, err.value.decode())


def batched_nms(boxes, scores, idxs, iou_threshold):
    """boxes.py:38-83: NMS per category via the reference's float32 coordinate-offset trick."""
    boxes = _cuda_f32(boxes, "boxes")
    dev = boxes.device
    if boxes.numel() == 0:  # boxes.py:66-67
        return torch.empty((0,), dtype=torch.int64, device=dev)
    scores = _cuda_f32(scores, "scores").to(dev)
    idxs = idxs.to(device=dev, dtype=torch.int64).contiguous()
    n = boxes.shape[0]
    keep = torch.empty((n,), dtype=torch.int64, device=dev)
    count = torch.zeros((1,), dtype=torch.int32, device=dev)
    _call(dev, lambda h: h.lib.isb_batched_nms(h.ptr, _ptr(boxes), _ptr(scores), _ptr(idxs), n, float(iou_threshold),
                                               _ptr(keep), _ptr(count), _stream(dev)))
    return keep[: int(count.item())]


# ---------------------------------------------------------------------------- det_utils.py
def decode_single(rel_codes, boxes, weights=(1.0, 1.0, 1.0, 1.0), bbox_xform_clip=BBOX_XFORM_CLIP):
    """det_utils.py:224-267 `BoxCoder.decode_single`: rel_codes [N,4k], boxes [N,4] -> [N,4k]."""
    rel = _cuda_f32(rel_codes, "rel_codes")
    dev = rel.device
    boxes = _cuda_f32(boxes, "boxes").to(dev)
    n, k = rel.shape[0], rel.shape[1] // 4
    out = torch.empty_like(rel)
    wx, wy, ww, wh = [float(w) for w in weights]
    _call(dev, lambda h: h.lib.isb_box_decode(h.ptr, _ptr(rel), _ptr(boxes), n, k, wx, wy, ww, wh, float(bbox_xform_clip),
                                              _ptr(out), _stream(dev)))
    return out


class BoxCoder(object):
    """Inference half of det_utils.py:140-267 (decode / decode_single); encode is training-only."""

    def __init__(self, weights, bbox_xform_clip=BBOX_XFORM_CLIP):
        self.weights = weights
        self.bbox_xform_clip = bbox_xform_clip

    def decode_single(self, rel_codes, boxes):
        return decode_single(rel_codes, boxes.to(rel_codes.dtype), self.weights, self.bbox_xform_clip)

    def decode(self, rel_codes, boxes):
        assert isinstance(boxes, (list, tuple))
        assert isinstance(rel_codes, torch.Tensor)
        concat = torch.cat(boxes, dim=0)
        box_sum = concat.shape[0]
        pred = self.decode_single(rel_codes, concat)
        if box_sum > 0:
            pred = pred.reshape(box_sum, -1, 4)
        return pred


# ---------------------------------------------------------------------------- rpn_function.py
def filter_proposals(proposals, objectness, image_shape, num_anchors_per_level, pre_nms_top_n=1000, post_nms_top_n=1000,
                     nms_thresh=0.7, score_thresh=0.0, min_size=1.0, out=None):
    """rpn_function.py:460-529 (+ _get_top_n_idx :434) for one image: proposals [A,4] decoded boxes,
    objectness [A] logits.  Returns (boxes [k,4], scores [k]).  Defaults: faster_rcnn_framework.py:261-270.
    `out` = (boxes [post,4] f32, scores [post] f32, count [1] i32) preallocated on the device: with the same input and
    output tensors call after call, the library replays the whole launch sequence as one CUDA graph."""
    proposals = _cuda_f32(proposals, "proposals").reshape(-1, 4)
    dev = proposals.device
    objectness = _cuda_f32(objectness, "objectness").to(dev).reshape(-1)
    levels = (C.c_int32 * len(num_anchors_per_level))(*[int(v) for v in num_anchors_per_level])
    post = int(post_nms_top_n)
    if out is None:
        out = (torch.empty((post, 4), dtype=torch.float32, device=dev), torch.empty((post,), dtype=torch.float32, device=dev),
               torch.zeros((1,), dtype=torch.int32, device=dev))
    out_b, out_s, count = out
    _call(dev, lambda h: h.lib.isb_rpn_filter(h.ptr, _ptr(proposals), _ptr(objectness), levels, len(num_anchors_per_level),
                                              int(image_shape[0]), int(image_shape[1]), int(pre_nms_top_n), post,
                                              float(nms_thresh), float(score_thresh), float(min_size), _ptr(out_b),
                                              _ptr(out_s), _ptr(count), _stream(dev)))
    k = int(count.item())
    return out_b[:k], out_s[:k]


# ---------------------------------------------------------------------------- roi_head.py
def postprocess_detections(class_logits, box_regression, proposals, image_shape, weights=(10.0, 10.0, 5.0, 5.0),
                           score_thresh=0.05, nms_thresh=0.5, detections_per_img=100, bbox_xform_clip=BBOX_XFORM_CLIP,
                           class_probs=None, out=None):
    """roi_head.py:366-457 for one image.  Returns (boxes [k,4], scores [k], labels int64 [k]).
    `class_probs` (softmaxed scores) may be given instead of logits.  `out` = preallocated (boxes [d,4] f32, scores [d] f32,
    labels [d] i64, count [1] i32), see filter_proposals."""
    reg = _cuda_f32(box_regression, "box_regression")
    dev = reg.device
    props = _cuda_f32(proposals, "proposals").to(dev)
    logits = _cuda_f32(class_logits, "class_logits").to(dev) if class_probs is None else None
    probs = _cuda_f32(class_probs, "class_probs").to(dev) if class_probs is not None else None
    r = props.shape[0]
    c = (probs if probs is not None else logits).shape[-1]
    d = int(detections_per_img)
    if out is None:
        out = (torch.empty((d, 4), dtype=torch.float32, device=dev), torch.empty((d,), dtype=torch.float32, device=dev),
               torch.empty((d,), dtype=torch.int64, device=dev), torch.zeros((1,), dtype=torch.int32, device=dev))
    out_b, out_s, out_l, count = out
    null = C.c_void_p(0)
    wx, wy, ww, wh = [float(w) for w in weights]
    _call(dev, lambda h: h.lib.isb_box_postprocess(h.ptr, _ptr(logits) if logits is not None else null,
                                                   _ptr(probs) if probs is not None else null, _ptr(reg), _ptr(props), r, c,
                                                   int(image_shape[0]), int(image_shape[1]), wx, wy, ww, wh,
                                                   float(bbox_xform_clip), float(score_thresh), float(nms_thresh), d,
                                                   _ptr(out_b), _ptr(out_s), _ptr(out_l), _ptr(count), _stream(dev)))
    k = int(count.item())
    if k < 0:
        raise RuntimeError("postprocess_detections: class_probs rows exceed the candidate bound (do they sum to 1?)")
    return out_b[:k], out_s[:k], out_l[:k]


# ---------------------------------------------------------------------------- method drop-ins
def boxcoder_decode_single(self, rel_codes, boxes):
    """`BoxCoder.decode_single(self, rel_codes, boxes)` (det_utils.py:224-267): reads `self.weights` and
    `self.bbox_xform_clip` like the reference; rel_codes [N,4k], boxes [N,4] -> [N,4k]."""
    return decode_single(rel_codes, boxes.to(rel_codes.dtype), self.weights, self.bbox_xform_clip)


def rpn_filter_proposals(self, proposals, objectness, image_shapes, num_anchors_per_level):
    """`RegionProposalNetwork.filter_proposals(self, proposals, objectness, image_shapes, num_anchors_per_level)`
    (rpn_function.py:460-529, with `_get_top_n_idx` :434-458 folded in): proposals [N,A,4] decoded boxes, objectness
    [N*A,1] or [N,A] logits -> (List[Tensor [k_i,4]], List[Tensor [k_i]]).  Reads `self.pre_nms_top_n()`,
    `self.post_nms_top_n()`, `self.nms_thresh`, `self.score_thresh`, `self.min_size` exactly as the reference does.
    One fused launch sequence per image into padded [N,post,..] buffers, then ONE 4N-byte read of the counts for the
    variable-length views the reference's return type needs (the reference synchronises several times per image)."""
    num_images = proposals.shape[0]
    proposals = _cuda_f32(proposals.detach(), "proposals").reshape(num_images, -1, 4)
    dev = proposals.device
    objectness = _cuda_f32(objectness.detach(), "objectness").to(dev).reshape(num_images, -1)      # :475-476
    post, pre = int(self.post_nms_top_n()), int(self.pre_nms_top_n())
    nlev = len(num_anchors_per_level)
    levels = (C.c_int32 * nlev)(*[int(v) for v in num_anchors_per_level])
    out_b = torch.empty((num_images, post, 4), dtype=torch.float32, device=dev)
    out_s = torch.empty((num_images, post), dtype=torch.float32, device=dev)
    counts = torch.zeros((num_images,), dtype=torch.int32, device=dev)
    h = get_handle(dev.index)
    with _lib.on_device(dev):
        for i in range(num_images):
            shp = image_shapes[i]
            h.check(h.lib.isb_rpn_filter(h.ptr, _ptr(proposals[i]), _ptr(objectness[i]), levels, nlev, int(shp[0]), int(shp[1]),
                                         pre, post, float(self.nms_thresh), float(self.score_thresh), float(self.min_size),
                                         _ptr(out_b[i]), _ptr(out_s[i]), C.c_void_p(counts[i:].data_ptr()), _stream(dev)))
    ks = counts.tolist()
    return [out_b[i, :k] for i, k in enumerate(ks)], [out_s[i, :k] for i, k in enumerate(ks)]


def roiheads_postprocess_detections(self, class_logits, box_regression, proposals, image_shapes):
    """`RoIHeads.postprocess_detections(self, class_logits, box_regression, proposals, image_shapes)`
    (roi_head.py:366-457): class_logits [sum R_i, C], box_regression [sum R_i, 4C], proposals List[[R_i,4]] ->
    (List[boxes [k_i,4]], List[scores [k_i]], List[labels int64 [k_i]]).  Reads `self.box_coder.weights`,
    `self.box_coder.bbox_xform_clip`, `self.score_thresh`, `self.nms_thresh`, `self.detection_per_img`.
    Decode, softmax, clip, filters, per-class NMS and top-k are one fused sequence per image; one read of the counts."""
    logits = _cuda_f32(class_logits.detach(), "class_logits")
    dev = logits.device
    reg = _cuda_f32(box_regression.detach(), "box_regression").to(dev)
    num_classes = logits.shape[-1]
    per_image = [int(p.shape[0]) for p in proposals]
    n_img, d = len(per_image), int(self.detection_per_img)
    wx, wy, ww, wh = [float(w) for w in self.box_coder.weights]
    clip = float(self.box_coder.bbox_xform_clip)
    out_b = torch.empty((n_img, d, 4), dtype=torch.float32, device=dev)
    out_s = torch.empty((n_img, d), dtype=torch.float32, device=dev)
    out_l = torch.empty((n_img, d), dtype=torch.int64, device=dev)
    counts = torch.zeros((max(n_img, 1),), dtype=torch.int32, device=dev)
    h = get_handle(dev.index)
    null = C.c_void_p(0)
    row = 0
    with _lib.on_device(dev):
        for i, r in enumerate(per_image):
            props = _cuda_f32(proposals[i].detach(), "proposals").to(dev)
            shp = image_shapes[i]
            if r > 0:
                h.check(h.lib.isb_box_postprocess(h.ptr, _ptr(logits[row:row + r]), null, _ptr(reg[row:row + r]), _ptr(props), r,
                                                  num_classes, int(shp[0]), int(shp[1]), wx, wy, ww, wh, clip,
                                                  float(self.score_thresh), float(self.nms_thresh), d, _ptr(out_b[i]),
                                                  _ptr(out_s[i]), _ptr(out_l[i]), C.c_void_p(counts[i:].data_ptr()), _stream(dev)))
            row += r
    ks = counts[:n_img].tolist()
    if any(k < 0 for k in ks):
        raise RuntimeError("postprocess_detections: candidate bound exceeded")
    return ([out_b[i, :k] for i, k in enumerate(ks)], [out_s[i, :k] for i, k in enumerate(ks)],
            [out_l[i, :k] for i, k in enumerate(ks)])


# ---------------------------------------------------------------------------- sync-free (padded) forms
# The reference's return types are lists of variable-length tensors, so the drop-ins above must read the counts on the
# host once per stage.  A pipeline that keeps everything padded never has to: proposals stay [N, post_nms_top_n, 4] with
# the valid rows first and ZERO rows after them (RoIAlign of a zero box is harmless), detections [N, detections_per_img, ..],
# and the counts stay on the device -- `isb_box_postprocess_rows` takes the number of valid proposal rows from device memory.
# No counterpart in the reference (its stages are separated by Python lists); results are the list forms' results, padded.
def rpn_filter_proposals_padded(proposals, objectness, image_shapes, num_anchors_per_level, pre_nms_top_n=1000,
                                post_nms_top_n=1000, nms_thresh=0.7, score_thresh=0.0, min_size=1.0):
    """`filter_proposals` (rpn_function.py:460-529) without a host read: proposals [N,A,4], objectness [N,A] ->
    (boxes float32 [N,post,4] zero-padded, scores [N,post], counts int32 [N]) on the device."""
    num_images = proposals.shape[0]
    proposals = _cuda_f32(proposals.detach(), "proposals").reshape(num_images, -1, 4)
    dev = proposals.device
    objectness = _cuda_f32(objectness.detach(), "objectness").to(dev).reshape(num_images, -1)
    post, pre = int(post_nms_top_n), int(pre_nms_top_n)
    nlev = len(num_anchors_per_level)
    levels = (C.c_int32 * nlev)(*[int(v) for v in num_anchors_per_level])
    out_b = torch.zeros((num_images, post, 4), dtype=torch.float32, device=dev)
    out_s = torch.zeros((num_images, post), dtype=torch.float32, device=dev)
    counts = torch.zeros((num_images,), dtype=torch.int32, device=dev)
    h = get_handle(dev.index)
    with _lib.on_device(dev):
        for i in range(num_images):
            shp = image_shapes[i]
            h.check(h.lib.isb_rpn_filter(h.ptr, _ptr(proposals[i]), _ptr(objectness[i]), levels, nlev, int(shp[0]), int(shp[1]),
                                         pre, post, float(nms_thresh), float(score_thresh), float(min_size),
                                         _ptr(out_b[i]), _ptr(out_s[i]), C.c_void_p(counts[i:].data_ptr()), _stream(dev)))
    return out_b, out_s, counts


def postprocess_detections_padded(class_logits, box_regression, proposals, proposal_counts, image_shapes,
                                  weights=(10.0, 10.0, 5.0, 5.0), bbox_xform_clip=BBOX_XFORM_CLIP, score_thresh=0.05,
                                  nms_thresh=0.5, detections_per_img=100):
    """`postprocess_detections` (roi_head.py:366-457) without a host read: class_logits [N*R,C], box_regression [N*R,4C],
    proposals [N,R,4] padded, proposal_counts int32 [N] on the device (rows past a count are ignored) ->
    (boxes [N,D,4] zero-padded, scores [N,D], labels int64 [N,D], counts int32 [N])."""
    props = _cuda_f32(proposals.detach(), "proposals")
    dev = props.device
    n_img, r = int(props.shape[0]), int(props.shape[1])
    logits = _cuda_f32(class_logits.detach(), "class_logits").to(dev)
    reg = _cuda_f32(box_regression.detach(), "box_regression").to(dev)
    num_classes = logits.shape[-1]
    if logits.shape[0] != n_img * r or proposal_counts.dtype != torch.int32 or not proposal_counts.is_cuda:
        raise ValueError("postprocess_detections_padded: class_logits must have N*R rows, proposal_counts must be CUDA int32")
    d = int(detections_per_img)
    wx, wy, ww, wh = [float(w) for w in weights]
    out_b = torch.zeros((n_img, d, 4), dtype=torch.float32, device=dev)
    out_s = torch.zeros((n_img, d), dtype=torch.float32, device=dev)
    out_l = torch.zeros((n_img, d), dtype=torch.int64, device=dev)
    counts = torch.zeros((n_img,), dtype=torch.int32, device=dev)
    h = get_handle(dev.index)
    null = C.c_void_p(0)
    with _lib.on_device(dev):
        for i in range(n_img):
            shp = image_shapes[i]
            h.check(h.lib.isb_box_postprocess_rows(h.ptr, _ptr(logits[i * r:(i + 1) * r]), null, _ptr(reg[i * r:(i + 1) * r]),
                                                   _ptr(props[i]), r, C.c_void_p(proposal_counts[i:].data_ptr()), num_classes,
                                                   int(shp[0]), int(shp[1]), wx, wy, ww, wh, float(bbox_xform_clip),
                                                   float(score_thresh), float(nms_thresh), d, _ptr(out_b[i]), _ptr(out_s[i]),
                                                   _ptr(out_l[i]), C.c_void_p(counts[i:].data_ptr()), _stream(dev)))
    return out_b, out_s, out_l, counts


def transform_postprocess(self, result, image_shapes, original_image_sizes):
    """`GeneralizedRCNNTransform.postprocess(self, result, image_shapes, original_image_sizes)` (transform.py:388-418):
    per image, boxes back to the original frame (resize_boxes :462) and the 28x28 mask probabilities pasted into
    [m,1,H,W] (paste_masks_in_image :212).  Training mode returns `result` untouched (:404-405)."""
    if getattr(self, "train_mode", None):
        return result
    for i, (pred, im_s, o_im_s) in enumerate(zip(result, image_shapes, original_image_sizes)):
        boxes = resize_boxes(pred["boxes"], im_s, o_im_s)                                 # :410
        result[i]["boxes"] = boxes
        if "masks" in pred:
            result[i]["masks"] = paste_masks_in_image(pred["masks"], boxes, o_im_s)       # :415
    return result


# ---------------------------------------------------------------------------- MultiScaleRoIAlign
def _infer_scale(feat_h, image_h):
    # poolers.py `_infer_scale`: 2 ** round(log2(feat / img)), float32 log2 as torch.tensor(...).log2()
    approx = float(feat_h) / float(image_h)
    return 2.0 ** float(np.round(np.log2(np.float32(approx))))


class MultiScaleRoIAlign(torchvision.ops.MultiScaleRoIAlign):
    """Same constructor and forward signature as torchvision's module (it subclasses it so that the
    reference's isinstance checks at faster_rcnn_framework.py:282 / mask_rcnn.py:138 pass); forward runs
    the level mapping, sampling and scatter in one CUDA kernel."""

    def forward(self, x, boxes, image_shapes):
        feats = [v for k, v in x.items() if k in self.featmap_names]
        feats = [_cuda_f32(f, "feature map") for f in feats]
        dev = feats[0].device
        if self.scales is None:
            max_h = max(s[0] for s in image_shapes)
            self.scales = [_infer_scale(f.shape[-2], max_h) for f in feats]
            self._k_min = int(-np.log2(np.float32(self.scales[0])))
            self._k_max = int(-np.log2(np.float32(self.scales[-1])))
        # (one image: no copy of the boxes; torch.cat of a single tensor still launches a kernel)
        concat = _cuda_f32(boxes[0] if len(boxes) == 1 else torch.cat(boxes, dim=0), "boxes").to(dev)
        k = concat.shape[0]
        batch_idx = None
        if len(boxes) > 1:
            batch_idx = torch.cat([torch.full((b.shape[0],), i, dtype=torch.int32, device=dev) for i, b in enumerate(boxes)])
        c = feats[0].shape[1]
        oh, ow = self.output_size
        out = torch.empty((k, c, oh, ow), dtype=torch.float32, device=dev)
        # the C struct is rebuilt only when the feature maps moved or changed shape (steady state of a loop: never)
        sig = tuple((f.data_ptr(), f.shape[-2], f.shape[-1]) for f in feats) + (c, oh, ow)
        cached = getattr(self, "_isb_cfg", None)
        if cached is not None and cached[0] == sig:
            cfg = cached[1]
        else:
            cfg = RoiAlignCfg()
            cfg.num_levels = len(feats)
            for i, f in enumerate(feats):
                cfg.feat[i] = f.data_ptr()
                cfg.feat_h[i], cfg.feat_w[i] = f.shape[-2], f.shape[-1]
                cfg.scale[i] = self.scales[i]
            cfg.k_min, cfg.k_max = self._k_min, self._k_max
            cfg.canonical_scale, cfg.canonical_level = int(self.canonical_scale), int(self.canonical_level)
            cfg.channels, cfg.out_h, cfg.out_w, cfg.sampling_ratio = c, oh, ow, int(self.sampling_ratio)
            self._isb_cfg = (sig, cfg)
        bptr = _ptr(batch_idx) if batch_idx is not None else C.c_void_p(0)
        _call(dev, lambda h: h.lib.isb_ms_roi_align(h.ptr, C.byref(cfg), _ptr(concat), bptr, k, _ptr(out), _stream(dev)))
        return out


# ---------------------------------------------------------------------------- roi_head.py
def maskrcnn_inference(x, labels):
    """roi_head.py:59-91: x [m,C,S,S] mask logits, labels list of int64 tensors -> list of [m_i,1,S,S]."""
    x = _cuda_f32(x, "mask logits")
    dev = x.device
    boxes_per_image = [label.shape[0] for label in labels]
    lab = (labels[0] if len(labels) == 1 else torch.cat(labels)).to(device=dev, dtype=torch.int64).contiguous()
    m, c, s = x.shape[0], x.shape[1], x.shape[-1]
    out = torch.empty((m, 1, s, s), dtype=torch.float32, device=dev)
    _call(dev, lambda h: h.lib.isb_mask_select_sigmoid(h.ptr, _ptr(x), _ptr(lab), m, c, s, _ptr(out), _stream(dev)))
    return out.split(boxes_per_image, dim=0)


# ---------------------------------------------------------------------------- transform.py
def resize_boxes(boxes, original_size, new_size):
    """transform.py:462-484."""
    boxes = _cuda_f32(boxes, "boxes")
    dev = boxes.device
    out = torch.empty_like(boxes)
    _call(dev, lambda h: h.lib.isb_resize_boxes(h.ptr, _ptr(boxes), boxes.shape[0], int(original_size[0]), int(original_size[1]),
                                                int(new_size[0]), int(new_size[1]), _ptr(out), _stream(dev)))
    return out


def paste_masks_in_image(masks, boxes, img_shape, padding=1):
    """transform.py:212-230: masks [m,1,S,S] probabilities, boxes [m,4] -> float32 [m,1,H,W]."""
    masks = _cuda_f32(masks, "masks")
    dev = masks.device
    boxes = _cuda_f32(boxes, "boxes").to(dev)
    im_h, im_w = int(img_shape[0]), int(img_shape[1])
    m, s = masks.shape[0], masks.shape[-1]
    out = torch.empty((m, 1, im_h, im_w), dtype=torch.float32, device=dev)
    if m > 0:
        _call(dev, lambda h: h.lib.isb_paste_masks(h.ptr, _ptr(masks), _ptr(boxes), m, s, int(padding), im_h, im_w, _ptr(out),
                                                   _stream(dev)))
    return out


# ---------------------------------------------------------------------------- mask_rcnn.py:174-198 (wrapper tail)
def masks_class_image(masks, boxes, labels, scores, confidence, paste_shape, image_shape, padding=1, return_arg=False):
    """`MASK_RCNN.detect_image` tail (inst_model/Mask_RCNN/mask_rcnn.py:174-198) fused with `paste_masks_in_image`:
    masks [m,1,S,S] probabilities, boxes [m,4] in the paste frame, labels int64 [m], scores [m]  ->  uint8 [H,W] class
    image (labels[arg] + 1 where the arg-max kept detection's resized mask is > 0.5, else 0).  Detections with
    score <= confidence are dropped first (:174); with nothing kept the image is all zeros (the wrapper returns the
    input picture unchanged in that case).  `return_arg` adds masks_arg (int64 [H,W], index into the kept detections)."""
    masks = _cuda_f32(masks, "masks")
    dev = masks.device
    boxes = _cuda_f32(boxes, "boxes").to(dev)
    scores = _cuda_f32(scores, "scores").to(dev)
    labels = labels.to(device=dev, dtype=torch.int64).contiguous()
    m, s = masks.shape[0], masks.shape[-1]
    ph, pw = int(paste_shape[0]), int(paste_shape[1])
    oh, ow = int(image_shape[0]), int(image_shape[1])
    out = torch.empty((oh, ow), dtype=torch.uint8, device=dev)
    arg = torch.empty((oh, ow), dtype=torch.int64, device=dev) if return_arg else None
    _call(dev, lambda h: h.lib.isb_paste_class_map(h.ptr, _ptr(masks), _ptr(boxes), _ptr(labels), _ptr(scores), m, s, int(padding),
                                                   ph, pw, oh, ow, float(confidence), _ptr(out),
                                                   _ptr(arg) if return_arg else None, None, _stream(dev)))
    return (out, arg) if return_arg else out


# ---------------------------------------------------------------------------- transform.py:232-459 (pre-processing)
def transform_output_size(h, w, min_size, max_size):
    """Size `GeneralizedRCNNTransform.resize` produces for an [.., h, w] image (transform.py:88-110): the scale is
    min(min_size / min(h, w), max_size / max(h, w)) evaluated the way torch evaluates `float / float32 tensor`
    (reciprocal, then a float32 product), and F.interpolate(recompute_scale_factor=True) takes int(size * scale)."""
    f32 = np.float32
    lo, hi = f32(min(h, w)), f32(max(h, w))
    scale = float(min((f32(1) / lo) * f32(min_size), (f32(1) / hi) * f32(max_size)))
    return int(h * scale), int(w * scale)


def rcnn_transform(images, min_size, max_size, image_mean, image_std, size_divisible=32):
    """`GeneralizedRCNNTransform.forward` for inference (transform.py:436-459): images = list of float32 [3,H,W] CUDA
    tensors -> (batch float32 [B,3,Hp,Wp], [(h_i, w_i)]): normalize, bilinear resize to the reference's size rule, zero
    canvas padded to a multiple of `size_divisible`, one fused kernel per image writing straight into the batch."""
    if isinstance(min_size, (list, tuple)):
        min_size = min_size[-1]                                    # transform.py:299: testing uses the largest scale
    imgs = [_cuda_f32(im, "image") for im in images]
    for im in imgs:
        if im.dim() != 3 or im.shape[0] != 3:
            raise ValueError(f"images is expected to be a list of 3d tensors of shape [3, H, W], got {tuple(im.shape)}")
    dev = imgs[0].device
    sizes = [transform_output_size(im.shape[1], im.shape[2], float(min_size), float(max_size)) for im in imgs]
    stride = float(size_divisible)
    hp = int(math.ceil(float(max(s[0] for s in sizes)) / stride) * stride)     # transform.py:372-376
    wp = int(math.ceil(float(max(s[1] for s in sizes)) / stride) * stride)
    batch = torch.empty((len(imgs), 3, hp, wp), dtype=torch.float32, device=dev)
    m3 = (C.c_float * 3)(*[float(v) for v in image_mean])
    s3 = (C.c_float * 3)(*[float(v) for v in image_std])
    for i, (im, (oh, ow)) in enumerate(zip(imgs, sizes)):
        im = im.to(dev)
        _call(dev, lambda h, im=im, oh=oh, ow=ow, i=i: h.lib.isb_rcnn_transform(
            h.ptr, _ptr(im), im.shape[1], im.shape[2], oh, ow, m3, s3, C.c_void_p(batch[i].data_ptr()), hp, wp, _stream(dev)))
    return batch, sizes
