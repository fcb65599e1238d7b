"""Drop-in for the reference's `BBoxUtility` (inst_model/yolact/utils/utils_bbox.py:7-243).

Same constructor, method names, argument order, defaults, return arity, dtypes and device
placement; the work is done by libisb.so on the tensor's CUDA device and stream.  Added on top:
`decode_nms_batch`, the batched form the B200 path is built around (one call per batch of
images, no host synchronisation, padded outputs + per-image counts).
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import YolactMapOut, ISB_NUM_COEF, YolactIn, YolactOut, get_handle

TOP_K = 200            # utils_bbox.py:50 default, never overridden by decode_nms (:207)
MAX_DETECTIONS = 100   # utils_bbox.py:50 default


def _cuda_f32(t, name):
    if not isinstance(t, torch.Tensor):
        raise TypeError(f"{name} must be a torch.Tensor")
    if not t.is_cuda:
        raise RuntimeError(f"{name} is on {t.device}: the B200 path has no CPU fallback, move it to a CUDA device")
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


def _stream(device):
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


def _ptr(t):
    return C.c_void_p(t.data_ptr())


class YolactBatchResult:
    """Padded batched outputs of `decode_nms_batch` (device tensors, nothing synchronised).

    boxes [B,D,4] f32, scores [B,D] f32, class_ids [B,D] i64, masks_arg [B,H,W] i64,
    masks_raw [B, slot] u8 (image b holds bool [H,W,m_b] densely at the start of its slot),
    counts [B] i32 (m_b; -1 = candidate overflow).
    """

    def __init__(self, boxes, scores, class_ids, masks_arg, masks_raw, counts, out_hw):
        self.boxes, self.scores, self.class_ids = boxes, scores, class_ids
        self.masks_arg, self.masks_raw, self.counts = masks_arg, masks_raw, counts
        self.out_hw = out_hw

    def image(self, b, m=None):
        """Reference-shaped 5-tuple for image b (reads counts[b] from the device if m is None)."""
        if m is None:
            m = int(self.counts[b].item())
        if m < 0:
            raise _lib.IsbError(_lib.ISB_ERR_OVERFLOW, f"image {b}: candidate buffer overflow (counts = -1): raise "
                                "cand_capacity (isb._lib.get_handle(device, cand_capacity=...)) and run the batch again")
        if m == 0:
            return (None,) * 5
        h, w = self.out_hw
        masks = self.masks_raw[b, : h * w * m].view(h, w, m).view(torch.bool)
        return self.boxes[b, :m], self.scores[b, :m], self.class_ids[b, :m], self.masks_arg[b], masks


class BBoxUtility(object):
    def __init__(self):
        pass

    # ------------------------------------------------------------------ utils_bbox.py:11
    def decode_boxes(self, pred_box, anchors, variances=[0.1, 0.2]):
        pred_box = _cuda_f32(pred_box, "pred_box")
        anchors = _cuda_f32(anchors, "anchors").to(pred_box.device)
        if pred_box.dim() != 2 or pred_box.shape[1] != 4 or anchors.shape != pred_box.shape:
            raise ValueError("decode_boxes expects pred_box [n,4] and anchors [n,4]")
        dev = pred_box.device
        out = torch.empty_like(pred_box)
        h = get_handle(dev.index)
        with _lib.on_device(dev):
            h.check(h.lib.isb_decode_boxes(h.ptr, _ptr(pred_box), _ptr(anchors), pred_box.shape[0], float(variances[0]),
                                           float(variances[1]), _ptr(out), _stream(dev)))
        return out

    # ------------------------------------------------------------------ utils_bbox.py:27
    def jaccard(self, box_a, box_b, iscrowd: bool = False):
        box_a = _cuda_f32(box_a, "box_a")
        box_b = _cuda_f32(box_b, "box_b").to(box_a.device)
        use_batch = box_a.dim() == 3
        if not use_batch:
            box_a, box_b = box_a[None], box_b[None]
        n, na, nb = box_a.shape[0], box_a.shape[1], box_b.shape[1]
        dev = box_a.device
        out = torch.empty((n, na, nb), dtype=torch.float32, device=dev)
        h = get_handle(dev.index)
        with _lib.on_device(dev):
            h.check(h.lib.isb_jaccard(h.ptr, _ptr(box_a), _ptr(box_b), n, na, nb, int(bool(iscrowd)), _ptr(out), _stream(dev)))
        return out if use_batch else out.squeeze(0)

    # ------------------------------------------------------------------ utils_bbox.py:50
    def fast_non_max_suppression(self, box_thre, class_thre, mask_thre, nms_iou=0.5, top_k=200, max_detections=100):
        box_thre = _cuda_f32(box_thre, "box_thre")
        class_thre = _cuda_f32(class_thre, "class_thre")
        mask_thre = _cuda_f32(mask_thre, "mask_thre")
        dev = box_thre.device
        n, ncls = class_thre.shape
        k = mask_thre.shape[1]
        top_k, max_det = int(top_k), int(max_detections)
        if max_det > _lib.ISB_MAX_DET_LIMIT:
            raise ValueError(f"max_detections {max_det} > {_lib.ISB_MAX_DET_LIMIT}")
        h = get_handle(dev.index, max_anchors=n, max_classes=ncls, max_det=max_det, max_top_k=top_k,
                       cand_capacity=max(1, n * ncls))
        out_box = torch.empty((max_det, 4), dtype=torch.float32, device=dev)
        out_score = torch.empty((max_det,), dtype=torch.float32, device=dev)
        out_cls = torch.empty((max_det,), dtype=torch.int64, device=dev)
        out_coef = torch.empty((max_det, k), dtype=torch.float32, device=dev)
        count = torch.zeros((1,), dtype=torch.int32, device=dev)
        with _lib.on_device(dev):
            h.check(h.lib.isb_fast_nms(h.ptr, _ptr(box_thre), _ptr(class_thre), _ptr(mask_thre), n, ncls, k, float(nms_iou),
                                       top_k, max_det, _ptr(out_box), _ptr(out_score), _ptr(out_cls), _ptr(out_coef),
                                       _ptr(count), _stream(dev)))
        m = int(count.item())  # the one 4-byte D2H a variable-length return needs
        return out_box[:m], out_score[:m], out_cls[:m], out_coef[:m]

    # ------------------------------------------------------------------ utils_bbox.py:102
    def traditional_non_max_suppression(self, box_thre, class_thre, mask_thre, pred_class_max, nms_iou, max_detections):
        """Per-class greedy NMS (class = argmax of the foreground scores, score = pred_class_max), detections of all
        classes ordered by score, first max_detections.  The reference builds its label tensor on the CPU (:129) and
        therefore fails on CUDA inputs; here the labels stay on the device."""
        box_thre = _cuda_f32(box_thre, "box_thre")
        class_thre = _cuda_f32(class_thre, "class_thre")
        mask_thre = _cuda_f32(mask_thre, "mask_thre")
        pred_class_max = _cuda_f32(pred_class_max, "pred_class_max")
        dev = box_thre.device
        n, ncls = class_thre.shape
        k = mask_thre.shape[1]
        max_det = int(max_detections)
        if max_det > _lib.ISB_MAX_DET_LIMIT:
            raise ValueError(f"max_detections {max_det} > {_lib.ISB_MAX_DET_LIMIT}")
        h = get_handle(dev.index)
        out_box = torch.empty((max_det, 4), dtype=torch.float32, device=dev)
        out_score = torch.empty((max_det,), dtype=torch.float32, device=dev)
        out_cls = torch.empty((max_det,), dtype=torch.int64, device=dev)
        out_coef = torch.empty((max_det, k), dtype=torch.float32, device=dev)
        count = torch.zeros((1,), dtype=torch.int32, device=dev)
        with _lib.on_device(dev):
            h.check(h.lib.isb_traditional_nms(h.ptr, _ptr(box_thre), _ptr(class_thre), _ptr(mask_thre), _ptr(pred_class_max),
                                              n, ncls, k, float(nms_iou), max_det, _ptr(out_box), _ptr(out_score),
                                              _ptr(out_cls), _ptr(out_coef), _ptr(count), _stream(dev)))
        m = int(count.item())
        return out_box[:m], out_score[:m], out_cls[:m], out_coef[:m]

    # ------------------------------------------------------------------ utils_bbox.py:142
    def yolact_correct_boxes(self, boxes, image_shape):
        boxes = _cuda_f32(boxes, "boxes")
        dev = boxes.device
        hgt, wid = int(image_shape[0]), int(image_shape[1])
        out = torch.empty_like(boxes)
        h = get_handle(dev.index)
        with _lib.on_device(dev):
            h.check(h.lib.isb_correct_boxes(h.ptr, _ptr(boxes), boxes.shape[0], hgt, wid, _ptr(out), _stream(dev)))
        return out

    # ------------------------------------------------------------------ utils_bbox.py:154
    def crop(self, masks, boxes):
        masks = _cuda_f32(masks, "masks")
        boxes = _cuda_f32(boxes, "boxes").to(masks.device)
        hgt, wid, m = masks.shape
        dev = masks.device
        out = torch.empty_like(masks)
        h = get_handle(dev.index)
        with _lib.on_device(dev):
            h.check(h.lib.isb_crop(h.ptr, _ptr(masks), _ptr(boxes), hgt, wid, m, _ptr(out), _stream(dev)))
        return out

    # ------------------------------------------------------------------ stage helpers (no reference method)
    def mask_logits(self, pred_proto, mask_coef):
        """proto [Ph,Pw,K] @ coef[m,K]^T -> [Ph,Pw,m]  (the matmul of utils_bbox.py:223)."""
        proto = _cuda_f32(pred_proto, "pred_proto")
        coef = _cuda_f32(mask_coef, "mask_coef").to(proto.device)
        ph, pw, k = proto.shape
        m = coef.shape[0]
        out = torch.empty((ph, pw, m), dtype=torch.float32, device=proto.device)
        h = get_handle(proto.device.index)
        with _lib.on_device(proto.device):
            h.check(h.lib.isb_mask_logits(h.ptr, _ptr(proto), _ptr(coef), ph, pw, m, _ptr(out), _stream(proto.device)))
        return out

    def mask_logits_batch(self, pred_proto, mask_coef, sigmoid=False):
        """utils_bbox.py:223 for a batch in one launch on the tensor cores: proto [B,Ph,Pw,K], coef [B,m,K] (m <= 128) ->
        [B,Ph,Pw,m] = proto @ coef^T (sigmoid=True: torch.sigmoid of it, the whole of :223)."""
        proto = _cuda_f32(pred_proto, "pred_proto")
        coef = _cuda_f32(mask_coef, "mask_coef").to(proto.device)
        b, ph, pw, k = proto.shape
        m = coef.shape[1]
        if k != ISB_NUM_COEF or coef.shape[0] != b or coef.shape[2] != k:
            raise ValueError("mask_logits_batch expects proto [B,Ph,Pw,32] and coef [B,m,32]")
        out = torch.empty((b, ph, pw, m), dtype=torch.float32, device=proto.device)
        h = get_handle(proto.device.index)
        with _lib.on_device(proto.device):
            h.check(h.lib.isb_mask_logits_batch(h.ptr, _ptr(proto), _ptr(coef), b, ph, pw, m, 1 if sigmoid else 0, _ptr(out),
                                                _stream(proto.device)))
        return out

    def assemble_masks(self, pred_proto, mask_coef, boxes_px, image_shape):
        """utils_bbox.py:223-241 for given detections: returns (masks_arg i64 [H,W], masks bool [H,W,m])."""
        proto = _cuda_f32(pred_proto, "pred_proto")
        dev = proto.device
        coef = _cuda_f32(mask_coef, "mask_coef").to(dev)
        boxes = _cuda_f32(boxes_px, "boxes_px").to(dev)
        ph, pw, k = proto.shape
        m = coef.shape[0]
        hgt, wid = int(image_shape[0]), int(image_shape[1])
        if k != ISB_NUM_COEF:
            raise ValueError(f"prototype channels must be {ISB_NUM_COEF}")
        nbytes = (hgt * wid * m + 15) // 16 * 16
        masks = torch.empty((max(nbytes, 16),), dtype=torch.uint8, device=dev)
        arg = torch.empty((hgt, wid), dtype=torch.int64, device=dev)
        h = get_handle(dev.index, max_det=max(m, 1))
        with _lib.on_device(dev):
            h.check(h.lib.isb_mask_assemble(h.ptr, _ptr(proto), _ptr(coef), _ptr(boxes), ph, pw, m, hgt, wid, _ptr(arg),
                                            _ptr(masks), _stream(dev)))
        return arg, masks[: hgt * wid * m].view(hgt, wid, m).view(torch.bool)

    # ------------------------------------------------------------------ batched fused path
    def decode_nms_batch(self, outputs, anchors, confidence, nms_iou, image_shape, top_k=TOP_K,
                         max_detections=MAX_DETECTIONS, out=None, logits=False):
        """Fused Detect + mask assembly for a batch: outputs = (box [B,A,4], conf [B,A,C],
        coef [B,A,K], proto [B,Ph,Pw,K]) on one CUDA device.  Enqueues three kernels on the
        current stream and returns a YolactBatchResult without synchronising.
        logits=True: outputs[1] holds the raw class logits of the prediction heads and the softmax of nets/yolact.py:204 is
        fused into the score filter."""
        box, conf, coef, proto = [_cuda_f32(t, n) for t, n in zip(outputs, ("box", "conf", "coef", "proto"))]
        dev = box.device
        anchors = _cuda_f32(anchors, "anchors").to(dev)
        if box.dim() != 3 or conf.dim() != 3 or coef.dim() != 3 or proto.dim() != 4:
            raise ValueError("decode_nms_batch expects [B,A,4], [B,A,C], [B,A,K], [B,Ph,Pw,K]")
        bsz, a, c = conf.shape
        ph, pw, k = proto.shape[1:]
        hgt, wid = int(image_shape[0]), int(image_shape[1])
        max_det = int(max_detections)
        slot = (hgt * wid * max_det + 15) // 16 * 16
        if out is None:
            out = YolactBatchResult(
                torch.empty((bsz, max_det, 4), dtype=torch.float32, device=dev),
                torch.empty((bsz, max_det), dtype=torch.float32, device=dev),
                torch.empty((bsz, max_det), dtype=torch.int64, device=dev),
                torch.empty((bsz, hgt, wid), dtype=torch.int64, device=dev),
                torch.empty((bsz, slot), dtype=torch.uint8, device=dev),
                torch.empty((bsz,), dtype=torch.int32, device=dev), (hgt, wid))
        h = get_handle(dev.index, max_batch=bsz, max_anchors=a, max_classes=c - 1, max_det=max_det, max_top_k=int(top_k))
        pin = YolactIn(bsz, a, c, k, ph, pw, hgt, wid, box.data_ptr(), conf.data_ptr(), coef.data_ptr(), proto.data_ptr(),
                       anchors.data_ptr(), float(confidence), float(nms_iou), int(top_k), max_det, 1 if logits else 0)
        pout = YolactOut(out.boxes.data_ptr(), out.scores.data_ptr(), out.class_ids.data_ptr(), out.masks_arg.data_ptr(),
                         out.masks_raw.data_ptr(), out.masks_raw.shape[1], out.counts.data_ptr())
        with _lib.on_device(dev):
            h.check(h.lib.isb_yolact_detect(h.ptr, C.byref(pin), C.byref(pout), _stream(dev)))
        return out

    # ------------------------------------------------------------------ yolact.py:162-170
    def decode_nms_class_map_batch(self, outputs, anchors, confidence, nms_iou, image_shape, top_k=TOP_K,
                                   max_detections=MAX_DETECTIONS, out=None, logits=False):
        """Detect + the wrapper's per-pixel class image (inst_model/yolact/yolact.py:162-170) for a batch, without ever
        materialising bool [H,W,m] / int64 [H,W]: returns (boxes [B,D,4], scores [B,D], class_ids [B,D] i64,
        class_map [B,H,W] u8, counts [B] i32) on the device, nothing synchronised.
        class_map[b,y,x] = class_ids[b, masks_arg[y,x]] + 1 where that detection's mask is set, else 0."""
        box, conf, coef, proto = [_cuda_f32(t, n) for t, n in zip(outputs, ("box", "conf", "coef", "proto"))]
        dev = box.device
        anchors = _cuda_f32(anchors, "anchors").to(dev)
        if box.dim() != 3 or conf.dim() != 3 or coef.dim() != 3 or proto.dim() != 4:
            raise ValueError("decode_nms_class_map_batch expects [B,A,4], [B,A,C], [B,A,K], [B,Ph,Pw,K]")
        bsz, a, c = conf.shape
        ph, pw, k = proto.shape[1:]
        hgt, wid = int(image_shape[0]), int(image_shape[1])
        max_det = int(max_detections)
        if out is None:
            out = (torch.empty((bsz, max_det, 4), dtype=torch.float32, device=dev),
                   torch.empty((bsz, max_det), dtype=torch.float32, device=dev),
                   torch.empty((bsz, max_det), dtype=torch.int64, device=dev),
                   torch.empty((bsz, hgt, wid), dtype=torch.uint8, device=dev),
                   torch.empty((bsz,), dtype=torch.int32, device=dev))
        h = get_handle(dev.index, max_batch=bsz, max_anchors=a, max_classes=c - 1, max_det=max_det, max_top_k=int(top_k))
        pin = YolactIn(bsz, a, c, k, ph, pw, hgt, wid, box.data_ptr(), conf.data_ptr(), coef.data_ptr(), proto.data_ptr(),
                       anchors.data_ptr(), float(confidence), float(nms_iou), int(top_k), max_det, 1 if logits else 0)
        pout = YolactMapOut(*[t.data_ptr() for t in out])
        with _lib.on_device(dev):
            h.check(h.lib.isb_yolact_class_map(h.ptr, C.byref(pin), C.byref(pout), _stream(dev)))
        return out

    def decode_nms_class_map(self, outputs, anchors, confidence, nms_iou, image_shape):
        """One image, reference-shaped: (box_thre [m,4], class_thre [m], class_ids [m], masks_class u8 [H,W]) -- the four
        things `YOLACT.detect_image` (yolact.py:158-170) uses after decode_nms -- or four `None`s."""
        box, conf, coef, proto = outputs[0], outputs[1], outputs[2], outputs[3]
        box = box.reshape(-1, 4)[None]
        conf = conf.reshape(box.shape[1], -1)[None]
        coef = coef.reshape(box.shape[1], -1)[None]
        proto = proto.reshape((1,) + tuple(proto.shape[-3:]))
        a, c = conf.shape[1], conf.shape[2]
        res = self.decode_nms_class_map_batch((box, conf, coef, proto), anchors, confidence, nms_iou, image_shape)
        m = int(res[4][0].item())
        while m < 0:  # candidate buffer overflow: grow the workspace and redo
            cap = int(get_handle(box.device.index).limits.cand_capacity)
            get_handle(box.device.index, cand_capacity=min(cap * 4, a * (c - 1)))
            res = self.decode_nms_class_map_batch((box, conf, coef, proto), anchors, confidence, nms_iou, image_shape)
            m = int(res[4][0].item())
        if m == 0:
            return (None,) * 4
        return res[0][0, :m], res[1][0, :m], res[2][0, :m], res[3][0]

    def decode_nms_from_logits(self, outputs, anchors, confidence, nms_iou, image_shape):
        """decode_nms (utils_bbox.py:170-243) fed with the heads' RAW class logits: outputs[1] is pred_classes before
        `F.softmax(pred_classes, -1)` (nets/yolact.py:204).  Same 5-tuple / five Nones."""
        box, conf, coef, proto = outputs[0], outputs[1], outputs[2], outputs[3]
        box = box.reshape(-1, 4)[None]
        conf = conf.reshape(box.shape[1], -1)[None]
        coef = coef.reshape(box.shape[1], -1)[None]
        proto = proto.reshape((1,) + tuple(proto.shape[-3:]))
        a, c = conf.shape[1], conf.shape[2]
        res = self.decode_nms_batch((box, conf, coef, proto), anchors, confidence, nms_iou, image_shape, logits=True)
        m = int(res.counts[0].item())
        while m < 0:
            cap = int(get_handle(box.device.index).limits.cand_capacity)
            get_handle(box.device.index, cand_capacity=min(cap * 4, a * (c - 1)))
            res = self.decode_nms_batch((box, conf, coef, proto), anchors, confidence, nms_iou, image_shape, logits=True)
            m = int(res.counts[0].item())
        return res.image(0, m)

    # ------------------------------------------------------------------ utils_map.py:33-34
    def rle_counts(self, masks):
        """Run lengths of every instance mask, computed on the device: masks bool/uint8 [H,W,m] (what decode_nms returns)
        -> list of m int64 numpy arrays, each the `counts` of pycocotools' uncompressed RLE (column-major, zeros first)."""
        if not masks.is_cuda:
            raise RuntimeError("isb.yolact: CUDA tensors only (no CPU fallback)")
        dev = masks.device
        hgt, wid, m = (int(v) for v in masks.shape)
        if m == 0:
            return []
        raw = masks.contiguous().view(torch.uint8)
        h = get_handle(dev.index)
        run_count = torch.empty((m,), dtype=torch.int32, device=dev)
        run_off = torch.empty((m,), dtype=torch.int64, device=dev)
        totals = torch.empty((2,), dtype=torch.int64, device=dev)
        cap = max(4096, 8 * wid * m)  # a blob crosses a column twice; grown below if the masks are busier
        while True:
            out = torch.empty((cap,), dtype=torch.int32, device=dev)
            with _lib.on_device(dev):
                h.check(h.lib.isb_rle_encode(h.ptr, raw.data_ptr(), raw.numel(), None, m, m, 1, hgt, wid, out.data_ptr(), cap,
                                             run_count.data_ptr(), run_off.data_ptr(), totals.data_ptr(), _stream(dev)))
            need = int(totals[1].item())
            if need <= cap:
                break
            cap = need
        flat = out[:need].cpu().numpy().view("uint32")
        cnt, off = run_count.cpu().numpy(), run_off.cpu().numpy()
        return [flat[off[j]:off[j] + cnt[j]].astype("int64") for j in range(m)]

    def rle_encode(self, masks):
        """`pycocotools.mask.encode(np.asfortranarray(masks))` (utils_map.py:34) for decode_nms' bool [H,W,m] masks without
        copying them to the host: list of {'size': [H, W], 'counts': bytes}, one per instance."""
        import numpy as np
        hgt, wid = int(masks.shape[0]), int(masks.shape[1])
        lib = get_handle(masks.device.index).lib
        res = []
        for cnts in self.rle_counts(masks):
            c32 = np.ascontiguousarray(cnts.astype(np.uint32))
            buf = C.create_string_buffer(7 * len(c32) + 8)
            n = lib.isb_rle_to_string(c32.ctypes.data_as(C.c_void_p), len(c32), buf, len(buf))
            if n < 0:
                raise RuntimeError("isb_rle_to_string failed (%d)" % n)
            res.append({"size": [hgt, wid], "counts": buf.raw[:n]})
        return res

    # ------------------------------------------------------------------ utils_bbox.py:170
    def decode_nms(self, outputs, anchors, confidence, nms_iou, image_shape, traditional_nms=False, max_detections=100):
        if traditional_nms:
            return self._decode_nms_traditional(outputs, anchors, confidence, nms_iou, image_shape, max_detections)
        # utils_bbox.py:177-180 squeeze the batch dimension: the reference handles one image.
        box, conf, coef, proto = outputs[0], outputs[1], outputs[2], outputs[3]
        box = box.reshape(-1, 4)[None]
        conf = conf.reshape(box.shape[1], -1)[None]
        coef = coef.reshape(box.shape[1], -1)[None]
        proto = proto.reshape((1,) + tuple(proto.shape[-3:]))
        a, c = conf.shape[1], conf.shape[2]
        # `max_detections` is accepted but, as in the reference (:207), not forwarded to Fast NMS.
        res = self.decode_nms_batch((box, conf, coef, proto), anchors, confidence, nms_iou, image_shape)
        m = int(res.counts[0].item())
        while m < 0:  # candidate buffer overflow: grow the workspace (still the CUDA path) and redo
            cur = get_handle(box.device.index)
            cap = int(cur.limits.cand_capacity)
            get_handle(box.device.index, cand_capacity=min(cap * 4, a * (c - 1)))
            res = self.decode_nms_batch((box, conf, coef, proto), anchors, confidence, nms_iou, image_shape)
            m = int(res.counts[0].item())
        return res.image(0, m)

    def _decode_nms_traditional(self, outputs, anchors, confidence, nms_iou, image_shape, max_detections):
        """utils_bbox.py:170-243 with traditional_nms=True (off by default in the reference): the stage kernels chained on
        the device; torch is used only to slice the surviving rows (:193-204)."""
        box = _cuda_f32(outputs[0], "box").reshape(-1, 4)
        conf = _cuda_f32(outputs[1], "conf").reshape(box.shape[0], -1)
        coef = _cuda_f32(outputs[2], "coef").reshape(box.shape[0], -1)
        proto = _cuda_f32(outputs[3], "proto")
        proto = proto.reshape(tuple(proto.shape[-3:]))
        boxes = self.decode_boxes(box, anchors)                                      # :186
        fg = conf[:, 1:]                                                             # :193
        cmax = fg.max(dim=1).values                                                  # :194
        keep = cmax > confidence                                                     # :195
        box_thre, class_thre, coef_thre = boxes[keep], fg[keep].contiguous(), coef[keep]
        if class_thre.shape[0] == 0:                                                 # :203
            return (None,) * 5
        b, s, c, k = self.traditional_non_max_suppression(box_thre, class_thre, coef_thre, cmax[keep], nms_iou, max_detections)
        b = self.yolact_correct_boxes(b, image_shape)                                # :216
        arg, masks = self.assemble_masks(proto, k, b, image_shape)                   # :223-241
        return b, s, c, arg, masks
