"""Generate tests/golden/maskrcnn_*.npz from the REAL reference functions + installed torchvision.

Build container only:   PYTHONDONTWRITEBYTECODE=1 python -m oracle.gen_golden_maskrcnn
"""
from __future__ import annotations

import functools
import sys
import types
from collections import OrderedDict

import numpy as np
import torch

from .gen_golden import REF, _save, digest

sys.dont_write_bytecode = True


def _ref():
    if REF not in sys.path:
        sys.path.insert(0, REF)
    from inst_model.Mask_RCNN.net.network_files import boxes, det_utils, roi_head, rpn_function, transform
    return types.SimpleNamespace(boxes=boxes, det_utils=det_utils, roi_head=roi_head, rpn=rpn_function, transform=transform)


def clustered_boxes(g, n, extent=(800.0, 600.0), clusters=40, jitter=6.0):
    ctr = torch.rand(clusters, 2, generator=g) * torch.tensor(extent)
    size = torch.rand(clusters, 2, generator=g) * 120 + 20
    pick = torch.randint(0, clusters, (n,), generator=g)
    c = ctr[pick] + torch.randn(n, 2, generator=g) * jitter
    s = size[pick] * (1 + torch.randn(n, 2, generator=g) * 0.08)
    return torch.cat([c - s / 2, c + s / 2], 1).contiguous()


def stages(name, seed):
    from torchvision.ops import MultiScaleRoIAlign
    R = _ref()
    g = torch.Generator().manual_seed(seed)
    out = {}
    # --- M1 decode_single: RPN weights (k=1) and box-head weights (k=6 classes), incl. clamp of dw/dh
    n = 400
    xy = torch.rand(n, 2, generator=g) * 700
    prop = torch.cat([xy, xy + torch.rand(n, 2, generator=g) * 250 + 4], 1)
    rel1 = torch.randn(n, 4, generator=g) * 0.6
    rel6 = torch.randn(n, 24, generator=g) * 3.0
    rel6[:7, 2::4] = 60.0   # hits bbox_xform_clip
    out.update(dec_boxes=prop.numpy(), dec_rel1=rel1.numpy(), dec_rel6=rel6.numpy(),
               dec_out1=R.det_utils.BoxCoder((1.0, 1.0, 1.0, 1.0)).decode_single(rel1, prop).numpy(),
               dec_out6=R.det_utils.BoxCoder((10.0, 10.0, 5.0, 5.0)).decode_single(rel6, prop).numpy())
    # --- clip / small boxes
    wild = torch.randn(300, 4, generator=g) * 500 + 300
    out.update(clip_in=wild.numpy(), clip_out=R.boxes.clip_boxes_to_image(wild, (600, 800)).numpy(),
               small_keep=R.boxes.remove_small_boxes(R.boxes.clip_boxes_to_image(wild, (600, 800)), 1.0).numpy())
    # --- nms / batched_nms on clustered boxes, with score ties, two thresholds
    nb = clustered_boxes(g, 2500)
    sc = torch.rand(2500, generator=g)
    sc[::7] = torch.round(sc[::7] * 20) / 20          # ties
    idx = torch.randint(0, 5, (2500,), generator=g)
    out.update(nms_boxes=nb.numpy(), nms_scores=sc.numpy(), nms_idxs=idx.numpy(),
               nms_keep_05=R.boxes.nms(nb, sc, 0.5).numpy(), nms_keep_07=R.boxes.nms(nb, sc, 0.7).numpy(),
               bnms_keep_05=R.boxes.batched_nms(nb, sc, idx, 0.5).numpy(),
               bnms_keep_07=R.boxes.batched_nms(nb, sc, idx, 0.7).numpy())
    # adversarial near-threshold pairs: IoU of (unit-height) boxes [0,w] and [d,w+d] is (w-d)/(w+d)
    adv = []
    for k in range(200):
        w = 100.0 + k
        d = w * (1 - 0.5) / (1 + 0.5) + (k % 5 - 2) * 1e-5 * w      # IoU within ~1e-5 of 0.5
        adv += [[0.0, 10.0 * k, w, 10.0 * k + 5.0], [d, 10.0 * k, w + d, 10.0 * k + 5.0]]
    adv = torch.tensor(adv, dtype=torch.float32)
    asc = torch.rand(adv.shape[0], generator=g)
    out.update(adv_boxes=adv.numpy(), adv_scores=asc.numpy(), adv_keep=R.boxes.nms(adv, asc, 0.5).numpy())
    # --- MultiScaleRoIAlign, small channel count, 7x7 and 14x14
    feats = [torch.randn(1, 8, 200 // 2 ** i, 336 // 2 ** i, generator=g) for i in range(4)]
    from . import synth
    rois = synth.roi_boxes(seed + 5, 120, (800, 1333))
    fd = OrderedDict((str(i), f) for i, f in enumerate(feats))
    for size in (7, 14):
        m = MultiScaleRoIAlign(["0", "1", "2", "3"], size, 2)
        out[f"roi_out{size}"] = m(fd, [rois], [(800, 1333)]).numpy()
    out.update(roi_boxes=rois.numpy(), **{f"roi_feat{i}": f.numpy() for i, f in enumerate(feats)})
    # --- M7 / M8 / M9
    logits = torch.randn(12, 6, 28, 28, generator=g) * 3
    labels = torch.randint(1, 6, (12,), generator=g)
    out.update(inf_logits=logits.numpy(), inf_labels=labels.numpy(),
               inf_out=R.roi_head.maskrcnn_inference(logits, [labels])[0].numpy())
    pb = clustered_boxes(g, 24, extent=(500.0, 400.0), clusters=24)
    pb[0] = torch.tensor([-3.2, -2.0, 40.5, 50.0])          # partly outside (top-left)
    pb[1] = torch.tensor([500.0, 480.0, 560.0, 551.0])      # partly outside (bottom-right)
    pb[2] = torch.tensor([100.2, 100.7, 100.9, 101.1])      # sub-pixel box -> 1x1 / 2x2 paste
    probs = torch.rand(24, 1, 28, 28, generator=g)
    out.update(paste_masks=probs.numpy(), paste_boxes=pb.numpy(),
               paste_out=R.transform.paste_masks_in_image(probs, pb, (544, 544)).numpy(),
               resize_out=R.transform.resize_boxes(pb, (800, 1216), (544, 544)).numpy())
    _save(name, **out)


def rpn_filter(name, seed, image_hw=(800, 1344), image_shape=(800, 1333)):
    from . import synth
    R = _ref()
    anchors, deltas, logits, per_level = synth.rpn_heads(seed, image_hw)
    coder = R.det_utils.BoxCoder((1.0, 1.0, 1.0, 1.0))
    proposals = coder.decode_single(deltas, anchors)
    stub = types.SimpleNamespace(min_size=1.0, score_thresh=0.0, nms_thresh=0.7,
                                 pre_nms_top_n=lambda: 1000, post_nms_top_n=lambda: 1000)
    stub._get_top_n_idx = functools.partial(R.rpn.RegionProposalNetwork._get_top_n_idx, stub)
    boxes, scores = R.rpn.RegionProposalNetwork.filter_proposals(stub, proposals[None], logits[None], [image_shape], per_level)
    _save(name, seed=seed, image_hw=np.array(image_hw), image_shape=np.array(image_shape),
          input_digest=digest(anchors.numpy(), deltas.numpy(), logits.numpy()),
          boxes=boxes[0].numpy(), scores=scores[0].numpy())


def box_post(name, seed, num_classes=21, image_shape=(800, 1333)):
    from . import synth
    R = _ref()
    proposals, logits, reg = synth.box_head_outputs(seed, 1000, num_classes, image_shape)
    stub = types.SimpleNamespace(box_coder=R.det_utils.BoxCoder((10.0, 10.0, 5.0, 5.0)), score_thresh=0.05,
                                 nms_thresh=0.5, detection_per_img=100)
    b, s, l = R.roi_head.RoIHeads.postprocess_detections(stub, logits, reg, [proposals], [image_shape])
    _save(name, seed=seed, num_classes=num_classes, image_shape=np.array(image_shape),
          input_digest=digest(proposals.numpy(), logits.numpy(), reg.numpy()),
          softmax=torch.softmax(logits, -1).numpy(), boxes=b[0].numpy(), scores=s[0].numpy(), labels=l[0].numpy())


def roi_full(name, seed):
    """Config-4 size: 256-channel FPN at 800x1344, 1000 RoIs 7x7 -> store 24 RoIs of the output."""
    from torchvision.ops import MultiScaleRoIAlign
    from . import synth
    feats = synth.fpn_features(seed, 256)
    rois = synth.roi_boxes(seed + 1, 1000)
    fd = OrderedDict((str(i), f) for i, f in enumerate(feats))
    out7 = MultiScaleRoIAlign(["0", "1", "2", "3"], 7, 2)(fd, [rois], [(800, 1333)])
    out14 = MultiScaleRoIAlign(["0", "1", "2", "3"], 14, 2)(fd, [rois[:100]], [(800, 1333)])
    pick = np.arange(0, 1000, 42)
    _save(name, seed=seed, pick=pick, out7=out7[pick].numpy().astype(np.float16 if False else np.float32)[:, ::8],
          out14=out14[::10, ::16].numpy(), sum7=float(out7.double().sum()), sum14=float(out14.double().sum()))


def class_image(name, seed, m=40, paste_hw=(272, 272), sizes=((272, 272), (360, 480), (200, 150))):
    """Wrapper tail mask_rcnn.py:174-198.  The paste is the reference's own `transform.paste_masks_in_image`; the tail
    lives inside `MASK_RCNN.detect_image` (not importable: its module pulls cv2 / pynvml-dependent configs), so its
    torch / numpy statements are re-enacted here one by one on the reference's pasted masks."""
    import torch.nn.functional as F
    R = _ref()
    g = torch.Generator().manual_seed(seed)
    boxes = clustered_boxes(g, m, extent=(float(paste_hw[1]) * 0.8, float(paste_hw[0]) * 0.8), clusters=12, jitter=5.0)
    boxes[:, :2] = boxes[:, :2].clamp(min=-8.0)
    probs = torch.sigmoid(torch.randn(m, 1, 28, 28, generator=g) * 2.5 + 0.5)
    # smooth the random masks a little so that they look like blobs rather than salt and pepper
    probs = F.avg_pool2d(probs, 5, stride=1, padding=2)
    probs = (probs - probs.amin(dim=(2, 3), keepdim=True)) / (probs.amax(dim=(2, 3), keepdim=True) - probs.amin(dim=(2, 3), keepdim=True))
    labels = torch.randint(1, 21, (m,), generator=g)
    scores = torch.rand(m, generator=g)
    confidence = 0.3
    pasted = R.transform.paste_masks_in_image(probs, boxes, paste_hw, padding=1)          # transform.py:212
    out = {}
    for k, (h, w) in enumerate(sizes):
        idxs = np.greater(scores.numpy(), confidence)                                      # :174
        classes = labels.numpy()[idxs]
        masks = pasted.squeeze(1)[torch.from_numpy(idxs)]
        ms = F.interpolate(masks.unsqueeze(0), (h, w), mode="bilinear", align_corners=False).squeeze(0)   # :186
        ms = ms.permute(1, 2, 0).contiguous()
        arg = torch.argmax(ms, dim=-1)
        ms = ms > 0.5
        mc = ms * (classes[None, None, :] + 1)                                             # :195
        mc = np.reshape(mc, [-1, np.shape(ms)[-1]])
        mc = np.reshape(mc[np.arange(np.shape(mc)[0]), np.reshape(arg, [-1])], [h, w])
        out[f"class_map_{k}"] = np.asarray(mc).astype(np.uint8)
        out[f"arg_{k}"] = arg.numpy().astype(np.int16)
    _save(name, seed=seed, boxes=boxes.numpy(), probs=probs.numpy(), labels=labels.numpy(), scores=scores.numpy(),
          confidence=np.float32(confidence), paste_hw=np.array(paste_hw), sizes=np.array(sizes), **out)


IMAGENET_MEAN, IMAGENET_STD = [0.485, 0.456, 0.406], [0.229, 0.224, 0.225]   # faster_rcnn_framework.py:255-258


def transform_golden(name, seed):
    """GeneralizedRCNNTransform.forward (transform.py:436-459) of the reference on seeded pictures: two small batches with
    their inputs, the wrapper's case (544 x 544 -> min 800 / max 1333) as a sub-sampled output, and the output size of
    400 random shapes (the size rule is host logic)."""
    R = _ref()
    out = dict(seed=seed, mean=np.array(IMAGENET_MEAN), std=np.array(IMAGENET_STD))
    g = torch.Generator().manual_seed(seed)
    for tag, mn, mx, shapes in (("a", 160, 200, [(97, 211), (150, 90), (33, 35)]), ("b", 48, 64, [(60, 80), (50, 45)])):
        t = R.transform.GeneralizedRCNNTransform(mn, mx, IMAGENET_MEAN, IMAGENET_STD, train_mode=False)
        imgs = [torch.rand(3, h, w, generator=g) for h, w in shapes]
        with torch.no_grad():
            il, _ = t(imgs)
        out[f"{tag}_min_max"] = np.array([mn, mx])
        out[f"{tag}_n"] = len(imgs)
        for i, im in enumerate(imgs):
            out[f"{tag}_in{i}"] = im.numpy()
        out[f"{tag}_shape"] = np.array(il.tensors.shape)
        out[f"{tag}_batch"] = il.tensors.numpy() if tag == "b" else il.tensors.numpy()[:, :, ::3, ::3]   # a: sub-sampled
        out[f"{tag}_sizes"] = np.array([tuple(s) for s in il.image_sizes])
    t = R.transform.GeneralizedRCNNTransform(800, 1333, IMAGENET_MEAN, IMAGENET_STD, train_mode=False)
    gf = torch.Generator().manual_seed(seed + 1)
    img = torch.rand(3, 544, 544, generator=gf)
    with torch.no_grad():
        il, _ = t([img])
    out["full_hw"] = np.array([544, 544])
    out["full_digest_in"] = digest(img.numpy())
    out["full_sizes"] = np.array([tuple(s) for s in il.image_sizes])
    out["full_shape"] = np.array(il.tensors.shape)
    out["full_sub"] = il.tensors.numpy()[:, :, ::13, ::11]
    out["full_sum"] = float(il.tensors.double().sum())
    rs = np.random.RandomState(seed)
    hw = np.stack([rs.randint(20, 1400, 400), rs.randint(20, 2000, 400)], 1)
    got = []
    for h, w in hw:
        with torch.no_grad():
            il, _ = t([torch.zeros(3, int(h), int(w))])
        got.append(tuple(il.image_sizes[0]))
    out["rule_hw"] = hw
    out["rule_sizes"] = np.array(got)
    _save(name, **out)


def chain(name, seed, num_classes=21, image_shape=(800, 1333), original_size=(600, 1000)):
    """The whole inference post-processing chain of the reference at configs[3] size, run by the reference's OWN objects:
    BoxCoder.decode (rpn_function.py:624) -> RegionProposalNetwork.filter_proposals (:460) -> RoIHeads.forward
    (roi_head.py:459-557: box_roi_pool 7x7, postprocess_detections, mask_roi_pool 14x14, maskrcnn_inference) ->
    GeneralizedRCNNTransform.postprocess (transform.py:388-418).  The network layers in between (box head, predictors,
    mask head) are the deterministic stand-ins of oracle/synth.py; everything else is reference / torchvision code."""
    from torchvision.ops import MultiScaleRoIAlign
    from . import synth
    R = _ref()
    feats, anchors, deltas, logits, per_level = synth.chain_inputs(seed)
    coder = R.det_utils.BoxCoder((1.0, 1.0, 1.0, 1.0))
    proposals = coder.decode(deltas, [anchors]).view(1, -1, 4)                       # rpn_function.py:624-625
    rpn = R.rpn.RegionProposalNetwork(None, None, 0.7, 0.3, 256, 0.5, dict(training=2000, testing=1000),
                                      dict(training=2000, testing=1000), 0.7, train_mode=False)
    boxes, scores = rpn.filter_proposals(proposals, logits[None], [image_shape], per_level)
    heads = R.roi_head.RoIHeads(MultiScaleRoIAlign(["0", "1", "2", "3"], 7, 2), synth.ChainBoxHead(),
                                synth.ChainBoxPredictor(num_classes), 0.5, 0.5, 512, 0.25, None, 0.05, 0.5, 100,
                                mask_roi_pool=MultiScaleRoIAlign(["0", "1", "2", "3"], 14, 2), mask_head=torch.nn.Identity(),
                                mask_predictor=synth.ChainMaskPredictor(num_classes), train_mode=False)
    fd = OrderedDict((str(i), f) for i, f in enumerate(feats))
    with torch.no_grad():
        det, _ = heads(fd, boxes, [image_shape])
    masks28 = det[0]["masks"].clone()
    boxes_net = det[0]["boxes"].clone()
    tf = R.transform.GeneralizedRCNNTransform(800, 1333, IMAGENET_MEAN, IMAGENET_STD, train_mode=False)
    out = tf.postprocess(det, [image_shape], [original_size])[0]
    pasted = out["masks"]
    nz = pasted != 0
    _save(name, seed=seed, num_classes=num_classes, image_shape=np.array(image_shape), original_size=np.array(original_size),
          input_digest=digest(*[f.numpy() for f in feats], anchors.numpy(), deltas.numpy(), logits.numpy()),
          rpn_boxes=boxes[0].numpy(), rpn_scores=scores[0].numpy(),
          det_boxes_net=boxes_net.numpy(), det_scores=out["scores"].numpy(), det_labels=out["labels"].numpy(),
          masks28=masks28.numpy(), det_boxes=out["boxes"].numpy(),
          pasted_sub=pasted[:, :, ::5, ::7].numpy(), pasted_sum=pasted.double().sum(dim=(1, 2, 3)).numpy(),
          pasted_nonzero=nz.sum(dim=(1, 2, 3)).numpy(),
          pasted_rows=np.packbits(nz.any(dim=3).numpy()), pasted_cols=np.packbits(nz.any(dim=2).numpy()))


def paste_full(name, seed, m=100, image_shape=(800, 1333)):
    """paste_masks_in_image (transform.py:212-230) at configs[3] size: 100 masks into (800, 1333) = 426.6 MB of float32.
    Stored: a sub-sampled copy, per-mask float64 sums, the exact zero pattern as row / column occupancy and counts."""
    R = _ref()
    g = torch.Generator().manual_seed(seed)
    boxes = clustered_boxes(g, m, extent=(1250.0, 760.0), clusters=30, jitter=9.0)
    boxes[0] = torch.tensor([-20.5, -11.0, 300.2, 410.0])
    boxes[1] = torch.tensor([1200.0, 700.0, 1400.0, 830.5])
    boxes[2] = torch.tensor([0.0, 0.0, 1333.0, 800.0])
    boxes[3] = torch.tensor([640.3, 400.2, 640.9, 400.6])
    probs = torch.rand(m, 1, 28, 28, generator=g)
    pasted = R.transform.paste_masks_in_image(probs, boxes, image_shape)
    nz = pasted != 0
    _save(name, seed=seed, image_shape=np.array(image_shape), boxes=boxes.numpy(), probs=probs.numpy(),
          sub=pasted[:, :, ::7, ::9].numpy(), sums=pasted.double().sum(dim=(1, 2, 3)).numpy(),
          nonzero=nz.sum(dim=(1, 2, 3)).numpy(), rows=np.packbits(nz.any(dim=3).numpy()), cols=np.packbits(nz.any(dim=2).numpy()),
          full2=pasted[[0, 57], 0].numpy())


def nan_inf(name, seed):
    """NaN / +-inf boxes and scores through the reference's own Fast NMS (utils_bbox.py:27-99: torch.min / max / clamp and
    the column max propagate NaN, so a NaN IoU suppresses) and through boxes.nms / batched_nms (boxes.py:7-83 ->
    torchvision's CPU kernel, std::max / std::min and `ovr > thr`, where a NaN IoU does NOT suppress)."""
    from .gen_golden import _ref as _yolact_ref
    R = _ref()
    bu = _yolact_ref()
    g = torch.Generator().manual_seed(seed)
    nan, inf = float("nan"), float("inf")
    # ---- Fast NMS: 3 classes x 64 boxes, clustered so that plenty of pairs overlap
    n, c = 64, 3
    box = clustered_boxes(g, n, extent=(1.0, 1.0), clusters=6, jitter=0.01)
    box = torch.cat([box[:, :2].clamp(0, 1), box[:, :2].clamp(0, 1) + (torch.rand(n, 2, generator=g) * 0.2 + 0.05)], 1)
    sc = torch.rand(n, c, generator=g)
    coef = torch.tanh(torch.randn(n, 32, generator=g))
    box[3, 0] = nan                      # NaN corner: NaN area -> suppresses every lower-ranked box of each class it is in
    box[9] = torch.tensor([0.2, 0.2, inf, 0.6])        # infinite width: IoU 0 with everything
    box[17] = torch.tensor([-inf, 0.1, inf, 0.9])      # infinite both ways
    box[25] = torch.tensor([inf, 0.3, inf, 0.5])       # inf - inf = NaN area
    box[31] = torch.tensor([0.4, 0.4, 0.4, inf])       # zero width x infinite height: 0 * inf = NaN intersection
    sc[40, 0] = inf
    sc[41, 1] = -inf
    sc[42, 2] = nan                      # torch.sort(descending=True) ranks NaN first
    sc[3, 0] = 0.99                      # the NaN box near the top of class 0 ...
    sc[3, 1] = 0.0001                    # ... and at the bottom of class 1
    out = dict(seed=seed, f_box=box.numpy(), f_scores=sc.numpy(), f_coef=coef.numpy())
    for tag, thr, top_k, max_det in (("a", 0.5, 200, 100), ("b", 0.3, 20, 15)):
        b, s, cl, k = bu.fast_non_max_suppression(box, sc, coef, thr, top_k, max_det)
        out.update({f"f{tag}_box": b.numpy(), f"f{tag}_score": s.numpy(), f"f{tag}_cls": cl.numpy(), f"f{tag}_coef": k.numpy(),
                    f"f{tag}_params": np.array([thr, top_k, max_det])})
    out["f_iou"] = bu.jaccard(box[:12], box[:36]).numpy()
    # ---- torchvision-style greedy NMS
    m = 90
    nb = clustered_boxes(g, m, extent=(300.0, 200.0), clusters=8, jitter=3.0)
    ns = torch.rand(m, generator=g)
    idx = torch.randint(0, 3, (m,), generator=g)
    nb[5, 2] = nan
    nb[11, 1] = nan
    nb[20] = torch.tensor([10.0, 10.0, inf, 80.0])
    nb[33] = torch.tensor([-inf, -inf, inf, inf])
    nb[47] = torch.tensor([inf, 5.0, inf, 50.0])
    ns[60] = inf
    ns[61] = -inf
    ns[62] = nan
    ns[5] = 0.97
    out.update(n_boxes=nb.numpy(), n_scores=ns.numpy(), n_idxs=idx.numpy(),
               n_keep=R.boxes.nms(nb, ns, 0.5).numpy(), n_keep_b=R.boxes.batched_nms(nb, ns, idx, 0.5).numpy())
    # finite boxes, non-finite scores only (batched_nms' max coordinate stays finite)
    fb = clustered_boxes(g, m, extent=(300.0, 200.0), clusters=8, jitter=3.0)
    out.update(n2_boxes=fb.numpy(), n2_keep=R.boxes.nms(fb, ns, 0.5).numpy(), n2_keep_b=R.boxes.batched_nms(fb, ns, idx, 0.5).numpy())
    # one infinite coordinate: boxes.py:78 max_coordinate = inf -> offsets inf / NaN
    ib = fb.clone()
    ib[7, 2] = inf
    fs = torch.rand(m, generator=g)
    out.update(n3_boxes=ib.numpy(), n3_scores=fs.numpy(), n3_keep=R.boxes.nms(ib, fs, 0.5).numpy(),
               n3_keep_b=R.boxes.batched_nms(ib, fs, idx, 0.5).numpy())
    _save(name, **out)


def main():
    nan_inf("nan_inf_cases", 4801)
    chain("maskrcnn_chain", 4601)
    paste_full("maskrcnn_paste_full", 4701)
    transform_golden("maskrcnn_transform", 4501)
    class_image("maskrcnn_class_image", 4401)
    stages("maskrcnn_stages", 31)
    rpn_filter("maskrcnn_rpn_filter", 4001)
    box_post("maskrcnn_box_post", 4101)
    roi_full("maskrcnn_roi_full", 4201)


if __name__ == "__main__":
    main()
