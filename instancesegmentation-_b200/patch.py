"""Install the B200 drop-ins into the reference's modules without editing the reference tree.

    import sys; sys.path.insert(0, "/path/to/InstanceSegmentation-")
    import isb.patch; isb.patch.install()
    # eval.py / predict.py now run the CUDA post-processing path

What is rebound:
  inst_model.yolact.yolact.BBoxUtility                      -> isb.yolact.BBoxUtility   (yolact.py:81,158,240,272)
  inst_model.yolact.yolact.get_anchors, ...utils.anchors.get_anchors -> isb.anchors.get_anchors (yolact.py:81, anchors.py:39)
  ...network_files.boxes.nms / batched_nms                  -> isb.maskrcnn.*           (rpn_function.py:521, roi_head.py:446)
  ...network_files.transform.paste_masks_in_image / resize_boxes -> isb.maskrcnn.*      (transform.py:410,415)
  ...network_files.roi_head.maskrcnn_inference              -> isb.maskrcnn.maskrcnn_inference (roi_head.py:555)
  METHODS, bound onto the reference's classes (same `(self, ...)` signatures, same attributes of self read):
  ...det_utils.BoxCoder.decode_single                       -> isb.maskrcnn.boxcoder_decode_single      (det_utils.py:224)
  ...rpn_function.RegionProposalNetwork.filter_proposals    -> isb.maskrcnn.rpn_filter_proposals        (rpn_function.py:460)
  ...roi_head.RoIHeads.postprocess_detections               -> isb.maskrcnn.roiheads_postprocess_detections (roi_head.py:366)
  ...transform.GeneralizedRCNNTransform.postprocess         -> isb.maskrcnn.transform_postprocess       (transform.py:388)
  torchvision MultiScaleRoIAlign instances of a built model -> isb.maskrcnn.MultiScaleRoIAlign (swap_roi_pools)
  model.transform.forward (GeneralizedRCNNTransform)         -> isb.maskrcnn.rcnn_transform (swap_transform)
"""
from __future__ import annotations

import importlib

from . import anchors, maskrcnn, yolact

_REBIND = [
    ("inst_model.yolact.yolact", "BBoxUtility", lambda: yolact.BBoxUtility),
    ("inst_model.yolact.yolact", "get_anchors", lambda: anchors.get_anchors),          # yolact.py:14,81
    ("inst_model.yolact.utils.anchors", "get_anchors", lambda: anchors.get_anchors),   # anchors.py:39
    ("inst_model.Mask_RCNN.net.network_files.boxes", "nms", lambda: maskrcnn.nms),
    ("inst_model.Mask_RCNN.net.network_files.boxes", "batched_nms", lambda: maskrcnn.batched_nms),
    ("inst_model.Mask_RCNN.net.network_files.transform", "paste_masks_in_image", lambda: maskrcnn.paste_masks_in_image),
    ("inst_model.Mask_RCNN.net.network_files.transform", "resize_boxes", lambda: maskrcnn.resize_boxes),
    ("inst_model.Mask_RCNN.net.network_files.roi_head", "maskrcnn_inference", lambda: maskrcnn.maskrcnn_inference),
    # methods: "Class.method" is resolved inside the module and set on the class, so every instance -- built before or
    # after install() -- dispatches to the drop-in
    ("inst_model.Mask_RCNN.net.network_files.det_utils", "BoxCoder.decode_single", lambda: maskrcnn.boxcoder_decode_single),
    ("inst_model.Mask_RCNN.net.network_files.rpn_function", "RegionProposalNetwork.filter_proposals",
     lambda: maskrcnn.rpn_filter_proposals),
    ("inst_model.Mask_RCNN.net.network_files.roi_head", "RoIHeads.postprocess_detections",
     lambda: maskrcnn.roiheads_postprocess_detections),
    ("inst_model.Mask_RCNN.net.network_files.transform", "GeneralizedRCNNTransform.postprocess",
     lambda: maskrcnn.transform_postprocess),
]


def _owner(mod, attr):
    """("Class.method") -> (class object, "method"); ("name") -> (module, "name")."""
    owner = mod
    parts = attr.split(".")
    for part in parts[:-1]:
        owner = getattr(owner, part)
    return owner, parts[-1]


def install(only=None, strict=False):
    """Rebind the reference's names to the drop-ins.  Returns the list of (module, name) rebound.
    Modules that cannot be imported (missing optional dependencies such as cv2) are skipped unless
    `strict`."""
    done = []
    for mod_name, attr, getter in _REBIND:
        if only is not None and attr not in only and attr.split(".")[-1] not in only:
            continue
        try:
            mod = importlib.import_module(mod_name)
        except Exception:
            if strict:
                raise
            continue
        if not hasattr(mod, "_isb_original"):
            mod._isb_original = {}
        owner, name = _owner(mod, attr)
        mod._isb_original.setdefault(attr, owner.__dict__[name] if isinstance(owner, type) else getattr(owner, name))
        setattr(owner, name, getter())
        done.append((mod_name, attr))
    return done


def uninstall():
    for mod_name, attr, _ in _REBIND:
        try:
            mod = importlib.import_module(mod_name)
        except Exception:
            continue
        orig = getattr(mod, "_isb_original", {})
        if attr in orig:
            owner, name = _owner(mod, attr)
            setattr(owner, name, orig.pop(attr))


def swap_roi_pools(model):
    """Replace the torchvision MultiScaleRoIAlign modules of a built Mask R-CNN (roi_heads.box_roi_pool,
    roi_heads.mask_roi_pool; faster_rcnn_framework.py:326-329, mask_rcnn.py:150) by the CUDA drop-in,
    keeping their configuration."""
    import torchvision
    swapped = []
    for name in ("box_roi_pool", "mask_roi_pool"):
        heads = getattr(model, "roi_heads", None)
        pool = getattr(heads, name, None)
        if isinstance(pool, torchvision.ops.MultiScaleRoIAlign) and not isinstance(pool, maskrcnn.MultiScaleRoIAlign):
            new = maskrcnn.MultiScaleRoIAlign(pool.featmap_names, pool.output_size, pool.sampling_ratio,
                                              canonical_scale=pool.canonical_scale, canonical_level=pool.canonical_level)
            setattr(heads, name, new)
            swapped.append(name)
    return swapped


def swap_transform(model):
    """Route `model.transform.forward` (GeneralizedRCNNTransform, transform.py:436-459) through the fused CUDA
    normalize + resize + batch kernel for inference (targets pass through untouched; training keeps the original)."""
    t = getattr(model, "transform", None)
    if t is None or getattr(t, "_isb_swapped", False):
        return False
    image_list = importlib.import_module(type(t).__module__.rsplit(".", 1)[0] + ".image_list")
    original = t.forward

    def forward(images, targets=None):
        if targets is not None or getattr(t, "train_mode", None):
            return original(images, targets)
        batch, sizes = maskrcnn.rcnn_transform(list(images), t.min_size, t.max_size, t.image_mean, t.image_std, t.size_divisible)
        return image_list.ImageList(batch, sizes), targets

    t.forward = forward
    t._isb_swapped = True
    return True
