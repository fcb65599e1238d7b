"""CPU: isb.patch rebinds the reference's names (only where the reference tree is present, i.e. in the
build container) and restores them."""
import os
import sys
import types

import pytest

REF = "/root/reference"


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree only exists in the build container")
def test_install_and_uninstall():
    sys.dont_write_bytecode = True
    if REF not in sys.path:
        sys.path.insert(0, REF)
    import isb.patch
    from isb import maskrcnn
    done = isb.patch.install()
    names = {a for _, a in done}
    assert {"nms", "batched_nms", "paste_masks_in_image", "resize_boxes", "maskrcnn_inference"} <= names
    from inst_model.Mask_RCNN.net.network_files import boxes, roi_head, transform
    assert boxes.batched_nms is maskrcnn.batched_nms and transform.paste_masks_in_image is maskrcnn.paste_masks_in_image
    assert roi_head.maskrcnn_inference is maskrcnn.maskrcnn_inference
    isb.patch.uninstall()
    assert boxes.batched_nms is not maskrcnn.batched_nms and boxes.batched_nms.__module__.endswith("boxes")


def test_swap_roi_pools():
    import torchvision
    import isb.patch
    from isb import maskrcnn
    heads = types.SimpleNamespace(box_roi_pool=torchvision.ops.MultiScaleRoIAlign(["0", "1"], 7, 2),
                                  mask_roi_pool=torchvision.ops.MultiScaleRoIAlign(["0", "1"], 14, 2))
    model = types.SimpleNamespace(roi_heads=heads)
    assert isb.patch.swap_roi_pools(model) == ["box_roi_pool", "mask_roi_pool"]
    assert isinstance(heads.box_roi_pool, maskrcnn.MultiScaleRoIAlign) and heads.mask_roi_pool.output_size == (14, 14)
    assert isinstance(heads.box_roi_pool, torchvision.ops.MultiScaleRoIAlign)   # reference isinstance checks still pass


def test_swap_transform_routes_inference_only(monkeypatch):
    """swap_transform: training calls and calls with targets keep the reference's forward; inference goes to
    isb.maskrcnn.rcnn_transform and comes back as the reference package's ImageList."""
    import sys
    import isb.patch
    from isb import maskrcnn
    pkg = types.ModuleType("fakeref")
    tmod, imod = types.ModuleType("fakeref.transform"), types.ModuleType("fakeref.image_list")

    class ImageList:
        def __init__(self, tensors, image_sizes):
            self.tensors, self.image_sizes = tensors, image_sizes

    class Transform:
        min_size, max_size, image_mean, image_std, size_divisible, train_mode = (800,), 1333, [0.5] * 3, [0.2] * 3, 32, False

        def forward(self, images, targets=None):
            return "original", targets
    Transform.__module__ = "fakeref.transform"
    imod.ImageList = ImageList
    for name, mod in (("fakeref", pkg), ("fakeref.transform", tmod), ("fakeref.image_list", imod)):
        monkeypatch.setitem(sys.modules, name, mod)
    calls = []
    monkeypatch.setattr(maskrcnn, "rcnn_transform", lambda images, *a: (calls.append(a) or ("batch", [(1, 2)])))
    model = types.SimpleNamespace(transform=Transform())
    assert isb.patch.swap_transform(model) and not isb.patch.swap_transform(model)
    out, tg = model.transform.forward(["img"])
    assert isinstance(out, ImageList) and out.tensors == "batch" and out.image_sizes == [(1, 2)] and tg is None
    assert calls == [((800,), 1333, [0.5] * 3, [0.2] * 3, 32)]
    assert model.transform.forward(["img"], targets=[{}])[0] == "original"
    model.transform.train_mode = True
    assert model.transform.forward(["img"])[0] == "original"


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree only exists in the build container")
def test_method_dropins_bind_with_reference_signatures():
    """the fused Mask R-CNN stages must be reachable from an UNEDITED reference model.  Build the
    reference's own RegionProposalNetwork / RoIHeads / BoxCoder / GeneralizedRCNNTransform objects, install the patch and
    check that the bound methods are ours and carry the reference methods' exact parameter lists."""
    import inspect
    sys.dont_write_bytecode = True
    if REF not in sys.path:
        sys.path.insert(0, REF)
    import isb.patch
    from isb import maskrcnn
    from inst_model.Mask_RCNN.net.network_files import det_utils, roi_head, rpn_function, transform
    isb.patch.uninstall()
    ref_sigs = {
        "decode_single": inspect.signature(det_utils.BoxCoder.decode_single),
        "filter_proposals": inspect.signature(rpn_function.RegionProposalNetwork.filter_proposals),
        "postprocess_detections": inspect.signature(roi_head.RoIHeads.postprocess_detections),
        "postprocess": inspect.signature(transform.GeneralizedRCNNTransform.postprocess),
    }
    # instances built BEFORE install() must be patched too (the methods are set on the classes)
    rpn = rpn_function.RegionProposalNetwork(anchor_generator=None, head=None, fg_iou_thresh=0.7, bg_iou_thresh=0.3,
                                             batch_size_per_image=256, positive_fraction=0.5,
                                             pre_nms_top_n=dict(training=2000, testing=1000),
                                             post_nms_top_n=dict(training=2000, testing=1000), nms_thresh=0.7)
    heads = roi_head.RoIHeads(None, None, None, 0.5, 0.5, 512, 0.25, None, 0.05, 0.5, 100)
    coder = det_utils.BoxCoder((10.0, 10.0, 5.0, 5.0))
    tfm = transform.GeneralizedRCNNTransform(800, 1333, [0.485, 0.456, 0.406], [0.229, 0.224, 0.225])
    try:
        done = {a for _, a in isb.patch.install()}
        assert {"BoxCoder.decode_single", "RegionProposalNetwork.filter_proposals", "RoIHeads.postprocess_detections",
                "GeneralizedRCNNTransform.postprocess"} <= done
        pairs = [(coder.decode_single, maskrcnn.boxcoder_decode_single, "decode_single"),
                 (rpn.filter_proposals, maskrcnn.rpn_filter_proposals, "filter_proposals"),
                 (heads.postprocess_detections, maskrcnn.roiheads_postprocess_detections, "postprocess_detections"),
                 (tfm.postprocess, maskrcnn.transform_postprocess, "postprocess")]
        for bound, ours, name in pairs:
            assert bound.__func__ is ours, name
            assert inspect.signature(ours) == ref_sigs[name], (name, inspect.signature(ours), ref_sigs[name])
        # the drop-ins read exactly these attributes of the reference objects
        assert rpn.pre_nms_top_n() == 1000 and rpn.post_nms_top_n() == 1000 and rpn.nms_thresh == 0.7
        assert rpn.score_thresh == 0.0 and rpn.min_size == 1.0
        assert heads.box_coder.weights == (10.0, 10.0, 5.0, 5.0) and heads.detection_per_img == 100
        assert heads.score_thresh == 0.05 and heads.nms_thresh == 0.5 and heads.box_coder.bbox_xform_clip > 4.13
        # no CPU fallback: CPU tensors raise instead of silently running the reference
        import torch
        with pytest.raises(RuntimeError):
            coder.decode_single(torch.zeros(2, 4), torch.zeros(2, 4))
    finally:
        isb.patch.uninstall()
    assert rpn_function.RegionProposalNetwork.filter_proposals.__module__.endswith("rpn_function")
    assert det_utils.BoxCoder.decode_single.__module__.endswith("det_utils")


def test_integration_md_structs_match_the_binding():
    """INTEGRATION.md shows the ctypes structs a maintainer would copy: they must have the layout of isb/_lib.py (and so of
    include/isb.h) -- a stale struct makes the library read past the caller's memory."""
    import ctypes as C
    import re
    from isb import _lib
    text = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "INTEGRATION.md")).read()
    blocks = re.findall(r"class (\w+)\(C\.Structure\):\n\s+_fields_ = \[(.*?)\][ \t]*(?:#[^\n]*)?\n", text, flags=re.S)
    seen = {}
    for name, body in blocks:
        body = re.sub(r"#.*", "", body)
        fields = re.findall(r'\("(\w+)",\s*C\.(\w+)\)', body)
        doc = type(name, (C.Structure,), {"_fields_": [(f, getattr(C, t)) for f, t in fields]})
        seen[name] = doc
        ours = getattr(_lib, name)
        assert C.sizeof(doc) == C.sizeof(ours), (name, C.sizeof(doc), C.sizeof(ours))
        assert [f for f, _ in doc._fields_] == [f for f, _ in ours._fields_], name
        for f, _ in doc._fields_:
            assert getattr(doc, f).offset == getattr(ours, f).offset, (name, f)
    assert {"Limits", "YolactIn", "YolactOut"} <= set(seen)
