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
