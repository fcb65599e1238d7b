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
