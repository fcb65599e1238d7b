"""Per-stage device timing of the Mask R-CNN side at BASELINE.json configs[3] sizes
(800x1333 image, padded 800x1344 FPN, 268569 anchors, 1000 proposals, 100 detections)."""
import json
import os
import sys
from collections import OrderedDict

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from isb import maskrcnn as R  # noqa: E402
from oracle import synth  # noqa: E402  (input generator only)


def timeit(fn, warm=10, iters=50):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3  # us


def main():
    dev = torch.device("cuda:0")
    res = OrderedDict()
    anchors, deltas, logits, per_level = synth.rpn_heads(4001)
    anchors, deltas, logits = anchors.to(dev), deltas.to(dev), logits.to(dev)
    a = anchors.shape[0]
    us = timeit(lambda: R.decode_single(deltas, anchors))
    res["decode_single_268569"] = {"us": us, "GBs": a * 48 / us / 1e3}
    proposals = R.decode_single(deltas, anchors)
    us = timeit(lambda: R.filter_proposals(proposals, logits, (800, 1333), per_level))
    res["rpn_filter_268569_to_1000"] = {"us": us}
    feats = OrderedDict((str(i), f.to(dev)) for i, f in enumerate(synth.fpn_features(4201, 256)))
    rois = synth.roi_boxes(4202, 1000).to(dev)
    res["roi_kernel"] = "staged (round 1, ISB_ROI_STAGED)" if os.environ.get("ISB_ROI_STAGED") else "register / shuffle rows kernel"
    p7, p14 = R.MultiScaleRoIAlign(["0", "1", "2", "3"], 7, 2), R.MultiScaleRoIAlign(["0", "1", "2", "3"], 14, 2)
    us = timeit(lambda: p7(feats, [rois], [(800, 1333)]))
    res["ms_roi_align_7x7_x1000"] = {"us": us, "out_GBs": 1000 * 256 * 49 * 4 / us / 1e3}
    us = timeit(lambda: p14(feats, [rois[:100]], [(800, 1333)]))
    res["ms_roi_align_14x14_x100"] = {"us": us, "out_GBs": 100 * 256 * 196 * 4 / us / 1e3}
    for c in (21, 81):
        props, lg, reg = [t.to(dev) for t in synth.box_head_outputs(4101, 1000, c)]
        us = timeit(lambda: R.postprocess_detections(lg, reg, props, (800, 1333)))
        res[f"box_postprocess_1000x{c}"] = {"us": us}
    mlog = torch.randn(100, 21, 28, 28, device=dev)
    labels = torch.randint(1, 21, (100,), device=dev)
    us = timeit(lambda: R.maskrcnn_inference(mlog, [labels]))
    res["mask_select_sigmoid_100"] = {"us": us}
    probs = R.maskrcnn_inference(mlog, [labels])[0]
    boxes = synth.roi_boxes(4301, 100).to(dev)
    for shape in ((800, 1333), (544, 544)):
        out = torch.empty((100, 1) + shape, device=dev)
        import ctypes as C
        from isb._lib import get_handle
        h = get_handle(0)

        def paste():
            h.check(h.lib.isb_paste_masks(h.ptr, C.c_void_p(probs.data_ptr()), C.c_void_p(boxes.data_ptr()), 100, 28, 1,
                                          shape[0], shape[1], C.c_void_p(out.data_ptr()),
                                          C.c_void_p(torch.cuda.current_stream().cuda_stream)))
        us = timeit(paste)
        res[f"paste_masks_100_to_{shape[0]}x{shape[1]}"] = {"us": us, "out_GBs": 100 * shape[0] * shape[1] * 4 / us / 1e3}
    # GeneralizedRCNNTransform.forward (transform.py:436-459): the wrapper's 544 x 544 tensor -> 800 x 800, and a 1080p picture
    for hw in ((544, 544), (1080, 1920)):
        img = torch.rand(3, *hw, device=dev)
        us = timeit(lambda: R.rcnn_transform([img], 800, 1333, [0.485, 0.456, 0.406], [0.229, 0.224, 0.225]))
        oh, ow = R.transform_output_size(hw[0], hw[1], 800.0, 1333.0)
        res[f"rcnn_transform_{hw[0]}x{hw[1]}_to_{oh}x{ow}"] = {"us": us, "GBs": (3 * hw[0] * hw[1] + 3 * ((oh + 31) // 32 * 32) * ((ow + 31) // 32 * 32)) * 4 / us / 1e3}
    # wrapper tail fused with the paste (mask_rcnn.py:174-198): 544x544 paste frame -> class image at the original size
    scores = torch.rand(100, device=dev)
    b544 = boxes * (544.0 / 1333.0)
    for shape in ((544, 544), (1080, 1920)):
        us = timeit(lambda: R.masks_class_image(probs, b544, labels, scores, 0.5, (544, 544), shape))
        res[f"class_image_100_544x544_to_{shape[0]}x{shape[1]}"] = {"us": us}
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
