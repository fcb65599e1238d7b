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
    # the C entry points alone (preallocated outputs, no count read-back): device time of the call itself
    import ctypes as C
    from isb._lib import get_handle
    h = get_handle(0)
    vp = lambda t: C.c_void_p(t.data_ptr())
    stream = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)
    ob, osc, ocnt = torch.empty(1000, 4, device=dev), torch.empty(1000, device=dev), torch.zeros(1, dtype=torch.int32, device=dev)
    lv = (C.c_int32 * len(per_level))(*[int(v) for v in per_level])
    objf = logits.reshape(-1).contiguous()

    def rpn_c():
        h.check(h.lib.isb_rpn_filter(h.ptr, vp(proposals), vp(objf), lv, len(per_level), 800, 1333, 1000, 1000, 0.7, 0.0, 1.0, vp(ob),
                                     vp(osc), vp(ocnt), stream()))
    res["isb_rpn_filter_device_only"] = {"us": timeit(rpn_c), "kept": int(ocnt.item()),
                                         "path": "chain" if os.environ.get("ISB_RCNN_CHAIN") else "one cooperative launch"}
    props, lg, reg = [t.to(dev) for t in synth.box_head_outputs(4101, 1000, 21)]
    db, ds, dl = torch.empty(100, 4, device=dev), torch.empty(100, device=dev), torch.empty(100, dtype=torch.int64, device=dev)

    def box_c():
        h.check(h.lib.isb_box_postprocess(h.ptr, vp(lg), C.c_void_p(0), vp(reg), vp(props), 1000, 21, 800, 1333, 10.0, 10.0, 5.0, 5.0,
                                          R.BBOX_XFORM_CLIP, 0.05, 0.5, 100, vp(db), vp(ds), vp(dl), vp(ocnt), stream()))
    res["isb_box_postprocess_device_only_1000x21"] = {"us": timeit(box_c), "kept": int(ocnt.item())}
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
