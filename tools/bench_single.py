"""Latency of ONE image through the reference-shaped call `BBoxUtility.decode_nms` (what predict.py / eval.py make per
picture: batch 1, five tensors back, one 4-byte read of the detection count) and through the class-image variant, for
the benchmark shape (550^2 COCO) and the repo's default (544^2 VOC head, 1080 x 1920 picture)."""
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from isb.yolact import BBoxUtility  # noqa: E402
from oracle import synth, yolact_np as O  # noqa: E402  (input generator only)


def wall_us(fn, warm=10, iters=100):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(iters):
        fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / iters * 1e6


def main():
    bu = BBoxUtility()
    res = {}
    for tag, size, classes, shape, conf, iou in (("550_coco", 550, 81, (550, 550), 0.05, 0.5),
                                                 ("544_voc_1080p", 544, 21, (1080, 1920), 0.5, 0.3)):
        a, p = synth.yolact_dims(size)
        heads = synth.yolact_heads(1234, a, num_classes=classes, proto=p)
        outputs = tuple(t[None].cuda() for t in heads)
        anchors = torch.from_numpy(O.get_anchors((size, size))).float().cuda()
        image_shape = np.array(shape)
        r = bu.decode_nms(outputs, anchors, conf, iou, image_shape)
        m = 0 if r[0] is None else int(r[0].shape[0])
        res[tag] = {"detections": m,
                    "decode_nms_us": wall_us(lambda: bu.decode_nms(outputs, anchors, conf, iou, image_shape)),
                    "decode_nms_class_map_us": wall_us(lambda: bu.decode_nms_class_map(outputs, anchors, conf, iou, image_shape)),
                    "output_MB": (shape[0] * shape[1] * (m + 8)) / 1e6}
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
