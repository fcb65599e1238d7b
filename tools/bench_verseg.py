"""BASELINE.json configs[4] -- VerSeg vertebra stress: ONE image, 5000 heavily overlapping boxes of one class through Fast
NMS (top_k = 5000, max_detections = 1024, called directly as SURVEY 8d says: decode_nms cannot raise top_k), then
yolact_correct_boxes + mask assembly of the kept detections (up to 1024) at 550x550.  Device times (CUDA events)."""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from isb.yolact import BBoxUtility  # noqa: E402
from oracle import synth  # noqa: E402  (input generator only)


def timeit(fn, warm=5, iters=20):
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
    bu = BBoxUtility()
    box, score, coef = [t.cuda() for t in synth.verseg_boxes(5001, 5000)]
    gen = torch.Generator().manual_seed(5002)
    proto = torch.relu(torch.randn(138, 138, 32, generator=gen)).cuda()
    res = {}
    nms = lambda: bu.fast_non_max_suppression(box, score, coef, 0.5, 5000, 1024)
    res["fast_nms_5000_top5000_max1024_us"] = timeit(nms)
    b, s, c, k = nms()
    # the same call without the Python mirror's allocations and its 4-byte D2H of the count (device time only)
    import ctypes as C
    from isb._lib import get_handle
    h = get_handle(0)
    ob, os_, oc = torch.empty(1024, 4, device="cuda"), torch.empty(1024, device="cuda"), torch.empty(1024, dtype=torch.int64, device="cuda")
    ok, cnt = torch.empty(1024, 32, device="cuda"), torch.empty(1, dtype=torch.int32, device="cuda")
    P = lambda t: C.c_void_p(t.data_ptr())
    raw = lambda: h.check(h.lib.isb_fast_nms(h.ptr, P(box), P(score), P(coef), 5000, 1, 32, 0.5, 5000, 1024, P(ob), P(os_), P(oc), P(ok),
                                              P(cnt), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    res["fast_nms_device_only_us"] = timeit(raw)
    m = int(b.shape[0])
    res["kept"] = m
    boxes_px = bu.yolact_correct_boxes(b, (550, 550))
    asm = lambda: bu.assemble_masks(proto, k, boxes_px, (550, 550))
    us = timeit(asm, warm=3, iters=10)
    res["assemble_masks_550x550_us"] = us
    res["assemble_out_GBs"] = (550 * 550 * (m + 8)) / us / 1e3
    res["reference_cpu_ms"] = {"fast_nms": 305, "mask_assembly": 1956, "source": "SURVEY.md 8(d) probe, torch CPU"}
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
