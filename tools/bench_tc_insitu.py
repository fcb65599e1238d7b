"""In-situ comparison asked for by VERDICT r1 item 10: the fused step (64 pictures, configs[1]) with K3a = FFMA sigmoid maps on the
footprints (default) against K3a = batched tcgen05 product scattering into the same footprint maps (ISB_TC_SMAP=1).  Run twice:
    python tools/bench_tc_insitu.py            # FFMA
    ISB_TC_SMAP=1 python tools/bench_tc_insitu.py
Prints the per-stage device times of isb_yolact_detect (isb_profile_*)."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from isb import _lib  # noqa: E402
from isb import anchors as A  # noqa: E402
from isb.yolact import BBoxUtility  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    anchors = torch.from_numpy(A.get_anchors((550, 550))).float().to(dev)
    a, p = anchors.shape[0], 138
    distinct = bench.synthetic_batch(2000, 16, a, num_classes=81, proto=p)
    ins = [t.repeat((4,) + (1,) * (t.dim() - 1)).contiguous().to(dev) for t in distinct]
    bu = BBoxUtility()
    out = None
    for _ in range(5):
        out = bu.decode_nms_batch(ins, anchors, 0.05, 0.5, (550, 550), out=out)
    torch.cuda.synchronize()
    h = _lib.get_handle(0)
    steps = 50
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):          # unprofiled: the two-half overlap is active (unless ISB_NO_SPLIT is set)
        bu.decode_nms_batch(ins, anchors, 0.05, 0.5, (550, 550), out=out)
    e1.record()
    torch.cuda.synchronize()
    ms_unprofiled = e0.elapsed_time(e1) / steps
    h.profile_begin(steps)
    e0.record()
    for _ in range(steps):          # profiled: four kernels back to back, stage events in between
        bu.decode_nms_batch(ins, anchors, 0.05, 0.5, (550, 550), out=out)
    e1.record()
    torch.cuda.synchronize()
    ms, calls = h.profile_read()
    print(json.dumps({"k3a": "tcgen05 batched product -> footprint scatter" if os.environ.get("ISB_TC_SMAP") else "FFMA on footprints (default)",
                      "ms_per_step": ms_unprofiled, "two_half_overlap": not os.environ.get("ISB_NO_SPLIT"),
                      "ms_per_step_profiled_sequential": e0.elapsed_time(e1) / steps,
                      "stage_us": {"candidates": ms[0] / calls * 1e3, "sort_nms": ms[1] / calls * 1e3,
                                   "sigmoid_maps": ms[2] / calls * 1e3, "mask_compose": ms[3] / calls * 1e3},
                      "mean_detections": float(out.counts.float().mean())}))


if __name__ == "__main__":
    main()
