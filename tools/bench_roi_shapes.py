"""RoIAlign timing by box population: square-ish boxes (every window fits the TMA kernel's 32 x 32 tile) vs the elongated
population of oracle.synth.roi_boxes.  Run once as is and once with ISB_ROI_NO_TMA=1 (rows kernel alone)."""
import json, os, sys
from collections import OrderedDict
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from isb import maskrcnn as R  # noqa: E402
from oracle import synth  # noqa: E402  (input generator only)
dev = torch.device("cuda:0")
feats = OrderedDict((str(i), f.to(dev)) for i, f in enumerate(synth.fpn_features(4201, 256)))


def square_boxes(seed, n, hw=(800, 1333), lo=16.0, hi=600.0, max_aspect=1.5):
    g = torch.Generator().manual_seed(seed)
    s = torch.exp(torch.rand(n, generator=g) * (torch.log(torch.tensor(hi)) - torch.log(torch.tensor(lo))) + torch.log(torch.tensor(lo)))
    a = torch.exp((torch.rand(n, generator=g) * 2 - 1) * torch.log(torch.tensor(max_aspect)) / 2)
    w, h = (s * a).clamp(max=hw[1] - 1), (s / a).clamp(max=hw[0] - 1)
    x = torch.rand(n, generator=g) * (hw[1] - w)
    y = torch.rand(n, generator=g) * (hw[0] - h)
    return torch.stack([x, y, x + w, y + h], 1).contiguous()


def time_us(fn, it=50):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1000.0 / it


res = {"kernel": "rows only" if os.environ.get("ISB_ROI_NO_TMA") else "tma"}
p7, p14 = R.MultiScaleRoIAlign(["0", "1", "2", "3"], 7, 2), R.MultiScaleRoIAlign(["0", "1", "2", "3"], 14, 2)
for name, boxes in (("square_small_16_112", square_boxes(1, 1000, hi=112.0)), ("square_16_600", square_boxes(2, 1000)),
                    ("elongated_synth", synth.roi_boxes(4202, 1000))):
    b = boxes.to(dev)
    res[name] = {"7x7_x1000_us": time_us(lambda: p7(feats, [b], [(800, 1333)])),
                 "14x14_x100_us": time_us(lambda: p14(feats, [b[:100]], [(800, 1333)]))}
print(json.dumps(res, indent=1))
