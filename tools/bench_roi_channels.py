import os, sys, json
from collections import OrderedDict
import torch
sys.path.insert(0, "/root/repo")
from isb import maskrcnn as R
from oracle import synth
sys.path.insert(0, "/root/repo/tools")
dev = torch.device("cuda:0")
def time_us(fn, it=30):
    for _ in range(5): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(it): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1000.0 / it
g = torch.Generator().manual_seed(3)
n = 1000
s = torch.exp(torch.rand(n, generator=g) * 2.0 + 3.0)
x = torch.rand(n, generator=g) * (1333 - s); y = torch.rand(n, generator=g) * (800 - s)
boxes = torch.stack([x, y, x + s, y + s], 1).contiguous().to(dev)
p7 = R.MultiScaleRoIAlign(["0", "1", "2", "3"], 7, 2)
res = {"kernel": "rows" if os.environ.get("ISB_ROI_NO_TMA") else "tma"}
for ch in (64, 128, 256, 512):
    feats = OrderedDict((str(i), f.to(dev)) for i, f in enumerate(synth.fpn_features(11, ch)))
    res[f"c{ch}"] = round(time_us(lambda: p7(feats, [boxes], [(800, 1333)])), 1)
    for nr in (250, 500):
        res[f"c{ch}_n{nr}"] = round(time_us(lambda: p7(feats, [boxes[:nr]], [(800, 1333)])), 1)
print(json.dumps(res))
