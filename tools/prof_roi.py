"""Runs MultiScaleRoIAlign 7x7 x1000 and 14x14 x100 a few times (ncu target)."""
import os, sys
from collections import OrderedDict
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from isb import maskrcnn as R  # noqa: E402
from oracle import synth  # noqa: E402  (input generator only)
dev = torch.device("cuda:0")
feats = OrderedDict((str(i), f.to(dev)) for i, f in enumerate(synth.fpn_features(4201, 256)))
rois = synth.roi_boxes(4202, 1000).to(dev)
p7, p14 = R.MultiScaleRoIAlign(["0", "1", "2", "3"], 7, 2), R.MultiScaleRoIAlign(["0", "1", "2", "3"], 14, 2)
for _ in range(3):
    p7(feats, [rois], [(800, 1333)])
    p14(feats, [rois[:100]], [(800, 1333)])
torch.cuda.synchronize()
print("ok")
