"""Runs the RPN filter and the box-head post-processing a few times (ncu target)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from isb import maskrcnn as R  # noqa: E402
from oracle import synth  # noqa: E402  (input generator only)
dev = torch.device("cuda:0")
anchors, deltas, logits, per_level = synth.rpn_heads(4001)
anchors, deltas, logits = anchors.to(dev), deltas.to(dev), logits.to(dev)
proposals = R.decode_single(deltas, anchors)
props, lg, reg = [t.to(dev) for t in synth.box_head_outputs(4101, 1000, 21)]
for _ in range(3):
    R.filter_proposals(proposals, logits, (800, 1333), per_level)
    R.postprocess_detections(lg, reg, props, (800, 1333))
torch.cuda.synchronize()
print("ok")
