"""One call of the tcgen05 mask-logit GEMM (isb_mask_logits) at the benchmark's shape, for ncu:
138*138 prototype pixels x 100 detections x K = 32 (utils_bbox.py:223)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from isb.yolact import BBoxUtility  # noqa: E402

g = torch.Generator().manual_seed(1)
proto = torch.relu(torch.randn(138, 138, 32, generator=g)).cuda()
coef = torch.tanh(torch.randn(100, 32, generator=g)).cuda()
bu = BBoxUtility()
for _ in range(5):
    out = bu.mask_logits(proto, coef)
torch.cuda.synchronize()
ref = proto.reshape(-1, 32).double() @ coef.double().t()
print("max abs err vs fp64:", float((out.reshape(-1, 100).double() - ref).abs().max()))
