import os, sys, torch
sys.path.insert(0, "/root/repo")
from isb import _lib
dev = torch.device("cuda:0"); g = torch.Generator().manual_seed(5); h = _lib.get_handle(0)
for n in (4819, 20000):
    ctr = torch.rand(n, 2, generator=g) * torch.tensor([1333.0, 800.0]); wh = torch.rand(n, 2, generator=g) * 120 + 8
    boxes = torch.cat([ctr - wh / 2, ctr + wh / 2], 1).contiguous().to(dev); scores = torch.rand(n, generator=g).to(dev)
    keep = torch.empty(n, dtype=torch.int64, device=dev); cnt = torch.zeros(1, dtype=torch.int32, device=dev)
    s = torch.cuda.current_stream().cuda_stream
    for _ in range(3):
        h.check(h.lib.isb_nms(h.ptr, boxes.data_ptr(), scores.data_ptr(), n, 0.5, keep.data_ptr(), cnt.data_ptr(), s))
    torch.cuda.synchronize(); print(n, int(cnt.item()))
