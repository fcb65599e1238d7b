"""Device time of the reference-shaped nms / batched_nms calls (boxes.py:7, :38) for a few sizes; ISB_NO_PDL=1 for the A/B."""
import json, os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from isb import _lib, maskrcnn as R  # noqa: E402
dev = torch.device("cuda:0")


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


res = {"pdl": not os.environ.get("ISB_NO_PDL")}
g = torch.Generator().manual_seed(5)
h = _lib.get_handle(0)
for n in (1000, 4819, 20000):
    ctr = torch.rand(n, 2, generator=g) * torch.tensor([1333.0, 800.0])
    wh = torch.rand(n, 2, generator=g) * 120 + 8
    boxes = torch.cat([ctr - wh / 2, ctr + wh / 2], 1).contiguous().to(dev)
    scores = torch.rand(n, generator=g).to(dev)
    idxs = torch.randint(0, 5, (n,), generator=g).to(dev)
    keep = torch.empty(n, dtype=torch.int64, device=dev)
    cnt = torch.zeros(1, dtype=torch.int32, device=dev)
    s = torch.cuda.current_stream().cuda_stream
    # device-only: the C entry points with preallocated outputs, no count read
    res[f"isb_nms_{n}_us"] = round(time_us(lambda: h.check(h.lib.isb_nms(h.ptr, boxes.data_ptr(), scores.data_ptr(), n, 0.5,
                                                                          keep.data_ptr(), cnt.data_ptr(), s))), 1)
    res[f"isb_batched_nms_{n}_us"] = round(time_us(lambda: h.check(h.lib.isb_batched_nms(h.ptr, boxes.data_ptr(), scores.data_ptr(),
                                                                                          idxs.data_ptr(), n, 0.7, keep.data_ptr(),
                                                                                          cnt.data_ptr(), s))), 1)
    res[f"python_nms_{n}_us"] = round(time_us(lambda: R.nms(boxes, scores, 0.5)), 1)
print(json.dumps(res))
