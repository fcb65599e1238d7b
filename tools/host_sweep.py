"""isb_yolact_detect_host at the benchmark workload for several worker-thread counts (ms per 64-image call, split)."""
import ctypes as C
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import bench  # noqa: E402
from isb import _lib  # noqa: E402
from isb import anchors as A  # noqa: E402


def main():
    bsz, d, shape = 64, 100, (550, 550)
    anchors = torch.from_numpy(A.get_anchors(shape)).float()
    a, p = anchors.shape[0], 138
    distinct = bench.synthetic_batch(2000, 16, a, num_classes=81, proto=p)
    host_in = [t.repeat((4,) + (1,) * (t.dim() - 1)).contiguous().pin_memory() for t in distinct]
    pa = anchors.pin_memory()
    slot = (shape[0] * shape[1] * d + 15) // 16 * 16
    hb, hs = torch.empty((bsz, d, 4)).pin_memory(), torch.empty((bsz, d)).pin_memory()
    hc = torch.empty((bsz, d), dtype=torch.int64).pin_memory()
    ha = torch.empty((bsz,) + shape, dtype=torch.int64).pin_memory()
    hm = torch.empty((bsz, slot), dtype=torch.uint8).pin_memory()
    hn = torch.empty((bsz,), dtype=torch.int32).pin_memory()
    h = _lib.get_handle(0, max_batch=bsz)
    pin = _lib.YolactIn(bsz, a, 81, 32, p, p, shape[0], shape[1], host_in[0].data_ptr(), host_in[1].data_ptr(), host_in[2].data_ptr(),
                        host_in[3].data_ptr(), pa.data_ptr(), 0.05, 0.5, 200, d, 0)
    pout = _lib.YolactOut(hb.data_ptr(), hs.data_ptr(), hc.data_ptr(), ha.data_ptr(), hm.data_ptr(), slot, hn.data_ptr())
    res = {}
    for threads in [int(v) for v in sys.argv[1:]] or [0, 2, 3, 4, 6, 8, 10, 12, 15]:
        h.set_host_threads(threads)
        for _ in range(2):
            h.check(h.lib.isb_yolact_detect_host(h.ptr, C.byref(pin), C.byref(pout), None))
        t0 = time.perf_counter()
        for _ in range(6):
            h.check(h.lib.isb_yolact_detect_host(h.ptr, C.byref(pin), C.byref(pout), None))
        ms = 1e3 * (time.perf_counter() - t0) / 6
        st = h.host_stats()
        res[threads] = {"ms": round(ms, 2), "img_s": round(bsz / ms * 1e3), "dense": st["images_copy_engine"], "pool": st["images_host_threads"]}
        print(threads, res[threads], file=sys.stderr)
    print(json.dumps(res))


if __name__ == "__main__":
    main()
