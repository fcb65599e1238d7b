"""GPU: `isb_nms_static`, the entry point that allocates nothing (fixed `static __device__` scratch), against `isb_nms`.

Collected last on purpose: the file name sorts after every other test file.  Calls go
through the C ABI with ctypes; must pass alone in a fresh process and after the other tests (the scratch array is shared by
every call on the GPU and keeps the previous call's contents)."""
import ctypes as C

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def lib():
    assert torch.cuda.is_available()
    from isb import _lib
    return _lib


def _boxes(seed, n, span=600.0):
    g = torch.Generator().manual_seed(seed)
    xy = torch.rand(n, 2, generator=g) * span
    wh = 8.0 + torch.rand(n, 2, generator=g) * 120.0
    scores = torch.rand(n, generator=g)
    scores[::7] = scores[min(3, n - 1)]                     # ties: order must fall back to the index, as in isb_nms
    return torch.cat([xy, xy + wh], 1).cuda().contiguous(), scores.cuda().contiguous()


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _existing(lib, boxes, scores, thr):
    n = boxes.shape[0]
    keep = torch.empty((max(n, 1),), dtype=torch.int64, device="cuda")
    count = torch.zeros((1,), dtype=torch.int32, device="cuda")
    h = lib.get_handle(torch.cuda.current_device())
    h.check(h.lib.isb_nms(h.ptr, _ptr(boxes), _ptr(scores), n, float(thr), _ptr(keep), _ptr(count), _stream()))
    return keep[: int(count.item())].cpu().numpy()


def _static(lib, boxes, scores, thr, keep, count):
    err = C.create_string_buffer(256)
    rc = lib.load().isb_nms_static(_ptr(boxes), _ptr(scores), boxes.shape[0], float(thr), _ptr(keep), _ptr(count), _stream(),
                                   err, len(err))
    return rc, err.value.decode()


def test_nms_static_matches_nms_on_two_calls_in_a_row(lib):
    nmax = lib.ISB_NMS_STATIC_MAX
    cases = [(11, nmax, 0.5), (12, 1000, 0.7)]            # the second, SMALLER call sees the first call's leftovers in the scratch
    want = [_existing(lib, *_boxes(seed, n), thr) for seed, n, thr in cases]
    inputs = [_boxes(seed, n) for seed, n, _ in cases]
    # outputs preallocated by the caller, one set per call
    keeps = [torch.full((nmax,), -7, dtype=torch.int64, device="cuda") for _ in cases]
    counts = [torch.full((1,), -7, dtype=torch.int32, device="cuda") for _ in cases]
    # warm-up call (first use on this device: looks the scratch address up, loads the kernels), then watch the memory
    rc, msg = _static(lib, *inputs[1], 0.3, keeps[0], counts[0])
    assert rc == 0, msg
    keeps[0].fill_(-7)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    torch_bytes0, torch_reserved0 = torch.cuda.memory_allocated(), torch.cuda.memory_reserved()
    for (boxes, scores), (_, n, thr), keep, count in zip(inputs, cases, keeps, counts):   # two calls in a row, nothing in between
        rc, msg = _static(lib, boxes, scores, thr, keep, count)
        assert rc == 0, msg
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    assert torch.cuda.memory_allocated() == torch_bytes0 and torch.cuda.memory_reserved() == torch_reserved0
    assert free1 >= free0, f"device memory shrank by {free0 - free1} bytes across two isb_nms_static calls"
    for keep, count, w, (_, n, thr) in zip(keeps, counts, want, cases):
        m = int(count.item())
        assert 0 < len(w) < n
        assert np.array_equal(keep[:m].cpu().numpy(), w), f"n={n} thr={thr}: kept indices differ from isb_nms (index work: bit-exact)"
        assert bool((keep[m:] == -7).all()), "wrote past the kept indices"


def test_nms_static_python_wrapper_edges_and_limit(lib):
    from isb import maskrcnn as R
    nmax = lib.ISB_NMS_STATIC_MAX
    keep = torch.empty((nmax + 1,), dtype=torch.int64, device="cuda")
    count = torch.full((1,), -7, dtype=torch.int32, device="cuda")
    for seed, n, thr in [(21, 1, 0.5), (22, 64, 0.0), (23, 65, 1.0), (24, 2049, 0.3)]:
        boxes, scores = _boxes(seed, n, span=200.0)
        R.nms_static(boxes, scores, thr, keep, count)
        got = keep[: int(count.item())].cpu().numpy()
        assert np.array_equal(got, _existing(lib, boxes, scores, thr)), (n, thr)
    # empty input: count 0, keep untouched
    R.nms_static(torch.empty((0, 4), device="cuda"), torch.empty((0,), device="cuda"), 0.5, keep, count)
    assert int(count.item()) == 0
    # beyond the fixed scratch: a clear error, nothing launched
    boxes, scores = _boxes(25, nmax + 1)
    count.fill_(-7)
    rc, msg = _static(lib, boxes, scores, 0.5, keep, count)
    assert rc == lib.ISB_ERR_LIMIT and "ISB_NMS_STATIC_MAX" in msg and str(nmax) in msg
    with pytest.raises(lib.IsbError, match="ISB_NMS_STATIC_MAX"):
        R.nms_static(boxes, scores, 0.5, keep, count)
    torch.cuda.synchronize()
    assert int(count.item()) == -7
    # null error buffer is allowed
    rc = lib.load().isb_nms_static(_ptr(boxes), _ptr(scores), nmax + 1, 0.5, _ptr(keep), _ptr(count), _stream(), None, 0)
    assert rc == lib.ISB_ERR_LIMIT
