"""ctypes binding of libisb.so (include/isb.h).  Fails loudly if the library is absent."""
from __future__ import annotations

import ctypes as C
import os
import threading

PKG = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(PKG, "libisb.so")

ISB_NUM_COEF = 32
ISB_MAX_DET_LIMIT = 1024
ISB_NMS_STATIC_MAX = 4096  # boxes per isb_nms_static call (fixed device scratch)
ISB_OK, ISB_ERR_BAD_ARG, ISB_ERR_CUDA, ISB_ERR_LIMIT, ISB_ERR_OVERFLOW = 0, -1, -2, -3, -4

# every exported symbol of include/isb.h (tests check the library exports exactly these)
SYMBOLS = (
    "isb_version", "isb_create", "isb_destroy", "isb_last_error", "isb_workspace_bytes", "isb_launch_count",
    "isb_profile_begin", "isb_profile_read",
    "isb_yolact_detect", "isb_yolact_detect_host", "isb_set_host_threads", "isb_host_stats",
    "isb_counts_export", "isb_counts_attach", "isb_counts_wait", "isb_counts_epoch", "isb_counts_detach", "isb_yolact_class_map", "isb_yolact_class_map_host", "isb_rle_encode", "isb_rle_to_string", "isb_decode_boxes", "isb_jaccard", "isb_fast_nms",
    "isb_traditional_nms", "isb_correct_boxes", "isb_crop", "isb_mask_logits", "isb_mask_logits_batch", "isb_mask_assemble",
    "isb_box_decode", "isb_nms", "isb_nms_static", "isb_batched_nms", "isb_ms_roi_align", "isb_mask_select_sigmoid", "isb_resize_boxes",
    "isb_paste_masks", "isb_paste_class_map", "isb_rpn_filter", "isb_box_postprocess", "isb_box_postprocess_rows", "isb_preprocess",
    "isb_rcnn_transform",
)


class IsbError(RuntimeError):
    def __init__(self, status, message):
        super().__init__(f"libisb status {status}: {message}")
        self.status = status


class Limits(C.Structure):
    _fields_ = [("max_batch", C.c_int32), ("max_anchors", C.c_int32), ("max_classes", C.c_int32),
                ("max_det", C.c_int32), ("max_top_k", C.c_int32), ("cand_capacity", C.c_int64)]


class YolactIn(C.Structure):
    _fields_ = [("batch", C.c_int32), ("num_anchors", C.c_int32), ("num_classes", C.c_int32), ("num_coef", C.c_int32),
                ("proto_h", C.c_int32), ("proto_w", C.c_int32), ("out_h", C.c_int32), ("out_w", C.c_int32),
                ("box", C.c_void_p), ("conf", C.c_void_p), ("coef", C.c_void_p), ("proto", C.c_void_p),
                ("anchors", C.c_void_p), ("confidence", C.c_float), ("nms_iou", C.c_float),
                ("top_k", C.c_int32), ("max_det", C.c_int32), ("conf_is_logits", C.c_int32)]


class YolactOut(C.Structure):
    _fields_ = [("boxes", C.c_void_p), ("scores", C.c_void_p), ("class_ids", C.c_void_p), ("masks_arg", C.c_void_p),
                ("masks", C.c_void_p), ("mask_slot_bytes", C.c_int64), ("counts", C.c_void_p)]


class YolactMapOut(C.Structure):
    _fields_ = [("boxes", C.c_void_p), ("scores", C.c_void_p), ("class_ids", C.c_void_p), ("class_map", C.c_void_p),
                ("counts", C.c_void_p)]


class RoiAlignCfg(C.Structure):
    _fields_ = [("num_levels", C.c_int32), ("feat", C.c_void_p * 8), ("feat_h", C.c_int32 * 8), ("feat_w", C.c_int32 * 8),
                ("scale", C.c_float * 8), ("k_min", C.c_int32), ("k_max", C.c_int32), ("canonical_scale", C.c_int32),
                ("canonical_level", C.c_int32), ("channels", C.c_int32), ("out_h", C.c_int32), ("out_w", C.c_int32),
                ("sampling_ratio", C.c_int32)]


_lib = None
_lock = threading.Lock()


def load():
    """Load libisb.so once.  No fallback: a missing library is an error."""
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "(nvcc, sm_100a).  This package has no CPU or PyTorch fallback.")
        lib = C.CDLL(LIB_PATH)
        vp, i32, i64, f32 = C.c_void_p, C.c_int32, C.c_int64, C.c_float
        lib.isb_version.restype = C.c_int
        lib.isb_create.argtypes = [C.c_int, C.POINTER(Limits), C.POINTER(vp)]
        lib.isb_destroy.argtypes = [vp]
        lib.isb_destroy.restype = None
        lib.isb_last_error.argtypes = [vp]
        lib.isb_last_error.restype = C.c_char_p
        lib.isb_workspace_bytes.argtypes = [vp]
        lib.isb_workspace_bytes.restype = C.c_size_t
        lib.isb_launch_count.argtypes = [vp]
        lib.isb_launch_count.restype = i64
        lib.isb_profile_begin.argtypes = [vp, i32]
        lib.isb_profile_read.argtypes = [vp, C.POINTER(C.c_double * 4), C.POINTER(i32)]
        lib.isb_yolact_detect.argtypes = [vp, C.POINTER(YolactIn), C.POINTER(YolactOut), vp]
        lib.isb_yolact_detect_host.argtypes = [vp, C.POINTER(YolactIn), C.POINTER(YolactOut), vp]
        lib.isb_counts_export.argtypes = [vp, i32, i32, i32, C.c_char_p]
        lib.isb_counts_attach.argtypes = [vp, i32, C.c_char_p]
        lib.isb_counts_wait.argtypes = [vp, i64, vp, vp, vp, i32, vp]
        lib.isb_counts_epoch.argtypes = [vp]
        lib.isb_counts_epoch.restype = i64
        lib.isb_counts_detach.argtypes = [vp]
        lib.isb_set_host_threads.argtypes = [vp, i32]
        lib.isb_host_stats.argtypes = [vp, C.POINTER(i64 * 8)]
        lib.isb_yolact_class_map.argtypes = [vp, C.POINTER(YolactIn), C.POINTER(YolactMapOut), vp]
        lib.isb_yolact_class_map_host.argtypes = [vp, C.POINTER(YolactIn), C.POINTER(YolactMapOut), vp]
        lib.isb_rle_encode.argtypes = [vp, vp, i64, vp, i32, i32, i32, i32, i32, vp, i64, vp, vp, vp, vp]
        lib.isb_rle_to_string.argtypes = [vp, i64, vp, i64]
        lib.isb_rle_to_string.restype = i64
        lib.isb_decode_boxes.argtypes = [vp, vp, vp, i64, f32, f32, vp, vp]
        lib.isb_jaccard.argtypes = [vp, vp, vp, i32, i32, i32, i32, vp, vp]
        lib.isb_fast_nms.argtypes = [vp, vp, vp, vp, i32, i32, i32, f32, i32, i32, vp, vp, vp, vp, vp, vp]
        lib.isb_traditional_nms.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, C.c_double, i32, vp, vp, vp, vp, vp, vp]
        lib.isb_correct_boxes.argtypes = [vp, vp, i64, i32, i32, vp, vp]
        lib.isb_crop.argtypes = [vp, vp, vp, i32, i32, i32, vp, vp]
        lib.isb_mask_logits.argtypes = [vp, vp, vp, i32, i32, i32, vp, vp]
        lib.isb_mask_logits_batch.argtypes = [vp, vp, vp, i32, i32, i32, i32, i32, vp, vp]
        lib.isb_mask_assemble.argtypes = [vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, vp, vp]
        f64 = C.c_double
        lib.isb_box_decode.argtypes = [vp, vp, vp, i64, i32, f32, f32, f32, f32, f32, vp, vp]
        lib.isb_nms.argtypes = [vp, vp, vp, i32, f64, vp, vp, vp]
        lib.isb_batched_nms.argtypes = [vp, vp, vp, vp, i32, f64, vp, vp, vp]
        lib.isb_nms_static.argtypes = [vp, vp, i32, f64, vp, vp, vp, C.c_char_p, i32]  # no handle: fixed device scratch
        lib.isb_ms_roi_align.argtypes = [vp, C.POINTER(RoiAlignCfg), vp, vp, i32, vp, vp]
        lib.isb_mask_select_sigmoid.argtypes = [vp, vp, vp, i32, i32, i32, vp, vp]
        lib.isb_resize_boxes.argtypes = [vp, vp, i64, i32, i32, i32, i32, vp, vp]
        lib.isb_paste_masks.argtypes = [vp, vp, vp, i32, i32, i32, i32, i32, vp, vp]
        lib.isb_paste_class_map.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, f32, vp, vp, vp, vp]
        lib.isb_rpn_filter.argtypes = [vp, vp, vp, C.POINTER(i32), i32, i32, i32, i32, i32, f64, f32, f32, vp, vp, vp, vp]
        lib.isb_box_postprocess.argtypes = [vp, vp, vp, vp, vp, i32, i32, i32, i32, f32, f32, f32, f32, f32, f32, f64, i32,
                                            vp, vp, vp, vp, vp]
        lib.isb_box_postprocess_rows.argtypes = [vp, vp, vp, vp, vp, i32, vp, i32, i32, i32, f32, f32, f32, f32, f32, f32, f64, i32,
                                                 vp, vp, vp, vp, vp]
        lib.isb_rcnn_transform.argtypes = [vp, vp, i32, i32, i32, i32, C.POINTER(f32), C.POINTER(f32), vp, i32, i32, vp]
        lib.isb_preprocess.argtypes = [vp, vp, i32, i32, i32, i32, i32, C.POINTER(C.c_double), C.POINTER(C.c_double), vp, vp, vp]
        for name in SYMBOLS:
            fn = getattr(lib, name)
            if name not in ("isb_destroy", "isb_last_error", "isb_workspace_bytes", "isb_launch_count", "isb_version",
                            "isb_counts_epoch", "isb_rle_to_string"):
                fn.restype = C.c_int
        _lib = lib
        return lib


class Handle:
    """Owns one isb_handle (per device).  `limits` are the sizes the workspace supports."""

    def __init__(self, device, **limits):
        self.lib = load()
        self.device = int(device)
        self.limits = Limits(**limits)
        ptr = C.c_void_p()
        rc = self.lib.isb_create(self.device, C.byref(self.limits), C.byref(ptr))
        self._h = ptr
        if rc != ISB_OK:
            msg = self.lib.isb_last_error(ptr).decode() if ptr else "allocation failed"
            if ptr:
                self.lib.isb_destroy(ptr)
            self._h = None
            raise IsbError(rc, msg)

    @property
    def ptr(self):
        if self._h is None:
            raise RuntimeError("isb handle already destroyed")
        return self._h

    def check(self, rc):
        if rc != ISB_OK:
            raise IsbError(rc, self.lib.isb_last_error(self.ptr).decode())

    @property
    def launch_count(self):
        return int(self.lib.isb_launch_count(self.ptr))

    @property
    def workspace_bytes(self):
        return int(self.lib.isb_workspace_bytes(self.ptr))

    def profile_begin(self, max_calls):
        self.check(self.lib.isb_profile_begin(self.ptr, int(max_calls)))

    def profile_read(self):
        """-> ([ms_candidates, ms_nms, ms_sigmoid_maps, ms_compose] summed over the recorded calls, calls)"""
        ms = (C.c_double * 4)()
        n = C.c_int32()
        self.check(self.lib.isb_profile_read(self.ptr, C.byref(ms), C.byref(n)))
        return [ms[0], ms[1], ms[2], ms[3]], int(n.value)

    def set_host_threads(self, threads):
        """Worker threads of the *_host entry points (-1 default, 0 = copy engine only); see include/isb.h."""
        self.check(self.lib.isb_set_host_threads(self.ptr, int(threads)))

    def host_stats(self):
        st = (C.c_int64 * 8)()
        self.check(self.lib.isb_host_stats(self.ptr, C.byref(st)))
        return dict(images_copy_engine=int(st[0]), images_host_threads=int(st[1]), d2h_bytes=int(st[2]), h2d_bytes=int(st[3]),
                    host_threads=int(st[4]))

    def close(self):
        if getattr(self, "_h", None) is not None:
            self.lib.isb_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_DEFAULTS = dict(max_batch=1, max_anchors=19248, max_classes=80, max_det=100, max_top_k=200, cand_capacity=16384)
_handles = {}


def get_handle(device, **needs):
    """Cached per-device handle whose limits cover `needs`; re-created (bigger) when they do not."""
    device = int(device)
    cur = _handles.get(device)
    want = dict(_DEFAULTS)
    if cur is not None:
        want = {k: getattr(cur.limits, k) for k in _DEFAULTS}
    grow = cur is None
    for k, v in needs.items():
        if int(v) > want[k]:
            want[k] = int(v)
            grow = True
    if grow:
        if cur is not None:
            cur.close()
        cur = Handle(device, **want)
        _handles[device] = cur
    return cur


class _NoGuard:
    def __enter__(self):
        return None

    def __exit__(self, *exc):
        return False


_NO_GUARD = _NoGuard()


def on_device(dev):
    """`torch.cuda.device(dev)` only when `dev` is not already current: the library selects its handle's device itself,
    the guard exists to put the caller's current device back, and entering / leaving it costs microseconds that show
    on every small call."""
    import torch
    return _NO_GUARD if torch.cuda.current_device() == dev.index else torch.cuda.device(dev)


def total_launches():
    return sum(h.launch_count for h in _handles.values())
