#!/usr/bin/env python
"""bench.py -- YOLACT-R50 550^2 COCO post-processing throughput (BASELINE.json metric).

A step = one pass of the hot path (Detect + mask assembly, utils_bbox.py:170-243) over one
batch of 64 synthetic images per GPU (BASELINE.json configs[1]; 64*N images at N GPUs = configs[2]
at N=8).  Prints ONE JSON line on rank 0.

  python bench.py [--gpus N] [--steps K] [--warmup W]           # this repo's CUDA path
  python bench.py --impl reference [--steps K] [--warmup W]     # reference algorithm on host cores
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "post-proc images/sec (YOLACT-R50 550^2 COCO Detect + mask assembly)"
UNIT = "images/s"
BATCH_PER_GPU = 64
INPUT_SIZE, NUM_CLASSES, CONFIDENCE, NMS_IOU, TOP_K, MAX_DET = 550, 81, 0.05, 0.5, 200, 100
DISTINCT_IMAGES = 16  # distinct synthetic images per rank, tiled to the batch (all resident, 733 MB > L2)
SEED_BASE = 2000      # seed = 1000 * config + image index


def workload_config(n_gpus):
    return {"workload": "configs[1]: YOLACT-ResNet50 550x550 COCO 81-class Detect + mask assembly, synthetic heads",
            "batch_per_gpu": BATCH_PER_GPU, "global_batch": BATCH_PER_GPU * n_gpus, "anchors": 19248, "classes": NUM_CLASSES,
            "protos": "138x138x32", "out_hw": [INPUT_SIZE, INPUT_SIZE], "confidence": CONFIDENCE, "nms_iou": NMS_IOU,
            "top_k": TOP_K, "max_det": MAX_DET, "objects_per_image": 12, "parallelism": f"per-image shard x{n_gpus}",
            "l2": "inputs (733 MB) and outputs (2.1 GB) per step exceed the 126 MB L2; no flush needed"}


# ----------------------------------------------------------------------------- CPU arm (oracle port)
def _cpu_worker(rank, n_img, barrier, go, queue):
    os.environ["OMP_NUM_THREADS"] = "1"
    import numpy as np
    import torch
    torch.set_num_threads(1)
    try:
        from threadpoolctl import threadpool_limits
        threadpool_limits(1)
    except Exception:
        pass
    from oracle import synth, yolact_np as O
    a, p = synth.yolact_dims(INPUT_SIZE)
    anchors = O.get_anchors((INPUT_SIZE, INPUT_SIZE)).astype(np.float32)
    imgs = [[t.numpy() for t in synth.yolact_heads(SEED_BASE + rank * n_img + i, a, num_classes=NUM_CLASSES, proto=p)]
            for i in range(n_img)]
    while True:
        barrier.wait()          # round start (the parent decides whether there is one)
        if not go.value:
            break
        for im in imgs:
            r = O.decode_nms(im, anchors, CONFIDENCE, NMS_IOU, (INPUT_SIZE, INPUT_SIZE))
            assert r[0] is not None
        barrier.wait()          # round end
    queue.put(rank)


def cpu_port_rounds(max_rounds, budget_s=150.0, n_img_per_worker=1, max_workers=32):
    """Run the numpy oracle (the reference algorithm restated) on `workers` single-thread host
    processes, each decoding `n_img_per_worker` images per round, for up to `max_rounds` rounds or
    until `budget_s` of wall clock is used.  Returns (per-round wall seconds, images per round, workers)."""
    import multiprocessing as mp
    workers = max(1, min(os.cpu_count() or 1, max_workers))
    ctx = mp.get_context("fork")
    barrier, go, queue = ctx.Barrier(workers + 1), ctx.Value("i", 1), ctx.Queue()
    procs = [ctx.Process(target=_cpu_worker, args=(r, n_img_per_worker, barrier, go, queue)) for r in range(workers)]
    for p in procs:
        p.start()
    secs, t_begin = [], None
    while True:
        more = len(secs) < max_rounds
        if more and len(secs) >= 2 and (time.time() - t_begin) + 1.2 * max(secs) > budget_s:
            more = False
        go.value = 1 if more else 0
        barrier.wait()
        if t_begin is None:
            t_begin = time.time()   # input generation (before the first barrier) is not counted
        if not more:
            break
        t0 = time.time()
        barrier.wait()
        secs.append(time.time() - t0)
    for _ in procs:
        queue.get()
    for p in procs:
        p.join()
    return secs, workers * n_img_per_worker, workers


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    secs, n_img, workers = cpu_port_rounds(args.warmup + args.steps)
    warm = min(args.warmup, max(1, len(secs) // 5)) if len(secs) > 1 else 0
    timed = secs[warm:]
    total = sum(timed)
    value = n_img * len(timed) / total
    sample = (f"{n_img} images per step ({workers} single-thread processes x 1 image); numpy restatement of the reference's "
              "torch code (the Python reference cannot travel to the GPU box)")
    if len(secs) < args.warmup + args.steps:
        sample += f"; bounded to {len(timed)} timed + {warm} warm-up steps by the 150 s budget"
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / len(timed), "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(args.gpus),
            "steps_timed": len(timed),
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": workers, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------- clocks
class ClockSampler(threading.Thread):
    """Polls SM clock and throttle reasons through NVML while the timed region runs."""

    def __init__(self, uuid):
        super().__init__(daemon=True)
        self.samples, self.reasons, self.max_mhz, self.stop_flag, self.err = [], set(), None, False, None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            try:
                self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid if isinstance(uuid, bytes) else uuid.encode())
            except Exception:
                self.h = pynvml.nvmlDeviceGetHandleByUUID(uuid)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception as e:  # pragma: no cover
            self.nv, self.err = None, repr(e)

    def run(self):
        if self.nv is None:
            return
        nv = self.nv
        names = {"hw_slowdown": getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8),
                 "hw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40),
                 "sw_thermal_slowdown": getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20),
                 "sw_power_cap": getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4)}
        while not self.stop_flag:
            try:
                self.samples.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                try:
                    mask = nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                except Exception:
                    mask = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for k, bit in names.items():
                    if mask & bit:
                        self.reasons.add(k)
            except Exception as e:  # pragma: no cover
                self.err = repr(e)
                return
            time.sleep(0.002)

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s), **({"error": self.err} if self.err else {})}


# ----------------------------------------------------------------------------- GPU arm
def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        with open(path) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def ncu_traffic():
    """dram bytes per launch of the dominant kernel from the committed ncu capture, if any."""
    try:
        with open(os.path.join(ROOT, "profiles", "mask_kernel_traffic.json")) as f:
            return json.load(f)
    except Exception:
        return None


def run_gpu(args):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        # before CUDA is initialised in this process (workers are forked)
        secs, n_img, workers = cpu_port_rounds(24, budget_s=20.0)
        cpu = {"value": n_img * (len(secs) - 1) / sum(secs[1:]), "unit": UNIT, "cores": workers, "kind": "port",
               "sample": f"{len(secs) - 1} timed rounds (+1 warm-up) of {n_img} images of the same workload ({workers} single-thread "
                         f"processes x 1 image, {sum(secs[1:]):.1f} s): numpy restatement of the reference's torch code"}

    import numpy as np
    import torch
    import torch.distributed as dist

    from isb import _lib, sharding
    from isb.yolact import BBoxUtility, YolactBatchResult
    from oracle import synth, yolact_np as O  # input generator only; nothing of the oracle is timed here

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    json_fd = None
    if world > 1:
        # NCCL writes its version banner to the process's stdout (fd 1) when NCCL_DEBUG is set on the box; stdout must
        # carry exactly ONE JSON line, so everything else that lands on fd 1 goes to stderr and the line is written to
        # a saved copy of the original descriptor
        sys.stdout.flush()
        json_fd = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", device_id=dev)

    a, p = synth.yolact_dims(INPUT_SIZE)
    anchors = torch.from_numpy(O.get_anchors((INPUT_SIZE, INPUT_SIZE))).float()
    bsz = BATCH_PER_GPU
    # rank r owns images [r*64, (r+1)*64) of the global batch (contiguous block partition)
    start, stop = sharding.partition(bsz * world, world, rank)
    assert stop - start == bsz
    distinct = synth.yolact_batch(SEED_BASE + (start % 1000), DISTINCT_IMAGES, a, num_classes=NUM_CLASSES, proto=p)
    reps = bsz // DISTINCT_IMAGES
    host_in = [t.repeat((reps,) + (1,) * (t.dim() - 1)).contiguous() for t in distinct]
    dev_in = [t.to(dev) for t in host_in]
    anchors_d = anchors.to(dev)
    bu = BBoxUtility()
    shape = (INPUT_SIZE, INPUT_SIZE)
    slot = (shape[0] * shape[1] * MAX_DET + 15) // 16 * 16
    out = YolactBatchResult(torch.empty((bsz, MAX_DET, 4), device=dev), torch.empty((bsz, MAX_DET), device=dev),
                            torch.empty((bsz, MAX_DET), dtype=torch.int64, device=dev),
                            torch.empty((bsz,) + shape, dtype=torch.int64, device=dev),
                            torch.empty((bsz, slot), dtype=torch.uint8, device=dev),
                            torch.empty((bsz,), dtype=torch.int32, device=dev), shape)

    xchg = sharding.CountExchange(bsz, dev) if world > 1 else None

    def step():
        bu.decode_nms_batch(dev_in, anchors_d, CONFIDENCE, NMS_IOU, shape, TOP_K, MAX_DET, out=out)
        if xchg is not None:
            xchg.post(out.counts)  # the path's one collective (int32 counts), overlapped with the next step

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    handle = _lib.get_handle(local_rank)
    handle.profile_begin(args.steps)
    sampler = ClockSampler("GPU-" + str(torch.cuda.get_device_properties(dev).uuid))
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    launches0 = handle.launch_count
    sampler.start()
    ev0.record()
    for _ in range(args.steps):
        step()
    if xchg is not None:
        xchg.flush()  # every posted gather has to finish inside the timed region
    ev1.record()
    torch.cuda.synchronize()
    sampler.stop_flag = True
    if world > 1:
        dist.barrier()
    launches = handle.launch_count - launches0
    ms = ev0.elapsed_time(ev1)
    stage_ms, calls = handle.profile_read()
    handle.profile_begin(0)
    counts = out.counts.cpu().numpy()
    assert (counts > 0).all(), "benchmark images must all produce detections"
    if xchg is not None:
        all_counts, all_offsets = xchg.result()
        assert all_counts.numel() == bsz * world and bool((all_counts[rank * bsz:(rank + 1) * bsz].cpu() == out.counts.cpu()).all())
        assert int(all_offsets[-1]) + int(all_counts[-1]) == int(all_counts.sum())

    # ---- end to end through the host-buffer C ABI entry (pinned host memory) -----------------
    import ctypes as C
    pin_in = [t.pin_memory() for t in host_in]
    pin_anchors = anchors.pin_memory()
    h_boxes = torch.empty((bsz, MAX_DET, 4)).pin_memory()
    h_scores = torch.empty((bsz, MAX_DET)).pin_memory()
    h_cls = torch.empty((bsz, MAX_DET), dtype=torch.int64).pin_memory()
    h_arg = torch.empty((bsz,) + shape, dtype=torch.int64).pin_memory()
    h_masks = torch.empty((bsz, slot), dtype=torch.uint8).pin_memory()
    h_counts = torch.empty((bsz,), dtype=torch.int32).pin_memory()
    pin = _lib.YolactIn(bsz, a, NUM_CLASSES, 32, p, p, shape[0], shape[1], pin_in[0].data_ptr(), pin_in[1].data_ptr(),
                        pin_in[2].data_ptr(), pin_in[3].data_ptr(), pin_anchors.data_ptr(), CONFIDENCE, NMS_IOU, TOP_K, MAX_DET)
    pout = _lib.YolactOut(h_boxes.data_ptr(), h_scores.data_ptr(), h_cls.data_ptr(), h_arg.data_ptr(), h_masks.data_ptr(),
                          slot, h_counts.data_ptr())
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    e2e_steps = max(3, min(args.steps, 10))
    for _ in range(2):
        handle.check(handle.lib.isb_yolact_detect_host(handle.ptr, C.byref(pin), C.byref(pout), stream))
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        handle.check(handle.lib.isb_yolact_detect_host(handle.ptr, C.byref(pin), C.byref(pout), stream))
    e2e_s = (time.perf_counter() - t0) / e2e_steps
    assert np.array_equal(h_counts.numpy(), counts)
    h2d = sum(t.numel() * t.element_size() for t in pin_in) + anchors.numel() * 4
    d2h = int(h_boxes.numel() * 4 + h_scores.numel() * 4 + h_cls.numel() * 8 + h_arg.numel() * 8 + h_counts.numel() * 4
              + shape[0] * shape[1] * int(counts.sum()))

    # ---- class-map variant (decode_nms + yolact.py:162-170 fused; uint8 [H,W] out) -- extra keys only,
    # the headline `value` / `e2e` above stay the full decode_nms outputs
    cm_out = bu.decode_nms_class_map_batch(dev_in, anchors_d, CONFIDENCE, NMS_IOU, shape, TOP_K, MAX_DET)
    for _ in range(3):
        bu.decode_nms_class_map_batch(dev_in, anchors_d, CONFIDENCE, NMS_IOU, shape, TOP_K, MAX_DET, out=cm_out)
    cm_steps = max(3, min(args.steps, 50))
    torch.cuda.synchronize()
    ev0.record()
    for _ in range(cm_steps):
        bu.decode_nms_class_map_batch(dev_in, anchors_d, CONFIDENCE, NMS_IOU, shape, TOP_K, MAX_DET, out=cm_out)
    ev1.record()
    torch.cuda.synchronize()
    cm_ms = ev0.elapsed_time(ev1) / cm_steps
    h_map = torch.empty((bsz,) + shape, dtype=torch.uint8).pin_memory()
    pmap = _lib.YolactMapOut(h_boxes.data_ptr(), h_scores.data_ptr(), h_cls.data_ptr(), h_map.data_ptr(), h_counts.data_ptr())
    for _ in range(2):
        handle.check(handle.lib.isb_yolact_class_map_host(handle.ptr, C.byref(pin), C.byref(pmap), stream))
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        handle.check(handle.lib.isb_yolact_class_map_host(handle.ptr, C.byref(pin), C.byref(pmap), stream))
    cm_e2e_s = (time.perf_counter() - t0) / e2e_steps
    assert np.array_equal(h_counts.numpy(), counts) and np.array_equal(h_map.numpy(), cm_out[3].cpu().numpy())
    cm_d2h = int(h_boxes.numel() * 4 + h_scores.numel() * 4 + h_cls.numel() * 8 + h_map.numel() + h_counts.numel() * 4)

    # ---- COCO RLE of the batch's masks on the device (utils_map.py:33-34 without the 2 GB D2H) -- extra keys only
    bu.decode_nms_batch(dev_in, anchors_d, CONFIDENCE, NMS_IOU, shape, TOP_K, MAX_DET, out=out)
    rle_cap = 8 * shape[1] * MAX_DET * bsz
    rle_out = torch.empty((rle_cap,), dtype=torch.int32, device=dev)
    rle_cnt = torch.empty((bsz, MAX_DET), dtype=torch.int32, device=dev)
    rle_off = torch.empty((bsz, MAX_DET), dtype=torch.int64, device=dev)
    rle_tot = torch.empty((2,), dtype=torch.int64, device=dev)

    def rle():
        handle.check(handle.lib.isb_rle_encode(handle.ptr, out.masks_raw.data_ptr(), out.masks_raw.shape[1], out.counts.data_ptr(), -1,
                                               MAX_DET, bsz, shape[0], shape[1], rle_out.data_ptr(), rle_cap, rle_cnt.data_ptr(),
                                               rle_off.data_ptr(), rle_tot.data_ptr(), stream))
    for _ in range(3):
        rle()
    torch.cuda.synchronize()
    ev0.record()
    for _ in range(10):
        rle()
    ev1.record()
    torch.cuda.synchronize()
    rle_ms = ev0.elapsed_time(ev1) / 10
    rle_runs = int(rle_tot[1].item())
    assert rle_runs <= rle_cap
    del rle_out

    # ---- raw-logits variant (F.softmax of nets/yolact.py:204 fused into the score filter) -- extra keys only
    lg = synth.yolact_batch(SEED_BASE + (start % 1000), DISTINCT_IMAGES, a, num_classes=NUM_CLASSES, proto=p, return_logits=True)[1]
    logits_d = lg.repeat((reps, 1, 1)).contiguous().to(dev)
    lg_in = [dev_in[0], logits_d, dev_in[2], dev_in[3]]
    for _ in range(3):
        bu.decode_nms_batch(lg_in, anchors_d, CONFIDENCE, NMS_IOU, shape, TOP_K, MAX_DET, out=out, logits=True)
    torch.cuda.synchronize()
    assert np.array_equal(out.counts.cpu().numpy(), counts)
    handle.profile_begin(cm_steps)
    ev0.record()
    for _ in range(cm_steps):
        bu.decode_nms_batch(lg_in, anchors_d, CONFIDENCE, NMS_IOU, shape, TOP_K, MAX_DET, out=out, logits=True)
    ev1.record()
    torch.cuda.synchronize()
    lg_ms = ev0.elapsed_time(ev1) / cm_steps
    lg_stage, lg_calls = handle.profile_read()
    handle.profile_begin(0)
    for _ in range(3):
        torch.softmax(logits_d, dim=-1)
    torch.cuda.synchronize()
    ev0.record()
    for _ in range(cm_steps):
        torch.softmax(logits_d, dim=-1)
    ev1.record()
    torch.cuda.synchronize()
    softmax_ms = ev0.elapsed_time(ev1) / cm_steps
    del logits_d, lg_in

    # ---- pre-processing variant (PIL bicubic resize + normalisation of yolact.py:137-148) -- extra keys only
    from isb import preprocess as pre
    pics = torch.randint(0, 256, (bsz, 1080, 1920, 3), dtype=torch.uint8, device=dev)
    for _ in range(3):
        pre.image_to_input(pics, shape)
    torch.cuda.synchronize()
    ev0.record()
    for _ in range(cm_steps):
        pre.image_to_input(pics, shape)
    ev1.record()
    torch.cuda.synchronize()
    pre_ms = ev0.elapsed_time(ev1) / cm_steps
    pre_bytes = bsz * (1080 * 1920 * 3 + shape[0] * shape[1] * 12)
    del pics

    if world > 1:
        t = torch.tensor([ms, e2e_s, cm_ms, cm_e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_s, cm_ms, cm_e2e_s = float(t[0]), float(t[1]), float(t[2]), float(t[3])
    if rank == 0:
        n_img = bsz * world * args.steps
        m_sum = float(counts.sum())
        hw = shape[0] * shape[1]
        # algorithmic bytes of ONE mask_compose_kernel launch: per-detection records in, int64 arg-max map
        # and bool [H,W,m] masks out (the sigmoid maps it samples are an L2-resident intermediate, not counted)
        mask_bytes = bsz * hw * 8 + m_sum * (hw + 32)
        mask_ms = stage_ms[3] / max(calls, 1)
        peak, peak_src = measured_peak()
        achieved = mask_bytes / (mask_ms * 1e-3) / 1e9
        path_bytes = bsz * 13_865_696 + 302_528 * m_sum  # SURVEY 8(d) bytes(m), whole path, per GPU
        tr = ncu_traffic()
        line = {"metric": METRIC, "value": n_img / (ms * 1e-3), "unit": UNIT, "n_gpus": world, "steps": args.steps,
                "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(world),
                "clocks": sampler.summary(),
                "e2e": {"value": bsz * world / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "steps": e2e_steps, "api": "isb_yolact_detect_host (pinned host buffers)"},
                "gpu_launches": launches,
                "roofline": {"bound": "hbm", "kernel": "mask_compose_kernel", "achieved": achieved, "peak": peak, "unit": "GB/s",
                             "frac": achieved / peak, "traffic": tr["dram_bytes_per_launch"] if tr else None,
                             "peak_source": peak_src, "algorithmic_bytes_per_launch": mask_bytes, "kernel_ms": mask_ms,
                             "stage_ms": {"candidates": stage_ms[0] / max(calls, 1), "sort_nms": stage_ms[1] / max(calls, 1),
                                          "sigmoid_maps": stage_ms[2] / max(calls, 1), "mask_compose": mask_ms},
                             "path_achieved_gbs": path_bytes / (ms / args.steps * 1e-3) / 1e9,
                             "path_frac": path_bytes / (ms / args.steps * 1e-3) / 1e9 / peak,
                             "mean_detections": m_sum / bsz},
                "class_map_variant": {"what": "decode_nms + yolact.py:162-170 fused, uint8 [H,W] class image instead of "
                                              "bool [H,W,m] + int64 [H,W] (SURVEY 8f rank 1); not the headline",
                                      "value": bsz * world / (cm_ms * 1e-3), "ms_per_step": cm_ms,
                                      "e2e": bsz * world / cm_e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d,
                                      "d2h_bytes_per_step": cm_d2h, "api": "isb_yolact_class_map[_host]"},
                "rle_variant": {"what": "COCO RLE (utils_map.py:33-34) of all instance masks of the step on the device "
                                        "(SURVEY 8f rank 3); not the headline", "ms_per_step": rle_ms, "runs": rle_runs,
                                "d2h_bytes_instead_of_masks": rle_runs * 4 + bsz * MAX_DET * 12,
                                "mask_bytes": int(shape[0] * shape[1] * counts.sum())},
                "from_logits_variant": {"what": "raw class logits in, softmax (nets/yolact.py:204) fused into the score filter "
                                                "(SURVEY 8f rank 2); not the headline",
                                        "value": bsz / (lg_ms * 1e-3), "ms_per_step": lg_ms, "unit": UNIT + " (this rank)",
                                        "candidates_ms": lg_stage[0] / max(lg_calls, 1),
                                        "torch_softmax_ms_it_replaces": softmax_ms},
                "preprocess_variant": {"what": "wrapper pre-processing on the device (SURVEY 8f rank 4): 1080x1920 uint8 pictures "
                                               "-> Pillow-exact bicubic resize -> normalised float32 CHW network input; "
                                               "not the headline", "value": bsz / (pre_ms * 1e-3), "unit": "pictures/s (this rank)",
                                       "ms_per_step": pre_ms, "algorithmic_GBs": pre_bytes / (pre_ms * 1e-3) / 1e9},
                "cpu_baseline": cpu}
        if json_fd is None:
            print(json.dumps(line), flush=True)
        else:
            os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="isb", choices=["isb", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
