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


def synthetic_heads(seed, num_anchors, num_classes=NUM_CLASSES, num_coef=32, proto=138, objects=12, run=40, return_logits=False):
    """One image of synthetic head outputs, CPU torch.Generator so every box draws the same bits:
    box ~ N(0, 0.3^2) [A,4]; logits ~ N(0,1) [A,C] with +7 on the background column and `objects` runs of `run`
    neighbouring anchors boosted by U(5,13) on one class; conf = softmax(logits); coef = tanh(N(0,1)); proto = relu(N(0,1)).
    Both arms of this benchmark draw their inputs here; the parity tests have their own generator under oracle/."""
    import torch
    g = torch.Generator().manual_seed(int(seed))
    a, c = num_anchors, num_classes
    box = torch.randn(a, 4, generator=g) * 0.3
    logits = torch.randn(a, c, generator=g)
    logits[:, 0] += 7.0
    for _ in range(objects):
        cls = int(torch.randint(1, c, (1,), generator=g))
        r = min(run, a)
        start = int(torch.randint(0, a - r + 1, (1,), generator=g))
        logits[start:start + r, cls] += torch.rand(r, generator=g) * 8.0 + 5.0
    conf = torch.softmax(logits, dim=1)
    coef = torch.tanh(torch.randn(a, num_coef, generator=g))
    prot = torch.relu(torch.randn(proto, proto, num_coef, generator=g))
    return box.contiguous(), (logits if return_logits else conf).contiguous(), coef.contiguous(), prot.contiguous()


def synthetic_batch(seed0, batch, num_anchors, **kw):
    import torch
    per = [synthetic_heads(seed0 + i, num_anchors, **kw) for i in range(batch)]
    return tuple(torch.stack([p[k] for p in per]).contiguous() for k in range(4))


# ----------------------------------------------------------------------------- configs[3] / configs[4] inputs
FPN_STRIDES = (4, 8, 16, 32, 64)


def rpn_inputs(seed, image_hw=(800, 1344), objects=25):
    """Synthetic RPN outputs for one 800 x 1333 image padded to 800 x 1344: anchors [A,4] (five levels,
    three ratios: 201600 / 50400 / 12600 / 3150 / 819 = 268569), deltas [A,4], objectness logits [A] = noise around -4 with
    `objects` boosted clusters, so the per-level top-1000 overlap and NMS has real work."""
    import torch
    g = torch.Generator().manual_seed(int(seed))
    per_level, out = [], []
    for stride, size in zip(FPN_STRIDES, (32, 64, 128, 256, 512)):
        fh, fw = -(-image_hw[0] // stride), -(-image_hw[1] // stride)
        hr = torch.sqrt(torch.tensor((0.5, 1.0, 2.0)))
        ws, hs = size / hr, size * hr
        base = torch.stack([-ws, -hs, ws, hs], 1) / 2
        ys, xs = torch.meshgrid(torch.arange(fh) * stride, torch.arange(fw) * stride, indexing="ij")
        a = (torch.stack([xs, ys, xs, ys], -1).reshape(-1, 1, 4).float() + base[None]).reshape(-1, 4)
        out.append(a)
        per_level.append(a.shape[0])
    anchors = torch.cat(out).contiguous()
    n = anchors.shape[0]
    deltas = torch.randn(n, 4, generator=g) * torch.tensor([0.1, 0.1, 0.2, 0.2])
    logits = torch.randn(n, generator=g) * 1.5 - 4.0
    ctr = (anchors[:, :2] + anchors[:, 2:]) / 2
    for _ in range(objects):
        c = torch.rand(2, generator=g) * torch.tensor([float(image_hw[1]), float(image_hw[0])])
        rad = float(torch.rand(1, generator=g)) * 120 + 30
        near = (ctr - c).abs().max(dim=1).values < rad
        logits[near] += torch.rand(int(near.sum()), generator=g) * 6.0 + 2.0
    return anchors, deltas.contiguous(), logits.contiguous(), per_level


def workload_config(n_gpus):
    return {"workload": "configs[1]: YOLACT-ResNet50 550x550 COCO 81-class Detect + mask assembly, synthetic heads",
            "batch_per_gpu": BATCH_PER_GPU, "global_batch": BATCH_PER_GPU * n_gpus, "anchors": 19248, "classes": NUM_CLASSES,
            "protos": "138x138x32", "out_hw": [INPUT_SIZE, INPUT_SIZE], "confidence": CONFIDENCE, "nms_iou": NMS_IOU,
            "top_k": TOP_K, "max_det": MAX_DET, "objects_per_image": 12, "parallelism": f"per-image shard x{n_gpus}",
            "count_exchange": "none (1 GPU)" if n_gpus == 1 else "see run",
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
    from oracle import yolact_np as O   # the CPU arm IS the oracle port: the one place bench.py may execute oracle/
    anchors = O.get_anchors((INPUT_SIZE, INPUT_SIZE)).astype(np.float32)
    a, p = anchors.shape[0], 2 * int(O.get_img_output_length(INPUT_SIZE, INPUT_SIZE)[0][0])
    imgs = [[t.numpy() for t in synthetic_heads(SEED_BASE + rank * n_img + i, a, num_classes=NUM_CLASSES, proto=p)]
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
    from isb import anchors as isb_anchors
    from isb.yolact import BBoxUtility, YolactBatchResult

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

    anchors = torch.from_numpy(isb_anchors.get_anchors((INPUT_SIZE, INPUT_SIZE))).float()   # yolact.py:81
    a, p = anchors.shape[0], isb_anchors.proto_size((INPUT_SIZE, INPUT_SIZE))[0]
    bsz = BATCH_PER_GPU
    # rank r owns images [r*64, (r+1)*64) of the global batch (contiguous block partition)
    start, stop = sharding.partition(bsz * world, world, rank)
    assert stop - start == bsz
    distinct = synthetic_batch(SEED_BASE + (start % 1000), DISTINCT_IMAGES, a, num_classes=NUM_CLASSES, proto=p)
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

    xchg = None
    if world > 1:
        # the path's one exchange (per-image detection counts): fused into the NMS kernel as NVLink peer stores when the
        # ranks can map each other's memory (isb_counts_*), else one NCCL all-gather per step posted off the critical path
        _lib.get_handle(local_rank, max_batch=bsz, max_anchors=a, max_classes=NUM_CLASSES - 1, max_det=MAX_DET, max_top_k=TOP_K)
        xchg = sharding.open_count_exchange(bsz, dev, prefer_fused=not args.nccl_counts)

    def step():
        bu.decode_nms_batch(dev_in, anchors_d, CONFIDENCE, NMS_IOU, shape, TOP_K, MAX_DET, out=out)
        if xchg is not None:
            xchg.post(out.counts)  # the path's one collective (int32 counts), overlapped with the next step

    for _ in range(max(args.warmup, 3)):
        step()
    torch.cuda.synchronize()
    handle = _lib.get_handle(local_rank)
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
    # per-stage device times (the roofline kernel's duration): a second, untimed pass with the library's event ring armed --
    # profiled calls run the four kernels back to back on one stream (the timed pass above overlaps the two halves' K1 / K2 / K3a)
    prof_steps = max(5, min(args.steps, 50))
    handle.profile_begin(prof_steps)
    for _ in range(prof_steps):
        bu.decode_nms_batch(dev_in, anchors_d, CONFIDENCE, NMS_IOU, shape, TOP_K, MAX_DET, out=out)
    torch.cuda.synchronize()
    stage_ms, calls = handle.profile_read()
    handle.profile_begin(0)
    counts = out.counts.cpu().numpy()
    assert (counts > 0).all(), "benchmark images must all produce detections"
    if xchg is not None:
        all_counts, all_offsets = xchg.result()
        assert all_counts.numel() == bsz * world and bool((all_counts[rank * bsz:(rank + 1) * bsz].cpu() == out.counts.cpu()).all())
        assert int(all_offsets[-1]) + int(all_counts[-1]) == int(all_counts.sum())
        if getattr(xchg, "fused", False):
            assert int(xchg.status.item()) == 0, "fused count exchange: a peer's counts did not arrive"
            ref_counts, ref_offsets = sharding.gather_counts(out.counts, pad_to=bsz)   # outside the timed region: NCCL as the checker
            assert torch.equal(ref_counts.cpu(), all_counts.cpu()) and torch.equal(ref_offsets.cpu(), all_offsets.cpu()), \
                "fused count exchange disagrees with the NCCL all-gather"

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
    # bytes that actually crossed PCIe in the last call, counted by the library copy by copy (isb_host_stats)
    delivery = handle.host_stats()
    h2d, d2h = delivery["h2d_bytes"], delivery["d2h_bytes"]
    assert h2d == sum(t.numel() * t.element_size() for t in pin_in) + anchors.numel() * 4
    # the host buffers must hold exactly what the device path produced (spot check: three images, every byte)
    for b in (0, bsz // 2, bsz - 1):
        m = int(counts[b])
        assert torch.equal(h_masks[b, : shape[0] * shape[1] * m], out.masks_raw[b, : shape[0] * shape[1] * m].cpu()), f"host masks of image {b}"
        assert torch.equal(h_arg[b], out.masks_arg[b].cpu()) and torch.equal(h_scores[b, :m], out.scores[b, :m].cpu())
    # same call with the copy engine alone (no host worker threads): what round 1 measured
    handle.set_host_threads(0)
    for _ in range(2):
        handle.check(handle.lib.isb_yolact_detect_host(handle.ptr, C.byref(pin), C.byref(pout), stream))
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        handle.check(handle.lib.isb_yolact_detect_host(handle.ptr, C.byref(pin), C.byref(pout), stream))
    e2e_dma_s = (time.perf_counter() - t0) / e2e_steps
    delivery_dma = handle.host_stats()
    handle.set_host_threads(-1)

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
    lg = synthetic_batch(SEED_BASE + (start % 1000), DISTINCT_IMAGES, a, num_classes=NUM_CLASSES, proto=p, return_logits=True)[1]
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

    # ---- sparse distribution (3 objects per image, the wrapper's own thresholds 0.5 / 0.3 of yolact.py:42,46 ->
    # a handful of detections per image; output bytes scale with m) -- extra keys only
    sp = synthetic_batch(SEED_BASE + 500 + (start % 1000), DISTINCT_IMAGES, a, num_classes=NUM_CLASSES, proto=p, objects=3)
    sp_host = [t.repeat((reps,) + (1,) * (t.dim() - 1)).contiguous() for t in sp]
    sp_in = [t.to(dev) for t in sp_host]
    for _ in range(3):
        bu.decode_nms_batch(sp_in, anchors_d, 0.5, 0.3, shape, TOP_K, MAX_DET, out=out)
    torch.cuda.synchronize()
    ev0.record()
    for _ in range(cm_steps):
        bu.decode_nms_batch(sp_in, anchors_d, 0.5, 0.3, shape, TOP_K, MAX_DET, out=out)
    ev1.record()
    torch.cuda.synchronize()
    sp_ms = ev0.elapsed_time(ev1) / cm_steps
    sp_counts = out.counts.cpu().numpy()
    sp_pin = [t.pin_memory() for t in sp_host]
    spin = _lib.YolactIn(bsz, a, NUM_CLASSES, 32, p, p, shape[0], shape[1], sp_pin[0].data_ptr(), sp_pin[1].data_ptr(),
                         sp_pin[2].data_ptr(), sp_pin[3].data_ptr(), pin_anchors.data_ptr(), 0.5, 0.3, TOP_K, MAX_DET)
    for _ in range(2):
        handle.check(handle.lib.isb_yolact_detect_host(handle.ptr, C.byref(spin), C.byref(pout), stream))
    t0 = time.perf_counter()
    for _ in range(e2e_steps):
        handle.check(handle.lib.isb_yolact_detect_host(handle.ptr, C.byref(spin), C.byref(pout), stream))
    sp_e2e_s = (time.perf_counter() - t0) / e2e_steps
    sp_stats = handle.host_stats()
    assert np.array_equal(h_counts.numpy(), sp_counts)
    sp_bytes = bsz * 13_865_696 + 302_528 * float(sp_counts.sum())
    del sp_in, sp_pin

    # ---- BASELINE configs[3] (Mask R-CNN 800x1333) and configs[4] (VerSeg stress): one GPU, extra keys only.  Last: the VerSeg
    # call grows the shared per-device handle (max_det 1024), which re-creates it.
    mr_variant = vs_variant = None
    if world == 1 and not args.no_variants:
        peak_gbs = measured_peak()[0]
        del dev_in, pin_in, h_masks, out
        torch.cuda.empty_cache()
        mr_variant = maskrcnn_variant(dev, peak_gbs, cpu=not args.no_cpu_baseline)
        vs_variant = verseg_variant(dev, peak_gbs, cpu=not args.no_cpu_baseline)

    if world > 1:
        t = torch.tensor([ms, e2e_s, cm_ms, cm_e2e_s, e2e_dma_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms, e2e_s, cm_ms, cm_e2e_s, e2e_dma_s = float(t[0]), float(t[1]), float(t[2]), float(t[3]), float(t[4])
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
                "vs_baseline": None, "dtype": "f32", "data": "synthetic",
                "config": {**workload_config(world),
                           "count_exchange": "none (1 GPU)" if xchg is None else
                           ("fused into the step's own kernels: NVLink peer stores of every image's count into every rank's ring (isb_counts_*), no collective launch"
                            if getattr(xchg, "fused", False) else "NCCL all_gather_into_tensor of int32 counts, posted asynchronously")},
                "clocks": sampler.summary(),
                "e2e": {"value": bsz * world / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "steps": e2e_steps, "api": "isb_yolact_detect_host (pinned host buffers in, pinned host bool [H,W,m] out)",
                        "delivery": {"what": "masks reach the host two ways at once: dense over the copy engine, or bit-packed "
                                             "(1/8 of the bytes) + expansion into the caller's buffer by host worker threads; "
                                             "split decided at run time (rank 0's last call shown)", **delivery},
                        "copy_engine_only": {"value": bsz * world / e2e_dma_s, "d2h_bytes_per_step": delivery_dma["d2h_bytes"]}},
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
                "sparse_variant": {"what": "same path, sparse images (3 objects, wrapper thresholds 0.5 / 0.3): output bytes scale with "
                                           "the detections (SURVEY 8d); this rank; not the headline",
                                   "mean_detections": float(sp_counts.mean()), "value": bsz / (sp_ms * 1e-3), "ms_per_step": sp_ms,
                                   "unit": UNIT + " (this rank)", "path_frac": sp_bytes / (sp_ms * 1e-3) / 1e9 / peak,
                                   "e2e": bsz / sp_e2e_s, "d2h_bytes_per_step": sp_stats["d2h_bytes"]},
                "maskrcnn_variant": mr_variant, "verseg_variant": vs_variant,
                "cpu_baseline": cpu}
        if json_fd is None:
            print(json.dumps(line), flush=True)
        else:
            os.write(json_fd, (json.dumps(line) + "\n").encode())
    if world > 1:
        dist.destroy_process_group()


# ----------------------------------------------------------------------------- configs[3]: Mask R-CNN 800 x 1333
class _StandInBoxHead:
    """Deterministic element-wise stand-ins for the network layers between the post-processing stages (box head, predictors,
    mask head): the path under test is what sits around them.  Not timed."""

    def __init__(self, num_classes):
        self.c = num_classes

    def box(self, f7):
        import torch
        x = f7.double().mean(dim=(2, 3)).float()
        c = self.c
        logits = torch.cat([x[:, :1] * 14.0 + 2.0, x[:, 1:c] * 14.0], 1)
        reg = x[:, c:5 * c] * torch.tensor([6.0, 6.0, 3.0, 3.0], device=x.device).repeat(c)
        return logits.contiguous(), reg.contiguous()

    def mask(self, f14):
        return (f14[:, :self.c].repeat_interleave(2, dim=2).repeat_interleave(2, dim=3) * 3.0).contiguous()


def _time_us(fn, warm=5, iters=30):
    import torch
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


def maskrcnn_variant(dev, peak, num_classes=21, cpu=True):
    """BASELINE.json configs[3]: one 800 x 1333 image (FPN padded to 800 x 1344, 268 569 anchors, 1000 proposals, 100
    detections) through the chained post-processing stages, each timed on the device (CUDA events, mean of 30 calls through
    the reference-signature Python entry points, i.e. including the one count read a variable-length return needs), with
    SURVEY 8(d)'s algorithmic bytes and the fraction of the measured HBM peak; beside it the reference's CPU path on this
    box's host cores (torchvision CPU ops for NMS / RoIAlign, the numpy oracle port for the rest)."""
    import types
    from collections import OrderedDict
    import torch
    from isb import maskrcnn as R
    shape, orig, c = (800, 1333), (800, 1333), num_classes
    g = torch.Generator().manual_seed(4601)
    feats_cpu = [torch.randn(1, 256, -(-800 // s_), -(-1344 // s_), generator=g) for s_ in FPN_STRIDES[:4]]
    anchors_c, deltas_c, logits_c, per_level = rpn_inputs(4602)
    feats = OrderedDict((str(i), f.to(dev)) for i, f in enumerate(feats_cpu))
    anchors, deltas, logits = anchors_c.to(dev), deltas_c.to(dev), logits_c.to(dev)
    coder = types.SimpleNamespace(weights=(1.0, 1.0, 1.0, 1.0), bbox_xform_clip=R.BBOX_XFORM_CLIP)
    rpn = types.SimpleNamespace(pre_nms_top_n=lambda: 1000, post_nms_top_n=lambda: 1000, nms_thresh=0.7, score_thresh=0.0, min_size=1.0)
    heads = types.SimpleNamespace(box_coder=types.SimpleNamespace(weights=(10.0, 10.0, 5.0, 5.0), bbox_xform_clip=R.BBOX_XFORM_CLIP),
                                  score_thresh=0.05, nms_thresh=0.5, detection_per_img=100)
    tfm = types.SimpleNamespace(train_mode=False)
    stand = _StandInBoxHead(c)
    p7, p14 = R.MultiScaleRoIAlign(["0", "1", "2", "3"], 7, 2), R.MultiScaleRoIAlign(["0", "1", "2", "3"], 14, 2)
    a = anchors.shape[0]
    feat_bytes = sum(f.numel() * 4 for f in feats_cpu)
    st = OrderedDict()

    def rec(name, us, nbytes):
        st[name] = {"us": us, "algorithmic_bytes": int(nbytes), "GBs": nbytes / us / 1e3, "frac": nbytes / us / 1e3 / peak}

    proposals = R.boxcoder_decode_single(coder, deltas, anchors)
    rec("decode_single (det_utils.py:224)", _time_us(lambda: R.boxcoder_decode_single(coder, deltas, anchors)), a * 48)
    boxes, scores = R.rpn_filter_proposals(rpn, proposals[None], logits[None], [shape], per_level)
    n_prop = int(boxes[0].shape[0])
    rec("filter_proposals (rpn_function.py:434-529)", _time_us(lambda: R.rpn_filter_proposals(rpn, proposals[None], logits[None], [shape], per_level)),
        a * 4 + 4819 * 36 + n_prop * 20)
    f7 = p7(feats, boxes, [shape])
    rec("MultiScaleRoIAlign 7x7 (roi_head.py:491)", _time_us(lambda: p7(feats, boxes, [shape])), feat_bytes + f7.numel() * 4)
    cl, reg = stand.box(f7)
    db, ds, dl = R.roiheads_postprocess_detections(heads, cl, reg, boxes, [shape])
    m = int(db[0].shape[0])
    rec("postprocess_detections (roi_head.py:366)", _time_us(lambda: R.roiheads_postprocess_detections(heads, cl, reg, boxes, [shape])),
        cl.numel() * 4 + reg.numel() * 4 + n_prop * 16 + m * 28)
    f14 = p14(feats, db, [shape])
    rec("MultiScaleRoIAlign 14x14 (roi_head.py:540)", _time_us(lambda: p14(feats, db, [shape])), feat_bytes + f14.numel() * 4)
    ml = stand.mask(f14)
    m28 = R.maskrcnn_inference(ml, dl)
    rec("maskrcnn_inference (roi_head.py:59)", _time_us(lambda: R.maskrcnn_inference(ml, dl)), 2 * m * 784 * 4)
    res = [dict(boxes=db[0], labels=dl[0], scores=ds[0], masks=m28[0])]

    def post():
        R.transform_postprocess(tfm, [dict(res[0])], [shape], [orig])
    rec("transform.postprocess: resize_boxes + paste_masks_in_image (transform.py:388)", _time_us(post), m * (784 * 4 + 32) + m * orig[0] * orig[1] * 4)
    total_us = sum(v["us"] for v in st.values())
    total_bytes = sum(v["algorithmic_bytes"] for v in st.values())

    # the same chain with nothing read on the host between the stages: padded proposals / detections and device-side counts
    # (isb.maskrcnn.*_padded, isb_box_postprocess_rows); the box / mask head outputs are the fixed stand-in tensors, padded
    def pad_rows(t, rows):
        return t if t.shape[0] == rows else torch.cat([t, t.new_zeros((rows - t.shape[0],) + tuple(t.shape[1:]))])
    cl_p, reg_p, ml_p = pad_rows(cl, 1000), pad_rows(reg, 1000), pad_rows(ml, 100)

    def pipeline():
        props = R.boxcoder_decode_single(coder, deltas, anchors)
        pb, _, pc = R.rpn_filter_proposals_padded(props[None], logits[None], [shape], per_level, 1000, 1000, 0.7, 0.0, 1.0)
        p7(feats, [pb[0]], [shape])
        qb, _, ql, _ = R.postprocess_detections_padded(cl_p, reg_p, pb, pc, [shape], score_thresh=0.05, nms_thresh=0.5,
                                                       detections_per_img=100)
        p14(feats, [qb[0]], [shape])
        mm = R.maskrcnn_inference(ml_p, [ql[0]])
        R.paste_masks_in_image(mm[0], R.resize_boxes(qb[0], shape, orig), orig)
    pipe_us = _time_us(pipeline)
    out = {"what": "BASELINE configs[3]: Mask R-CNN R50-FPN 800x1333 post-processing chain, one image, synthetic FPN / RPN outputs; "
                   "not the headline", "num_classes": c, "proposals": n_prop, "detections": m, "stages": st,
           "total_us": total_us, "images_per_s": 1e6 / total_us, "algorithmic_bytes": int(total_bytes),
           "frac_of_hbm_peak": total_bytes / total_us / 1e3 / peak,
           "sync_free_pipeline": {"what": "the seven stages enqueued back to back with padded tensors and device-side counts, "
                                          "no host read in between (isb.maskrcnn.*_padded); CUDA events around the whole chain",
                                  "us": pipe_us, "images_per_s": 1e6 / pipe_us,
                                  "frac_of_hbm_peak": total_bytes / pipe_us / 1e3 / peak}}
    if cpu:
        out["cpu_baseline"] = maskrcnn_cpu_baseline(feats_cpu, anchors_c, deltas_c, logits_c, per_level, c, shape, orig, stand)
        out["speedup_vs_cpu"] = out["cpu_baseline"]["total_ms"] * 1e3 / total_us
    return out


def maskrcnn_cpu_baseline(feats, anchors, deltas, logits, per_level, c, shape, orig, stand):
    """The reference's CPU path for the same image on this box's host cores: torchvision's CPU kernels where the reference
    calls them (boxes.py:35 torch.ops.torchvision.nms, roi_head.py:491/540 MultiScaleRoIAlign) with all host threads, the
    numpy oracle port of the reference's torch code for the rest (one thread).  One pass after one warm-up pass."""
    from collections import OrderedDict
    import numpy as np
    import torch
    import torchvision
    from oracle import maskrcnn_np as M
    fd = OrderedDict((str(i), f) for i, f in enumerate(feats))
    tv7 = torchvision.ops.MultiScaleRoIAlign(["0", "1", "2", "3"], 7, 2)
    tv14 = torchvision.ops.MultiScaleRoIAlign(["0", "1", "2", "3"], 14, 2)
    real_nms = M.batched_nms

    def tv_batched_nms(boxes, scores, idxs, thr):       # boxes.py:38-83 with torchvision's kernel, as the reference runs it
        b = torch.from_numpy(np.ascontiguousarray(boxes))
        if b.numel() == 0:
            return np.zeros((0,), np.int64)
        off = torch.from_numpy(np.asarray(idxs)).to(b) * (b.max() + 1)
        return torchvision.ops.nms(b + off[:, None], torch.from_numpy(np.ascontiguousarray(scores)), thr).numpy()

    times = OrderedDict()

    def run():
        t = time.perf_counter()
        props = M.decode_single(deltas.numpy(), anchors.numpy())
        times["decode_single"] = time.perf_counter() - t
        t = time.perf_counter()
        rb, rs = M.filter_proposals(props, logits.numpy(), shape, per_level)
        times["filter_proposals"] = time.perf_counter() - t
        t = time.perf_counter()
        with torch.no_grad():
            f7 = tv7(fd, [torch.from_numpy(rb)], [shape])
        times["MultiScaleRoIAlign 7x7"] = time.perf_counter() - t
        cl, reg = stand.box(f7)
        t = time.perf_counter()
        db, ds, dl = M.postprocess_detections(cl.numpy(), reg.numpy(), rb, shape)
        times["postprocess_detections"] = time.perf_counter() - t
        t = time.perf_counter()
        with torch.no_grad():
            f14 = tv14(fd, [torch.from_numpy(db)], [shape])
        times["MultiScaleRoIAlign 14x14"] = time.perf_counter() - t
        ml = stand.mask(f14)
        t = time.perf_counter()
        m28 = M.maskrcnn_inference(ml.numpy(), dl)
        times["maskrcnn_inference"] = time.perf_counter() - t
        t = time.perf_counter()
        M.paste_masks_in_image(m28, M.resize_boxes(db, shape, orig), orig)
        times["transform.postprocess"] = time.perf_counter() - t

    M.batched_nms = tv_batched_nms
    try:
        run()
        run()
    finally:
        M.batched_nms = real_nms
    total = sum(times.values())
    return {"kind": "reference ops (torchvision CPU nms / MultiScaleRoIAlign) + port (numpy oracle of the reference's torch code)",
            "cores": os.cpu_count(), "torch_threads": torch.get_num_threads(), "stage_ms": {k: v * 1e3 for k, v in times.items()},
            "total_ms": total * 1e3, "images_per_s": 1.0 / total, "sample": "one image, second of two passes"}


# ----------------------------------------------------------------------------- configs[4]: VerSeg stress
def verseg_variant(dev, peak, cpu=True):
    """BASELINE.json configs[4]: ONE image, 5000 heavily overlapping boxes of one class through Fast NMS (top_k 5000,
    max_detections 1024, called directly: decode_nms cannot raise top_k, utils_bbox.py:207), then yolact_correct_boxes and
    mask assembly of the kept detections at 550 x 550 (bool [550,550,m], ~310 MB)."""
    import ctypes as C
    import numpy as np
    import torch
    from isb import _lib
    from isb.yolact import BBoxUtility
    bu = BBoxUtility()
    g = torch.Generator().manual_seed(5001)
    n = 5000
    ctr = torch.rand(n, 2, generator=g)
    w, hh = torch.rand(n, generator=g) * 0.08 + 0.01, torch.rand(n, generator=g) * 0.05 + 0.01
    box_c = torch.stack([ctr[:, 0] - w / 2, ctr[:, 1] - hh / 2, ctr[:, 0] + w / 2, ctr[:, 1] + hh / 2], 1).contiguous()
    score_c = (torch.rand(n, 1, generator=g) * 0.9 + 0.05).contiguous()
    coef_c = torch.tanh(torch.randn(n, 32, generator=g)).contiguous()
    proto_c = torch.relu(torch.randn(138, 138, 32, generator=g))
    box, score, coef, proto = box_c.to(dev), score_c.to(dev), coef_c.to(dev), proto_c.to(dev)
    h = _lib.get_handle(dev.index, max_anchors=n, max_classes=1, max_det=1024, max_top_k=5000, cand_capacity=n)
    ob, os_ = torch.empty(1024, 4, device=dev), torch.empty(1024, device=dev)
    oc, ok = torch.empty(1024, dtype=torch.int64, device=dev), torch.empty(1024, 32, device=dev)
    cnt = torch.empty(1, dtype=torch.int32, device=dev)
    P = lambda t: C.c_void_p(t.data_ptr())
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    nms = lambda: h.check(h.lib.isb_fast_nms(h.ptr, P(box), P(score), P(coef), n, 1, 32, 0.5, 5000, 1024, P(ob), P(os_), P(oc), P(ok),
                                             P(cnt), stream))
    nms_us = _time_us(nms, 3, 20)
    nms_api_us = _time_us(lambda: bu.fast_non_max_suppression(box, score, coef, 0.5, 5000, 1024), 3, 20)
    m = int(cnt.item())
    boxes_px = bu.yolact_correct_boxes(ob[:m].contiguous(), (550, 550))
    kept_coef = ok[:m].contiguous()
    asm_us = _time_us(lambda: bu.assemble_masks(proto, kept_coef, boxes_px, (550, 550)), 3, 10)
    out_bytes = 550 * 550 * (m + 8)
    out = {"what": "BASELINE configs[4]: VerSeg one-class stress, 5000 boxes -> Fast NMS (top_k 5000, max_det 1024) -> masks at "
                   "550x550; one image (does not shard: replicas only); not the headline", "kept": m,
           "fast_nms_us": nms_us, "fast_nms_python_api_us": nms_api_us,
           "assemble_masks_us": asm_us, "assemble_algorithmic_bytes": out_bytes, "assemble_GBs": out_bytes / asm_us / 1e3,
           "assemble_frac": out_bytes / asm_us / 1e3 / peak, "images_per_s": 1e6 / (nms_us + asm_us)}
    if cpu:
        from oracle import yolact_np as O
        t = time.perf_counter()
        rb, rs, rc, rk = O.fast_non_max_suppression(box_c.numpy(), score_c.numpy(), coef_c.numpy(), 0.5, 5000, 1024)
        t_nms = time.perf_counter() - t
        bpx = O.yolact_correct_boxes(rb, (550, 550))
        sub = min(128, len(rs))
        t = time.perf_counter()
        O.assemble_masks(proto_c.numpy(), rk[:sub], bpx[:sub], (550, 550))
        t_asm = (time.perf_counter() - t) * len(rs) / sub
        out["cpu_baseline"] = {"kind": "port", "cores": 1, "fast_nms_ms": t_nms * 1e3, "assemble_masks_ms": t_asm * 1e3,
                               "images_per_s": 1.0 / (t_nms + t_asm),
                               "sample": f"numpy oracle, one thread: Fast NMS of all 5000 boxes once; mask assembly of {sub} of the "
                                         f"{len(rs)} kept detections, scaled by {len(rs) / sub:.1f}"}
        out["speedup_vs_cpu"] = (t_nms + t_asm) * 1e6 / (nms_us + asm_us)
        assert len(rs) == m
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=10)
    ap.add_argument("--impl", default="isb", choices=["isb", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-variants", action="store_true", help="skip the configs[3] / configs[4] extra keys")
    ap.add_argument("--nccl-counts", action="store_true", help="multi-GPU: use the NCCL all-gather for the count exchange "
                                                               "instead of the fused NVLink peer stores")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
