"""Generate tests/golden/*.npz by running the REAL reference (imported from /root/reference).

Run in the build container only (the reference tree does not exist on the GPU box):

    PYTHONDONTWRITEBYTECODE=1 python -m oracle.gen_golden

The fixtures pin the numpy oracle (tests/test_oracle_golden.py) and, on the GPU, the CUDA
path (tests/test_yolact_gpu.py).  Library versions used are recorded in every file.
Full-size cases store only outputs (inputs are re-drawn from the recorded seed and checked
against a stored checksum); tiny cases store inputs too.
"""
from __future__ import annotations

import hashlib
import os
import sys

import numpy as np
import torch
import torchvision

sys.dont_write_bytecode = True
REF = "/root/reference"
HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")


def _ref():
    if REF not in sys.path:
        sys.path.insert(0, REF)
    from inst_model.yolact.utils.utils_bbox import BBoxUtility
    return BBoxUtility()


def digest(*arrays):
    h = hashlib.sha1()
    for a in arrays:
        h.update(np.ascontiguousarray(a).tobytes())
    return h.hexdigest()


def _meta():
    return dict(torch_version=torch.__version__, torchvision_version=torchvision.__version__)


def _save(name, **kw):
    os.makedirs(OUT, exist_ok=True)
    path = os.path.join(OUT, name + ".npz")
    np.savez_compressed(path, **_meta(), **kw)
    print(f"{name}: {os.path.getsize(path) / 1024:.1f} KiB")


def _pack_result(res):
    boxes, scores, cls, arg, masks = [r.numpy() for r in res]
    return dict(boxes=boxes, scores=scores, class_ids=cls, masks_arg=arg.astype(np.int32),
                masks_packed=np.packbits(masks), masks_shape=np.array(masks.shape))


def yolact_full(name, seed, input_size, num_classes, image_shape, confidence, nms_iou, **kw):
    from . import synth, yolact_np
    bu = _ref()
    a, p = synth.yolact_dims(input_size)
    anchors = torch.from_numpy(yolact_np.get_anchors((input_size, input_size))).float()
    box, conf, coef, proto = synth.yolact_heads(seed, a, num_classes=num_classes, proto=p, **kw)
    res = bu.decode_nms((box[None], conf[None], coef[None], proto[None]), anchors, confidence, nms_iou,
                        np.array(image_shape))
    assert res[0] is not None
    _save(name, seed=seed, input_size=input_size, num_classes=num_classes, image_shape=np.array(image_shape),
          confidence=confidence, nms_iou=nms_iou, objects=kw.get("objects", 12),
          tie_free=synth.tie_free(conf.numpy(), confidence),
          input_digest=digest(box.numpy(), conf.numpy(), coef.numpy(), proto.numpy()), **_pack_result(res))


def yolact_tiny(name, seed, a, c, k, ph, pw, image_shape, confidence, nms_iou, objects=4, run=12, boost=None):
    """Small fully stored case with made-up anchors (the kernels are size-generic)."""
    from . import synth
    bu = _ref()
    g = torch.Generator().manual_seed(seed + 7)
    anchors = torch.cat([torch.rand(a, 2, generator=g), torch.rand(a, 2, generator=g) * 0.3 + 0.05], 1)
    box, conf, coef, proto = synth.yolact_heads(seed, a, num_classes=c, num_coef=k, proto_hw=(ph, pw),
                                                objects=objects, run=run)
    if boost is not None:
        conf = boost(conf)
    res = bu.decode_nms((box[None], conf[None], coef[None], proto[None]), anchors, confidence, nms_iou,
                        np.array(image_shape))
    out = dict(anchors=anchors.numpy(), box=box.numpy(), conf=conf.numpy(), coef=coef.numpy(), proto=proto.numpy(),
               image_shape=np.array(image_shape), confidence=confidence, nms_iou=nms_iou,
               is_none=res[0] is None)
    if res[0] is not None:
        out.update(_pack_result(res))
    _save(name, **out)


def yolact_stages(name, seed):
    """Stage-wise known answers: each reference method on its own inputs."""
    from . import synth
    bu = _ref()
    g = torch.Generator().manual_seed(seed)
    n, c, k = 260, 7, 32
    anchors = torch.cat([torch.rand(n, 2, generator=g), torch.rand(n, 2, generator=g) * 0.3 + 0.02], 1)
    pred = torch.randn(n, 4, generator=g) * 0.5
    dec = bu.decode_boxes(pred, anchors)
    iou = bu.jaccard(dec[:90], dec[40:150])
    scores = torch.rand(n, c, generator=g)                 # all classes for all rows, as Fast NMS gets them
    coef = torch.tanh(torch.randn(n, k, generator=g))
    # clustered boxes so suppression happens
    ctr = torch.rand(12, 2, generator=g)[torch.randint(0, 12, (n,), generator=g)] + torch.randn(n, 2, generator=g) * 0.02
    wh = torch.rand(n, 2, generator=g) * 0.1 + 0.1
    nbox = torch.cat([ctr - wh / 2, ctr + wh / 2], 1)
    f1 = bu.fast_non_max_suppression(nbox, scores, coef, 0.5, 200, 100)
    f2 = bu.fast_non_max_suppression(nbox, scores, coef, 0.3, 50, 20)
    raw = torch.randn(40, 4, generator=g) * 0.6 + 0.5      # includes flipped / out-of-range corners
    corr = bu.yolact_correct_boxes(raw, np.array([123, 77]))
    maps = torch.rand(37, 53, 9, generator=g)
    cb = torch.rand(9, 4, generator=g) * torch.tensor([53.0, 37.0, 53.0, 37.0])
    cb = torch.cat([torch.min(cb[:, :2], cb[:, 2:]), torch.max(cb[:, :2], cb[:, 2:])], 1)
    cropped = bu.crop(maps, cb)
    small = torch.rand(5, 11, 13, generator=g)
    up = torch.nn.functional.interpolate(small[None], (47, 29), mode="bilinear", align_corners=False)[0]
    down = torch.nn.functional.interpolate(small[None], (7, 5), mode="bilinear", align_corners=False)[0]
    _save(name, anchors=anchors.numpy(), pred=pred.numpy(), decoded=dec.numpy(), iou=iou.numpy(),
          nms_box=nbox.numpy(), nms_scores=scores.numpy(), nms_coef=coef.numpy(),
          f1_box=f1[0].numpy(), f1_score=f1[1].numpy(), f1_cls=f1[2].numpy(), f1_coef=f1[3].numpy(),
          f2_box=f2[0].numpy(), f2_score=f2[1].numpy(), f2_cls=f2[2].numpy(), f2_coef=f2[3].numpy(),
          raw=raw.numpy(), corrected=corr.numpy(), maps=maps.numpy(), crop_boxes=cb.numpy(), cropped=cropped.numpy(),
          small=small.numpy(), up=up.numpy(), down=down.numpy())


def yolact_traditional(name, seed):
    """utils_bbox.py:102-140 and the traditional_nms=True branch of decode_nms (CPU only in the reference)."""
    from . import synth
    bu = _ref()
    g = torch.Generator().manual_seed(seed + 7)
    a, c, k = 500, 6, 32
    anchors = torch.cat([torch.rand(a, 2, generator=g), torch.rand(a, 2, generator=g) * 0.3 + 0.05], 1)
    box, conf, coef, proto = synth.yolact_heads(seed, a, num_classes=c, num_coef=k, proto_hw=(20, 24), objects=5, run=25)
    res = bu.decode_nms((box[None], conf[None], coef[None], proto[None]), anchors, 0.05, 0.5, np.array((45, 70)),
                        traditional_nms=True, max_detections=40)
    # stage call on clustered boxes
    n = 300
    ctr = torch.rand(10, 2, generator=g)[torch.randint(0, 10, (n,), generator=g)] + torch.randn(n, 2, generator=g) * 0.02
    wh = torch.rand(n, 2, generator=g) * 0.1 + 0.1
    nbox = torch.cat([ctr - wh / 2, ctr + wh / 2], 1)
    scores = torch.rand(n, 5, generator=g)
    ncoef = torch.tanh(torch.randn(n, k, generator=g))
    t = bu.traditional_non_max_suppression(nbox, scores, ncoef, scores.max(dim=1).values, 0.4, 50)
    _save(name, anchors=anchors.numpy(), box=box.numpy(), conf=conf.numpy(), coef=coef.numpy(), proto=proto.numpy(),
          image_shape=np.array((45, 70)), confidence=0.05, nms_iou=0.5, max_detections=40, **_pack_result(res),
          t_box=nbox.numpy(), t_scores=scores.numpy(), t_coef=ncoef.numpy(), t_iou=0.4, t_max=50,
          t_out_box=t[0].numpy(), t_out_score=t[1].numpy(), t_out_cls=t[2].numpy(), t_out_coef=t[3].numpy())


def verseg(name, seed, n=5000, top_k=5000, max_det=1024):
    from . import synth
    bu = _ref()
    box, score, coef = synth.verseg_boxes(seed, n)
    r = bu.fast_non_max_suppression(box, score, coef, 0.5, top_k, max_det)
    _save(name, seed=seed, n=n, top_k=top_k, max_det=max_det,
          input_digest=digest(box.numpy(), score.numpy(), coef.numpy()),
          box=r[0].numpy(), score=r[1].numpy(), cls=r[2].numpy(), coef_digest=digest(r[3].numpy()))


def main():
    torch.manual_seed(0)
    yolact_stages("yolact_stages", 11)
    yolact_tiny("yolact_tiny_a", 21, a=600, c=6, k=32, ph=20, pw=24, image_shape=(45, 70), confidence=0.05, nms_iou=0.5)
    yolact_tiny("yolact_tiny_b", 22, a=333, c=3, k=32, ph=17, pw=17, image_shape=(64, 64), confidence=0.3, nms_iou=0.3,
                objects=6, run=30)
    # nothing passes the threshold -> five Nones (utils_bbox.py:203-204)
    yolact_tiny("yolact_tiny_none", 23, a=200, c=4, k=32, ph=8, pw=8, image_shape=(20, 20), confidence=0.9999,
                nms_iou=0.5, objects=0)
    # exactly one survivor
    def one(conf):
        conf = conf.clone() * 0.01
        conf[57, 2] = 0.9
        return conf
    yolact_tiny("yolact_tiny_one", 24, a=200, c=4, k=32, ph=8, pw=8, image_shape=(20, 31), confidence=0.5,
                nms_iou=0.5, objects=0, boost=one)
    yolact_full("yolact_550_coco", 2001, 550, 81, (550, 550), 0.05, 0.5)
    yolact_full("yolact_550_coco_sparse", 2101, 550, 81, (550, 550), 0.05, 0.5, objects=3)
    yolact_full("yolact_544_voc_1080p", 1001, 544, 21, (1080, 1920), 0.5, 0.3)
    verseg("verseg_fastnms", 5001)
    yolact_traditional("yolact_traditional", 31)


if __name__ == "__main__":
    main()
