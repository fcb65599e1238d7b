"""CPU: the numpy oracle against the fixtures produced by the real reference
(oracle/gen_golden.py).  This is the pin SURVEY.md section 8(c) asks for."""
import numpy as np
import pytest

from conftest import assert_close_f32, assert_masks_close, load_golden, unpack_masks
from oracle import synth, yolact_np as O
from oracle.gen_golden import digest


def test_anchor_counts():
    assert O.get_anchors((550, 550)).shape == (19248, 4)
    assert O.get_anchors((544, 544)).shape == (18525, 4)
    assert synth.yolact_dims(550) == (19248, 138)
    assert synth.yolact_dims(544) == (18525, 136)


def test_stage_known_answers():
    g = load_golden("yolact_stages")
    assert_close_f32(O.decode_boxes(g["pred"], g["anchors"]), g["decoded"], 1e-6, "decode_boxes")
    dec = g["decoded"]
    iou = O.jaccard(dec[:90], dec[40:150])
    assert np.array_equal(iou, g["iou"]), "jaccard must be bit-exact (separate fp32 ops)"
    for tag, (thr, top_k, max_det) in {"f1": (0.5, 200, 100), "f2": (0.3, 50, 20)}.items():
        b, s, c, k = O.fast_non_max_suppression(g["nms_box"], g["nms_scores"], g["nms_coef"], thr, top_k, max_det)
        assert np.array_equal(s, g[tag + "_score"]) and np.array_equal(c, g[tag + "_cls"])
        assert np.array_equal(b, g[tag + "_box"]) and np.array_equal(k, g[tag + "_coef"])
    assert np.array_equal(O.yolact_correct_boxes(g["raw"], (123, 77)), g["corrected"])
    assert np.array_equal(O.crop(g["maps"], g["crop_boxes"]), g["cropped"])
    assert_close_f32(O.upsample_bilinear(g["small"], 47, 29), g["up"], 1e-6, "upsample")
    assert_close_f32(O.upsample_bilinear(g["small"], 7, 5), g["down"], 1e-6, "downsample")


def _check_decode_nms(g, res):
    boxes, scores, cls, arg, masks = res
    assert np.array_equal(scores, g["scores"]), "scores (keep set) must be bit-exact"
    assert np.array_equal(cls, g["class_ids"]) and cls.dtype == np.int64
    assert_close_f32(boxes, g["boxes"], 1e-5, "boxes")
    assert_masks_close(masks, unpack_masks(g), 1e-3, "masks")
    assert float(np.mean(arg != g["masks_arg"])) <= 1e-3
    assert arg.dtype == np.int64 and masks.dtype == np.bool_


@pytest.mark.parametrize("name", ["yolact_tiny_a", "yolact_tiny_b", "yolact_tiny_one"])
def test_tiny_end_to_end(name):
    g = load_golden(name)
    res = O.decode_nms((g["box"], g["conf"], g["coef"], g["proto"]), g["anchors"], float(g["confidence"]),
                       float(g["nms_iou"]), g["image_shape"])
    assert not bool(g["is_none"])
    _check_decode_nms(g, res)


def test_no_survivor_returns_nones():
    g = load_golden("yolact_tiny_none")
    assert bool(g["is_none"])
    res = O.decode_nms((g["box"], g["conf"], g["coef"], g["proto"]), g["anchors"], float(g["confidence"]),
                       float(g["nms_iou"]), g["image_shape"])
    assert res == (None,) * 5


@pytest.mark.parametrize("name", ["yolact_550_coco", "yolact_550_coco_sparse", "yolact_544_voc_1080p"])
def test_full_size_end_to_end(name):
    g = load_golden(name)
    size, c = int(g["input_size"]), int(g["num_classes"])
    a, p = synth.yolact_dims(size)
    box, conf, coef, proto = [t.numpy() for t in synth.yolact_heads(int(g["seed"]), a, num_classes=c, proto=p,
                                                                     objects=int(g["objects"]))]
    assert digest(box, conf, coef, proto) == str(g["input_digest"]), "synthetic generator drifted"
    anchors = O.get_anchors((size, size)).astype(np.float32)
    res = O.decode_nms((box, conf, coef, proto), anchors, float(g["confidence"]), float(g["nms_iou"]), g["image_shape"])
    _check_decode_nms(g, res)


def test_verseg_fast_nms():
    g = load_golden("verseg_fastnms")
    box, score, coef = [t.numpy() for t in synth.verseg_boxes(int(g["seed"]), int(g["n"]))]
    assert digest(box, score, coef) == str(g["input_digest"])
    b, s, c, k = O.fast_non_max_suppression(box, score, coef, 0.5, int(g["top_k"]), int(g["max_det"]))
    assert np.array_equal(s, g["score"]) and np.array_equal(b, g["box"]) and np.array_equal(c, g["cls"])
    assert digest(k) == str(g["coef_digest"])


@pytest.mark.parametrize("case", range(6))
def test_fast_nms_list_structures(case):
    """Many short class lists, lists longer than top_k, a few long lists, a full 4096-candidate buffer, a handful of
    candidates: the reference's own fast_non_max_suppression outputs (oracle/gen_golden_fnms.py) pin the oracle."""
    from oracle.gen_golden_fnms import fnms_case_inputs
    g = load_golden("fnms_lists")
    n, c, top_k, max_det = [int(v) for v in g[f"c{case}_params"]]
    box, score, coef = [t.numpy() for t in fnms_case_inputs(n, c)]
    assert digest(box, score, coef) == str(g[f"c{case}_digest"])
    b, s, cl, k = O.fast_non_max_suppression(box, score, coef, 0.5, top_k, max_det)
    assert np.array_equal(s, g[f"c{case}_score"]) and np.array_equal(cl, g[f"c{case}_cls"])
    assert np.array_equal(b, g[f"c{case}_box"]) and np.array_equal(k, g[f"c{case}_coef"])


def test_traditional_nms_known_answers():
    """utils_bbox.py:102-140 and the traditional_nms=True branch of decode_nms, pinned to the reference run on CPU."""
    g = load_golden("yolact_traditional")
    t = O.traditional_non_max_suppression(g["t_box"], g["t_scores"], g["t_coef"], g["t_scores"].max(axis=1),
                                          float(g["t_iou"]), int(g["t_max"]))
    assert np.array_equal(t[0], g["t_out_box"]) and np.array_equal(t[1], g["t_out_score"])
    assert np.array_equal(t[2], g["t_out_cls"]) and np.array_equal(t[3], g["t_out_coef"])
    r = O.decode_nms((g["box"], g["conf"], g["coef"], g["proto"]), g["anchors"], float(g["confidence"]), float(g["nms_iou"]),
                     g["image_shape"], traditional_nms=True, max_detections=int(g["max_detections"]))
    assert np.array_equal(r[1], g["scores"]) and np.array_equal(r[2], g["class_ids"])
    assert_close_f32(r[0], g["boxes"], 1e-5, "boxes")
    assert_masks_close(r[4], unpack_masks(g), 1e-3, "masks")


def test_masks_class_against_the_wrapper_formula():
    """yolact.py:167-169 evaluated literally (broadcast multiply, flatten, fancy-index) on reference-generated decode_nms
    outputs vs the oracle's gather-first formulation."""
    from conftest import unpack_masks
    for name in ("yolact_tiny_a", "yolact_tiny_b", "yolact_tiny_one"):
        g = load_golden(name)
        masks, cls, arg = unpack_masks(g), g["class_ids"], g["masks_arg"]
        h, w, m = masks.shape
        full = masks * (cls[None, None, :] + 1)
        flat = full.reshape(-1, m)
        literal = flat[np.arange(h * w), arg.reshape(-1)].reshape(h, w)
        got = O.masks_class(masks, cls, arg)
        assert got.shape == (h, w) and np.array_equal(got, literal)
        assert got.max() <= cls.max() + 1 and (got[~masks.any(axis=2)] == 0).all()


def test_nan_inf_cases():
    """NaN / +-inf boxes and scores, oracle against the reference's own Fast NMS (a NaN IoU
    suppresses: torch.min / max / clamp / max(dim) propagate NaN) and torchvision-backed nms / batched_nms (std::max /
    std::min, `ovr > thr`: a NaN IoU does not suppress)."""
    from oracle import maskrcnn_np as M
    g = load_golden("nan_inf_cases")
    assert np.array_equal(O.jaccard(g["f_box"][:12], g["f_box"][:36]), g["f_iou"], equal_nan=True)
    for tag in ("a", "b"):
        thr, top_k, max_det = g[f"f{tag}_params"]
        b, s, c, k = O.fast_non_max_suppression(g["f_box"], g["f_scores"], g["f_coef"], float(thr), int(top_k), int(max_det))
        assert np.array_equal(s, g[f"f{tag}_score"], equal_nan=True) and np.array_equal(c, g[f"f{tag}_cls"])
        assert np.array_equal(b, g[f"f{tag}_box"], equal_nan=True) and np.array_equal(k, g[f"f{tag}_coef"])
    assert np.array_equal(M.nms(g["n_boxes"], g["n_scores"], 0.5), g["n_keep"])
    assert np.array_equal(M.batched_nms(g["n_boxes"], g["n_scores"], g["n_idxs"], 0.5), g["n_keep_b"])
    assert np.array_equal(M.nms(g["n2_boxes"], g["n_scores"], 0.5), g["n2_keep"])
    assert np.array_equal(M.batched_nms(g["n2_boxes"], g["n_scores"], g["n_idxs"], 0.5), g["n2_keep_b"])
    assert np.array_equal(M.nms(g["n3_boxes"], g["n3_scores"], 0.5), g["n3_keep"])
    assert np.array_equal(M.batched_nms(g["n3_boxes"], g["n3_scores"], g["n_idxs"], 0.5), g["n3_keep_b"])


def test_product_anchors_equal_oracle_and_reference():
    """Y1: isb.anchors.get_anchors (the product's prior table, vectorised) is bit-equal to the oracle's restatement of
    anchors.py:7-48 (itself pinned to the reference) and, in the build container, to the reference's own function."""
    import os
    import sys
    from isb import anchors as A
    for shape in ((550, 550), (544, 544), (300, 420)):
        ours = A.get_anchors(list(shape))
        assert ours.dtype == np.float64 and np.array_equal(ours, O.get_anchors(shape))
    assert A.get_anchors().shape == (19248, 4) and A.proto_size((550, 550)) == (138, 138) and A.proto_size((544, 544)) == (136, 136)
    if os.path.isdir("/root/reference"):
        sys.dont_write_bytecode = True
        if "/root/reference" not in sys.path:
            sys.path.insert(0, "/root/reference")
        from inst_model.yolact.utils.anchors import get_anchors as ref
        for shape in ([550, 550], [544, 544], [320, 480]):
            assert np.array_equal(A.get_anchors(shape), ref(shape))
        assert np.array_equal(A.get_anchors([550, 550], [16, 32, 64, 128, 256]), ref([550, 550], [16, 32, 64, 128, 256]))
