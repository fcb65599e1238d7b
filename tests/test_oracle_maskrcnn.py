"""CPU: the numpy Mask R-CNN oracle against fixtures produced by the real reference functions and the
installed torchvision ops (oracle/gen_golden_maskrcnn.py)."""
import numpy as np

from conftest import assert_close_f32, load_golden
from oracle import maskrcnn_np as M, synth
from oracle.gen_golden import digest


def test_stage_known_answers():
    g = load_golden("maskrcnn_stages")
    assert_close_f32(M.decode_single(g["dec_rel1"], g["dec_boxes"]), g["dec_out1"], 1e-5, "decode k=1")
    ext = np.abs(g["dec_out6"]).reshape(-1, 6, 4).max(axis=2, keepdims=True).repeat(4, axis=2).reshape(-1, 24)
    assert_close_f32(M.decode_single(g["dec_rel6"], g["dec_boxes"], (10, 10, 5, 5)), g["dec_out6"], 1e-5, "decode k=6", ext)
    clipped = M.clip_boxes_to_image(g["clip_in"], (600, 800))
    assert np.array_equal(clipped, g["clip_out"])
    assert np.array_equal(M.remove_small_boxes(clipped, 1.0), g["small_keep"])
    for thr, tag in ((0.5, "05"), (0.7, "07")):
        assert np.array_equal(M.nms(g["nms_boxes"], g["nms_scores"], thr), g["nms_keep_" + tag])
        assert np.array_equal(M.batched_nms(g["nms_boxes"], g["nms_scores"], g["nms_idxs"], thr), g["bnms_keep_" + tag])
    assert np.array_equal(M.nms(g["adv_boxes"], g["adv_scores"], 0.5), g["adv_keep"]), "near-threshold pairs"
    feats = [g[f"roi_feat{i}"] for i in range(4)]
    for size in (7, 14):
        out = M.multiscale_roi_align(feats, g["roi_boxes"], (800, 1333), size, 2)
        assert_close_f32(out, g[f"roi_out{size}"], 1e-5, f"ms roi align {size}")
    assert_close_f32(M.maskrcnn_inference(g["inf_logits"], g["inf_labels"]), g["inf_out"], 1e-5, "mask inference")
    assert np.array_equal(M.resize_boxes(g["paste_boxes"], (800, 1216), (544, 544)), g["resize_out"])
    assert_close_f32(M.paste_masks_in_image(g["paste_masks"], g["paste_boxes"], (544, 544)), g["paste_out"], 1e-5, "paste")


def test_rpn_filter_full_size():
    g = load_golden("maskrcnn_rpn_filter")
    anchors, deltas, logits, per_level = synth.rpn_heads(int(g["seed"]), tuple(g["image_hw"]))
    assert digest(anchors.numpy(), deltas.numpy(), logits.numpy()) == str(g["input_digest"])
    proposals = M.decode_single(deltas.numpy(), anchors.numpy())
    boxes, scores = M.filter_proposals(proposals, logits.numpy(), tuple(g["image_shape"]), per_level)
    assert boxes.shape == g["boxes"].shape
    assert_close_f32(scores, g["scores"], 1e-6, "rpn scores")
    assert_close_f32(boxes, g["boxes"], 1e-5, "rpn boxes")


def test_box_postprocess():
    g = load_golden("maskrcnn_box_post")
    proposals, logits, reg = synth.box_head_outputs(int(g["seed"]), 1000, int(g["num_classes"]), tuple(g["image_shape"]))
    assert digest(proposals.numpy(), logits.numpy(), reg.numpy()) == str(g["input_digest"])
    b, s, l = M.postprocess_detections(logits.numpy(), reg.numpy(), proposals.numpy(), tuple(g["image_shape"]),
                                       scores=g["softmax"])
    assert np.array_equal(l, g["labels"]) and np.array_equal(s, g["scores"])
    assert_close_f32(b, g["boxes"], 1e-5, "detections")
    b2, s2, l2 = M.postprocess_detections(logits.numpy(), reg.numpy(), proposals.numpy(), tuple(g["image_shape"]))
    assert np.array_equal(l2, g["labels"])
    assert_close_f32(s2, g["scores"], 1e-5, "softmax scores")


def test_class_image_wrapper_tail():
    """mask_rcnn.py:174-198 (next row): oracle vs the reference's paste + the wrapper's own statements."""
    g = load_golden("maskrcnn_class_image")
    pasted = M.paste_masks_in_image(g["probs"], g["boxes"], tuple(int(v) for v in g["paste_hw"]))
    for k, (h, w) in enumerate(g["sizes"]):
        mc, arg = M.masks_class_image(pasted, g["labels"], g["scores"], float(g["confidence"]), (int(h), int(w)))
        assert float(np.mean(mc != g[f"class_map_{k}"])) <= 1e-3
        assert float(np.mean(arg != g[f"arg_{k}"])) <= 1e-3
    none, _ = M.masks_class_image(pasted, g["labels"], g["scores"], 2.0, (50, 50))
    assert none is None


def test_rcnn_transform_vs_reference_golden():
    """GeneralizedRCNNTransform.forward (transform.py:436-459): normalize + bilinear resize + batch padding, and the
    output-size rule on 400 random shapes (exact), against outputs of the reference itself."""
    g = load_golden("maskrcnn_transform")
    mean, std = g["mean"], g["std"]
    for tag in ("a", "b"):
        mn, mx = [int(v) for v in g[f"{tag}_min_max"]]
        imgs = [g[f"{tag}_in{i}"] for i in range(int(g[f"{tag}_n"]))]
        batch, sizes = M.rcnn_transform(imgs, mn, mx, mean, std)
        assert batch.shape == tuple(g[f"{tag}_shape"]) and [tuple(s) for s in g[f"{tag}_sizes"]] == sizes
        got = batch if tag == "b" else batch[:, :, ::3, ::3]
        assert_close_f32(got, g[f"{tag}_batch"], 1e-5, f"transform {tag}")
        assert np.array_equal(got == 0, g[f"{tag}_batch"] == 0)
    for (h, w), want in zip(g["rule_hw"], g["rule_sizes"]):
        assert M.transform_output_size(int(h), int(w), 800, 1333) == tuple(want), (h, w)


def test_product_size_rule_matches_reference_golden():
    from isb import maskrcnn as R
    g = load_golden("maskrcnn_transform")
    for (h, w), want in zip(g["rule_hw"], g["rule_sizes"]):
        assert R.transform_output_size(int(h), int(w), 800, 1333) == tuple(want), (h, w)
