"""CPU: the numpy Mask R-CNN oracle against fixtures produced by the real reference functions and the
installed torchvision ops (oracle/gen_golden_maskrcnn.py)."""
import numpy as np

from conftest import assert_close_f32, check_pasted, load_golden
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


def test_paste_full_size():
    """configs[3]: 100 masks pasted into (800, 1333) -- oracle vs the reference's own paste_masks_in_image."""
    g = load_golden("maskrcnn_paste_full")
    out = M.paste_masks_in_image(g["probs"], g["boxes"], tuple(int(v) for v in g["image_shape"]))
    check_pasted(out, g, "", (7, 9), what="oracle paste 800x1333x100")
    assert_close_f32(out[[0, 57], 0], g["full2"], 1e-5, "two whole masks")


CHAIN_TOL = 1e-4


def oracle_chain(feats, anchors, deltas, logits, per_level, num_classes, image_shape, original_size):
    """The inference chain of oracle functions, stage k fed stage k-1's output (same stand-in heads as the golden)."""
    import torch
    props = M.decode_single(deltas, anchors)
    rb, rs = M.filter_proposals(props, logits, image_shape, per_level)
    f7 = M.multiscale_roi_align(feats, rb, image_shape, 7, 2)
    with torch.no_grad():
        cl, reg = synth.ChainBoxPredictor(num_classes)(synth.ChainBoxHead()(torch.from_numpy(f7)))
    db, ds, dl = M.postprocess_detections(cl.numpy(), reg.numpy(), rb, image_shape)
    f14 = M.multiscale_roi_align(feats, db, image_shape, 14, 2)
    with torch.no_grad():
        ml = synth.ChainMaskPredictor(num_classes)(torch.from_numpy(f14)).numpy()
    m28 = M.maskrcnn_inference(ml, dl)
    ob = M.resize_boxes(db, image_shape, original_size)
    pasted = M.paste_masks_in_image(m28, ob, original_size)
    return dict(rpn_boxes=rb, rpn_scores=rs, det_boxes_net=db, det_scores=ds, det_labels=dl, masks28=m28, det_boxes=ob, pasted=pasted)


def test_chain_vs_reference_golden():
    """The whole chain through the oracle against the chain run by the reference's own objects (gen_golden_maskrcnn.chain)."""
    g = load_golden("maskrcnn_chain")
    feats, anchors, deltas, logits, per_level = synth.chain_inputs(int(g["seed"]))
    assert digest(*[f.numpy() for f in feats], anchors.numpy(), deltas.numpy(), logits.numpy()) == str(g["input_digest"])
    shape, orig = tuple(int(v) for v in g["image_shape"]), tuple(int(v) for v in g["original_size"])
    o = oracle_chain([f.numpy() for f in feats], anchors.numpy(), deltas.numpy(), logits.numpy(), per_level,
                     int(g["num_classes"]), shape, orig)
    assert o["rpn_boxes"].shape == g["rpn_boxes"].shape
    assert_close_f32(o["rpn_scores"], g["rpn_scores"], 1e-6, "chain rpn scores")
    assert_close_f32(o["rpn_boxes"], g["rpn_boxes"], 1e-5, "chain rpn boxes")
    assert np.array_equal(o["det_labels"], g["det_labels"]), "chain labels"
    assert_close_f32(o["det_scores"], g["det_scores"], 1e-5, "chain scores")
    assert_close_f32(o["det_boxes_net"], g["det_boxes_net"], 1e-5, "chain boxes (network frame)")
    assert_close_f32(o["det_boxes"], g["det_boxes"], 1e-5, "chain boxes (original frame)")
    # the mask values sit behind three chained fp32 stages (decode -> RoIAlign at box positions that already differ by
    # 1e-5 relative -> x3 gain of the stand-in predictor -> sigmoid): end to end the bar is 1e-4; stage by stage, each
    # fed the same input, it stays 1e-5 (test_stage_known_answers, tests/test_maskrcnn_gpu.py::test_chain_*)
    assert_close_f32(o["masks28"], g["masks28"], CHAIN_TOL, "chain 28x28 masks")
    check_pasted(o["pasted"], g, "pasted_", (5, 7), tol=CHAIN_TOL, what="chain pasted masks")
