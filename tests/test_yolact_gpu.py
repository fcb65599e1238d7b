"""GPU parity tests for the YOLACT path: the CUDA kernels (through the C ABI, via the
BBoxUtility drop-in) against the reference-generated golden fixtures and the numpy oracle."""
import numpy as np
import pytest
import torch

from conftest import assert_close_f32, assert_masks_close, load_golden, unpack_masks
from oracle import synth, yolact_np as O

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def bu():
    assert torch.cuda.is_available()
    from isb import _lib
    _lib.load()  # must be the in-tree libisb.so; raises if missing
    from isb.yolact import BBoxUtility
    return BBoxUtility()


def dev(x):
    return torch.as_tensor(np.asarray(x)).cuda()


def host(t):
    return t.cpu().numpy()


# ------------------------------------------------------------------------------- stage KATs
def test_decode_boxes(bu):
    g = load_golden("yolact_stages")
    out = host(bu.decode_boxes(dev(g["pred"]), dev(g["anchors"])))
    assert_close_f32(out, g["decoded"], 1e-5, "decode_boxes")


def test_jaccard_bit_exact(bu):
    g = load_golden("yolact_stages")
    dec = g["decoded"]
    out = host(bu.jaccard(dev(dec[:90]), dev(dec[40:150])))
    assert np.array_equal(out, g["iou"])
    # batched form
    a = np.stack([dec[:50], dec[50:100]])
    out3 = host(bu.jaccard(dev(a), dev(a)))
    assert np.array_equal(out3, O.jaccard(a, a))


@pytest.mark.parametrize("tag,thr,top_k,max_det", [("f1", 0.5, 200, 100), ("f2", 0.3, 50, 20)])
def test_fast_nms_stage(bu, tag, thr, top_k, max_det):
    g = load_golden("yolact_stages")
    b, s, c, k = bu.fast_non_max_suppression(dev(g["nms_box"]), dev(g["nms_scores"]), dev(g["nms_coef"]), thr, top_k, max_det)
    assert np.array_equal(host(s), g[tag + "_score"])
    assert np.array_equal(host(c), g[tag + "_cls"]) and c.dtype == torch.int64
    assert np.array_equal(host(b), g[tag + "_box"])
    assert np.array_equal(host(k), g[tag + "_coef"])


def test_traditional_nms_stage_and_decode(bu):
    """Y15: per-class greedy NMS (utils_bbox.py:102-140) and decode_nms(traditional_nms=True), against the reference's
    CPU outputs: keep set, order, class ids and coefficient rows bit-exact."""
    g = load_golden("yolact_traditional")
    smax = g["t_scores"].max(axis=1)
    b, s, c, k = bu.traditional_non_max_suppression(dev(g["t_box"]), dev(g["t_scores"]), dev(g["t_coef"]), dev(smax),
                                                    float(g["t_iou"]), int(g["t_max"]))
    assert np.array_equal(host(s), g["t_out_score"])
    assert np.array_equal(host(c), g["t_out_cls"]) and c.dtype == torch.int64
    assert np.array_equal(host(b), g["t_out_box"]) and np.array_equal(host(k), g["t_out_coef"])
    res = bu.decode_nms((dev(g["box"])[None], dev(g["conf"])[None], dev(g["coef"])[None], dev(g["proto"])[None]),
                        dev(g["anchors"]), float(g["confidence"]), float(g["nms_iou"]), g["image_shape"],
                        traditional_nms=True, max_detections=int(g["max_detections"]))
    _check(g, res)


def test_correct_boxes_and_crop_bit_exact(bu):
    g = load_golden("yolact_stages")
    assert np.array_equal(host(bu.yolact_correct_boxes(dev(g["raw"]), np.array([123, 77]))), g["corrected"])
    assert np.array_equal(host(bu.crop(dev(g["maps"]), dev(g["crop_boxes"]))), g["cropped"])


def test_mask_logits_within_1e5(bu):
    gen = torch.Generator().manual_seed(3)
    proto = torch.relu(torch.randn(138, 138, 32, generator=gen))
    coef = torch.tanh(torch.randn(100, 32, generator=gen))
    ref = torch.matmul(proto.double(), coef.double().t()).numpy()
    out = host(bu.mask_logits(proto.cuda(), coef.cuda()))
    assert_close_f32(out, ref, 1e-5, "mask logits")


def test_assemble_masks_vs_oracle(bu):
    gen = torch.Generator().manual_seed(4)
    proto = torch.relu(torch.randn(23, 31, 32, generator=gen))
    coef = torch.tanh(torch.randn(13, 32, generator=gen))
    shape = (61, 90)
    raw = torch.rand(13, 4, generator=gen)
    boxes = O.yolact_correct_boxes(raw.numpy(), shape)
    boxes[3] = [0, 0, shape[1], shape[0]]          # a full-image detection
    boxes[5] = [10.5, 7.25, 10.5, 30.0]            # empty box
    arg_ref, m_ref = O.assemble_masks(proto.numpy(), coef.numpy(), boxes, shape)
    arg, m = bu.assemble_masks(proto.cuda(), coef.cuda(), dev(boxes), shape)
    assert m.dtype == torch.bool and arg.dtype == torch.int64
    assert_masks_close(host(m), m_ref, 1e-3, "masks")
    assert float(np.mean(host(arg) != arg_ref)) <= 1e-3


def test_assemble_masks_downsample_and_many(bu):
    # H < P (tile shrinks to keep the prototype footprint bounded) and m > 256 (several list rounds)
    gen = torch.Generator().manual_seed(6)
    proto = torch.relu(torch.randn(40, 36, 32, generator=gen))
    shape = (17, 23)
    m = 300
    coef = torch.tanh(torch.randn(m, 32, generator=gen))
    boxes = O.yolact_correct_boxes(torch.rand(m, 4, generator=gen).numpy(), shape)
    arg_ref, m_ref = O.assemble_masks(proto.numpy(), coef.numpy(), boxes, shape)
    arg, mk = bu.assemble_masks(proto.cuda(), coef.cuda(), dev(boxes), shape)
    assert_masks_close(host(mk), m_ref, 1e-3, "masks")
    assert float(np.mean(host(arg) != arg_ref)) <= 2e-3


# ------------------------------------------------------------------------------- end to end
def _check(g, res):
    boxes, scores, cls, arg, masks = res
    assert boxes.dtype == torch.float32 and scores.dtype == torch.float32 and cls.dtype == torch.int64
    assert arg.dtype == torch.int64 and masks.dtype == torch.bool and masks.is_contiguous()
    assert np.array_equal(host(scores), g["scores"]), "keep set / scores must be bit-exact"
    assert np.array_equal(host(cls), g["class_ids"])
    assert_close_f32(host(boxes), g["boxes"], 1e-5, "boxes")
    assert_masks_close(host(masks), unpack_masks(g), 1e-3, "masks")
    assert float(np.mean(host(arg) != g["masks_arg"])) <= 1e-3


@pytest.mark.parametrize("name", ["yolact_tiny_a", "yolact_tiny_b", "yolact_tiny_one"])
def test_tiny_end_to_end(bu, name):
    g = load_golden(name)
    res = bu.decode_nms((dev(g["box"])[None], dev(g["conf"])[None], dev(g["coef"])[None], dev(g["proto"])[None]),
                        dev(g["anchors"]), float(g["confidence"]), float(g["nms_iou"]), g["image_shape"])
    _check(g, res)


def test_no_survivor_returns_nones(bu):
    g = load_golden("yolact_tiny_none")
    res = bu.decode_nms((dev(g["box"])[None], dev(g["conf"])[None], dev(g["coef"])[None], dev(g["proto"])[None]),
                        dev(g["anchors"]), float(g["confidence"]), float(g["nms_iou"]), g["image_shape"])
    assert res == (None,) * 5


def _full_inputs(g):
    """Re-draw the full-size inputs from the golden's seed and check them against the digest recorded when the reference
    produced the golden: a different CPU (softmax / RNG code path) would otherwise show up as an opaque keep-set failure."""
    from oracle.gen_golden import digest
    size, c = int(g["input_size"]), int(g["num_classes"])
    a, p = synth.yolact_dims(size)
    heads = synth.yolact_heads(int(g["seed"]), a, num_classes=c, proto=p, objects=int(g["objects"]))
    anchors = torch.from_numpy(O.get_anchors((size, size))).float()
    if "input_digest" in g.files:
        got = digest(*[t.numpy() for t in heads])
        assert got == str(g["input_digest"]), ("seeded inputs differ from the ones the reference golden was generated from "
                                               f"(digest {got} vs {g['input_digest']}): this host's torch CPU RNG / softmax "
                                               "does not reproduce the build container's bits")
    return heads, anchors


@pytest.mark.parametrize("name", ["yolact_550_coco", "yolact_550_coco_sparse", "yolact_544_voc_1080p"])
def test_full_size_end_to_end(bu, name):
    g = load_golden(name)
    (box, conf, coef, proto), anchors = _full_inputs(g)
    res = bu.decode_nms((box[None].cuda(), conf[None].cuda(), coef[None].cuda(), proto[None].cuda()), anchors.cuda(),
                        float(g["confidence"]), float(g["nms_iou"]), g["image_shape"])
    _check(g, res)


def test_batch_equals_single_and_oracle(bu):
    a, p = synth.yolact_dims(550)
    anchors = torch.from_numpy(O.get_anchors((550, 550))).float()
    bsz = 5
    box, conf, coef, proto = synth.yolact_batch(2000, bsz, a, num_classes=81, proto=p)
    conf[3] = conf[3] * 0.0 + 0.001                        # image 3: nothing passes -> count 0
    res = bu.decode_nms_batch((box.cuda(), conf.cuda(), coef.cuda(), proto.cuda()), anchors.cuda(), 0.05, 0.5, (550, 550))
    counts = host(res.counts)
    assert counts[3] == 0 and (counts[[0, 1, 2, 4]] == 100).all()
    for b in (0, 1, 4):
        single = bu.decode_nms((box[b:b + 1].cuda(), conf[b:b + 1].cuda(), coef[b:b + 1].cuda(), proto[b:b + 1].cuda()),
                               anchors.cuda(), 0.05, 0.5, np.array([550, 550]))
        batched = res.image(b)
        for x, y in zip(single, batched):
            assert torch.equal(x, y)
    ref = O.decode_nms((box[1].numpy(), conf[1].numpy(), coef[1].numpy(), proto[1].numpy()), anchors.numpy(), 0.05, 0.5, (550, 550))
    got = res.image(1)
    assert np.array_equal(host(got[1]), ref[1]) and np.array_equal(host(got[2]), ref[2])
    assert_masks_close(host(got[4]), ref[4], 1e-3, "masks")
    assert res.image(3) == (None,) * 5


def test_verseg_fast_nms(bu):
    g = load_golden("verseg_fastnms")
    box, score, coef = synth.verseg_boxes(int(g["seed"]), int(g["n"]))
    from oracle.gen_golden import digest
    assert digest(box.numpy(), score.numpy(), coef.numpy()) == str(g["input_digest"]), "seeded VerSeg inputs differ from the golden's"
    b, s, c, k = bu.fast_non_max_suppression(box.cuda(), score.cuda(), coef.cuda(), 0.5, int(g["top_k"]), int(g["max_det"]))
    assert np.array_equal(host(s), g["score"]) and np.array_equal(host(b), g["box"]) and np.array_equal(host(c), g["cls"])


@pytest.mark.parametrize("case", range(6))
def test_fast_nms_list_structures_vs_reference_golden(bu, case):
    """The single-CTA Fast NMS kernel on the list structures it treats differently (sort-free class walk, rank counting for
    lists longer than top_k, bitonic fallback, full buffer) against the REFERENCE's outputs (oracle/gen_golden_fnms.py)."""
    from oracle.gen_golden import digest
    from oracle.gen_golden_fnms import fnms_case_inputs
    g = load_golden("fnms_lists")
    n, c, top_k, max_det = [int(v) for v in g[f"c{case}_params"]]
    box, score, coef = fnms_case_inputs(n, c)
    assert digest(box.numpy(), score.numpy(), coef.numpy()) == str(g[f"c{case}_digest"]), "seeded inputs differ from the golden's"
    b, s, cl, k = bu.fast_non_max_suppression(box.cuda(), score.cuda(), coef.cuda(), 0.5, top_k, max_det)
    assert np.array_equal(host(s), g[f"c{case}_score"]) and np.array_equal(host(cl), g[f"c{case}_cls"])
    assert np.array_equal(host(b), g[f"c{case}_box"]) and np.array_equal(host(k), g[f"c{case}_coef"])


def test_candidate_overflow_grows_workspace(bu, monkeypatch):
    # confidence 0 makes every (anchor, class) entry a candidate: 600*5 = 3000 per image, more than
    # the 1024 the workspace is created with here -> counts = -1 -> the shim grows it and reruns
    from isb import _lib
    for hd in _lib._handles.values():
        hd.close()
    _lib._handles.clear()
    monkeypatch.setitem(_lib._DEFAULTS, "cand_capacity", 1024)
    g = load_golden("yolact_tiny_a")
    args = ((dev(g["box"])[None], dev(g["conf"])[None], dev(g["coef"])[None], dev(g["proto"])[None]), dev(g["anchors"]))
    res = bu.decode_nms(*args, 0.0, 0.5, g["image_shape"])
    ref = O.decode_nms((g["box"], g["conf"], g["coef"], g["proto"]), g["anchors"], 0.0, 0.5, g["image_shape"])
    assert np.array_equal(host(res[1]), ref[1]) and np.array_equal(host(res[2]), ref[2])
    assert_masks_close(host(res[4]), ref[4], 1e-3, "masks")
    assert _lib.get_handle(0).limits.cand_capacity >= 3000


# ------------------------------------------------------------------------------- full-size properties
def test_benchmark_size_properties(bu):
    """B = 64 at COCO 550^2 (BASELINE.json configs[1]): properties that need no oracle."""
    a, p = synth.yolact_dims(550)
    anchors = torch.from_numpy(O.get_anchors((550, 550))).float().cuda()
    bsz = 64
    one = synth.yolact_batch(3000, 4, a, num_classes=81, proto=p)
    box, conf, coef, proto = [t.cuda().repeat((bsz // 4,) + (1,) * (t.dim() - 1)) for t in one]
    r1 = bu.decode_nms_batch((box, conf, coef, proto), anchors, 0.05, 0.5, (550, 550))
    r2 = bu.decode_nms_batch((box, conf, coef, proto), anchors, 0.05, 0.5, (550, 550))
    torch.cuda.synchronize()
    counts = host(r1.counts)
    assert (counts > 0).all() and (counts <= 100).all()
    assert torch.equal(r1.counts, r2.counts)
    for b in (0, 17, 63):
        m = int(counts[b])
        bx, sc, cl, arg, mk = r1.image(b, m)
        bx2, sc2, cl2, arg2, mk2 = r2.image(b, m)
        assert torch.equal(mk, mk2) and torch.equal(arg, arg2) and torch.equal(sc, sc2)      # idempotent
        assert torch.equal(mk, r1.image(b % 4, m)[4])                                       # replicas agree
        s = host(sc)
        assert (np.diff(s) <= 0).all() and (s > 0.05).all()                                 # sorted, thresholded
        bxh = host(bx)
        assert (bxh[:, 0] >= 0).all() and (bxh[:, 2] <= 550).all() and (bxh[:, 1] >= 0).all() and (bxh[:, 3] <= 550).all()
        mkh, argh = host(mk), host(arg)
        ys, xs, js = np.nonzero(mkh)
        assert (xs >= bxh[js, 0]).all() and (xs < bxh[js, 2]).all() and (ys >= bxh[js, 1]).all() and (ys < bxh[js, 3]).all()
        assert ((argh >= 0) & (argh < m)).all()
        cy, cx = np.nonzero(mkh.any(axis=2))
        assert mkh[cy, cx, argh[cy, cx]].all()            # the arg-max detection is > 0.5 wherever any is


@pytest.mark.parametrize("bsz", [1, 7, 26])
def test_detect_host_matches_device(bu, bsz):
    """isb_yolact_detect_host (pinned host buffers, 3-lane H2D/compute/D2H pipeline) == device path."""
    import ctypes as C
    from isb import _lib
    g = load_golden("yolact_tiny_a")
    gen = torch.Generator().manual_seed(bsz)
    a, c = g["conf"].shape
    ph, pw, k = g["proto"].shape
    box = torch.from_numpy(g["box"])[None].repeat(bsz, 1, 1) + torch.randn(bsz, a, 4, generator=gen) * 0.05
    conf = torch.from_numpy(g["conf"])[None].repeat(bsz, 1, 1) * (torch.rand(bsz, a, 1, generator=gen) * 0.5 + 0.5)
    coef = torch.from_numpy(g["coef"])[None].repeat(bsz, 1, 1)
    proto = torch.from_numpy(g["proto"])[None].repeat(bsz, 1, 1, 1) + torch.rand(bsz, ph, pw, k, generator=gen) * 0.1
    if bsz > 2:
        conf[2] = 0.0                              # an image with no detections
    anchors = torch.from_numpy(g["anchors"])
    hgt, wid = int(g["image_shape"][0]), int(g["image_shape"][1])
    d = 100
    ref = bu.decode_nms_batch((box.cuda(), conf.cuda(), coef.cuda(), proto.cuda()), anchors.cuda(), 0.05, 0.5, (hgt, wid))
    torch.cuda.synchronize()
    slot = (hgt * wid * d + 15) // 16 * 16
    pin = [t.contiguous().pin_memory() for t in (box, conf, coef, proto, anchors)]
    hb, hs = torch.zeros((bsz, d, 4)).pin_memory(), torch.zeros((bsz, d)).pin_memory()
    hc = torch.zeros((bsz, d), dtype=torch.int64).pin_memory()
    ha = torch.zeros((bsz, hgt, wid), dtype=torch.int64).pin_memory()
    hm = torch.zeros((bsz, slot), dtype=torch.uint8).pin_memory()
    hn = torch.zeros((bsz,), dtype=torch.int32).pin_memory()
    h = _lib.get_handle(0, max_batch=bsz, max_anchors=a, max_classes=c - 1)
    yin = _lib.YolactIn(bsz, a, c, k, ph, pw, hgt, wid, pin[0].data_ptr(), pin[1].data_ptr(), pin[2].data_ptr(),
                        pin[3].data_ptr(), pin[4].data_ptr(), 0.05, 0.5, 200, d)
    yout = _lib.YolactOut(hb.data_ptr(), hs.data_ptr(), hc.data_ptr(), ha.data_ptr(), hm.data_ptr(), slot, hn.data_ptr())
    h.check(h.lib.isb_yolact_detect_host(h.ptr, C.byref(yin), C.byref(yout), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
    counts = ref.counts.cpu()
    assert torch.equal(hn, counts)
    for b in range(bsz):
        m = int(counts[b])
        if m == 0:
            continue
        assert torch.equal(hb[b, :m], ref.boxes[b, :m].cpu()) and torch.equal(hs[b, :m], ref.scores[b, :m].cpu())
        assert torch.equal(hc[b, :m], ref.class_ids[b, :m].cpu()) and torch.equal(ha[b], ref.masks_arg[b].cpu())
        assert torch.equal(hm[b, : hgt * wid * m], ref.masks_raw[b, : hgt * wid * m].cpu())


# ------------------------------------------------------------------------------- class map
@pytest.mark.parametrize("name", ["yolact_tiny_a", "yolact_tiny_b", "yolact_tiny_one"])
def test_class_map_tiny_vs_golden(bu, name):
    g = load_golden(name)
    res = bu.decode_nms_class_map((dev(g["box"])[None], dev(g["conf"])[None], dev(g["coef"])[None], dev(g["proto"])[None]),
                                  dev(g["anchors"]), float(g["confidence"]), float(g["nms_iou"]), g["image_shape"])
    boxes, scores, cls, cmap = res
    assert cmap.dtype == torch.uint8 and tuple(cmap.shape) == tuple(int(v) for v in g["image_shape"])
    assert np.array_equal(host(scores), g["scores"]) and np.array_equal(host(cls), g["class_ids"])
    assert_close_f32(host(boxes), g["boxes"], 1e-5, "boxes")
    expect = O.masks_class(unpack_masks(g), g["class_ids"], g["masks_arg"])
    assert float(np.mean(host(cmap).astype(np.int64) != expect)) <= 1e-3


def test_class_map_none(bu):
    g = load_golden("yolact_tiny_none")
    res = bu.decode_nms_class_map((dev(g["box"])[None], dev(g["conf"])[None], dev(g["coef"])[None], dev(g["proto"])[None]),
                                  dev(g["anchors"]), float(g["confidence"]), float(g["nms_iou"]), g["image_shape"])
    assert res == (None,) * 4


def test_class_map_batch_matches_full_outputs_and_host_entry(bu):
    """The class map must be exactly what yolact.py:167-169 computes from this library's own decode_nms outputs, the
    empty image gets zeros, and the host-buffer entry point returns the same bytes."""
    import ctypes as C
    from isb import _lib
    a, p = synth.yolact_dims(550)
    anchors = torch.from_numpy(O.get_anchors((550, 550))).float()
    bsz = 4
    box, conf, coef, proto = synth.yolact_batch(2100, bsz, a, num_classes=81, proto=p)
    conf[2] = conf[2] * 0.0 + 0.001                        # image 2: nothing passes
    ins = (box.cuda(), conf.cuda(), coef.cuda(), proto.cuda())
    full = bu.decode_nms_batch(ins, anchors.cuda(), 0.05, 0.5, (550, 550))
    boxes, scores, cls, cmap, counts = bu.decode_nms_class_map_batch(ins, anchors.cuda(), 0.05, 0.5, (550, 550))
    assert np.array_equal(host(counts), host(full.counts)) and host(counts)[2] == 0
    for b in range(bsz):
        m = int(host(counts)[b])
        if m == 0:
            assert not host(cmap[b]).any()
            continue
        fb, fs, fc, farg, fmask = full.image(b, m)
        assert torch.equal(boxes[b, :m], fb) and torch.equal(scores[b, :m], fs) and torch.equal(cls[b, :m], fc)
        expect = O.masks_class(host(fmask), host(fc), host(farg))
        assert np.array_equal(host(cmap[b]).astype(np.int64), expect)
    # host-buffer entry
    h = _lib.get_handle(0)
    pins = [t.contiguous().pin_memory() for t in (box, conf, coef, proto)]
    panc = anchors.pin_memory()
    hb, hs = torch.empty((bsz, 100, 4)).pin_memory(), torch.empty((bsz, 100)).pin_memory()
    hc = torch.empty((bsz, 100), dtype=torch.int64).pin_memory()
    hm = torch.empty((bsz, 550, 550), dtype=torch.uint8).pin_memory()
    hn = torch.empty((bsz,), dtype=torch.int32).pin_memory()
    pin = _lib.YolactIn(bsz, a, 81, 32, p, p, 550, 550, pins[0].data_ptr(), pins[1].data_ptr(), pins[2].data_ptr(),
                        pins[3].data_ptr(), panc.data_ptr(), 0.05, 0.5, 200, 100)
    pout = _lib.YolactMapOut(hb.data_ptr(), hs.data_ptr(), hc.data_ptr(), hm.data_ptr(), hn.data_ptr())
    h.check(h.lib.isb_yolact_class_map_host(h.ptr, C.byref(pin), C.byref(pout), None))
    assert np.array_equal(hn.numpy(), host(counts)) and np.array_equal(hm.numpy(), host(cmap))
    for b in range(bsz):
        m = int(hn[b])
        assert np.array_equal(hs[b, :m].numpy(), host(scores[b, :m])) and np.array_equal(hc[b, :m].numpy(), host(cls[b, :m]))


# ------------------------------------------------------------------------------- detect from logits
def _check_logits(g, res):
    """Like _check, but the scores went through a different (fused) softmax: they agree to 2e-6 relative, not bit for bit,
    so detections whose scores are that close (saturated scores near 1.0 are one ulp apart) may swap places.  Match the
    detections by (class, box), require the matched scores to agree and the order to be score-descending, and compare
    masks / arg-max through the permutation."""
    boxes, scores, cls, arg, masks = [host(t) for t in res]
    m = len(g["scores"])
    assert len(scores) == m
    perm = np.full(m, -1)
    used = np.zeros(m, bool)
    for i in range(m):
        tol = 1e-5 * np.maximum(np.abs(g["boxes"][i]), 1.0)
        cand = [j for j in range(m) if not used[j] and cls[j] == g["class_ids"][i] and (np.abs(boxes[j] - g["boxes"][i]) <= tol).all()]
        assert cand, f"golden detection {i} has no counterpart"
        j = min(cand, key=lambda j_: abs(j_ - i))
        perm[i], used[j] = j, True
    np.testing.assert_allclose(scores[perm], g["scores"], rtol=2e-6, atol=0)
    assert (np.diff(scores) <= 0).all()
    moved = np.nonzero(perm != np.arange(m))[0]
    assert all(abs(scores[perm[i]] - scores[i]) <= 4e-6 * scores[i] for i in moved), "only near-ties may swap"
    assert_masks_close(masks[:, :, perm], unpack_masks(g), 1e-3, "masks")
    inv = np.empty(m, np.int64)
    inv[perm] = np.arange(m)
    assert float(np.mean(inv[arg] != g["masks_arg"])) <= 1e-3


@pytest.mark.parametrize("name", ["yolact_550_coco", "yolact_550_coco_sparse", "yolact_544_voc_1080p"])
def test_detect_from_logits_vs_golden(bu, name):
    """Raw logits in, softmax (nets/yolact.py:204) fused into the score filter: same detections as the reference run on
    softmax(logits)."""
    g = load_golden(name)
    size, c = int(g["input_size"]), int(g["num_classes"])
    a, p = synth.yolact_dims(size)
    box, logits, coef, proto = synth.yolact_heads(int(g["seed"]), a, num_classes=c, proto=p, objects=int(g["objects"]),
                                                  return_logits=True)
    anchors = torch.from_numpy(O.get_anchors((size, size))).float()
    res = bu.decode_nms_from_logits((box[None].cuda(), logits[None].cuda(), coef[None].cuda(), proto[None].cuda()), anchors.cuda(),
                                    float(g["confidence"]), float(g["nms_iou"]), g["image_shape"])
    _check_logits(g, res)


@pytest.mark.parametrize("a,c,bsz", [(601, 21, 5), (130, 7, 3), (19248, 81, 3)])
def test_detect_from_logits_matches_softmax_path(bu, a, c, bsz):
    """Odd sizes: rows that are not a multiple of the 128-row tile, images whose byte offset is not 16-byte aligned,
    a tensor whose last floats lie past the last aligned boundary.  Candidate sets and detections must equal the
    softmax-input path run on torch.softmax(logits)."""
    g = torch.Generator().manual_seed(77 + a)
    per = [synth.yolact_heads(900 + i, a, num_classes=c, proto=40, objects=4, run=20, return_logits=True) for i in range(bsz)]
    box, logits, coef, proto = (torch.stack([p_[k] for p_ in per]).contiguous().cuda() for k in range(4))
    anchors = torch.rand(a, 4, generator=g).mul(0.5).add(0.1).cuda()
    conf = torch.softmax(logits, dim=-1)
    ref = bu.decode_nms_batch((box, conf, coef, proto), anchors, 0.05, 0.5, (120, 90))
    got = bu.decode_nms_batch((box, logits, coef, proto), anchors, 0.05, 0.5, (120, 90), logits=True)
    assert np.array_equal(host(ref.counts), host(got.counts)) and (host(got.counts) > 0).all()
    for b in range(bsz):
        m = int(host(ref.counts)[b])
        rb, rs, rc, rarg, rmask = ref.image(b, m)
        gb, gs, gc, garg, gmask = got.image(b, m)
        assert torch.equal(rc, gc)
        np.testing.assert_allclose(host(gs), host(rs), rtol=2e-6, atol=0)
        assert torch.equal(rb, gb)
        assert float((rmask != gmask).float().mean()) <= 1e-3


@pytest.mark.parametrize("a,c,scale", [(333, 81, 2.0), (333, 21, 3.0), (210, 100, 1.0), (210, 100, 3.0), (97, 84, 2.5)])
def test_detect_from_logits_undecided_heads_and_wide_rows(bu, a, c, scale):
    """Heads without a dominant background: most rows pass the max-only pre-test, so the register-resident lane-group
    pass of the score filter runs instead of the warp-wide one; 100 classes take the shared-memory path for rows wider
    than the register quarter rows.  Same candidates and detections as softmax-then-filter."""
    g = torch.Generator().manual_seed(1234 + a + c)
    bsz = 3
    logits = (torch.randn(bsz, a, c, generator=g) * scale).cuda()
    box = (torch.randn(bsz, a, 4, generator=g) * 0.3).cuda()
    coef = torch.tanh(torch.randn(bsz, a, 32, generator=g)).cuda()
    proto = torch.relu(torch.randn(bsz, 40, 40, 32, generator=g)).cuda()
    anchors = torch.rand(a, 4, generator=g).mul(0.5).add(0.1).cuda()
    conf = torch.softmax(logits, dim=-1)
    for thr in (0.05, 0.3):
        ref = bu.decode_nms_batch((box, conf, coef, proto), anchors, thr, 0.5, (96, 80))
        got = bu.decode_nms_batch((box, logits, coef, proto), anchors, thr, 0.5, (96, 80), logits=True)
        assert np.array_equal(host(ref.counts), host(got.counts))
        for b in range(bsz):
            m = int(host(ref.counts)[b])
            if m == 0:  # 100 undecided classes at unit scale: no probability reaches the threshold, in either path
                continue
            rb, rs, rc, rarg, rmask = ref.image(b, m)
            gb, gs, gc, garg, gmask = got.image(b, m)
            assert torch.equal(rc, gc)
            np.testing.assert_allclose(host(gs), host(rs), rtol=2e-6, atol=0)
            assert torch.equal(rb, gb)
    assert int(host(got.counts).sum()) > 0 or scale == 1.0


# ------------------------------------------------------------------------------- COCO RLE
def test_rle_encode_vs_oracle(bu):
    """utils_map.py:33-34 on the device: counts and strings equal the restated cocoapi algorithm, and decode back to the
    masks (round trip) -- on real decode_nms output, on degenerate masks and on a busy random one (capacity regrowth)."""
    from oracle import rle_np as RL
    g = load_golden("yolact_550_coco_sparse")
    (box, conf, coef, proto), anchors = _full_inputs(g)
    res = bu.decode_nms((box[None].cuda(), conf[None].cuda(), coef[None].cuda(), proto[None].cuda()), anchors.cuda(),
                        float(g["confidence"]), float(g["nms_iou"]), g["image_shape"])
    masks = res[4]
    hm = host(masks)
    enc = bu.rle_encode(masks)
    assert len(enc) == hm.shape[2]
    for j, e in enumerate(enc):
        assert e["size"] == [hm.shape[0], hm.shape[1]]
        assert e["counts"] == RL.rle_to_string(RL.rle_counts(hm[:, :, j]))
        assert np.array_equal(RL.rle_decode(RL.rle_from_string(e["counts"]), hm.shape[0], hm.shape[1]), hm[:, :, j])
    rng = np.random.default_rng(5)
    odd = np.zeros((37, 53, 5), bool)
    odd[:, :, 1] = True                                  # all ones -> [0, HW]
    odd[:, :, 2] = rng.random((37, 53)) > 0.5            # salt and pepper: ~HW/2 runs, forces the capacity to grow
    odd[0, 0, 3] = True                                  # first pixel set -> leading zero-length run
    odd[-1, -1, 4] = True                                # last pixel
    got = bu.rle_counts(torch.from_numpy(odd).cuda())
    for j in range(5):
        assert got[j].tolist() == RL.rle_counts(odd[:, :, j]).tolist()
    assert got[0].tolist() == [37 * 53] and got[1].tolist() == [0, 37 * 53]


def test_rle_encode_batch_with_device_counts(bu):
    """The batched entry point on decode_nms_batch's slots and device-side counts (ragged m per image, one empty image)."""
    import ctypes as C
    from isb import _lib
    from oracle import rle_np as RL
    a, p = synth.yolact_dims(550)
    anchors = torch.from_numpy(O.get_anchors((550, 550))).float()
    box, conf, coef, proto = synth.yolact_batch(2300, 3, a, num_classes=81, proto=p, objects=2)
    conf[1] = conf[1] * 0.0 + 0.001
    res = bu.decode_nms_batch((box.cuda(), conf.cuda(), coef.cuda(), proto.cuda()), anchors.cuda(), 0.3, 0.5, (550, 550), max_detections=40)
    counts = host(res.counts)
    assert counts[1] == 0 and counts[0] > 0 and counts[2] > 0
    bsz, d, cap = 3, 40, 1 << 20
    out = torch.empty((cap,), dtype=torch.int32, device="cuda")
    rc = torch.empty((bsz, d), dtype=torch.int32, device="cuda")
    ro = torch.empty((bsz, d), dtype=torch.int64, device="cuda")
    tot = torch.empty((2,), dtype=torch.int64, device="cuda")
    h = _lib.get_handle(0)
    h.check(h.lib.isb_rle_encode(h.ptr, res.masks_raw.data_ptr(), res.masks_raw.shape[1], res.counts.data_ptr(), -1, d, bsz, 550, 550,
                                 out.data_ptr(), cap, rc.data_ptr(), ro.data_ptr(), tot.data_ptr(), None))
    torch.cuda.synchronize()
    flat, rc, ro = host(out).view(np.uint32), host(rc), host(ro)
    assert int(host(tot)[1]) == int(rc.sum()) <= cap
    for b in range(bsz):
        m = int(counts[b])
        assert (rc[b, m:] == 0).all()
        if m == 0:
            continue
        masks = host(res.image(b, m)[4])
        for j in range(m):
            got = flat[ro[b, j]:ro[b, j] + rc[b, j]]
            assert got.tolist() == RL.rle_counts(masks[:, :, j]).tolist()


def test_assemble_masks_verseg_many_detections(bu):
    """configs[4] shape of the problem at a size the oracle finishes quickly: ~1000 small overlapping detections (the compose
    kernel then runs with a long detection list and few warps) on full-size prototypes."""
    box, score, coef = synth.verseg_boxes(5003, 1024)
    gen = torch.Generator().manual_seed(5004)
    proto = torch.relu(torch.randn(138, 138, 32, generator=gen))
    shape = (160, 130)
    boxes = O.yolact_correct_boxes(box.numpy(), shape)
    arg_ref, m_ref = O.assemble_masks(proto.numpy(), coef.numpy(), boxes, shape)
    arg, mk = bu.assemble_masks(proto.cuda(), coef.cuda(), dev(boxes), shape)
    assert tuple(mk.shape) == (160, 130, 1024)
    assert_masks_close(host(mk), m_ref, 1e-3, "masks")
    assert float(np.mean(host(arg) != arg_ref)) <= 2e-3


@pytest.mark.parametrize("n,c,top_k,max_det", [(5000, 1, 5000, 1024), (3000, 4, 200, 100), (40000, 1, 300, 64)])
def test_fast_nms_wide_path_vs_oracle(bu, n, c, top_k, max_det):
    """More candidates than the single-CTA kernel holds in shared memory: the multi-CTA path (rank-sort variant up to 32768
    keys, bitonic above) must give the oracle's detections bit for bit."""
    g = torch.Generator().manual_seed(n + c)
    ctr = torch.rand(n, 2, generator=g)
    wh = torch.rand(n, 2, generator=g) * 0.05 + 0.01
    box = torch.cat([ctr - wh / 2, ctr + wh / 2], 1).contiguous()
    score = torch.rand(n, c, generator=g).contiguous()
    coef = torch.tanh(torch.randn(n, 32, generator=g))
    rb, rs, rc_, rk = O.fast_non_max_suppression(box.numpy(), score.numpy(), coef.numpy(), 0.5, top_k, max_det)
    b, s, cls, k = bu.fast_non_max_suppression(box.cuda(), score.cuda(), coef.cuda(), 0.5, top_k, max_det)
    assert np.array_equal(host(s), rs) and np.array_equal(host(cls), rc_)
    assert np.array_equal(host(b), rb) and np.array_equal(host(k), rk)


@pytest.mark.parametrize("n,c,top_k,max_det,path", [(40, 80, 200, 100, "rank"), (60, 30, 20, 50, "rank, lists longer than top_k"),
                                                     (1000, 4, 200, 100, "bitonic"), (300, 10, 50, 20, "bitonic"),
                                                     (4096, 1, 100, 100, "bitonic, full buffer"), (7, 3, 200, 100, "tiny")])
def test_fast_nms_single_cta_sort_paths_vs_oracle(bu, n, c, top_k, max_det, path):
    """Inside one CTA the class-major order comes either from a counting sort by class + in-class ranking (many short
    lists) or from the bitonic network (few long lists); both, with lists shorter and longer than top_k and heavily
    overlapping boxes, must give the oracle's detections bit for bit."""
    g = torch.Generator().manual_seed(7 * n + c)
    ctr = torch.rand(n, 2, generator=g)
    wh = torch.rand(n, 2, generator=g) * 0.2 + 0.1
    box = torch.cat([ctr - wh / 2, ctr + wh / 2], 1).contiguous()
    score = torch.rand(n, c, generator=g).contiguous()
    coef = torch.tanh(torch.randn(n, 32, generator=g))
    rb, rs, rc_, rk = O.fast_non_max_suppression(box.numpy(), score.numpy(), coef.numpy(), 0.5, top_k, max_det)
    b, s, cls, k = bu.fast_non_max_suppression(box.cuda(), score.cuda(), coef.cuda(), 0.5, top_k, max_det)
    assert np.array_equal(host(s), rs) and np.array_equal(host(cls), rc_), path
    assert np.array_equal(host(b), rb) and np.array_equal(host(k), rk), path


# ------------------------------------------------------------------------------- error convention
def test_error_convention(bu):
    """int status < 0 + isb_last_error text, nothing launched, handle usable afterwards; "no detections" is a value."""
    import ctypes as C
    from isb import _lib
    h = _lib.Handle(0, max_batch=2, max_anchors=64, max_classes=4, max_det=8, max_top_k=16, cand_capacity=256)
    try:
        f = torch.zeros(64, device="cuda")
        P = lambda t: C.c_void_p(t.data_ptr())
        box, conf, coef = torch.zeros(1, 64, 4, device="cuda"), torch.zeros(1, 64, 5, device="cuda"), torch.zeros(1, 64, 32, device="cuda")
        proto, anchors = torch.zeros(1, 8, 8, 32, device="cuda"), torch.rand(64, 4, device="cuda")
        ob, os_, oc = torch.empty(1, 8, 4, device="cuda"), torch.empty(1, 8, device="cuda"), torch.empty(1, 8, dtype=torch.int64, device="cuda")
        arg = torch.empty(1, 16, 16, dtype=torch.int64, device="cuda")
        slot = 16 * 16 * 8
        masks, cnt = torch.empty(1, slot, dtype=torch.uint8, device="cuda"), torch.full((1,), 7, dtype=torch.int32, device="cuda")

        def call(**kw):
            a = dict(batch=1, a=64, c=5, k=32, ph=8, pw=8, oh=16, ow=16, box=P(box), conf=P(conf), coef=P(coef), proto=P(proto),
                     anchors=P(anchors), thr=0.5, iou=0.5, top_k=16, max_det=8, masks=P(masks), slot=slot)
            a.update(kw)
            pin = _lib.YolactIn(a["batch"], a["a"], a["c"], a["k"], a["ph"], a["pw"], a["oh"], a["ow"], a["box"], a["conf"], a["coef"],
                                a["proto"], a["anchors"], a["thr"], a["iou"], a["top_k"], a["max_det"], 0)
            pout = _lib.YolactOut(P(ob), P(os_), P(oc), P(arg), a["masks"], a["slot"], P(cnt))
            return h.lib.isb_yolact_detect(h.ptr, C.byref(pin), C.byref(pout), None)

        n0 = h.launch_count
        cases = [(dict(box=None), _lib.ISB_ERR_BAD_ARG if hasattr(_lib, "ISB_ERR_BAD_ARG") else -1),      # null input
                 (dict(k=16), -1),                                                                             # wrong coefficient count
                 (dict(slot=slot - 16), -1),                                                                   # slot too small
                 (dict(proto=C.c_void_p(proto.data_ptr() + 4)), -1),                                           # misaligned
                 (dict(batch=3), -3), (dict(a=65), -3), (dict(c=6), -3), (dict(max_det=9), -3), (dict(top_k=17), -3)]
        for kw, want in cases:
            rc = call(**kw)
            assert rc == want, (kw, rc)
            assert len(h.lib.isb_last_error(h.ptr)) > 0
        assert h.launch_count == n0, "a rejected call must not launch anything"
        # all-zero scores: a valid call, zero detections, status OK
        assert call() == 0
        torch.cuda.synchronize()
        assert int(cnt[0]) == 0
        # stage entry points
        assert h.lib.isb_decode_boxes(h.ptr, None, P(f), 4, 0.1, 0.2, P(f), None) == -1
        assert h.lib.isb_fast_nms(h.ptr, P(f), P(f), P(f), 4, 1, 16, 0.5, 16, 8, P(f), P(f), P(f), P(f), P(cnt), None) == -1
        assert h.lib.isb_nms(h.ptr, None, P(f), 4, 0.5, P(f), P(cnt), None) < 0
    finally:
        h.close()


# ------------------------------------------------------------------------------- VERDICT r1 weak 1
def test_nan_inf_fast_nms_and_jaccard_vs_reference(bu):
    """NaN / +-inf boxes and scores: the reference's Fast NMS propagates NaN through torch.min / max / clamp and the column
    max, so a NaN IoU suppresses, and torch.sort ranks a NaN score first (golden from utils_bbox.py:27-99 itself)."""
    g = load_golden("nan_inf_cases")
    assert np.array_equal(host(bu.jaccard(dev(g["f_box"][:12]), dev(g["f_box"][:36]))), g["f_iou"], equal_nan=True)
    for tag in ("a", "b"):
        thr, top_k, max_det = g[f"f{tag}_params"]
        b, s, c, k = bu.fast_non_max_suppression(dev(g["f_box"]), dev(g["f_scores"]), dev(g["f_coef"]), float(thr), int(top_k),
                                                 int(max_det))
        assert np.array_equal(host(s), g[f"f{tag}_score"], equal_nan=True), tag
        assert np.array_equal(host(c), g[f"f{tag}_cls"])
        assert np.array_equal(host(b), g[f"f{tag}_box"], equal_nan=True) and np.array_equal(host(k), g[f"f{tag}_coef"])


def test_verseg_full_size_1024_masks_at_550(bu):
    """BASELINE configs[4] at FULL size: 1024 detections assembled at 550 x 550 (310 MB of bool, the compose kernel's
    large-m branch) against the oracle, evaluated in chunks of 64 detections with a running max / first arg-max so that
    the [H,W,m] float tensors of the reference (1.2 GB each) never exist.  <= 0.1 % pixels, arg-max <= 0.1 %."""
    box, score, coef = synth.verseg_boxes(5005, 1024)
    gen = torch.Generator().manual_seed(5006)
    proto = torch.relu(torch.randn(138, 138, 32, generator=gen))
    shape = (550, 550)
    boxes = O.yolact_correct_boxes(box.numpy(), shape)
    arg, mk = bu.assemble_masks(proto.cuda(), coef.cuda(), dev(boxes), shape)
    assert tuple(mk.shape) == (550, 550, 1024) and mk.is_contiguous()
    mk_h, arg_h = host(mk), host(arg)
    best = np.zeros(shape, np.float32)
    ref_arg = np.zeros(shape, np.int64)
    bad = 0
    for j0 in range(0, 1024, 64):
        s = O.sigmoid(O.mask_logits(proto.numpy(), coef.numpy()[j0:j0 + 64]))
        up = O.upsample_bilinear(np.ascontiguousarray(np.transpose(s, (2, 0, 1))), *shape)
        up = O.crop(np.ascontiguousarray(np.transpose(up, (1, 2, 0))), boxes[j0:j0 + 64])
        bad += int(np.count_nonzero(mk_h[:, :, j0:j0 + 64] != (up > np.float32(0.5))))
        cmax, carg = up.max(axis=-1), up.argmax(axis=-1)
        better = cmax > best                      # strict: the first maximum wins across chunks, as torch.argmax
        ref_arg[better] = carg[better] + j0
        best[better] = cmax[better]
    assert bad <= 1e-3 * mk_h.size, f"{bad} of {mk_h.size} mask pixels differ"
    assert float(np.mean(arg_h != ref_arg)) <= 1e-3
    # every set pixel lies inside its detection's crop box; columns / rows outside all boxes are empty
    ys, xs, js = np.nonzero(mk_h[::3, ::3])
    assert (xs * 3 >= boxes[js, 0]).all() and (xs * 3 < boxes[js, 2]).all() and (ys * 3 >= boxes[js, 1]).all() and (ys * 3 < boxes[js, 3]).all()


@pytest.mark.parametrize("max_det,grow_first", [(50, False), (50, True), (100, True)])
def test_class_map_batch_with_other_max_detections(bu, max_det, grow_first):
    """the class map read class ids with the WORKSPACE stride (the handle's max_det limit) while Fast
    NMS wrote them with the call's max_detections -- wrong classes for every image b >= 1 whenever the two differ (a call
    with max_detections=50 on the default handle, or any call after the shared handle grew to 1024 for VerSeg)."""
    from isb import _lib
    if grow_first:
        _lib.get_handle(0, max_det=1024)                 # what a VerSeg call leaves behind
    a, p = synth.yolact_dims(550)
    anchors = torch.from_numpy(O.get_anchors((550, 550))).float().cuda()
    box, conf, coef, proto = synth.yolact_batch(2400, 3, a, num_classes=81, proto=p)
    ins = (box.cuda(), conf.cuda(), coef.cuda(), proto.cuda())
    full = bu.decode_nms_batch(ins, anchors, 0.05, 0.5, (550, 550), max_detections=max_det)
    boxes, scores, cls, cmap, counts = bu.decode_nms_class_map_batch(ins, anchors, 0.05, 0.5, (550, 550), max_detections=max_det)
    assert np.array_equal(host(counts), host(full.counts)) and (host(counts) == max_det).all()
    assert _lib.get_handle(0).limits.max_det != max_det or not grow_first
    for b in range(3):
        fb, fs, fc, farg, fmask = full.image(b, max_det)
        assert torch.equal(cls[b, :max_det], fc)
        expect = O.masks_class(host(fmask), host(fc), host(farg))
        assert np.array_equal(host(cmap[b]).astype(np.int64), expect), f"image {b}"


def test_batch_result_overflow_raises(bu, monkeypatch):
    """counts = -1 (candidate buffer overflow) must not read as "no detections" in the batched API."""
    from isb import _lib
    for hd in _lib._handles.values():
        hd.close()
    _lib._handles.clear()
    monkeypatch.setitem(_lib._DEFAULTS, "cand_capacity", 1024)
    g = load_golden("yolact_tiny_a")
    res = bu.decode_nms_batch((dev(g["box"])[None], dev(g["conf"])[None], dev(g["coef"])[None], dev(g["proto"])[None]),
                              dev(g["anchors"]), 0.0, 0.5, g["image_shape"])
    assert int(res.counts[0]) == -1
    with pytest.raises(_lib.IsbError):
        res.image(0)
    for hd in _lib._handles.values():
        hd.close()
    _lib._handles.clear()


@pytest.mark.parametrize("threads", [0, 3, -1])
def test_detect_host_full_size_matches_device(bu, threads):
    """isb_yolact_detect_host at the benchmark's image size (550 x 550, up to 100 detections) with the
    hybrid delivery -- dense copies by the copy engine and bit-packed copies expanded by host threads, split at run time --
    must leave the caller's host buffers byte-identical to the device path: every mask byte, every image, whichever way
    it travelled.  threads = 0 is the copy engine alone, 3 a small pool, -1 the default."""
    import ctypes as C
    from isb import _lib
    from isb import anchors as A
    anchors = torch.from_numpy(A.get_anchors([550, 550])).float()
    a, p = anchors.shape[0], 138
    bsz, d, hgt, wid = 11, 100, 550, 550
    box, conf, coef, proto = synth.yolact_batch(2600, bsz, a, num_classes=81, proto=p)
    conf[4] = conf[4] * 0.0 + 0.001                      # an image without detections in the middle of a chunk
    conf[7, :, 1:] *= (torch.rand(a, 1) < 0.02)           # an image with few detections (ragged m)
    ref = bu.decode_nms_batch((box.cuda(), conf.cuda(), coef.cuda(), proto.cuda()), anchors.cuda(), 0.05, 0.5, (hgt, wid))
    torch.cuda.synchronize()
    counts = ref.counts.cpu()
    assert int(counts[4]) == 0 and 0 < int(counts[7]) < 100 and int(counts[0]) == 100
    slot = (hgt * wid * d + 15) // 16 * 16
    pin = [t.contiguous().pin_memory() for t in (box, conf, coef, proto, anchors)]
    hb, hs = torch.zeros((bsz, d, 4)).pin_memory(), torch.zeros((bsz, d)).pin_memory()
    hc = torch.zeros((bsz, d), dtype=torch.int64).pin_memory()
    ha = torch.zeros((bsz, hgt, wid), dtype=torch.int64).pin_memory()
    hm = torch.full((bsz, slot), 7, dtype=torch.uint8).pin_memory()      # poisoned: every byte must be overwritten
    hn = torch.zeros((bsz,), dtype=torch.int32).pin_memory()
    h = _lib.get_handle(0, max_batch=bsz, max_anchors=a, max_classes=80)
    h.set_host_threads(threads)
    try:
        yin = _lib.YolactIn(bsz, a, 81, 32, p, p, hgt, wid, pin[0].data_ptr(), pin[1].data_ptr(), pin[2].data_ptr(),
                            pin[3].data_ptr(), pin[4].data_ptr(), 0.05, 0.5, 200, d, 0)
        yout = _lib.YolactOut(hb.data_ptr(), hs.data_ptr(), hc.data_ptr(), ha.data_ptr(), hm.data_ptr(), slot, hn.data_ptr())
        for rep in range(2):                                              # second call: recycled slots / events / jobs
            hm.fill_(7)
            h.check(h.lib.isb_yolact_detect_host(h.ptr, C.byref(yin), C.byref(yout), C.c_void_p(torch.cuda.current_stream().cuda_stream)))
            assert torch.equal(hn, counts)
            for b in range(bsz):
                m = int(counts[b])
                if m == 0:
                    continue
                assert torch.equal(hb[b, :m], ref.boxes[b, :m].cpu()) and torch.equal(hs[b, :m], ref.scores[b, :m].cpu())
                assert torch.equal(hc[b, :m], ref.class_ids[b, :m].cpu()) and torch.equal(ha[b], ref.masks_arg[b].cpu())
                assert torch.equal(hm[b, : hgt * wid * m], ref.masks_raw[b, : hgt * wid * m].cpu()), f"masks of image {b} (call {rep})"
                assert bool((hm[b, hgt * wid * m:] == 7).all()), "bytes past H*W*m of a slot must not be touched"
            st = h.host_stats()
            assert st["images_copy_engine"] + st["images_host_threads"] == int((counts > 0).sum())
            if threads == 0:
                assert st["images_host_threads"] == 0 and st["host_threads"] == 0
            if threads == 3:
                assert st["host_threads"] == 3
            dense_bytes = hgt * wid * int(counts.sum())
            assert st["d2h_bytes"] <= dense_bytes + bsz * (4 + d * 28 + hgt * wid * 8)
    finally:
        h.set_host_threads(-1)


def test_fused_count_exchange_single_rank(bu):
    """isb_counts_* (the multi-GPU count exchange fused into fast_nms_kernel as peer stores) with world = 1: the rank's own
    ring is its only peer.  Every publishing detect call advances the epoch; the wait kernel returns the counts and their
    exclusive scan; an epoch that left the ring is refused; the wait is bounded when a word never arrives."""
    import ctypes as C
    from isb import _lib
    g = load_golden("yolact_tiny_a")
    bsz = 3
    ins = tuple(dev(g[k])[None].repeat((bsz,) + (1,) * g[k].ndim).contiguous() for k in ("box", "conf", "coef", "proto"))
    ins[1][1] *= 0.0                                   # image 1: no detections
    anchors = dev(g["anchors"])
    h = _lib.get_handle(0, max_batch=bsz, max_anchors=g["conf"].shape[0], max_classes=g["conf"].shape[1] - 1)
    mine = C.create_string_buffer(64)
    h.check(h.lib.isb_counts_export(h.ptr, bsz, 1, 2, mine))
    h.check(h.lib.isb_counts_attach(h.ptr, 0, mine.raw))
    try:
        counts = torch.full((bsz,), -7, dtype=torch.int32, device="cuda")
        offsets = torch.zeros((bsz,), dtype=torch.int64, device="cuda")
        status = torch.full((1,), -1, dtype=torch.int32, device="cuda")
        P = lambda t: C.c_void_p(t.data_ptr())
        for call in range(1, 4):
            res = bu.decode_nms_batch(ins, anchors, float(g["confidence"]), float(g["nms_iou"]), g["image_shape"])
            assert int(h.lib.isb_counts_epoch(h.ptr)) == call
            h.check(h.lib.isb_counts_wait(h.ptr, -1, P(counts), P(offsets), P(status), 500, None))
            torch.cuda.synchronize()
            assert int(status[0]) == 0 and torch.equal(counts, res.counts)
            c = res.counts.cpu().to(torch.int64).clamp(min=0)
            assert torch.equal(offsets.cpu(), torch.cumsum(c, 0) - c) and int(counts[1]) == 0 and int(counts[0]) > 0
        # a batch of another size does not publish
        bu.decode_nms_batch(tuple(t[:2] for t in ins), anchors, float(g["confidence"]), float(g["nms_iou"]), g["image_shape"])
        assert int(h.lib.isb_counts_epoch(h.ptr)) == 3
        assert h.lib.isb_counts_wait(h.ptr, 1, P(counts), P(offsets), P(status), 500, None) == _lib.ISB_ERR_LIMIT   # depth 2
        assert h.lib.isb_counts_wait(h.ptr, 9, P(counts), P(offsets), P(status), 500, None) == _lib.ISB_ERR_BAD_ARG
    finally:
        h.check(h.lib.isb_counts_detach(h.ptr))
    assert h.lib.isb_counts_wait(h.ptr, -1, P(counts), P(offsets), P(status), 500, None) == _lib.ISB_ERR_BAD_ARG


@pytest.mark.parametrize("m,bsz", [(100, 5), (37, 2), (128, 1), (8, 3)])
def test_mask_logits_batch_tensor_cores(bu, m, bsz):
    """isb_mask_logits_batch (batched tcgen05 + TMEM kernel, bulk-TMA prototype tiles): within 1e-5 of the float64 product
    for every picture, including the partial last tile (138 * 138 = 148 * 128 + 100 pixels), a detection count that is not
    a multiple of 4 or 16, and the sigmoid epilogue (utils_bbox.py:223)."""
    gen = torch.Generator().manual_seed(30 + m)
    proto = torch.relu(torch.randn(bsz, 138, 138, 32, generator=gen))
    coef = torch.tanh(torch.randn(bsz, m, 32, generator=gen))
    ref = torch.matmul(proto.double(), coef.double().transpose(1, 2)[:, None])
    out = bu.mask_logits_batch(proto.cuda(), coef.cuda())
    assert tuple(out.shape) == (bsz, 138, 138, m)
    assert_close_f32(host(out), ref.numpy(), 1e-5, "batched mask logits")
    sg = bu.mask_logits_batch(proto.cuda(), coef.cuda(), sigmoid=True)
    assert_close_f32(host(sg), torch.sigmoid(ref).numpy(), 1e-5, "batched sigmoid(mask logits)")
    # the per-picture stage kernel agrees
    one = bu.mask_logits(proto[0].cuda(), coef[0].cuda())
    assert_close_f32(host(one), ref[0].numpy(), 1e-5, "single-picture mask logits")


def test_tensor_core_sigmoid_maps_on_the_fused_path():
    """ISB_TC_SMAP=1 swaps K3a (FFMA, footprints only) for the batched tcgen05 product scattering into the same packed
    footprint maps: the fused path must still reproduce the reference goldens (run in a child process: the switch is read
    once per process)."""
    import os
    import subprocess
    import sys
    env = dict(os.environ, ISB_TC_SMAP="1")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    r = subprocess.run([sys.executable, "-m", "pytest", os.path.join(root, "tests", "test_yolact_gpu.py"), "-q", "-x", "-m", "gpu", "-k",
                        "test_full_size_end_to_end or test_batch_equals_single_and_oracle or test_tiny_end_to_end"],
                       env=env, cwd=root, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-2000:]


# ------------------------------------------------------------------ memory-safety stand-in (compute-sanitizer is closed on this pool)
def test_guard_bands_and_repeatability(bu):
    """Every output of the fused call sits between 4 KB bands of a known byte: none may change (out-of-bounds writes), and two
    runs into differently poisoned buffers must agree bit for bit (reads of uninitialised memory, races)."""
    from conftest import BandedAllocator
    from isb.yolact import YolactBatchResult
    a, p = synth.yolact_dims(550)
    anchors = torch.from_numpy(O.get_anchors((550, 550))).float().cuda()
    bsz, max_det, hgt, wid = 5, 100, 550, 550
    box, conf, coef, proto = synth.yolact_batch(2100, bsz, a, num_classes=81, proto=p)
    conf[2] = conf[2] * 0.0 + 0.001                      # an empty image in the middle of the batch
    outs = [t.cuda() for t in (box, conf, coef, proto)]
    slot = (hgt * wid * max_det + 15) // 16 * 16
    runs = []
    for poison in (0x00, 0xFF):
        alloc = BandedAllocator()
        d = "cuda"
        res = YolactBatchResult(alloc.empty((bsz, max_det, 4), dtype=torch.float32, device=d),
                                alloc.empty((bsz, max_det), dtype=torch.float32, device=d),
                                alloc.empty((bsz, max_det), dtype=torch.int64, device=d),
                                alloc.empty((bsz, hgt, wid), dtype=torch.int64, device=d),
                                alloc.empty((bsz, slot), dtype=torch.uint8, device=d),
                                alloc.empty((bsz,), dtype=torch.int32, device=d), (hgt, wid))
        for t in (res.boxes, res.scores, res.class_ids, res.masks_arg, res.masks_raw, res.counts):
            t.view(torch.uint8).fill_(poison)
        bu.decode_nms_batch(outs, anchors, 0.05, 0.5, (hgt, wid), out=res)
        torch.cuda.synchronize()
        alloc.check()
        runs.append(res)
    c0, c1 = host(runs[0].counts), host(runs[1].counts)
    assert np.array_equal(c0, c1) and c0[2] == 0 and (c0[[0, 1, 3, 4]] > 0).all()
    for b in range(bsz):
        if c0[b] == 0:
            continue
        for x, y in zip(runs[0].image(b, int(c0[b])), runs[1].image(b, int(c1[b]))):
            assert torch.equal(x, y), f"image {b}: two runs differ"
