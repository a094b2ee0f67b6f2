"""Generate tests/golden/*.npz by executing the REAL reference code (via oracle/ref_shim.py).

TEST INFRASTRUCTURE ONLY.  Run in the build container, where /root/reference is mounted:

    python -m oracle.make_golden

The fixtures hold OUTPUTS of the reference on seeded synthetic inputs (inputs are
re-derived from the seeds by oracle/synth.py, so the files stay small).  They pin
oracle/restate.py (tests/test_oracle_golden.py, CPU) and the CUDA path (tests -m gpu).
Each file records numpy's version: float32 exp/log last-ulp results and NEP-50 scalar
promotion are properties of the numpy build that produced it.
"""
import hashlib
import os
import platform
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))

from oracle import ref_shim, restate, synth  # noqa: E402

OUT = os.path.join(os.path.dirname(HERE), "tests", "golden")


def sha(a):
    return np.frombuffer(hashlib.sha256(np.ascontiguousarray(a).tobytes()).digest(), dtype=np.uint8)


def meta():
    return dict(numpy_version=np.array(np.__version__), machine=np.array(platform.machine()),
                generator=np.array("oracle/make_golden.py via oracle/ref_shim.py (real reference code)"))


def tuple_str(t):
    return "(" + ", ".join(str(x) for x in t) + ")"


def run_reference_proposal(R, cls, box, info, cfg, train, seed):
    """The reference's own ProposalOperator.forward (rcnn/rpn/proposal.py:39) through the mx stub."""
    key = "TRAIN" if train else "TEST"
    R.config[key].RPN_PRE_NMS_TOP_N = cfg["pre"]
    R.config[key].RPN_POST_NMS_TOP_N = cfg["post"]
    R.config[key].RPN_NMS_THRESH = cfg["thresh"]
    R.config[key].RPN_MIN_SIZE = cfg["min_size"]
    op = R.ProposalOperator(str(cfg["feat_stride"]), tuple_str(cfg["scales"]), tuple_str(cfg["ratios"]),
                            is_train=train, output_score=True)
    post = cfg["post"]
    out = [ref_shim.NDArrayStub(np.zeros((post, 5), np.float32)), ref_shim.NDArrayStub(np.zeros((post, 1), np.float32))]
    np.random.seed(seed)
    op.forward(False, ["write", "write"], [ref_shim.NDArrayStub(cls), ref_shim.NDArrayStub(box),
                                           ref_shim.NDArrayStub(info)], out, [])
    return out[0].arr, out[1].arr


PROPOSAL_CASES = [("tiny", [0, 1, 2, 3], False, None), ("tiny", [4, 5], True, (8, 11)),
                  ("voc_test", [0, 1, 2], False, None), ("voc_alt", [0], False, None),
                  ("wider", [0], False, None),
                  # the reference's TRAIN defaults at full size (rcnn/config.py:47-50: pre 12000, post 6000)
                  ("voc_train", [3, 6], False, None)]  # seeds with a greedy-path IoU margin > 1e-5


def golden_proposal(R, only=None):
    cases = []
    for name, seeds, train, pad in PROPOSAL_CASES:
        if only and name not in only:
            continue
        cfg = synth.CONFIGS[name]
        A = len(cfg["scales"]) * len(cfg["ratios"])
        for s in seeds:
            cls, box, info = synth.proposal_inputs(s, H=cfg["H"], W=cfg["W"], A=A, im_info=cfg["im_info"], pad_hw=pad)
            rois, scores = run_reference_proposal(R, cls, box, info, cfg, train, s)
            # stage values by calling the reference's helpers in the op's own order (proposal.py:110-131)
            st = restate.proposal_stages(cls, box, info, cfg["feat_stride"], cfg["scales"], cfg["ratios"], train,
                                         cfg["pre"], cfg["post"], cfg["thresh"], cfg["min_size"],
                                         tie="stable", rng=np.random.RandomState(s), audit=True)
            ref_dec = R.clip_boxes(R.bbox_pred(st["anchors"], _rows(box, st, A)), info[0, :2])
            cases.append((name, s, train, pad))
            tag = "%s_s%d%s" % (name, s, "_train" if train else "")
            np.savez_compressed(
                os.path.join(OUT, "proposal_%s.npz" % tag), config=np.array(name), seed=s, train=train,
                pad_hw=np.array(pad if pad else (0, 0)), rois=rois, scores=scores,
                decoded_sha256=sha(ref_dec), decoded_head=ref_dec[:64], decoded_tail=ref_dec[-64:],
                keep=st["keep"], keep_anchor_index=st["keep_anchor_index"], nkeep=st["nkeep"],
                nms_keep_full_len=st["nms_keep_full_len"], n_filter=len(st["filter_keep"]),
                margin_filter=st["margin_filter"], margin_nms=st["margin_nms"], nms_pairs=st["nms_pairs"], **meta())
            assert np.array_equal(ref_dec, st["decoded"]), "restatement decode differs from reference helpers"
            assert np.array_equal(rois, st["rois"]) and np.array_equal(scores, st["scores"]), tag
            print("proposal", tag, "nkeep", st["nkeep"], "full", st["nms_keep_full_len"],
                  "margins %.2e %.2e" % (st["margin_filter"], st["margin_nms"]))
    if only:
        return
    # a padded case: few boxes survive NMS, so npr.choice pads (proposal.py:135-139)
    cfg = dict(synth.CONFIGS["tiny"], post=120, thresh=0.3)
    for s in [7, 8]:
        cls, box, info = synth.proposal_inputs(s, H=cfg["H"], W=cfg["W"], A=9, im_info=cfg["im_info"])
        rois, scores = run_reference_proposal(R, cls, box, info, cfg, False, s)
        st = restate.proposal_stages(cls, box, info, 16, cfg["scales"], cfg["ratios"], False, cfg["pre"], cfg["post"],
                                     cfg["thresh"], cfg["min_size"], rng=np.random.RandomState(s), audit=True)
        assert st["nkeep"] < cfg["post"]
        assert np.array_equal(rois, st["rois"]) and np.array_equal(scores, st["scores"])
        np.savez_compressed(os.path.join(OUT, "proposal_pad_s%d.npz" % s), seed=s, post=cfg["post"], thresh=cfg["thresh"],
                            rois=rois, scores=scores, keep=st["keep"], nkeep=st["nkeep"], pad_draw=st["pad_draw"],
                            margin_filter=st["margin_filter"], margin_nms=st["margin_nms"], **meta())
        print("proposal pad", s, "nkeep", st["nkeep"])


def _rows(box, st, A):
    """deltas in (h,w,a) row order for the (possibly cropped) map, as proposal.py:99-100 builds them."""
    n = st["anchors"].shape[0]
    # recover crop from the anchors' extent
    K = n // A
    xs = np.unique(st["anchors"][:, 0] - st["anchors"][0, 0])
    w = len(np.unique(st["anchors"][::A, 0]))
    h = K // w
    d = box[:, :, :h, :w]
    return d.transpose((0, 2, 3, 1)).reshape((-1, 4))


def golden_helpers(R):
    d = {}
    d["anchors_voc"] = R.generate_anchors(base_size=16.0, ratios=[0.5, 1, 2], scales=np.array([8, 16, 32], dtype=float))
    d["anchors_wider"] = R.generate_anchors(base_size=16.0, ratios=[0.5, 1, 2], scales=np.array([4, 8, 16, 32], dtype=float))
    d["anchors_assign"] = R.generate_anchors(base_size=16, ratios=[0.5, 1, 2], scales=np.array((8, 16, 32), dtype=np.float32))
    # bbox_pred / clip_boxes at the detector's shape (rcnn/detector.py:73-74): 300 rois x 21 classes
    rs = np.random.RandomState(11)
    gt = synth.gt_boxes(11, G=8)
    rois = synth.jittered_rois(11, gt, R=300)[:, 1:5]
    deltas = (rs.normal(0, 0.2, size=(300, 84))).astype(np.float32)
    pred = R.bbox_pred(rois, deltas)
    d["bp_pred"] = pred.copy()
    d["bp_clip"] = R.clip_boxes(pred.copy(), (600, 1000))
    d["bp_empty_shape"] = np.array(R.bbox_pred(np.zeros((0, 4)), np.zeros((0, 84), np.float32)).shape)
    # bbox_transform f64 and f32
    ex = rois.astype(np.float64)
    g = gt[rs.randint(0, 8, 300), :4]
    d["bt_f64"] = R.bbox_transform(ex, g.astype(np.float64))
    d["bt_f32"] = R.bbox_transform(rois, g)
    # bbox_overlaps (real python double loop): 300 x 8 and degenerate cases
    d["ov_300x8"] = R.bbox_overlaps(ex, gt.astype(np.float64))
    anc = restate.proposal_stages(*synth.proposal_inputs(0, H=6, W=9, A=9, im_info=(96, 144, 1.0)), pre_nms_topN=200,
                                  post_nms_topN=30)["anchors"]
    gts = synth.gt_boxes(5, G=6, im_h=96, im_w=144)
    d["ov_anchor_gt"] = R.bbox_overlaps(anc, gts.astype(np.float64))
    # expand targets
    data = np.zeros((128, 5), np.float32)
    data[:32, 0] = rs.randint(1, 21, 32)
    data[:, 1:] = rs.normal(0, 1, size=(128, 4))
    t, w = R.expand_bbox_regression_targets(data, 21)
    d["ex_in"], d["ex_t"], d["ex_w"] = data, t, w
    # nms: f64 (tester.py:54-55) and f32 (demo.py:57-58) dets, thresholds 0.3 / 0.7
    for i, (n, dt, th) in enumerate([(300, np.float64, 0.3), (300, np.float32, 0.3), (2000, np.float64, 0.3),
                                     (2000, np.float32, 0.7), (1, np.float64, 0.3), (97, np.float32, 0.5)]):
        dets = synth.class_dets(20 + i, n=n, dtype=dt)
        d["nms%d_keep" % i] = np.asarray(R.nms(dets, th), dtype=np.int64)
        d["nms%d_spec" % i] = np.array([20 + i, n, 8 if dt == np.float64 else 4, th])
    d["nms_empty"] = np.asarray(R.nms(np.zeros((0, 5)), 0.3), dtype=np.int64)
    dets = synth.class_dets(40, n=60, dtype=np.float32)
    d["nest_keep"] = np.asarray(R.nest(dets, 0.9), dtype=np.int64)
    np.savez_compressed(os.path.join(OUT, "helpers.npz"), **d, **meta())
    print("helpers ok")


def golden_anchor_target(R):
    R.config.END2END = 0
    for tag, H, W, im, G, seed in [("tiny_g6", 14, 20, (224, 320, 1.0), 6, 0), ("tiny_g0", 14, 20, (224, 320, 1.0), 0, 1),
                                   ("tiny_g1", 14, 20, (224, 320, 1.0), 1, 2), ("mid_g8", 19, 31, (300, 500, 0.8), 8, 3),
                                   ("voc_g64", 38, 63, (600, 1000, 1.6), 64, 4), ("voc_g8", 38, 63, (600, 1000, 1.6), 8, 5)]:
        gt = synth.gt_boxes(seed, G=G, im_h=im[0], im_w=im[1])
        info = np.array([im], dtype=np.float32)
        np.random.seed(seed)
        lab = R.assign_anchor((1, 18, H, W), gt, info)
        rec = {}
        mine = restate.assign_anchor((1, 18, H, W), gt, info, rng=np.random.RandomState(seed), record=rec)
        for k in ("label", "bbox_target", "bbox_inside_weight", "bbox_outside_weight"):
            assert np.array_equal(lab[k], mine[k], equal_nan=True), (tag, k)
        np.savez_compressed(os.path.join(OUT, "anchor_target_%s.npz" % tag), H=H, W=W, im_info=info, G=G, seed=seed,
                            label=lab["label"], bbox_target=lab["bbox_target"],
                            bbox_inside_weight=lab["bbox_inside_weight"], bbox_outside_weight=lab["bbox_outside_weight"],
                            n_fg_pre=rec["n_fg_pre"], n_bg_pre=rec["n_bg_pre"], **meta())
        print("anchor_target", tag, "fg/bg pre", rec["n_fg_pre"], rec["n_bg_pre"])


def golden_proposal_target(R):
    T = R.config.TRAIN
    for tag, G, Rn, lo, seed, pre in [("g8", 8, 2000, 0.0, 0, True), ("g64", 64, 2000, 0.0, 1, True),
                                      ("g2_fewfg", 2, 300, 0.1, 2, False), ("g1", 1, 64, 0.0, 3, True)]:
        T.BG_THRESH_LO = lo
        T.BBOX_NORMALIZATION_PRECOMPUTED = pre
        gt = synth.gt_boxes(seed, G=G)
        rois = synth.jittered_rois(seed, gt, R=Rn, sigma=40.0 if tag != "g2_fewfg" else 150.0)
        gtp = synth.pad_gt(gt)
        op = R.ProposalTargetOperator("21", is_train=True)
        outs = [ref_shim.NDArrayStub(np.zeros(s, np.float32)) for s in [(128, 5), (128,), (128, 84), (128, 84), (128, 84)]]
        np.random.seed(seed)
        op.forward(True, ["write"] * 5, [ref_shim.NDArrayStub(rois), ref_shim.NDArrayStub(gtp)], outs, [])
        cfg = restate.default_config()
        cfg.TRAIN.BG_THRESH_LO = lo
        cfg.TRAIN.BBOX_NORMALIZATION_PRECOMPUTED = pre
        rec = {}
        mine = restate.proposal_target_forward(rois, gtp, 21, cfg=cfg, rng=np.random.RandomState(seed), record=rec)
        for a, b in zip(outs, mine):
            assert np.array_equal(a.arr, b), tag
        np.savez_compressed(os.path.join(OUT, "proposal_target_%s.npz" % tag), G=G, R=Rn, bg_lo=lo, seed=seed,
                            precomputed=pre, sigma=40.0 if tag != "g2_fewfg" else 150.0,
                            rois=outs[0].arr, labels=outs[1].arr, bbox_targets=outs[2].arr, inside=outs[3].arr,
                            outside=outs[4].arr, keep_inds=rec["keep_inds"], fg_this=rec["fg_this"], **meta())
        print("proposal_target", tag, "fg_this", rec["fg_this"])
    T.BG_THRESH_LO = 0.1
    T.BBOX_NORMALIZATION_PRECOMPUTED = False
    # Fast R-CNN sampling from a precomputed roidb (rcnn/minibatch.py:151)
    for seed in [0, 1]:
        db = synth.roidb_entry(seed, R=2000 if seed == 0 else 90)
        np.random.seed(seed)
        out = R.sample_rois(db, 16, 64, 21)
        rec = {}
        mine = restate.sample_rois(db, 16, 64, 21, rng=np.random.RandomState(seed), record=rec)
        for a, b in zip(out, mine):
            assert np.array_equal(a, b)
        np.savez_compressed(os.path.join(OUT, "sample_rois_s%d.npz" % seed), seed=seed, R=len(db["boxes"]),
                            rois=out[0], labels=out[1], bbox_targets=out[2], inside=out[3], overlaps=out[4],
                            keep_indexes=rec["keep_indexes"], **meta())
        print("sample_rois", seed)


class _FakeSymbol(object):
    """What rcnn/detector.py:38-58 asks of an mx Symbol: shapes, names, and an executor whose outputs are the
    network outputs we supply (the network itself is outside the path)."""

    def __init__(self, rois5, scores, deltas, end2end):
        self._names = ["rpn_rois_output" if end2end else "rois_output", "cls_prob_reshape_output",
                       "bbox_pred_reshape_output"]
        self._outs = [ref_shim.NDArrayStub(rois5), ref_shim.NDArrayStub(scores[None]), ref_shim.NDArrayStub(deltas[None])]

    def list_arguments(self):
        return ["data", "im_info", "cls_prob_label"]

    def infer_shape(self, **kw):
        return [kw["data"], kw["im_info"], (1,)], [o.shape for o in self._outs], []

    def list_outputs(self):
        return self._names

    def bind(self, ctx, args, args_grad=None, grad_req="null", aux_states=None):
        sym = self

        class _Exec(object):
            outputs = sym._outs

            def forward(self, is_train=False):
                pass
        return _Exec()


class _Batch(object):
    def __init__(self, data):
        self.data = data


class _TestData(list):
    shuffle = False


class _FakeImdb(object):
    def __init__(self, n_images, C, cache):
        self.num_images, self.num_classes = n_images, C
        self.classes = ["c%d" % j for j in range(C)]
        self.cache_path, self.name = cache, "synthetic"
        self.image_set_index = list(range(n_images))
        self.all_boxes = None

    def evaluate_detections(self, all_boxes):
        self.all_boxes = all_boxes


def run_reference_box_head(R, cases, nms_thresh=0.3):
    """The reference's own Detector.im_detect (rcnn/detector.py:17-81) + pred_eval (rcnn/tester.py:10-84) over a
    list of (rois, deltas, scores, im_shape, scale); returns all_boxes[j][i] as pred_eval hands it to the imdb."""
    import tempfile
    T = R.config.TEST
    saved = (T.HAS_RPN, T.NMS, R.config.END2END)
    T.HAS_RPN, T.NMS = True, nms_thresh
    try:
        C = cases[0][2].shape[1]
        batches, syms = _TestData(), []
        for rois, deltas, scores, (im_h, im_w), scale in cases:
            rois5 = np.hstack((np.zeros((rois.shape[0], 1), np.float32), rois))
            syms.append(_FakeSymbol(rois5, scores, deltas, R.config.END2END))
            batches.append(_Batch({"data": np.zeros((1, 3, im_h, im_w), np.float32),
                                   "im_info": np.array([[im_h, im_w, scale]], dtype=np.float32)}))

        class _Det(R.Detector):  # one fake symbol per image, everything else is the reference's Detector
            def im_detect(self, im_array, im_info=None, roi_array=None):
                self.symbol = syms[self.calls]
                self.calls += 1
                return R.Detector.im_detect(self, im_array, im_info=im_info, roi_array=roi_array)
        det = _Det(None, ctx="cpu", arg_params={}, aux_params={})
        det.calls = 0
        with tempfile.TemporaryDirectory() as tmp:
            imdb = _FakeImdb(len(cases), C, tmp)
            R.pred_eval(det, batches, imdb, vis=False)
        return imdb.all_boxes
    finally:
        T.HAS_RPN, T.NMS, R.config.END2END = saved


def golden_box_head(R):
    # (tag, seed, kwargs of synth.box_head_inputs, nms threshold)
    specs = [("voc_s0", 0, {}, 0.3), ("voc_s1", 1, {}, 0.3), ("dense_s2", 2, dict(G=2, peaked=0.9), 0.3),
             ("few_s3", 3, dict(R=40, G=1, peaked=0.1), 0.3), ("big_s4", 4, dict(R=1500, G=12, peaked=0.5), 0.5),
             ("wider_s5", 5, dict(R=2000, C=2, G=40, peaked=0.7, im_h=1200, im_w=1600, scale=1.0), 0.3)]
    for tag, seed, kw, thr in specs:
        case = synth.box_head_inputs(seed, **kw)
        all_boxes = run_reference_box_head(R, [case], nms_thresh=thr)
        C = case[2].shape[1]
        dets = [np.zeros((0, 5))] + [np.asarray(all_boxes[j][0], dtype=np.float64).reshape(-1, 5) for j in range(1, C)]
        mine, rows = restate.box_head(*case, thresh=0.05, nms_thresh=thr, max_per_image=100)
        for j in range(C):
            assert np.array_equal(dets[j], mine[j]), (tag, j)
        offsets = np.cumsum([0] + [len(d) for d in dets]).astype(np.int32)
        np.savez_compressed(os.path.join(OUT, "box_head_%s.npz" % tag), seed=seed, kw=np.array(repr(kw)), nms_thresh=thr,
                            dets=np.concatenate(dets), rows=np.concatenate(rows), offsets=offsets, **meta())
        print("box_head", tag, "dets", int(offsets[-1]), "of", int((case[2][:, 1:] > 0.05).sum()), "above threshold")


def run_reference_roidb(R, gt_roidb, box_list, C, precomputed):
    """The reference's offline chain on one synthetic dataset: IMDB.create_roidb_from_box_list + merge_roidbs
    (helper/dataset/imdb.py:32-79), the max_overlaps / max_classes lines of prepare_roidb (helper/processing/roidb.py:27-31;
    the image-path and cv2.imread lines around them are dataset IO), add_bbox_regression_targets (roidb.py:42-90) on
    the images that have ground truth, and the report printed by IMDB.evaluate_recall (imdb.py:109-188)."""
    import contextlib
    import copy
    import io
    imdb = R.IMDB("synthetic")
    imdb.num_classes, imdb.num_images = C, len(box_list)
    imdb.classes = ["c%d" % j for j in range(C)]
    imdb.image_set_index = list(range(len(box_list)))
    sink = io.StringIO()
    with contextlib.redirect_stdout(sink):
        ss = imdb.create_roidb_from_box_list(box_list, gt_roidb)
        ss_overlaps = [e["gt_overlaps"].toarray() for e in ss]
        roidb = R.IMDB.merge_roidbs(copy.deepcopy(gt_roidb), ss)
        for e in roidb:
            dense = e["gt_overlaps"].toarray()
            e["max_overlaps"], e["max_classes"] = dense.max(axis=1), dense.argmax(axis=1)
        with_gt = [e for e in roidb if (e["max_overlaps"] == 1).any()]
        saved = R.config.TRAIN.BBOX_NORMALIZATION_PRECOMPUTED
        R.config.TRAIN.BBOX_NORMALIZATION_PRECOMPUTED = precomputed
        try:
            means, stds = R.add_bbox_regression_targets(with_gt)
        finally:
            R.config.TRAIN.BBOX_NORMALIZATION_PRECOMPUTED = saved
    report = io.StringIO()
    with contextlib.redirect_stdout(report):
        imdb.evaluate_recall(roidb)
        imdb.evaluate_recall(roidb, area="medium", limit=100)
    return ss_overlaps, roidb, with_gt, means, stds, report.getvalue()


def golden_roidb(R):
    for seed, pre, kw in [(0, False, {}), (1, True, dict(n_images=3, R=120, C=5, max_gt=3))]:
        C = kw.get("C", 21)
        gt_roidb, box_list = synth.gt_roidb_and_box_list(seed, **kw)
        ss_overlaps, roidb, with_gt, means, stds, report = run_reference_roidb(R, gt_roidb, box_list, C, pre)
        cfg = restate.default_config()
        cfg.TRAIN.BBOX_NORMALIZATION_PRECOMPUTED = pre
        for i, want in enumerate(ss_overlaps):
            mine = restate.box_list_overlaps(box_list[i], gt_roidb[i]["boxes"], gt_roidb[i]["gt_classes"], C)
            assert np.array_equal(mine, want), i
        mine_db = [dict(boxes=e["boxes"], max_overlaps=e["max_overlaps"], max_classes=e["max_classes"],
                        gt_overlaps=e["gt_overlaps"]) for e in with_gt]
        m2, s2 = restate.add_bbox_regression_targets(mine_db, cfg)
        assert np.array_equal(m2, means) and np.array_equal(s2, stds)
        for a, b in zip(mine_db, with_gt):
            assert np.array_equal(a["bbox_targets"], b["bbox_targets"])
        np.savez_compressed(os.path.join(OUT, "roidb_s%d.npz" % seed), seed=seed, kw=np.array(repr(kw)), precomputed=pre,
                            ss_overlaps=np.concatenate(ss_overlaps), bbox_targets=np.concatenate([e["bbox_targets"] for e in with_gt]),
                            means=means, stds=stds, recall_report=np.array(report), **meta())
        print("roidb", seed, "images", len(roidb), "with gt", len(with_gt), "report lines", report.count("\n"))


GROUPS = {"roidb": golden_roidb, "helpers": golden_helpers, "proposal": golden_proposal, "anchor_target": golden_anchor_target,
          "proposal_target": golden_proposal_target, "box_head": golden_box_head}


def main():
    os.makedirs(OUT, exist_ok=True)
    R = ref_shim.ref()
    for name in (sys.argv[1:] or list(GROUPS)):  # python -m oracle.make_golden [group ...]
        GROUPS[name](R)


if __name__ == "__main__":
    main()
