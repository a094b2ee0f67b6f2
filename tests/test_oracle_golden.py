"""CPU: oracle/restate.py against the fixtures produced by the REAL reference (oracle/make_golden.py)."""
import numpy as np
import pytest

from oracle import restate, synth
from tests.util import golden, golden_names, proposal_case


def test_anchor_table_known_answer():
    # the one hand-checkable KAT of the path
    want = np.array([[-84, -40, 99, 55], [-176, -88, 191, 103], [-360, -184, 375, 199], [-56, -56, 71, 71],
                     [-120, -120, 135, 135], [-248, -248, 263, 263], [-36, -80, 51, 95], [-80, -168, 95, 183],
                     [-168, -344, 183, 359]], dtype=np.float64)
    got = restate.generate_anchors(16.0, [0.5, 1, 2], np.array([8.0, 16.0, 32.0]))
    assert got.dtype == np.float64 and np.array_equal(got, want)
    h = golden("helpers")
    assert np.array_equal(h["anchors_voc"], want)
    assert np.array_equal(restate.generate_anchors(16.0, [0.5, 1, 2], np.array([4.0, 8, 16, 32])), h["anchors_wider"])
    assert np.array_equal(restate.generate_anchors(16, [0.5, 1, 2], np.array((8, 16, 32), np.float32)), h["anchors_assign"])


def test_helpers_golden():
    h = golden("helpers")
    rs = np.random.RandomState(11)
    gt = synth.gt_boxes(11, G=8)
    rois = synth.jittered_rois(11, gt, R=300)[:, 1:5]
    deltas = rs.normal(0, 0.2, size=(300, 84)).astype(np.float32)
    pred = restate.bbox_pred(rois, deltas)
    assert np.array_equal(pred, h["bp_pred"])
    assert np.array_equal(restate.clip_boxes(pred.copy(), (600, 1000)), h["bp_clip"])
    assert tuple(h["bp_empty_shape"]) == restate.bbox_pred(np.zeros((0, 4)), np.zeros((0, 84), np.float32)).shape
    ex = rois.astype(np.float64)
    g = gt[rs.randint(0, 8, 300), :4]
    assert np.array_equal(restate.bbox_transform(ex, g.astype(np.float64)), h["bt_f64"])
    assert np.array_equal(restate.bbox_transform(rois, g), h["bt_f32"])
    assert np.array_equal(restate.bbox_overlaps(ex, gt.astype(np.float64)), h["ov_300x8"])
    assert np.array_equal(restate.bbox_overlaps_loop(ex[:40], gt.astype(np.float64)), h["ov_300x8"][:40])
    t, w = restate.expand_bbox_regression_targets(h["ex_in"], 21)
    assert np.array_equal(t, h["ex_t"]) and np.array_equal(w, h["ex_w"])


def test_nms_golden():
    h = golden("helpers")
    for i in range(6):
        seed, n, size, th = h["nms%d_spec" % i]
        dets = synth.class_dets(int(seed), n=int(n), dtype=np.float64 if size == 8 else np.float32)
        keep = restate.nms(dets, float(th))
        assert isinstance(keep, list)
        assert np.array_equal(np.asarray(keep, dtype=np.int64), h["nms%d_keep" % i]), i
    assert restate.nms(np.zeros((0, 5)), 0.3) == []
    dets = synth.class_dets(40, n=60, dtype=np.float32)
    assert np.array_equal(np.asarray(restate.nest(dets, 0.9), dtype=np.int64), h["nest_keep"])


def test_nms_early_stop_is_prefix():
    dets = synth.class_dets(3, n=800)
    full = restate.nms(dets, 0.5)
    assert restate.nms(dets, 0.5, stop_after=50) == full[:50]


@pytest.mark.parametrize("name", golden_names("proposal_tiny") + golden_names("proposal_voc_test")
                         + golden_names("proposal_voc_alt"))
def test_proposal_golden(name):
    g = golden(name)
    cls, box, info, cfg, train = proposal_case(g)
    st = restate.proposal_stages(cls, box, info, cfg["feat_stride"], cfg["scales"], cfg["ratios"], train,
                                 cfg["pre"], cfg["post"], cfg["thresh"], cfg["min_size"],
                                 rng=np.random.RandomState(int(g["seed"])))
    assert np.array_equal(st["rois"], g["rois"]) and np.array_equal(st["scores"], g["scores"])
    assert np.array_equal(st["keep"], g["keep"])
    assert np.array_equal(st["decoded"][:64], g["decoded_head"]) and np.array_equal(st["decoded"][-64:], g["decoded_tail"])
    assert st["rois"].dtype == np.float32 and st["rois"].shape == (cfg["post"], 5)


@pytest.mark.parametrize("seed", [7, 8])
def test_proposal_pad_golden(seed):
    g = golden("proposal_pad_s%d" % seed)
    cfg = dict(synth.CONFIGS["tiny"], post=int(g["post"]), thresh=float(g["thresh"]))
    cls, box, info = synth.proposal_inputs(seed, H=cfg["H"], W=cfg["W"], A=9, im_info=cfg["im_info"])
    st = restate.proposal_stages(cls, box, info, 16, cfg["scales"], cfg["ratios"], False, cfg["pre"], cfg["post"],
                                 cfg["thresh"], cfg["min_size"], rng=np.random.RandomState(seed))
    assert st["nkeep"] == int(g["nkeep"]) < cfg["post"]
    assert np.array_equal(st["rois"], g["rois"]) and np.array_equal(st["scores"], g["scores"])


def test_proposal_nan_raises():
    cls, box, info = synth.proposal_inputs(0, H=6, W=9, A=9, im_info=(96, 144, 1.0))
    bad = cls.copy()
    bad[0, 12, 2, 3] = np.nan
    with pytest.raises(ValueError):
        restate.proposal_stages(bad, box, info)
    bad = box.copy()
    bad[0, 5, 1, 1] = np.nan
    with pytest.raises(ValueError):
        restate.proposal_stages(cls, bad, info)
    # NaN in the unused background half is NOT an error (rcnn/rpn/proposal.py:57-58)
    ok = cls.copy()
    ok[0, 0, 0, 0] = np.nan
    restate.proposal_stages(ok, box, info, pre_nms_topN=200, post_nms_topN=5)


@pytest.mark.parametrize("name", golden_names("anchor_target_"))
def test_anchor_target_golden(name):
    g = golden(name)
    H, W, G, seed = int(g["H"]), int(g["W"]), int(g["G"]), int(g["seed"])
    info = g["im_info"]
    gt = synth.gt_boxes(seed, G=G, im_h=info[0, 0], im_w=info[0, 1])
    out = restate.assign_anchor((1, 18, H, W), gt, info, rng=np.random.RandomState(seed))
    for k in ("label", "bbox_target", "bbox_inside_weight", "bbox_outside_weight"):
        assert out[k].dtype == np.float32
        assert np.array_equal(out[k], g[k], equal_nan=True), k


@pytest.mark.parametrize("name", golden_names("proposal_target_"))
def test_proposal_target_golden(name):
    g = golden(name)
    seed = int(g["seed"])
    gt = synth.gt_boxes(seed, G=int(g["G"]))
    rois = synth.jittered_rois(seed, gt, R=int(g["R"]), sigma=float(g["sigma"]))
    cfg = restate.default_config()
    cfg.TRAIN.BG_THRESH_LO = float(g["bg_lo"])
    cfg.TRAIN.BBOX_NORMALIZATION_PRECOMPUTED = bool(g["precomputed"])
    rec = {}
    out = restate.proposal_target_forward(rois, synth.pad_gt(gt), 21, cfg=cfg, rng=np.random.RandomState(seed), record=rec)
    for a, k in zip(out, ("rois", "labels", "bbox_targets", "inside", "outside")):
        assert np.array_equal(a, g[k]), k
    assert np.array_equal(rec["keep_inds"], g["keep_inds"])


@pytest.mark.parametrize("seed", [0, 1])
def test_sample_rois_golden(seed):
    g = golden("sample_rois_s%d" % seed)
    db = synth.roidb_entry(seed, R=int(g["R"]))
    out = restate.sample_rois(db, 16, 64, 21, rng=np.random.RandomState(seed))
    for a, k in zip(out, ("rois", "labels", "bbox_targets", "inside", "overlaps")):
        assert np.array_equal(a, g[k]), k


@pytest.mark.parametrize("name", golden_names("box_head_"))
def test_box_head_golden(name):
    """rcnn/detector.py:73-74 + rcnn/tester.py:50-65 as run by the reference's own Detector / pred_eval."""
    from tests.util import box_head_case
    case, thr, dets, rows = box_head_case(golden(name))
    mine, mrows = restate.box_head(*case, thresh=0.05, nms_thresh=thr, max_per_image=100)
    assert len(mine) == len(dets) == case[2].shape[1]
    for j in range(len(dets)):
        assert mine[j].dtype == np.float64 and np.array_equal(mine[j], dets[j]), j
        assert np.array_equal(mrows[j], rows[j]), j


@pytest.mark.parametrize("seed", [0, 1])
def test_roidb_chain_golden(seed):
    """IMDB.create_roidb_from_box_list + add_bbox_regression_targets as run by the reference (oracle/make_golden.py)."""
    import ast
    import scipy.sparse
    g = golden("roidb_s%d" % seed)
    kw = ast.literal_eval(str(g["kw"]))
    C = kw.get("C", 21)
    gt_roidb, box_list = synth.gt_roidb_and_box_list(seed, **kw)
    ss = [restate.box_list_overlaps(box_list[i], gt_roidb[i]["boxes"], gt_roidb[i]["gt_classes"], C) for i in range(len(box_list))]
    assert np.array_equal(np.concatenate(ss), g["ss_overlaps"])
    db = []
    for gt, boxes, ov in zip(gt_roidb, box_list, ss):
        dense = np.vstack((gt["gt_overlaps"].toarray(), ov))
        if not (dense.max(axis=1) == 1).any():
            continue
        db.append(dict(boxes=np.vstack((gt["boxes"], boxes)), gt_overlaps=scipy.sparse.csr_matrix(dense),
                       max_overlaps=dense.max(axis=1), max_classes=dense.argmax(axis=1)))
    cfg = restate.default_config()
    cfg.TRAIN.BBOX_NORMALIZATION_PRECOMPUTED = bool(g["precomputed"])
    means, stds = restate.add_bbox_regression_targets(db, cfg)
    assert np.array_equal(means, g["means"]) and np.array_equal(stds, g["stds"])
    assert np.array_equal(np.concatenate([e["bbox_targets"] for e in db]), g["bbox_targets"])
