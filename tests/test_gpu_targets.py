"""GPU parity: assign_anchor, proposal_target (_sample_rois) and sample_rois against fixtures from
the real reference.  Labels, sampled indices and weights are bit-exact; regression targets agree
to the last ulp of log() (rtol written below)."""
import numpy as np
import pytest

from oracle import restate, synth
from tests.util import golden, golden_names

pytestmark = pytest.mark.gpu

LOG_RTOL = 2e-6   # float32 targets: CUDA logf/log vs numpy's SIMD log differ in the last ulp only


@pytest.mark.parametrize("name", golden_names("anchor_target_"))
def test_assign_anchor_vs_reference_fixture(name):
    from mxrcnn_b200.config import config
    from mxrcnn_b200.rcnn.minibatch import assign_anchor
    g = golden(name)
    H, W, G, seed = int(g["H"]), int(g["W"]), int(g["G"]), int(g["seed"])
    info = g["im_info"]
    gt = synth.gt_boxes(seed, G=G, im_h=info[0, 0], im_w=info[0, 1])
    config.END2END = 0
    np.random.seed(seed)
    out = assign_anchor((1, 18, H, W), gt, info)
    assert set(out) == {"label", "bbox_target", "bbox_inside_weight", "bbox_outside_weight"}
    for k in out:
        assert out[k].dtype == np.float32 and out[k].shape == g[k].shape, k
    assert np.array_equal(out["label"], g["label"])                       # incl. the RNG subsampling
    assert np.array_equal(out["bbox_inside_weight"], g["bbox_inside_weight"])
    assert np.array_equal(out["bbox_outside_weight"], g["bbox_outside_weight"])
    np.testing.assert_allclose(out["bbox_target"], g["bbox_target"], rtol=LOG_RTOL, atol=1e-7)
    t = out["bbox_target"].reshape(9, 4, H, W)
    tg = g["bbox_target"].reshape(9, 4, H, W)
    assert np.array_equal(t[:, :2], tg[:, :2])                            # dx, dy: no transcendental, bit-exact
    # the RNG stream must have been consumed exactly like the reference consumed it
    ref_state = np.random.RandomState(seed)
    restate.assign_anchor((1, 18, H, W), gt, info, rng=ref_state)
    assert np.random.randint(1 << 30) == ref_state.randint(1 << 30)


def test_assign_anchor_end2end_and_edge_cases():
    from mxrcnn_b200.config import config
    from mxrcnn_b200.rcnn.minibatch import assign_anchor
    info = np.array([[300, 500, 0.8]], np.float32)
    gt = synth.gt_boxes(31, G=5, im_h=300, im_w=500)
    config.END2END = 1
    try:
        np.random.seed(31)
        out = assign_anchor((1, 18, 19, 31), gt, info)
        assert "gt_boxes" in out and out["gt_boxes"] is gt
    finally:
        config.END2END = 0
    # every setting the reference reads: clobber, positive weight, allowed border, other anchors
    config.TRAIN.RPN_CLOBBER_POSITIVES, config.TRAIN.RPN_POSITIVE_WEIGHT = True, 0.25
    try:
        cfg = restate.default_config()
        cfg.TRAIN.RPN_CLOBBER_POSITIVES, cfg.TRAIN.RPN_POSITIVE_WEIGHT = True, 0.25
        np.random.seed(32)
        out = assign_anchor((1, 24, 19, 31), gt, info, scales=(4, 8, 16, 32), allowed_border=9)
        want = restate.assign_anchor((1, 24, 19, 31), gt, info, scales=(4, 8, 16, 32), allowed_border=9, cfg=cfg,
                                     rng=np.random.RandomState(32))
        for k in ("label", "bbox_inside_weight", "bbox_outside_weight"):
            assert np.array_equal(out[k], want[k]), k
        np.testing.assert_allclose(out["bbox_target"], want["bbox_target"], rtol=LOG_RTOL, atol=1e-7)
    finally:
        config.TRAIN.RPN_CLOBBER_POSITIVES, config.TRAIN.RPN_POSITIVE_WEIGHT = False, -1.0
    # no anchor inside a small image: the reference raises (argmax of an empty array)
    small = np.array([[96, 144, 1.0]], np.float32)
    with pytest.raises(ValueError):
        assign_anchor((1, 18, 6, 9), synth.gt_boxes(1, G=3, im_h=96, im_w=144), small)
    # a gt that overlaps no inside anchor makes EVERY zero-overlap anchor positive (quirk Q4)
    far = np.array([[5000, 5000, 5100, 5100, 1]], np.float32)
    np.random.seed(3)
    out = assign_anchor((1, 18, 19, 31), far, info)
    want = restate.assign_anchor((1, 18, 19, 31), far, info, rng=np.random.RandomState(3))
    assert np.array_equal(out["label"], want["label"]) and (out["label"] == 1).sum() == 128


def test_assign_anchor_wider_many_gt():
    """WIDER-face scale (C5): 75x100 map, 12 anchors, 1968 gt boxes (the dataset's maximum) -- more than
    one shared-memory chunk of gt boxes.  Checked against the oracle on labels only (bounded CPU time)."""
    from mxrcnn_b200.rcnn.minibatch import assign_anchor
    info = np.array([[1200, 1600, 1.0]], np.float32)
    rs = np.random.RandomState(9)
    G = 1968
    w = rs.uniform(8, 60, G)
    hgt = rs.uniform(8, 60, G)
    x1 = rs.uniform(0, 1599 - w)
    y1 = rs.uniform(0, 1199 - hgt)
    gt = np.stack([x1, y1, x1 + w, y1 + hgt, np.ones(G)], axis=1).astype(np.float32)
    np.random.seed(9)
    out = assign_anchor((1, 24, 75, 100), gt, info, scales=(4, 8, 16, 32))
    want = restate.assign_anchor((1, 24, 75, 100), gt, info, scales=(4, 8, 16, 32), rng=np.random.RandomState(9))
    assert np.array_equal(out["label"], want["label"])
    assert np.array_equal(out["bbox_outside_weight"], want["bbox_outside_weight"])
    np.testing.assert_allclose(out["bbox_target"], want["bbox_target"], rtol=LOG_RTOL, atol=1e-7)


@pytest.mark.parametrize("name", golden_names("proposal_target_"))
def test_proposal_target_operator_vs_reference_fixture(name):
    from mxrcnn_b200 import operator as mxop
    from mxrcnn_b200.config import config
    import mxrcnn_b200.rcnn.rpn.proposal_target  # noqa: F401 (registers 'proposal_target')
    g = golden(name)
    seed = int(g["seed"])
    gt = synth.gt_boxes(seed, G=int(g["G"]))
    rois = synth.jittered_rois(seed, gt, R=int(g["R"]), sigma=float(g["sigma"]))
    old = (config.TRAIN.BG_THRESH_LO, config.TRAIN.BBOX_NORMALIZATION_PRECOMPUTED)
    config.TRAIN.BG_THRESH_LO, config.TRAIN.BBOX_NORMALIZATION_PRECOMPUTED = float(g["bg_lo"]), bool(g["precomputed"])
    try:
        prop, op = mxop.create("proposal_target", num_classes=21, is_train=True)
        assert prop.list_arguments() == ["rpn_roi", "gt_boxes"]
        assert prop.list_outputs() == ["roi", "label", "bbox_target", "bbox_inside_weight", "bbox_outside_weight"]
        _, out_shapes = prop.infer_shape([rois.shape, (100, 5)])
        assert out_shapes == [(128, 5), (128,), (128, 84), (128, 84), (128, 84)]
        outs = [np.zeros(s, np.float32) for s in out_shapes]
        np.random.seed(seed)
        op.forward(True, ["write"] * 5, [rois, synth.pad_gt(gt)], outs, [])
    finally:
        config.TRAIN.BG_THRESH_LO, config.TRAIN.BBOX_NORMALIZATION_PRECOMPUTED = old
    assert np.array_equal(outs[0], g["rois"])                 # same sampled rois, same order: same RNG draws
    assert np.array_equal(outs[1], g["labels"])
    assert np.array_equal(outs[3], g["inside"]) and np.array_equal(outs[4], g["outside"])
    np.testing.assert_allclose(outs[2], g["bbox_targets"], rtol=LOG_RTOL, atol=2e-6)
    assert np.array_equal(outs[2] != 0, g["bbox_targets"] != 0)


def test_sample_rois_e2e_vs_oracle_and_empty_gt():
    from mxrcnn_b200.config import config
    from mxrcnn_b200.rcnn.rpn.proposal_target import _sample_rois
    gt = synth.gt_boxes(5, G=4)
    rois = np.vstack((synth.jittered_rois(5, gt, R=500), np.hstack((np.zeros((4, 1), np.float32), gt[:, :4]))))
    old = config.TRAIN.BG_THRESH_LO
    config.TRAIN.BG_THRESH_LO = 0.0
    try:
        np.random.seed(5)
        labels, out_rois, bt, bw = _sample_rois(rois, gt, 32, 128, 21, "TRAIN")
    finally:
        config.TRAIN.BG_THRESH_LO = old
    cfg = restate.default_config()
    cfg.TRAIN.BG_THRESH_LO = 0.0
    want = restate.sample_rois_e2e(rois, gt, 32, 128, 21, "TRAIN", cfg=cfg, rng=np.random.RandomState(5))
    assert np.array_equal(labels, want[0]) and np.array_equal(out_rois, want[1]) and np.array_equal(bw, want[3])
    np.testing.assert_allclose(bt, want[2], rtol=LOG_RTOL, atol=2e-6)
    with pytest.raises(ValueError):
        _sample_rois(rois, np.zeros((0, 5), np.float32), 32, 128, 21, "TRAIN")


@pytest.mark.parametrize("seed", [0, 1])
def test_sample_rois_roidb_vs_reference_fixture(seed):
    from mxrcnn_b200.rcnn.minibatch import sample_rois
    g = golden("sample_rois_s%d" % seed)
    db = synth.roidb_entry(seed, R=int(g["R"]))
    np.random.seed(seed)
    out = sample_rois(db, 16, 64, 21)
    for a, k in zip(out, ("rois", "labels", "bbox_targets", "inside", "overlaps")):
        assert np.array_equal(a, g[k]), k
        assert a.dtype == g[k].dtype, k
