"""GPU parity: bbox_pred / clip_boxes / clip_pad / bbox_transform / bbox_overlaps /
expand_bbox_regression_targets through the C ABI, against fixtures from the real reference."""
import numpy as np
import pytest

from oracle import restate, synth
from tests.util import golden

pytestmark = pytest.mark.gpu


def _inputs():
    rs = np.random.RandomState(11)
    gt = synth.gt_boxes(11, G=8)
    rois = synth.jittered_rois(11, gt, R=300)[:, 1:5]
    deltas = rs.normal(0, 0.2, size=(300, 84)).astype(np.float32)
    g = gt[rs.randint(0, 8, 300), :4]
    return gt, rois, deltas, g


def test_bbox_pred_and_clip_vs_reference_fixture():
    from mxrcnn_b200.helper.processing.bbox_transform import bbox_pred, clip_boxes
    h = golden("helpers")
    gt, rois, deltas, _ = _inputs()
    pred = bbox_pred(rois, deltas)
    assert pred.dtype == np.float64 and pred.shape == (300, 84)
    # exp() is CUDA's expf vs numpy's SIMD exp: last-ulp differences only
    np.testing.assert_allclose(pred, h["bp_pred"], rtol=1e-6, atol=1e-4)
    # x/y centres involve no exp: they match to f64 rounding
    cx = 0.5 * (pred[:, 0::4] + pred[:, 2::4])
    cx_ref = 0.5 * (h["bp_pred"][:, 0::4] + h["bp_pred"][:, 2::4])
    np.testing.assert_allclose(cx, cx_ref, rtol=1e-12, atol=1e-9)
    clipped = clip_boxes(pred, (600, 1000))
    assert clipped is pred                                  # in place, returns the same array
    np.testing.assert_allclose(clipped, h["bp_clip"], rtol=1e-6, atol=1e-4)
    # stage-level and bit-exact: clip the reference's own unclipped boxes
    ref = h["bp_pred"].copy()
    assert np.array_equal(clip_boxes(ref, (600, 1000)), h["bp_clip"])
    # float32 limits as the proposal op passes them (im_info[:2])
    ref = h["bp_pred"].copy()
    info = np.array([600, 1000], np.float32)
    assert np.array_equal(clip_boxes(ref, info), restate.clip_boxes(h["bp_pred"].copy(), info))
    # empty input (bbox_transform.py:43-44)
    e = bbox_pred(np.zeros((0, 4)), np.zeros((0, 84), np.float32))
    assert e.shape == tuple(h["bp_empty_shape"]) and e.dtype == np.float64


def test_bbox_pred_dtypes_and_is_train():
    from mxrcnn_b200.helper.processing.bbox_transform import bbox_pred
    gt, rois, deltas, _ = _inputs()
    for boxes in (rois, rois.astype(np.float64), rois.astype(np.uint16)):
        for d in (deltas, deltas.astype(np.float64)):
            np.testing.assert_allclose(bbox_pred(boxes, d), restate.bbox_pred(boxes, d), rtol=1e-6, atol=1e-4)
    big = deltas[:, :4] * 100
    got = bbox_pred(rois, big, is_train=True)
    clamped = big.copy()
    for k, lim in ((0, 10), (1, 10), (2, 8), (3, 8)):
        clamped[:, k::4] = np.where(np.abs(big[:, k::4]) > lim, np.sign(big[:, k::4]) * lim, big[:, k::4])
    np.testing.assert_allclose(got, restate.bbox_pred(rois, clamped), rtol=1e-6, atol=1e-3)


def test_clip_pad():
    from mxrcnn_b200.helper.processing.bbox_transform import clip_pad
    a = np.arange(2 * 3 * 5 * 7, dtype=np.float32).reshape(2, 3, 5, 7)
    for shape in ((5, 7), (4, 7), (5, 3), (2, 2), (9, 9)):
        assert np.array_equal(clip_pad(a, shape), restate.clip_pad(a, shape))


def test_bbox_transform_vs_reference_fixture():
    from mxrcnn_b200.helper.processing.bbox_transform import bbox_transform
    h = golden("helpers")
    gt, rois, _, g = _inputs()
    t64 = bbox_transform(rois.astype(np.float64), g.astype(np.float64))
    assert t64.dtype == np.float64
    assert np.array_equal(t64[:, :2], h["bt_f64"][:, :2])          # no transcendental: bit-exact
    np.testing.assert_allclose(t64[:, 2:], h["bt_f64"][:, 2:], rtol=1e-13, atol=1e-15)   # log: last ulp
    t32 = bbox_transform(rois, g)
    assert t32.dtype == np.float32
    assert np.array_equal(t32[:, :2], h["bt_f32"][:, :2])
    np.testing.assert_allclose(t32[:, 2:], h["bt_f32"][:, 2:], rtol=2e-6, atol=1e-7)
    # mixed dtypes (rcnn/minibatch.py:338: float64 anchors, float32 gt): each side in its own dtype first
    for e, q in ((rois.astype(np.float64), g), (rois, g.astype(np.float64))):
        mixed, want = bbox_transform(e, q), restate.bbox_transform(e, q)
        assert mixed.dtype == np.float64 and np.array_equal(mixed[:, :2], want[:, :2])
        np.testing.assert_allclose(mixed[:, 2:], want[:, 2:], rtol=1e-13, atol=1e-15)


def test_bbox_overlaps_bit_exact():
    from mxrcnn_b200.helper.processing.bbox_regression import bbox_overlaps, bbox_overlaps_reduced
    h = golden("helpers")
    gt, rois, _, _ = _inputs()
    ov = bbox_overlaps(rois.astype(np.float64), gt.astype(np.float64))   # gt has 5 columns: only 0..3 are read
    assert ov.dtype == np.float64 and np.array_equal(ov, h["ov_300x8"])
    anc = restate.proposal_stages(*synth.proposal_inputs(0, H=6, W=9, A=9, im_info=(96, 144, 1.0)), pre_nms_topN=200,
                                  post_nms_topN=30)["anchors"]
    gts = synth.gt_boxes(5, G=6, im_h=96, im_w=144)
    assert np.array_equal(bbox_overlaps(anc, gts.astype(np.float64)), h["ov_anchor_gt"])
    # larger random case incl. more queries than one shared-memory chunk, and the fused reductions
    big_gt = synth.gt_boxes(9, G=700).astype(np.float64)
    big = synth.jittered_rois(9, big_gt.astype(np.float32), R=3000)[:, 1:5].astype(np.float64)
    want = restate.bbox_overlaps(big, big_gt)
    got = bbox_overlaps(big, big_gt)
    assert np.array_equal(got, want)
    arg, rmax, cmax = bbox_overlaps_reduced(big, big_gt)
    assert np.array_equal(arg, want.argmax(axis=1)) and np.array_equal(rmax, want.max(axis=1))
    assert np.array_equal(cmax, want.max(axis=0))
    # degenerate shapes
    assert bbox_overlaps(np.zeros((0, 4)), big_gt).shape == (0, 700)
    assert bbox_overlaps(big, np.zeros((0, 5))).shape == (3000, 0)
    # touching / disjoint / identical boxes
    b = np.array([[0, 0, 9, 9], [10, 0, 19, 9], [9, 9, 20, 20], [0, 0, 9, 9.]])
    assert np.array_equal(bbox_overlaps(b, b), restate.bbox_overlaps_loop(b, b))


def test_expand_targets_bit_exact():
    from mxrcnn_b200.helper.processing.bbox_regression import expand_bbox_regression_targets
    h = golden("helpers")
    t, w = expand_bbox_regression_targets(h["ex_in"], 21)
    assert t.dtype == np.float32 and np.array_equal(t, h["ex_t"]) and np.array_equal(w, h["ex_w"])
    t, w = expand_bbox_regression_targets(np.zeros((0, 5), np.float32), 21)
    assert t.shape == (0, 84) and w.shape == (0, 84)


def test_helpers_on_device_arrays():
    import torch
    from mxrcnn_b200.helper.processing.bbox_regression import bbox_overlaps, expand_bbox_regression_targets
    from mxrcnn_b200.helper.processing.bbox_transform import bbox_pred, bbox_transform, clip_boxes
    h = golden("helpers")
    gt, rois, deltas, g = _inputs()
    dev = "cuda"
    pred = bbox_pred(torch.from_numpy(rois).to(dev), torch.from_numpy(deltas).to(dev))
    assert pred.is_cuda and np.array_equal(pred.cpu().numpy(), bbox_pred(rois, deltas))
    c = clip_boxes(pred, (600, 1000))
    assert c is pred
    np.testing.assert_allclose(c.cpu().numpy(), h["bp_clip"], rtol=1e-6, atol=1e-4)
    ov = bbox_overlaps(torch.from_numpy(rois).to(dev), torch.from_numpy(gt).to(dev))
    assert np.array_equal(ov.cpu().numpy(), h["ov_300x8"])
    t = bbox_transform(torch.from_numpy(rois).to(dev), torch.from_numpy(g).to(dev))
    assert np.array_equal(t.cpu().numpy(), bbox_transform(rois, g))
    et, ew = expand_bbox_regression_targets(torch.from_numpy(h["ex_in"]).to(dev), 21)
    assert np.array_equal(et.cpu().numpy(), h["ex_t"]) and np.array_equal(ew.cpu().numpy(), h["ex_w"])
