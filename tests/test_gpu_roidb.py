"""GPU parity for the offline roidb users of bbox_overlaps: class-overlap tables,
bbox regression targets + normalisation, recall report -- against the fixtures produced by the reference's own
IMDB / add_bbox_regression_targets, and against oracle/restate.py."""
import ast
import contextlib
import copy
import io

import numpy as np
import pytest

from oracle import restate, synth
from tests.util import golden

pytestmark = pytest.mark.gpu


def build(seed, kw):
    from mxrcnn_b200.helper.dataset.imdb import IMDB
    C = kw.get("C", 21)
    gt_roidb, box_list = synth.gt_roidb_and_box_list(seed, **kw)
    imdb = IMDB("synthetic")
    imdb.num_classes, imdb.num_images = C, len(box_list)
    return imdb, gt_roidb, box_list, C


def merged(gt_roidb, ss):
    import scipy.sparse
    out = []
    for a, b in zip(copy.deepcopy(gt_roidb), ss):  # what IMDB.merge_roidbs + prepare_roidb leave in each entry
        e = {"boxes": np.vstack((a["boxes"], b["boxes"])), "gt_classes": np.hstack((a["gt_classes"], b["gt_classes"])),
             "gt_overlaps": scipy.sparse.vstack([a["gt_overlaps"], b["gt_overlaps"]])}
        dense = e["gt_overlaps"].toarray()
        e["max_overlaps"], e["max_classes"] = dense.max(axis=1), dense.argmax(axis=1)
        out.append(e)
    return out


@pytest.mark.parametrize("seed", [0, 1])
def test_roidb_chain_vs_reference_fixture(seed):
    from mxrcnn_b200.config import config
    from mxrcnn_b200.helper.processing.roidb import add_bbox_regression_targets
    g = golden("roidb_s%d" % seed)
    imdb, gt_roidb, box_list, C = build(seed, ast.literal_eval(str(g["kw"])))
    ss = imdb.create_roidb_from_box_list(box_list, gt_roidb)
    got = np.concatenate([e["gt_overlaps"].toarray() for e in ss])
    assert got.dtype == np.float32 and np.array_equal(got, g["ss_overlaps"])  # IoU: no transcendental, bit-exact
    roidb = merged(gt_roidb, ss)
    with_gt = [e for e in roidb if (e["max_overlaps"] == 1).any()]
    saved = config.TRAIN.BBOX_NORMALIZATION_PRECOMPUTED
    config.TRAIN.BBOX_NORMALIZATION_PRECOMPUTED = bool(g["precomputed"])
    try:
        with contextlib.redirect_stdout(io.StringIO()):
            means, stds = add_bbox_regression_targets(with_gt)
    finally:
        config.TRAIN.BBOX_NORMALIZATION_PRECOMPUTED = saved
    t = np.concatenate([e["bbox_targets"] for e in with_gt])
    assert t.dtype == np.float32 and np.array_equal(t[:, 0], g["bbox_targets"][:, 0])
    # dw, dh go through log(): CUDA's log and numpy's may differ in the last ulp of the float64 result, which the
    # float32 store and the division by a small std can amplify to a few float32 ulps
    np.testing.assert_allclose(means, g["means"], rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(stds, g["stds"], rtol=1e-6, atol=1e-9)
    np.testing.assert_allclose(t[:, 1:], g["bbox_targets"][:, 1:], rtol=2e-5, atol=2e-6)
    report = io.StringIO()
    with contextlib.redirect_stdout(report):
        res = imdb.evaluate_recall(roidb)
        imdb.evaluate_recall(roidb, area="medium", limit=100)
    assert report.getvalue() == str(g["recall_report"])
    assert 0.0 <= res["ar"] <= 1.0 and res["gt_overlaps"].size > 0


@pytest.mark.parametrize("seed", [5, 6])
def test_compute_targets_and_cover_vs_oracle(seed):
    from mxrcnn_b200.helper.dataset.imdb import class_overlaps, greedy_cover
    from mxrcnn_b200.helper.processing.bbox_regression import bbox_overlaps, compute_bbox_regression_targets
    gt_roidb, box_list = synth.gt_roidb_and_box_list(seed, n_images=3, R=500, max_gt=12)
    for i in (0, 2):
        gt, boxes = gt_roidb[i], box_list[i]
        want = restate.box_list_overlaps(boxes, gt["boxes"], gt["gt_classes"], 21)
        got = class_overlaps(boxes, gt["boxes"], gt["gt_classes"], 21)
        assert np.array_equal(got, want)
        rois = np.vstack((gt["boxes"], boxes))
        ov = np.hstack((np.ones(len(gt["boxes"])), want.max(axis=1)))
        lab = np.hstack((gt["gt_classes"], want.argmax(axis=1)))
        a = compute_bbox_regression_targets(rois, ov, lab)
        b = restate.compute_bbox_regression_targets(rois, ov, lab)
        assert np.array_equal(a[:, 0], b[:, 0]) and np.array_equal(a[:, 1:3], b[:, 1:3])  # dx, dy: no transcendental
        np.testing.assert_allclose(a[:, 3:], b[:, 3:], rtol=2e-6, atol=1e-7)
        m = bbox_overlaps(boxes.astype(np.float64), gt["boxes"].astype(np.float64))
        assert np.array_equal(greedy_cover(m), restate.recall_gt_overlaps(boxes, gt["boxes"]))
        from mxrcnn_b200.helper.dataset.imdb import recall_cover
        assert np.array_equal(recall_cover(boxes, gt["boxes"]), restate.recall_gt_overlaps(boxes, gt["boxes"]))
    assert class_overlaps(box_list[1], gt_roidb[1]["boxes"], gt_roidb[1]["gt_classes"], 21).sum() == 0  # image without gt


def test_recall_cover_ties_many_gt_and_too_few_proposals():
    """The matching kernel against the reference loop (imdb.py:153-170, restated in the oracle) where ties decide:
    duplicated proposals and duplicated gt boxes (first maximum wins, by column then by row), integer-grid boxes whose
    IoUs coincide, more gt boxes than one warp pass (700), and fewer proposals than gt boxes (the reference asserts)."""
    from mxrcnn_b200.helper.dataset.imdb import recall_cover
    rs = np.random.RandomState(3)

    def grid(n, lo=0, hi=30):
        x1 = rs.randint(lo, hi, n) * 8.0
        y1 = rs.randint(lo, hi, n) * 8.0
        w = rs.choice([16, 32, 64], n).astype(np.float64)
        h = rs.choice([16, 32, 64], n).astype(np.float64)
        return np.stack([x1, y1, x1 + w - 1, y1 + h - 1], axis=1)

    for n, k in [(300, 40), (1500, 700), (64, 64), (5, 3)]:
        boxes, gt = grid(n), grid(k)
        boxes[n // 2:] = boxes[:n - n // 2]            # duplicated proposals
        gt[k // 2:] = gt[:k - k // 2]                  # duplicated gt boxes
        assert np.array_equal(recall_cover(boxes, gt), restate.recall_gt_overlaps(boxes, gt)), (n, k)
    with pytest.raises(AssertionError):
        recall_cover(grid(3), grid(5))
    assert recall_cover(grid(4), np.zeros((0, 4))).shape == (0,)
