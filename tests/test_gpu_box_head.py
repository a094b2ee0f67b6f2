"""GPU parity: the fused test-time box head (mxr_box_head*) against the fixtures produced by
the reference's own Detector.im_detect + pred_eval, and against oracle/restate.py on seeded inputs."""
import numpy as np
import pytest

from oracle import restate, synth
from tests.util import box_head_case, golden, golden_names

pytestmark = pytest.mark.gpu

# bbox_pred evaluates exp() in float32: CUDA's expf and numpy's differ in the last ulp now and then, which moves a
# decoded coordinate by ~1e-5 relative (the tolerance north_star states); scores are gathered: bit-exact
RTOL = 1e-5


def check(got, got_rows, want, want_rows):
    assert len(got) == len(want)
    assert got[0] == [] or len(got[0]) == 0
    for j in range(1, len(want)):
        assert got[j].dtype == np.float64 and got[j].shape == want[j].shape, (j, got[j].shape, want[j].shape)
        assert np.array_equal(got_rows[j], want_rows[j]), j
        assert np.array_equal(got[j][:, 4], want[j][:, 4]), j
        np.testing.assert_allclose(got[j][:, :4], want[j][:, :4], rtol=RTOL, atol=1e-4)


@pytest.mark.parametrize("name", golden_names("box_head_"))
def test_box_head_vs_reference_fixture(name):
    from mxrcnn_b200.rcnn.tester import detections
    (rois, deltas, scores, im_shape, scale), thr, dets, rows = box_head_case(golden(name))
    got, got_rows = detections(scores, rois, deltas, im_shape, scale, thresh=0.05, nms_thresh=thr, max_per_image=100,
                               with_rows=True)
    check(got, got_rows, dets, rows)


@pytest.mark.parametrize("seed,kw,cap,thr", [
    (40, {}, 0, 0.3), (41, {}, 100, 0.3), (42, dict(R=1, G=1, peaked=1.0), 100, 0.3),
    (43, dict(R=1024, G=8, peaked=0.6), 0, 0.5), (44, dict(R=1025, C=3, G=30, peaked=0.8), 0, 0.3),
    (45, dict(R=4096, C=2, G=60, peaked=0.5, im_h=1200, im_w=1600, scale=1.0), 2000, 0.3),
    (46, dict(R=300, C=81, G=10), 100, 0.3), (47, dict(R=65, G=2, peaked=0.9), 7, 0.7)])
def test_box_head_vs_oracle(seed, kw, cap, thr):
    from mxrcnn_b200.rcnn.tester import detections
    rois, deltas, scores, im_shape, scale = synth.box_head_inputs(seed, **kw)
    want, want_rows = restate.box_head(rois, deltas, scores, im_shape, scale, thresh=0.05, nms_thresh=thr, max_per_image=cap)
    got, got_rows = detections(scores, rois, deltas, im_shape, scale, thresh=0.05, nms_thresh=thr, max_per_image=cap,
                               with_rows=True)
    check(got, got_rows, want, want_rows)


def test_box_head_cap_keeps_ties_and_python_float_scale():
    """tester.py:60-65 keeps every detection whose score EQUALS the max_per_image-th largest; the Fast R-CNN
    branch divides by a python float scale (tester.py:47)."""
    from mxrcnn_b200.rcnn.tester import detections
    rois, deltas, scores, im_shape, _ = synth.box_head_inputs(48, R=200, G=20, peaked=0.9)
    scores = np.round(scores * 16) / 16  # heavy score ties, also across classes
    scale = 600.0 / 375.0
    for cap in (5, 20, 60):
        want, want_rows = restate.box_head(rois, deltas, scores, im_shape, scale, max_per_image=cap)
        got, got_rows = detections(scores, rois, deltas, im_shape, scale, max_per_image=cap, with_rows=True)
        check(got, got_rows, want, want_rows)
    assert sum(len(d) for d in want[1:]) > 60  # the ties did push the total over the cap


def test_box_head_nothing_above_threshold_and_empty_inputs():
    from mxrcnn_b200.rcnn.tester import detections
    rois, deltas, scores, im_shape, scale = synth.box_head_inputs(49, R=50)
    got = detections(np.full_like(scores, 0.01), rois, deltas, im_shape, scale)
    assert got[0] == [] and all(d.shape == (0, 5) for d in got[1:])
    got = detections(scores[:0], rois[:0], deltas[:0], im_shape, scale)
    assert all(d.shape == (0, 5) for d in got[1:]) and len(got) == scores.shape[1]
    nan = scores.copy()
    nan[::3] = np.nan  # NaN > thresh is False: those rows are simply not detections
    want, want_rows = restate.box_head(rois, deltas, nan, im_shape, scale)
    got, got_rows = detections(nan, rois, deltas, im_shape, scale, with_rows=True)
    check(got, got_rows, want, want_rows)


def test_box_head_batch_and_device_arrays():
    import torch
    from mxrcnn_b200.rcnn.tester import detections_batch, detections_device
    cases = [synth.box_head_inputs(50 + i) for i in range(5)]
    scales = [np.float32(1.6), np.float32(1.0), 2.0, np.float32(0.8), np.float32(1.6)]
    shapes = [(600, 1000), (600, 800), (480, 640), (600, 1000), (562, 1000)]
    rois, deltas, scores = (np.stack([c[k] for c in cases]) for k in range(3))
    got, got_rows = detections_batch(scores, rois, deltas, shapes, scales, with_rows=True)
    wants = []
    for i, c in enumerate(cases):
        want, want_rows = restate.box_head(c[0], c[1], c[2], shapes[i], scales[i])
        check(got[i], got_rows[i], want, want_rows)
        wants.append((want, want_rows))
    dev = torch.device("cuda:0")
    im = torch.tensor([[s[0], s[1], float(sc)] for s, sc in zip(shapes, scales)], dtype=torch.float64, device=dev)
    dets, rows, count = detections_device(torch.from_numpy(scores).to(dev), torch.from_numpy(rois).to(dev),
                                          torch.from_numpy(deltas).to(dev), im)
    dets, rows, count = dets.cpu().numpy(), rows.cpu().numpy(), count.cpu().numpy()
    assert (count[:, 0] == 0).all() and (count[:, :, 0] <= count[:, :, 1]).all()
    for i, (want, want_rows) in enumerate(wants):
        g = [[]] + [dets[i, j, :count[i, j, 0]] for j in range(1, scores.shape[2])]
        r = [[]] + [rows[i, j, :count[i, j, 0]] for j in range(1, scores.shape[2])]
        check(g, r, want, want_rows)


def test_box_head_rejects_bad_shapes():
    from mxrcnn_b200 import _lib
    from mxrcnn_b200.rcnn.tester import detections
    rois, deltas, scores, im_shape, scale = synth.box_head_inputs(51, R=20)
    with pytest.raises(ValueError):
        detections(scores, rois, deltas[:, :-4], im_shape, scale)
    big = np.zeros((4097, 2), np.float32)
    with pytest.raises((ValueError, _lib.MxrError)):
        detections(big, np.zeros((4097, 4), np.float32), np.zeros((4097, 8), np.float32), im_shape, scale)
