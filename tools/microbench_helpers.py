"""Latency of the helper entry points through their Python mirrors (numpy in / numpy out), next to the
CPU restatement.  Not part of bench.py's contract; used to steer optimisation."""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

from oracle import restate, synth  # noqa: E402


def timeit(fn, n=30, warm=3):
    for _ in range(warm):
        fn()
    t0 = time.perf_counter()
    for _ in range(n):
        fn()
    return (time.perf_counter() - t0) / n * 1e6


def main():
    from mxrcnn_b200.helper.processing.bbox_regression import bbox_overlaps
    from mxrcnn_b200.helper.processing.bbox_transform import bbox_pred, clip_boxes
    from mxrcnn_b200.helper.processing.nms import nms
    from mxrcnn_b200.rcnn.minibatch import assign_anchor
    from mxrcnn_b200.rcnn.rpn.proposal_target import _sample_rois
    rows = []
    for n in (300, 2000, 6000):
        d = synth.class_dets(1, n=n)
        rows.append(("nms f64 n=%d" % n, timeit(lambda: nms(d, 0.3)), timeit(lambda: restate.nms(d, 0.3), n=3, warm=1)))
    gt = synth.gt_boxes(11, G=8)
    rois = synth.jittered_rois(11, gt, R=300)[:, 1:5]
    deltas = np.random.RandomState(0).normal(0, 0.2, size=(300, 84)).astype(np.float32)
    rows.append(("bbox_pred+clip 300x84", timeit(lambda: clip_boxes(bbox_pred(rois, deltas), (600, 1000))),
                 timeit(lambda: restate.clip_boxes(restate.bbox_pred(rois, deltas), (600, 1000)))))
    gt64 = synth.gt_boxes(4, G=64)
    info = np.array([[600, 1000, 1.6]], np.float32)
    rows.append(("assign_anchor 38x63 G=64", timeit(lambda: assign_anchor((1, 18, 38, 63), gt64, info), n=10),
                 timeit(lambda: restate.assign_anchor((1, 18, 38, 63), gt64, info), n=3, warm=1)))
    anc = np.random.RandomState(1).uniform(0, 900, size=(21546, 4))
    anc[:, 2:] += anc[:, :2]
    rows.append(("bbox_overlaps 21546x64", timeit(lambda: bbox_overlaps(anc, gt64.astype(np.float64)), n=10),
                 timeit(lambda: restate.bbox_overlaps(anc, gt64.astype(np.float64)), n=3, warm=1)))
    allr = np.vstack((synth.jittered_rois(5, gt, R=2000), np.hstack((np.zeros((8, 1), np.float32), gt[:, :4]))))
    rows.append(("_sample_rois 2008x8", timeit(lambda: _sample_rois(allr, gt, 32, 128, 21, "TRAIN"), n=10),
                 timeit(lambda: restate.sample_rois_e2e(allr, gt, 32, 128, 21, "TRAIN"), n=3, warm=1)))
    from mxrcnn_b200 import _lib
    from mxrcnn_b200.rcnn.tester import detections, detections_batch
    for tag, kw in (("box head 300x21", {}), ("box head 2000x2 (wider)", dict(R=2000, C=2, G=40, peaked=0.7, im_h=1200, im_w=1600, scale=1.0))):
        r, d, sc, shp, scl = synth.box_head_inputs(7, **kw)
        rows.append((tag, timeit(lambda: detections(sc, r, d, shp, scl)), timeit(lambda: restate.box_head(r, d, sc, shp, scl), n=3, warm=1)))
    cases = [synth.box_head_inputs(60 + i) for i in range(8)]
    r8, d8, s8 = (np.stack([c[k] for c in cases]) for k in range(3))
    rows.append(("box head 8 x 300x21", timeit(lambda: detections_batch(s8, r8, d8, [(600, 1000)] * 8, [np.float32(1.6)] * 8)),
                 8 * rows[-2][2]))
    h = _lib.default_handle()
    h.profile_enable(True)
    r, d, sc, shp, scl = synth.box_head_inputs(7, R=2000, C=2, G=40, peaked=0.7, im_h=1200, im_w=1600, scale=1.0)
    for tag, fn in (("1 image 300x21", lambda: detections(cases[0][2], cases[0][0], cases[0][1], (600, 1000), np.float32(1.6))),
                    ("8 images 300x21", lambda: detections_batch(s8, r8, d8, [(600, 1000)] * 8, [np.float32(1.6)] * 8)),
                    ("1 image 2000x2", lambda: detections(sc, r, d, shp, scl))):
        for _ in range(10):
            fn()
        ms, cnt = h.profile_read()["box_head_kernel"]
        print("box_head_kernel alone, %s: %.1f us" % (tag, ms / cnt * 1e3))
    h.profile_enable(False)
    print("%-28s %12s %16s" % ("call (host arrays)", "GPU path us", "numpy restate us"))
    for name, a, b in rows:
        print("%-28s %12.1f %16.1f" % (name, a, b))


if __name__ == "__main__":
    main()
