"""Wall time of the training-target generators through their Python mirrors (host arrays in and out), the batched
loader hand-off (device tensors out), and the kernels' own time from the library's per-launch CUDA events.
Usage: python tools/targets_probe.py"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mxrcnn_b200 import _lib  # noqa: E402
from mxrcnn_b200.rcnn.loader import anchor_labels  # noqa: E402
from mxrcnn_b200.rcnn.minibatch import assign_anchor  # noqa: E402
from mxrcnn_b200.rcnn.rpn.proposal_target import _sample_rois  # noqa: E402
from oracle import synth  # noqa: E402

info = np.array([[600, 1000, 1.6]], np.float32)
gt64 = synth.gt_boxes(4, G=64)
gt8 = synth.gt_boxes(5, G=8)
allr = np.vstack((synth.jittered_rois(5, gt8, R=2000), np.hstack((np.zeros((8, 1), np.float32), gt8[:, :4]))))
gts8 = [synth.gt_boxes(40 + i, G=g) for i, g in enumerate((64, 6, 1, 40, 12, 3, 25, 7))]
infos8 = [info] * 8
h = _lib.default_handle() if hasattr(_lib, "default_handle") else _lib.handle_for(None)
cases = ((lambda: assign_anchor((1, 18, 38, 63), gt64, info), "assign_anchor (64 gt, host arrays)", 1),
         (lambda: _sample_rois(allr, gt8, 32, 128, 21, "TRAIN"), "_sample_rois (2008 rois, host arrays)", 1),
         (lambda: anchor_labels((1, 18, 38, 63), gts8, infos8, device="cuda:0"), "anchor_labels (8 images, device tensors)", 8))
for fn, name, per in cases:
    for _ in range(5):
        fn()
    t0 = time.perf_counter()
    for _ in range(50):
        fn()
    wall = (time.perf_counter() - t0) / 50 * 1e6
    h.profile_enable(True)
    for _ in range(10):
        fn()
    prof = h.profile_read()
    h.profile_enable(False)
    kern = {k: (round(ms * 1e3 / 10, 1), cnt // 10) for k, (ms, cnt) in prof.items()}
    print("%s: %.1f us per call (%.1f us per image); kernels per call (us, launches): %s" % (name, wall, wall / per, kern))
