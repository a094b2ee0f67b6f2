import sys, os, time
sys.path.insert(0, os.getcwd())
import numpy as np
from oracle import synth
from mxrcnn_b200.rcnn.minibatch import assign_anchor
from mxrcnn_b200.rcnn.rpn.proposal_target import _sample_rois
info = np.array([[600, 1000, 1.6]], np.float32)
gt64 = synth.gt_boxes(4, G=64)
gt8 = synth.gt_boxes(5, G=8)
allr = np.vstack((synth.jittered_rois(5, gt8, R=2000), np.hstack((np.zeros((8, 1), np.float32), gt8[:, :4]))))
for _ in range(5):
    assign_anchor((1, 18, 38, 63), gt64, info)
    _sample_rois(allr, gt8, 32, 128, 21, "TRAIN")
for fn, name in ((lambda: assign_anchor((1, 18, 38, 63), gt64, info), "assign_anchor"), (lambda: _sample_rois(allr, gt8, 32, 128, 21, "TRAIN"), "_sample_rois")):
    t0 = time.perf_counter()
    for _ in range(50):
        fn()
    print(name, "%.1f us per call (host arrays)" % ((time.perf_counter() - t0) / 50 * 1e6))
