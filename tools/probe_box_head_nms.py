import os, sys
sys.path.insert(0, os.getcwd())
import numpy as np
from oracle import synth
from mxrcnn_b200.rcnn.tester import detections
from mxrcnn_b200.helper.processing.nms import nms
r, d, sc, shp, scl = synth.box_head_inputs(7)
d6000 = synth.class_dets(905, n=6000, dtype=np.float64, clusters=150)
for _ in range(6):
    detections(sc, r, d, shp, scl)
    nms(d6000, 0.7)
