"""Mean phase times of the fused proposal kernel (its own %globaltimer stamps) on one batch of 8 L2-resident images.
Used with temporary extra stamps to time sub-phases while tuning; python tools/phase_probe.py"""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import synth
from mxrcnn_b200 import _lib
from mxrcnn_b200.helper.processing.generate_anchor import generate_anchors
from mxrcnn_b200.rcnn.rpn.proposal import proposal_forward
cfg = synth.CONFIGS["voc_test"]
anchors = generate_anchors(16.0, list(cfg["ratios"]), np.array(cfg["scales"], dtype=float))
h = _lib.Handle(0)
h.profile_enable(True)
c, b, i = (torch.from_numpy(x).cuda() for x in synth.proposal_batch(range(3, 11), cfg))
acc = {}
for it in range(30):
    try:
        proposal_forward(c, b, i, anchors, 16, cfg["pre"], cfg["post"], cfg["thresh"], cfg["min_size"], handle=h, rng=np.random.RandomState(0))
    except Exception as e:
        pass
    if it >= 5:
        for p_ in h.proposal_phase_times():
            for k, v in p_.items():
                acc.setdefault(k, []).append(v)
print({k: round(float(np.mean(v)), 2) for k, v in acc.items()})
