"""Per-kernel times of one proposal call for a named config (default voc_alt).  python tools/phase_probe_alt.py [config]"""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import synth
from mxrcnn_b200 import _lib
from mxrcnn_b200.helper.processing.generate_anchor import generate_anchors
from mxrcnn_b200.rcnn.rpn.proposal import proposal_forward
name = sys.argv[1] if len(sys.argv) > 1 else "voc_alt"
cfg = synth.CONFIGS[name]
anchors = generate_anchors(float(cfg["feat_stride"]), list(cfg["ratios"]), np.array(cfg["scales"], dtype=float))
h = _lib.Handle(0)
c, b, i = (torch.from_numpy(x).cuda() for x in synth.proposal_batch(range(3, 11), cfg))
for it in range(3):
    out = proposal_forward(c, b, i, anchors, cfg["feat_stride"], cfg["pre"], cfg["post"], cfg["thresh"], cfg["min_size"], handle=h, rng=np.random.RandomState(0))
h.profile_enable(True)
for it in range(5):
    out = proposal_forward(c, b, i, anchors, cfg["feat_stride"], cfg["pre"], cfg["post"], cfg["thresh"], cfg["min_size"], handle=h, rng=np.random.RandomState(0))
print({k: (round(v[0] / v[1] * 1e3, 1), v[1]) for k, v in h.profile_read().items()})
print("nkeep", out["nkeep"], "phases", h.proposal_phase_times()[:2])
