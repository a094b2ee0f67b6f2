"""Kernel time of one fused proposal call as a function of the batch and of the cluster size
(MXRCNN_B200_CLUSTER), to pick the per-call choice in csrc/proposal.cu.  python tools/cluster_sweep.py [config]"""
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import synth  # noqa: E402  (synthetic inputs only)
from mxrcnn_b200 import _lib  # noqa: E402
from mxrcnn_b200.helper.processing.generate_anchor import generate_anchors  # noqa: E402
from mxrcnn_b200.rcnn.rpn.proposal import proposal_forward  # noqa: E402


def main():
    cfgname = sys.argv[1] if len(sys.argv) > 1 else "voc_test"
    cfg = synth.CONFIGS[cfgname]
    anchors = generate_anchors(float(cfg["feat_stride"]), list(cfg["ratios"]), np.array(cfg["scales"], dtype=float))
    batches = tuple(int(b) for b in os.environ.get("SWEEP_BATCHES", "1,4,8,16,18,32,37,64,74,128,256").split(","))
    base = synth.proposal_batch(range(3, 3 + 32), cfg)
    sizes = tuple(os.environ.get("SWEEP_SIZES", "0,2,4,8,16").split(","))
    table = {}
    for cs in sizes:
        os.environ["MXRCNN_B200_CLUSTER"] = cs
        h = _lib.Handle(0)
        for B in batches:
            reps = (B + 31) // 32
            c, b, i = (torch.from_numpy(np.concatenate([x] * reps)[:B]).cuda() for x in base)

            def call():
                proposal_forward(c, b, i, anchors, cfg["feat_stride"], cfg["pre"], cfg["post"], cfg["thresh"],
                                 cfg["min_size"], handle=h, rng=np.random.RandomState(0))
            for _ in range(3):
                call()
            h.profile_enable(True)
            for _ in range(20):
                call()
            prof = h.profile_read()
            h.profile_enable(False)
            ms, cnt = prof.get("proposal_fused_kernel") or prof["proposal_cluster_kernel"]
            table[(cs, B)] = ms / cnt * 1e3
        h.close()
    print("| images per call | " + " | ".join("cluster %s" % ("auto" if s == "0" else s) for s in sizes) + " |  (us per call)")
    print("|---|" + "---|" * len(sizes))
    for B in batches:
        print("| %d | " % B + " | ".join("%.1f" % table[(s, B)] for s in sizes) + " |")


if __name__ == "__main__":
    main()
