"""Kernel time of one proposal call for small batches, with 16-CTA clusters allowed (default) and with
MXRCNN_B200_CLUSTER8=1.  Run on a GPU box:  python tools/latency_by_batch.py"""
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
    A = len(cfg["scales"]) * len(cfg["ratios"])
    anchors = generate_anchors(float(cfg["feat_stride"]), list(cfg["ratios"]), np.array(cfg["scales"], dtype=float))
    print("| images per call | 16-CTA clusters allowed (us) | 8-CTA clusters only (us) |")
    print("|---|---|---|")
    rows = {}
    for env in ("0", "1"):
        os.environ["MXRCNN_B200_CLUSTER8"] = env
        h = _lib.Handle(0)
        for B in (1, 2, 3, 4, 5, 6, 8):
            cls, box, info = synth.proposal_batch(range(3, 3 + B), cfg)
            c, b, i = (torch.from_numpy(x).cuda() for x in (cls, box, info))

            def call():
                proposal_forward(c, b, i, anchors, cfg["feat_stride"], cfg["pre"], cfg["post"], cfg["thresh"],
                                 cfg["min_size"], handle=h, rng=np.random.RandomState(0))
            for _ in range(10):
                call()
            h.profile_enable(True)
            for _ in range(100):
                call()
            prof = h.profile_read()
            h.profile_enable(False)
            ms, cnt = prof.get("proposal_fused_kernel") or prof["proposal_cluster_kernel"]
            rows.setdefault(B, []).append(ms / cnt * 1e3)
        h.close()
    for B, (w, n) in sorted(rows.items()):
        print("| %d | %.1f | %.1f |" % (B, w, n))


if __name__ == "__main__":
    main()
