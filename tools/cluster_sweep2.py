"""Back-to-back time per fused proposal call (CUDA events over 50 calls on one stream, no profiling) as a function of
the batch and the cluster size (MXRCNN_B200_CLUSTER pins it): the table behind kWaveCost in csrc/proposal.cu.

    SWEEP_BATCHES=8,16 SWEEP_SIZES=0,16,10,8 python tools/cluster_sweep2.py [config]
"""
import ctypes
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mxrcnn_b200 import _lib  # noqa: E402
from mxrcnn_b200.helper.processing.generate_anchor import generate_anchors  # noqa: E402
from oracle import synth  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "voc_test"
cfg = synth.CONFIGS[name]
A = len(cfg["scales"]) * len(cfg["ratios"])
batches = [int(x) for x in os.environ.get("SWEEP_BATCHES", "1,4,8,16,32,64,128,256").split(",")]
sizes = os.environ.get("SWEEP_SIZES", "0,16,14,12,10,8,6,4,2").split(",")
dev = torch.device("cuda:0")
base = synth.proposal_batch(list(range(700, 732)), cfg)
anchors = generate_anchors(float(cfg["feat_stride"]), list(cfg["ratios"]), np.array(cfg["scales"], dtype=float))
anc = anchors.ctypes.data_as(_lib.c_double_p)
lib = _lib.load()
post = cfg["post"]
stream = torch.cuda.Stream(dev)
sp = int(stream.cuda_stream)
table = {}
for cs in sizes:
    os.environ["MXRCNN_B200_CLUSTER"] = cs
    h = _lib.Handle(0)
    for B in batches:
        reps = (B + 31) // 32
        c, b, i = (torch.from_numpy(np.concatenate([x] * reps)[:B]).to(dev) for x in base)
        p = _lib.ProposalParams(B, A, cfg["H"], cfg["W"], 0, 0, float(cfg["feat_stride"]), cfg["pre"], post,
                                cfg["thresh"], float(cfg["min_size"]), 0)
        rois = torch.zeros((B * post, 5), dtype=torch.float32, device=dev)
        nk = torch.zeros((B, 2), dtype=torch.int32, device=dev)

        def step():
            _lib.check(lib.mxr_proposal_forward(h.ptr, sp, ctypes.byref(p), c.data_ptr(), b.data_ptr(), i.data_ptr(), anc,
                                                post, rois.data_ptr(), None, None, None, nk.data_ptr()))
        try:
            for _ in range(5):
                step()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            n = 50 if B <= 64 else 20
            e0.record(stream)
            for _ in range(n):
                step()
            e1.record(stream)
            torch.cuda.synchronize()
            table[(cs, B)] = 1e3 * e0.elapsed_time(e1) / n
        except Exception as ex:  # a pinned size that does not fit this map
            table[(cs, B)] = float("nan")
            print("skip", cs, B, ex, file=sys.stderr)
    h.close()
print("| images per call | " + " | ".join("cluster %s" % ("auto" if s == "0" else s) for s in sizes) + " |  (us per call, %s)" % name)
print("|---|" + "---|" * len(sizes))
for B in batches:
    print("| %d | " % B + " | ".join("%.1f" % table[(s, B)] for s in sizes) + " |")
