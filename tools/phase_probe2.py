"""Per-phase stamps of the fused proposal kernel (proposal_fused.cuh) for one workload: mean over the images of a batch.

    python tools/phase_probe2.py [config] [batch]
"""
import ctypes
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from mxrcnn_b200 import _lib  # noqa: E402
from mxrcnn_b200.helper.processing.generate_anchor import generate_anchors  # noqa: E402
from oracle import synth  # noqa: E402

name = sys.argv[1] if len(sys.argv) > 1 else "voc_test"
B = int(sys.argv[2]) if len(sys.argv) > 2 else 8
cfg = synth.CONFIGS[name]
A = len(cfg["scales"]) * len(cfg["ratios"])
dev = torch.device("cuda:0")
cls, box, info = synth.proposal_batch(list(range(500, 500 + B)), cfg)
cls_d, box_d, info_d = (torch.from_numpy(x).to(dev) for x in (cls, box, info))
lib = _lib.load()
h = _lib.Handle(0)
anchors = generate_anchors(float(cfg["feat_stride"]), list(cfg["ratios"]), np.array(cfg["scales"], dtype=float))
anc = anchors.ctypes.data_as(_lib.c_double_p)
post = cfg["post"]
p = _lib.ProposalParams(B, A, cfg["H"], cfg["W"], 0, 0, float(cfg["feat_stride"]), cfg["pre"], post, cfg["thresh"],
                        float(cfg["min_size"]), 0)
rois = torch.zeros((B * post, 5), dtype=torch.float32, device=dev)
nk = torch.zeros((B, 2), dtype=torch.int32, device=dev)
stream = torch.cuda.Stream(dev)
sp = int(stream.cuda_stream)


def step():
    _lib.check(lib.mxr_proposal_forward(h.ptr, sp, ctypes.byref(p), cls_d.data_ptr(), box_d.data_ptr(), info_d.data_ptr(),
                                        anc, post, rois.data_ptr(), None, None, None, nk.data_ptr()))


for _ in range(5):
    step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record(stream)
for _ in range(50):
    step()
e1.record(stream)
torch.cuda.synchronize()
print("%s batch %d: %.2f us per call, nkeep min %d" % (name, B, 1e3 * e0.elapsed_time(e1) / 50, int(nk[:, 0].min())))
h.profile_enable(True)
acc = []
for _ in range(5):
    step()
    acc.append(h.proposal_stamps(B).astype(np.int64))
h.profile_enable(False)
s = np.mean(np.stack(acc[1:]), axis=0)  # (B,16)
names = ["start", "start2", "decode+barrier", "coarse table", "window 0 (fine table, scatter, barrier)", "-", "-",
         "NMS total", "outputs"]
m = s.mean(axis=0)
for i in range(1, 9):
    print("  %-42s %7.2f us" % (names[i], (m[i] - m[i - 1]) * 1e-3))
d = s[:, [10, 11, 12, 13]] - s[:, [6, 10, 11, 12]]
for nm, v in zip(("first stripe: staged", "bit blocks", "barrier", "resolve"), d.mean(axis=0)):
    print("    %-40s %7.2f us" % (nm, v * 1e-3))
clk = s[:, 16:48]
cn = ["loop top", "planned", "loads issued", "sync1 (comps in smem)", "ranked", "decoded+stored", "sync2", "my blocks done",
      "sync (CTA blocks done)", "barrier A passed", "keep init sync", "nz masks", "round0", "round1", "round2", "round3", "round4",
      "round5", "round6", "round7", "rounds done", "kept stored", "status pushed", "barrier B passed"]
prev = None
for i, nm in enumerate(cn):
    v = clk[:, i]
    if (v == 0).all():
        continue
    if prev is not None:
        print("      %-28s +%7.0f cycles" % (nm, (v - clk[:, prev]).mean()))
    prev = i
print("  stripes %.1f windows %.1f fixed-point rounds (first stripe) %.1f" % (m[9], m[14], m[15]))
h.count_pairs(True)
step()
h.count_pairs(False)
pr = h.pairs_read(B)
print("  IoU pairs evaluated per image: %.0f, of which settled in float64: %.1f" % (pr[:, 0].mean(), pr[:, 1].mean()))
