"""Time of K back-to-back proposal calls (CUDA events, stream idle at the start) as a function of K: T(K) = a + b K.
a = what a short timed region pays once (first launch reaching an idle GPU, no predecessor to overlap with)."""
import ctypes
import os
import sys

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from mxrcnn_b200 import _lib  # noqa: E402
from oracle import synth  # noqa: E402

cfg = synth.CONFIGS["voc_test"]
dev = torch.device("cuda:0")
c, b, i = synth.proposal_batch(list(range(8)), cfg)
lib = _lib.load()
h = _lib.Handle(0)
r = bench.ProposalRunner(lib, _lib, h, dev, cfg, 8, torch.from_numpy(c).to(dev)[None], torch.from_numpy(b).to(dev)[None],
                         torch.from_numpy(i).to(dev)[None])
for _ in range(2000):
    r.step(0)
torch.cuda.synchronize()
for K in (1, 2, 5, 10, 20, 50, 200):
    ts = []
    for rep in range(7):
        torch.cuda.synchronize()
        ts.append(r.timed(0, K) * K * 1e3)
    spaced = []
    for rep in range(7):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(r.stream):
            torch.cuda._sleep(200000)  # ~100 us of busy GPU: the launches below queue up behind it
        e0.record(r.stream)
        for j in range(K):
            r.step(j)
        e1.record(r.stream)
        r.stream.synchronize()
        spaced.append(e0.elapsed_time(e1) * 1e3)
    print("K=%3d  from idle: median %.1f us total, %.2f us/step   | queued behind a busy GPU: %.1f us total, %.2f us/step"
          % (K, np.median(ts), np.median(ts) / K, np.median(spaced), np.median(spaced) / K))
