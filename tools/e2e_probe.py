"""End-to-end (host buffers in, host results out) rate of the proposal op through the two-slot host API, with the inputs
as the reference lays them out (three arrays, foreground half strided) and packed by the caller into one block; next to
what the host link gives a plain pinned copy.   python tools/e2e_probe.py [steps]"""
import ctypes
import os
import sys
import time

import numpy as np
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from mxrcnn_b200 import _lib  # noqa: E402
from mxrcnn_b200.helper.processing.generate_anchor import generate_anchors  # noqa: E402
from oracle import synth  # noqa: E402

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 300
cfg = synth.CONFIGS["voc_test"]
A, B, post = 9, 8, cfg["post"]
H, W = cfg["H"], cfg["W"]
dev = torch.device("cuda:0")
lib = _lib.load()
h = _lib.Handle(0)
anchors = generate_anchors(16.0, [0.5, 1, 2], np.array([8.0, 16.0, 32.0]))
anc = anchors.ctypes.data_as(_lib.c_double_p)
p = _lib.ProposalParams(B, A, H, W, 0, 0, 16.0, cfg["pre"], post, cfg["thresh"], float(cfg["min_size"]), 0)
pool = 40
cls_h = torch.empty((pool, B, 2 * A, H, W), dtype=torch.float32).pin_memory()
box_h = torch.empty((pool, B, 4 * A, H, W), dtype=torch.float32).pin_memory()
info_h = torch.empty((pool, B, 3), dtype=torch.float32).pin_memory()
n_fg, n_box = B * A * H * W, B * 4 * A * H * W
pack_h = torch.empty((pool, n_fg + n_box + B * 3), dtype=torch.float32).pin_memory()
for i in range(pool):
    c, b_, inf = synth.proposal_batch([5000 + i * B + j for j in range(B)], cfg)
    cls_h[i].copy_(torch.from_numpy(c))
    box_h[i].copy_(torch.from_numpy(b_))
    info_h[i].copy_(torch.from_numpy(inf))
    pack_h[i, :n_fg] = torch.from_numpy(np.ascontiguousarray(c[:, A:])).reshape(-1)
    pack_h[i, n_fg:n_fg + n_box] = torch.from_numpy(b_).reshape(-1)
    pack_h[i, n_fg + n_box:] = torch.from_numpy(inf).reshape(-1)
cls_np, box_np, info_np, pack_np = cls_h.numpy(), box_h.numpy(), info_h.numpy(), pack_h.numpy()
outs = [(np.empty((B * post, 5), np.float32), np.empty((B * post,), np.float32), np.empty((B, post), np.int32),
         np.zeros((B, 2), np.int32)) for _ in range(2)]
lib.mxr_reserve_proposal(h.ptr, ctypes.byref(p))


def submit(i, packed):
    j = i % pool
    if packed:
        st = lib.mxr_proposal_submit_host_packed(h.ptr, i % 2, ctypes.byref(p), _lib.hptr(pack_np[j]), anc, post)
    else:
        st = lib.mxr_proposal_submit_host(h.ptr, i % 2, ctypes.byref(p), _lib.hptr(cls_np[j]), _lib.hptr(box_np[j]),
                                          _lib.hptr(info_np[j]), anc, post)
    _lib.check(st)


def wait(i):
    o = outs[i % 2]
    _lib.check(lib.mxr_proposal_wait_host(h.ptr, i % 2, _lib.hptr(o[0]), _lib.hptr(o[1]), _lib.hptr(o[2]), None, _lib.hptr(o[3])))
    assert o[3][0, 0] == post


def run(first, count, packed):
    submit(first, packed)
    for i in range(first + 1, first + count):
        submit(i, packed)
        wait(i - 1)
    wait(first + count - 1)


h2d = (n_fg + n_box + B * 3) * 4
keeps = {}
for name, packed in (("three arrays (reference layout)", False), ("caller-packed block", True), ("three arrays again", False),
                     ("caller-packed again", True)):
    run(0, 10, packed)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    run(10, steps, packed)
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    keeps[packed] = outs[(10 + steps - 1) % 2][2].copy()
    print("%-34s %8.0f images/s  %6.2f us/step  H2D %.1f GB/s" % (name, B * steps / dt, 1e6 * dt / steps, h2d * steps / dt / 1e9))
assert np.array_equal(keeps[False], keeps[True])
st = torch.cuda.Stream(dev)
for nbytes in (h2d, 16 << 20, 64 << 20):
    pin = torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    dst = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    with torch.cuda.stream(st):
        for _ in range(3):
            dst.copy_(pin, non_blocking=True)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = max(4, int(400e6 / nbytes))
        e0.record(st)
        for _ in range(n):
            dst.copy_(pin, non_blocking=True)
        e1.record(st)
    st.synchronize()
    print("plain pinned copy of %9d bytes: %.1f GB/s" % (nbytes, nbytes * n / (e0.elapsed_time(e1) * 1e-3) / 1e9))
