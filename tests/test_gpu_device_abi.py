"""GPU: the device-pointer flavours of the C ABI (the ones a framework with its own device tensors
binds), called through ctypes with torch tensors, against the host-array flavours."""
import ctypes

import numpy as np
import pytest

from oracle import synth
from tests.util import golden, proposal_case

pytestmark = pytest.mark.gpu


def test_proposal_operator_with_device_tensors():
    import torch
    from mxrcnn_b200 import operator as mxop
    from mxrcnn_b200.config import config
    import mxrcnn_b200.rcnn.rpn.proposal  # noqa: F401
    g = golden("proposal_voc_test_s1")
    cls, box, info, cfg, _ = proposal_case(g)
    config.TEST.RPN_PRE_NMS_TOP_N, config.TEST.RPN_POST_NMS_TOP_N = cfg["pre"], cfg["post"]
    _, op = mxop.create("proposal", feat_stride=16, scales=(8, 16, 32), ratios=(0.5, 1, 2), output_score=True)
    dev = torch.device("cuda:0")
    ins = [torch.from_numpy(a).to(dev) for a in (cls, box, info)]
    outs = [torch.zeros((300, 5), device=dev), torch.zeros((300, 1), device=dev)]
    op.forward(False, ["write", "write"], ins, outs, [])      # written in place, no host round trip of the maps
    assert np.array_equal(outs[1].cpu().numpy(), g["scores"])
    np.testing.assert_allclose(outs[0].cpu().numpy(), g["rois"], rtol=1e-5, atol=1e-4)
    # req='add' goes through assign()
    acc = [torch.ones((300, 5), device=dev), torch.ones((300, 1), device=dev)]
    op.forward(False, ["add", "add"], ins, acc, [])
    assert np.array_equal(acc[1].cpu().numpy(), g["scores"] + 1)


def test_anchor_target_phase2_device_outputs():
    import torch
    from mxrcnn_b200 import _lib
    from mxrcnn_b200.helper.processing.generate_anchor import generate_anchors
    from mxrcnn_b200.rcnn.minibatch import assign_anchor
    lib, h = _lib.load(), _lib.Handle(0)
    info = np.array([[300, 500, 0.8]], np.float32)
    gt = synth.gt_boxes(33, G=7, im_h=300, im_w=500)
    np.random.seed(33)
    want = assign_anchor((1, 18, 19, 31), gt, info)           # host flavour, consumes the RNG
    base = generate_anchors(16, [0.5, 1, 2], np.array((8, 16, 32), np.float32))
    p = _lib.AnchorTargetParams(19, 31, 9, 16, 0, 300.0, 500.0, 7, 0.3, 0.7, 0)
    counts = np.zeros(3, np.int32)
    _lib.check(lib.mxr_anchor_target_phase1(h.ptr, ctypes.byref(p), base.ctypes.data_as(_lib.c_double_p), _lib.hptr(gt),
                                            counts.ctypes.data_as(_lib.c_int_p)))
    n_in, n_fg, n_bg = (int(c) for c in counts)
    np.random.seed(33)                                        # same draws as the host flavour made
    empty = np.zeros(0, np.int32)                             # a draw happens only when over quota (minibatch.py:321,330)
    fg = np.ascontiguousarray(np.random.permutation(n_fg)[:n_fg - 128], dtype=np.int32) if n_fg > 128 else empty
    nfg = n_fg - fg.size
    bg = np.ascontiguousarray(np.random.permutation(n_bg)[:n_bg - (256 - nfg)], dtype=np.int32) if n_bg > 256 - nfg else empty
    w = 1.0 / (nfg + n_bg - bg.size)
    iw = np.ones(4, np.float32)
    N = 9 * 19 * 31
    dev = torch.device("cuda:0")
    lab = torch.empty((1, N), device=dev)
    maps = [torch.empty((1, 36, 19, 31), device=dev) for _ in range(3)]
    _lib.check(lib.mxr_anchor_target_phase2(h.ptr, None, fg.ctypes.data_as(_lib.c_int_p), fg.size,
                                            bg.ctypes.data_as(_lib.c_int_p), bg.size, w, w, iw.ctypes.data_as(_lib.c_float_p),
                                            lab.data_ptr(), maps[0].data_ptr(), maps[1].data_ptr(), maps[2].data_ptr()))
    torch.cuda.synchronize()
    assert np.array_equal(lab.cpu().numpy(), want["label"])
    for t, k in zip(maps, ("bbox_target", "bbox_inside_weight", "bbox_outside_weight")):
        assert np.array_equal(t.cpu().numpy(), want[k]), k
    h.close()


def test_device_flavours_of_lists_gather_and_nest():
    import torch
    from mxrcnn_b200 import _lib
    from mxrcnn_b200.helper.processing.nms import nest
    lib, h = _lib.load(), _lib.Handle(0)
    dev = torch.device("cuda:0")
    ov = np.random.RandomState(3).uniform(0, 1, 5000)
    for dt, code in ((np.float64, 1), (np.float32, 0)):
        o = ov.astype(dt)
        t = torch.from_numpy(o).to(dev)
        fg = torch.empty(5000, dtype=torch.int32, device=dev)
        bg = torch.empty(5000, dtype=torch.int32, device=dev)
        cnt = torch.zeros(2, dtype=torch.int32, device=dev)
        _lib.check(lib.mxr_fg_bg_lists(h.ptr, None, t.data_ptr(), code, 5000, 0.5, 0.5, 0.1, fg.data_ptr(), bg.data_ptr(),
                                       cnt.data_ptr()))
        c = cnt.cpu().numpy()
        th = np.array([0.5, 0.1], dtype=dt)                   # numpy compares in the array's dtype
        assert np.array_equal(fg[:c[0]].cpu().numpy(), np.where(o >= th[0])[0])
        assert np.array_equal(bg[:c[1]].cpu().numpy(), np.where((o < th[0]) & (o >= th[1]))[0])
    dets = synth.class_dets(8, n=400, dtype=np.float64)
    flags = torch.zeros(400, dtype=torch.int32, device=dev)
    _lib.check(lib.mxr_nest(h.ptr, None, torch.from_numpy(dets).to(dev).data_ptr(), 400, 1, 0.9, flags.data_ptr()))
    assert [int(i) for i in np.where(flags.cpu().numpy() != 0)[0]] == nest(dets, 0.9)
    h.close()
