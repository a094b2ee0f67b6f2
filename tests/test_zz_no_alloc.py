"""GPU: mxr_nms_noalloc -- the NMS entry point for callers that must not allocate device memory.

Some users of this library cannot call cudaMalloc: a plug-in running inside a framework whose
memory pool already owns the whole GPU, or a serving loop that cannot take the device-wide stall
an allocation causes.  For them the library offers one NMS entry point whose scratch is a fixed
array inside the library itself (2048 boxes at most).  This file checks, through the C ABI, that it
returns exactly what the regular entry point (mxr_nms) returns -- keep lists are index lists, so
the bar is bit-exact equality, as in tests/test_gpu_nms.py -- on two calls in a row, the second one
smaller than the first so that anything the first call left behind in the shared array would show.

The file name sorts last on purpose: with `pytest -x` every other test has already run.
"""
import ctypes

import numpy as np
import pytest

from oracle import restate, synth

pytestmark = pytest.mark.gpu


def _regular(lib, handle, dets_t, dtype_code, thresh, max_keep):
    import torch
    n = dets_t.shape[0]
    keep = torch.full((max(n, 1),), -7, dtype=torch.int32, device=dets_t.device)
    nk = torch.zeros((2,), dtype=torch.int32, device=dets_t.device)
    st = lib.mxr_nms(handle.ptr, None, dets_t.data_ptr(), n, dtype_code, thresh, max_keep, keep.data_ptr(), nk.data_ptr())
    assert st == 0, lib.mxr_last_error_string()
    torch.cuda.synchronize()
    return keep.cpu().numpy()[:int(nk[0])], nk.cpu().numpy()


def _noalloc(lib, dets_t, dtype_code, thresh, max_keep):
    import torch
    n = dets_t.shape[0]
    keep = torch.full((max(n, 1),), -7, dtype=torch.int32, device=dets_t.device)
    nk = torch.zeros((2,), dtype=torch.int32, device=dets_t.device)
    st = lib.mxr_nms_noalloc(None, dets_t.data_ptr(), n, dtype_code, thresh, max_keep, keep.data_ptr(), nk.data_ptr())
    assert st == 0, lib.mxr_last_error_string()
    torch.cuda.synchronize()
    return keep.cpu().numpy()[:int(nk[0])], nk.cpu().numpy()


def test_noalloc_matches_regular_entry_point_on_consecutive_calls():
    import torch
    from mxrcnn_b200 import _lib
    lib = _lib.load()
    torch.cuda.set_device(0)          # the entry point works on the current device
    handle = _lib.Handle(0)           # only the regular entry point needs a handle
    # two calls in a row through the no-alloc entry point: a full-size one, then a small one
    # (f64 as rcnn/tester.py:54-55 produces, f32 as demo.py:57-58 does)
    calls = [(2048, np.float64, 0.3, 0), (300, np.float32, 0.7, 0), (1500, np.float64, 0.5, 100), (1, np.float32, 0.3, 0)]
    for i, (n, dt, thresh, max_keep) in enumerate(calls):
        dets = synth.class_dets(500 + i, n=n, dtype=dt, clusters=max(2, n // 40))
        t = torch.from_numpy(dets).cuda()
        code = 1 if dt == np.float64 else 0
        got, got_nk = _noalloc(lib, t, code, thresh, max_keep)
        want, want_nk = _regular(lib, handle, t, code, thresh, max_keep)
        assert np.array_equal(got_nk, want_nk), (i, got_nk, want_nk)
        assert np.array_equal(got, want), "call %d: no-alloc keep list differs from mxr_nms" % i
        # and both equal the oracle (helper/processing/nms.py restated)
        ref = [int(k) for k in restate.nms(dets, thresh)]
        assert list(got) == (ref[:max_keep] if max_keep else ref)


def test_noalloc_rejects_sizes_beyond_its_fixed_scratch():
    import torch
    from mxrcnn_b200 import _lib
    lib = _lib.load()
    torch.cuda.set_device(0)
    n = 2049
    t = torch.zeros((n, 5), dtype=torch.float64, device="cuda")
    keep = torch.zeros((n,), dtype=torch.int32, device="cuda")
    nk = torch.zeros((2,), dtype=torch.int32, device="cuda")
    st = lib.mxr_nms_noalloc(None, t.data_ptr(), n, 1, 0.5, 0, keep.data_ptr(), nk.data_ptr())
    assert st == -4  # MXR_ERR_LIMIT
    msg = lib.mxr_last_error_string().decode()
    assert "2048" in msg and "mxr_nms" in msg
    with pytest.raises(ValueError):
        _lib.check(st)
    # an empty problem is fine and keeps nothing
    st = lib.mxr_nms_noalloc(None, None, 0, 1, 0.5, 0, None, nk.data_ptr())
    assert st == 0
    torch.cuda.synchronize()
    assert int(nk[0]) == 0
