"""helper/processing/nms.py of the reference: ``nms(dets, thresh)`` -> python list of kept rows."""
import ctypes

import numpy as np

from ... import _lib


def nms(dets, thresh, max_keep=0, handle=None):
    """Greedy NMS (reference: helper/processing/nms.py:4-38).

    dets: (n,5) [x1,y1,x2,y2,score], numpy float32/float64 (arithmetic in that dtype, like
    numpy) or a device array of those dtypes.  Returns a python list of int row indices in
    keep order (quirk Q6).  Score ties are processed higher-index-first
    (``argsort(kind='stable')[::-1]``); the reference leaves tie order to numpy's unstable sort.
    ``max_keep`` > 0 (extension) stops after that many boxes.
    """
    lib = _lib.load()
    h = _lib.handle_for(handle, dets)
    if _lib.is_device_array(dets):
        v = _lib.device_view(dets)
        if v.dtype not in (np.float32, np.float64) or len(v.shape) != 2 or v.shape[1] != 5:
            raise ValueError("dets must be (n,5) float32/float64")
        n = v.shape[0]
        if n == 0:
            return []
        import torch
        keep = torch.empty((n,), dtype=torch.int32, device=dets.device)
        nk = torch.empty((2,), dtype=torch.int32, device=dets.device)
        stream = _lib.current_stream_ptr(dets)
        _lib.check(lib.mxr_nms(h.ptr, stream, v.ptr, n, 1 if v.dtype == np.float64 else 0, float(thresh),
                               int(max_keep), keep.data_ptr(), nk.data_ptr()))
        nk_h = nk.cpu()
        if int(nk_h[1]) & 1:
            raise ValueError("nms: NaN in dets")
        return [int(i) for i in keep[:int(nk_h[0])].cpu().numpy()]
    dets = np.asarray(dets)
    if dets.ndim != 2 or dets.shape[1] < 5:
        raise ValueError("dets must be (n,5)")
    if dets.dtype != np.float32:
        dets = dets.astype(np.float64, copy=False)
    d = np.ascontiguousarray(dets[:, :5])
    n = d.shape[0]
    if n == 0:
        return []
    keep = np.empty((n,), dtype=np.int32)
    nk = np.zeros((2,), dtype=np.int32)
    _lib.check(lib.mxr_nms_host(h.ptr, _lib.hptr(d), n, 1 if d.dtype == np.float64 else 0, float(thresh),
                                int(max_keep), _lib.hptr(keep), _lib.hptr(nk)))
    return [int(i) for i in keep[:int(nk[0])]]


def nms_batched(dets_list, thresh, max_keep=0, handle=None):
    """Extension: the per-class loop of rcnn/tester.py:50-56 (``nms(cls_dets, TEST.NMS)`` for every class)
    as ONE launch.  ``dets_list``: list of (n_c,5) numpy arrays of one dtype (float32 or float64).
    Returns a list of python lists, element c equal to ``nms(dets_list[c], thresh)``."""
    lib = _lib.load()
    h = _lib.handle_for(handle, dets_list)
    if len(dets_list) == 0:
        return []
    dt = np.float32 if all(np.asarray(d).dtype == np.float32 for d in dets_list) else np.float64
    arrs = [np.ascontiguousarray(np.asarray(d)[:, :5], dtype=dt).reshape(-1, 5) for d in dets_list]
    offsets = np.zeros((len(arrs) + 1,), dtype=np.int32)
    offsets[1:] = np.cumsum([a.shape[0] for a in arrs])
    total = int(offsets[-1])
    if total == 0:
        return [[] for _ in arrs]
    flat = np.ascontiguousarray(np.concatenate(arrs, axis=0))
    keep_h = np.empty((total,), dtype=np.int32)
    nk_h = np.zeros((len(arrs), 2), dtype=np.int32)
    _lib.check(lib.mxr_nms_batched_host(h.ptr, _lib.hptr(flat), offsets.ctypes.data_as(_lib.c_int_p), len(arrs),
                                        1 if dt == np.float64 else 0, float(thresh), int(max_keep), _lib.hptr(keep_h),
                                        _lib.hptr(nk_h)))
    if (nk_h[:, 1] & 1).any():
        raise ValueError("nms: NaN in dets")
    return [[int(i) for i in keep_h[offsets[c]:offsets[c] + nk_h[c, 0]]] for c in range(len(arrs))]


def nest(dets, thresh=0.90):
    """Drop every box that lies (almost) inside another one: box i goes when inter(i,j)/area(i) > thresh
    for some j != i (reference: helper/processing/nms.py:40-70, called from predict.py:75).
    Returns the python list of kept row indices, ascending."""
    lib, h = _lib.load(), _lib.handle_for(None, dets)
    d = np.asarray(dets)
    if d.ndim != 2 or d.shape[1] < 4:
        raise ValueError("dets must be (n, >=4)")
    if d.dtype != np.float32:
        d = d.astype(np.float64, copy=False)
    n = d.shape[0]
    if n == 0:
        return []
    buf = np.zeros((n, 5), dtype=d.dtype)
    buf[:, :4] = d[:, :4]
    flags = np.zeros((n,), dtype=np.int32)
    _lib.check(lib.mxr_nest_host(h.ptr, _lib.hptr(buf), n, 1 if buf.dtype == np.float64 else 0, float(thresh),
                                 _lib.hptr(flags)))
    return [int(i) for i in np.where(flags != 0)[0]]
