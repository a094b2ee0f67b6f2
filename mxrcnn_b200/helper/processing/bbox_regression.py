"""helper/processing/bbox_regression.py of the reference (hot-path functions), same signatures."""
import numpy as np

from ... import _lib
from ...config import config
from .bbox_transform import bbox_transform  # noqa: F401  (the reference re-exports it the same way)


def _overlaps_call(boxes, query_boxes, want_matrix, want_reductions):
    lib, h = _lib.load(), _lib.handle_for(None, boxes)
    if _lib.is_device_array(boxes):
        import torch
        b = boxes.double().contiguous()
        q = query_boxes.double().contiguous()
        n, k = b.shape[0], q.shape[0]
        dev = b.device
        out = torch.zeros((n, k), dtype=torch.float64, device=dev) if want_matrix else None
        arg = torch.zeros((n,), dtype=torch.int32, device=dev) if want_reductions else None
        rmax = torch.zeros((n,), dtype=torch.float64, device=dev) if want_reductions else None
        cmax = torch.zeros((k,), dtype=torch.float64, device=dev) if want_reductions else None
        _lib.check(lib.mxr_bbox_overlaps(
            h.ptr, _lib.current_stream_ptr(b), b.data_ptr(), n, b.shape[1], q.data_ptr(), k, q.shape[1],
            out.data_ptr() if want_matrix else None, arg.data_ptr() if want_reductions else None,
            rmax.data_ptr() if want_reductions else None, cmax.data_ptr() if want_reductions else None))
        return out, arg, rmax, cmax
    b = np.ascontiguousarray(boxes, dtype=np.float64)
    q = np.ascontiguousarray(query_boxes, dtype=np.float64)
    n, k = b.shape[0], q.shape[0]
    out = np.zeros((n, k), dtype=np.float64) if want_matrix else None
    arg = np.zeros((n,), dtype=np.int32) if want_reductions else None
    rmax = np.zeros((n,), dtype=np.float64) if want_reductions else None
    cmax = np.zeros((k,), dtype=np.float64) if want_reductions else None
    if n and k:
        _lib.check(lib.mxr_bbox_overlaps_host(h.ptr, _lib.hptr(b), n, b.shape[1], _lib.hptr(q), k, q.shape[1],
                                              _lib.hptr(out), _lib.hptr(arg), _lib.hptr(rmax), _lib.hptr(cmax)))
    return out, arg, rmax, cmax


def bbox_overlaps(boxes, query_boxes):
    """(n,k) float64 IoU with +1 widths; only columns 0..3 of either input are read
    (reference: bbox_regression.py:11-31)."""
    return _overlaps_call(boxes, query_boxes, True, False)[0]


def bbox_overlaps_reduced(boxes, query_boxes):
    """Extension: (row_argmax, row_max, col_max) without materialising the (n,k) matrix -- what
    rcnn/minibatch.py:296-300 and rcnn/rpn/proposal_target.py:143-144 take from it."""
    return _overlaps_call(boxes, query_boxes, False, True)[1:]


def expand_bbox_regression_targets(bbox_targets_data, num_classes):
    """(k,5) [cls,dx,dy,dw,dh] -> ((k,4C) targets, (k,4C) inside weights), float32
    (reference: bbox_regression.py:67-85)."""
    lib, h = _lib.load(), _lib.handle_for(None, bbox_targets_data)
    iw = np.ascontiguousarray(config.TRAIN.BBOX_INSIDE_WEIGHTS, dtype=np.float32)
    if _lib.is_device_array(bbox_targets_data):
        import torch
        d = bbox_targets_data.float().contiguous()
        k = d.shape[0]
        t = torch.zeros((k, 4 * num_classes), dtype=torch.float32, device=d.device)
        w = torch.zeros((k, 4 * num_classes), dtype=torch.float32, device=d.device)
        _lib.check(lib.mxr_expand_bbox_targets(h.ptr, _lib.current_stream_ptr(d), d.data_ptr(), k, int(num_classes),
                                               iw.ctypes.data_as(_lib.c_float_p), t.data_ptr(), w.data_ptr()))
        return t, w
    d = np.ascontiguousarray(bbox_targets_data, dtype=np.float32)
    k = d.shape[0]
    t = np.zeros((k, 4 * num_classes), dtype=np.float32)
    w = np.zeros((k, 4 * num_classes), dtype=np.float32)
    if k:
        _lib.check(lib.mxr_expand_bbox_targets_host(h.ptr, _lib.hptr(d), k, int(num_classes),
                                                    iw.ctypes.data_as(_lib.c_float_p), _lib.hptr(t), _lib.hptr(w)))
    return t, w


def compute_bbox_regression_targets(rois, overlaps, labels):
    """(k,5) float32 [class, dx, dy, dw, dh] of one image's roidb entry (reference: bbox_regression.py:34-64):
    every roi with max overlap >= TRAIN.BBOX_REGRESSION_THRESH regresses onto the ground-truth roi (overlap == 1)
    it overlaps most; IoU + argmax and the transform run on the GPU."""
    rois = np.asarray(rois).astype(np.float64, copy=False)
    gt_inds = np.where(overlaps == 1)[0]
    if len(gt_inds) == 0:
        print('something wrong : zero ground truth rois')
    ex_inds = np.where(overlaps >= config.TRAIN.BBOX_REGRESSION_THRESH)[0]
    targets = np.zeros((rois.shape[0], 5), dtype=np.float32)
    if len(ex_inds) and len(gt_inds) == 0:
        raise ValueError("attempt to get argmax of an empty sequence")  # what the reference's argmax raises (:57)
    if len(ex_inds) == 0:
        return targets
    ex_rois = np.ascontiguousarray(rois[ex_inds, :4])
    gt_assignment = bbox_overlaps_reduced(ex_rois, np.ascontiguousarray(rois[gt_inds, :4]))[0]
    gt_rois = np.ascontiguousarray(rois[gt_inds[gt_assignment], :4])
    targets[ex_inds, 0] = labels[ex_inds]
    targets[ex_inds, 1:] = bbox_transform(ex_rois, gt_rois)
    return targets
