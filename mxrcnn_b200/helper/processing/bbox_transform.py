"""helper/processing/bbox_transform.py of the reference: same functions, same signatures.

numpy in -> numpy out (the library stages the copies); device arrays (torch / DLPack) in ->
device arrays out.  All arithmetic runs in libmxrcnn_b200.so.
"""
import numpy as np

from ... import _lib


def _code(dt):
    return _lib.MXR_DTYPE_F64 if np.dtype(dt) == np.float64 else _lib.MXR_DTYPE_F32


def _float_array(a):
    """f32 stays f32, everything else becomes f64 (what numpy's promotion with python floats gives)."""
    a = np.asarray(a)
    return np.ascontiguousarray(a if a.dtype in (np.float32, np.float64) else a.astype(np.float64))


def bbox_transform(ex_rois, gt_rois):
    """Regression targets from ex_rois to gt_rois, (N,4) each (reference: bbox_transform.py:8-32).
    Like numpy, each side is evaluated in its own dtype and the result in the promoted one
    (float32 only when both inputs are float32)."""
    lib, h = _lib.load(), _lib.handle_for(None, ex_rois)
    if _lib.is_device_array(ex_rois):
        import torch
        ex = (ex_rois if ex_rois.dtype in (torch.float32, torch.float64) else ex_rois.double())[:, :4].contiguous()
        gt = (gt_rois if gt_rois.dtype in (torch.float32, torch.float64) else gt_rois.double())[:, :4].contiguous()
        dt = torch.float32 if (ex.dtype == torch.float32 and gt.dtype == torch.float32) else torch.float64
        out = torch.empty((ex.shape[0], 4), dtype=dt, device=ex.device)
        _lib.check(lib.mxr_bbox_transform(h.ptr, _lib.current_stream_ptr(ex), ex.data_ptr(),
                                          0 if ex.dtype == torch.float32 else 1, gt.data_ptr(),
                                          0 if gt.dtype == torch.float32 else 1, ex.shape[0], out.data_ptr()))
        return out
    ex = _float_array(np.asarray(ex_rois)[:, :4])
    gt = _float_array(np.asarray(gt_rois)[:, :4])
    if ex.shape[0] != gt.shape[0]:
        raise ValueError("bbox_transform: %d ex_rois vs %d gt_rois" % (ex.shape[0], gt.shape[0]))
    dt = np.float32 if (ex.dtype == np.float32 and gt.dtype == np.float32) else np.float64
    out = np.empty((ex.shape[0], 4), dtype=dt)
    _lib.check(lib.mxr_bbox_transform_host(h.ptr, _lib.hptr(ex), _code(ex.dtype), _lib.hptr(gt), _code(gt.dtype),
                                           ex.shape[0], _lib.hptr(out)))
    return out


def bbox_pred(boxes, box_deltas, is_train=False):
    """Apply (N, 4*num_classes) deltas to (N,4) boxes -> (N, 4*num_classes) float64
    (reference: bbox_transform.py:35-78).  Empty input returns zeros((0, 4C)) (:43-44).

    ``is_train`` (never enabled by any caller of the reference, and under py3 its ``map`` form does
    not even run) clamps |dx|,|dy| to 10 and |dw|,|dh| to 8 elementwise before the transform (:56-64).
    """
    lib, h = _lib.load(), _lib.handle_for(None, boxes)
    if _lib.is_device_array(boxes):
        import torch
        n, c4 = box_deltas.shape
        out = torch.zeros((n, c4), dtype=torch.float64, device=box_deltas.device)
        if n == 0:
            return out
        b = boxes if boxes.dtype in (torch.float32, torch.float64) else boxes.double()
        b = b[:, :4].contiguous()
        d = box_deltas if box_deltas.dtype in (torch.float32, torch.float64) else box_deltas.double()
        d = _clamp_train(d, torch) if is_train else d
        d = d.contiguous()
        _lib.check(lib.mxr_bbox_pred(h.ptr, _lib.current_stream_ptr(d), b.data_ptr(), 0 if b.dtype == torch.float32 else 1,
                                     d.data_ptr(), 0 if d.dtype == torch.float32 else 1, n, c4, out.data_ptr()))
        return out
    boxes = np.asarray(boxes)
    box_deltas = np.asarray(box_deltas)
    if boxes.shape[0] == 0:
        return np.zeros((0, box_deltas.shape[1]))
    b = _float_array(boxes[:, :4])
    d = _float_array(box_deltas)
    if is_train:
        d = _clamp_train(d, np)
    n, c4 = d.shape
    out = np.empty((n, c4), dtype=np.float64)
    _lib.check(lib.mxr_bbox_pred_host(h.ptr, _lib.hptr(b), _code(b.dtype), _lib.hptr(d), _code(d.dtype), n, c4,
                                      _lib.hptr(out)))
    return out


def _clamp_train(d, xp):
    d = d.clone() if xp is not np else d.copy()
    for k, lim in ((0, 10), (1, 10), (2, 8), (3, 8)):
        v = d[:, k::4]
        d[:, k::4] = xp.where(abs(v) > lim, xp.sign(v) * lim, v)
    return d


def clip_boxes(boxes, im_shape):
    """Clip (N, 4*num_classes) boxes to the image, IN PLACE, and return the same array
    (reference: bbox_transform.py:81-96).  x limits come from im_shape[1], y from im_shape[0]."""
    lib, h = _lib.load(), _lib.handle_for(None, boxes)
    lim_x, lim_y = float(im_shape[1] - 1), float(im_shape[0] - 1)  # numpy scalar rules apply to the subtraction
    if _lib.is_device_array(boxes):
        import torch
        if boxes.dtype not in (torch.float32, torch.float64) or not boxes.is_contiguous():
            raise ValueError("clip_boxes: device boxes must be contiguous float32/float64")
        n, c4 = boxes.shape
        _lib.check(lib.mxr_clip_boxes(h.ptr, _lib.current_stream_ptr(boxes), boxes.data_ptr(),
                                      0 if boxes.dtype == torch.float32 else 1, n, c4, lim_x, lim_y))
        return boxes
    if boxes.size == 0:
        return boxes
    if boxes.dtype in (np.float32, np.float64) and boxes.flags.c_contiguous:
        n, c4 = boxes.shape
        _lib.check(lib.mxr_clip_boxes_host(h.ptr, _lib.hptr(boxes), _code(boxes.dtype), n, c4, lim_x, lim_y))
        return boxes
    tmp = np.ascontiguousarray(boxes, dtype=np.float64)
    n, c4 = tmp.shape
    _lib.check(lib.mxr_clip_boxes_host(h.ptr, _lib.hptr(tmp), _lib.MXR_DTYPE_F64, n, c4, lim_x, lim_y))
    boxes[...] = tmp
    return boxes


def clip_pad(boxes, pad_shape):
    """Crop an [n, c, H, W] map to [n, c, h, w] (reference: bbox_transform.py:98-111).  Pure data
    movement: a slice copy, for numpy and device arrays alike."""
    H, W = boxes.shape[2:]
    h, w = pad_shape
    dev = _lib.is_device_array(boxes)
    if h < H:
        boxes = boxes[:, :, :h, :].contiguous() if dev else boxes[:, :, :h, :].copy()
    if w < W:
        boxes = boxes[:, :, :, :w].contiguous() if dev else boxes[:, :, :, :w].copy()
    return boxes
