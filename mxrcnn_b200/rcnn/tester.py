"""Test-time box head of the reference: the post-processing of ``Detector.im_detect``
(rcnn/detector.py:73-74) and the class loop of ``pred_eval`` (rcnn/tester.py:50-65) as one GPU launch.

    scores, boxes = detector.im_detect(...)            # reference: bbox_pred + clip_boxes on the host
    for j in range(1, num_classes): ... nms(...)       # reference: python loop, one nms per class
    ... cap at max_per_image ...

becomes ``boxes_this_image = detections(scores, rois, bbox_deltas, im_shape, scale)``.
"""
import ctypes

import numpy as np

from .. import _lib
from ..config import config


def _params(B, R, C, thresh, nms_thresh, max_per_image):
    nms_thresh = config.TEST.NMS if nms_thresh is None else nms_thresh
    return _lib.BoxHeadParams(int(B), int(R), int(C), int(max_per_image), float(thresh), float(nms_thresh))


def detections(scores, rois, bbox_deltas, im_shape, scale, thresh=0.05, nms_thresh=None, max_per_image=100,
               handle=None, with_rows=False):
    """Detections of ONE image, as ``pred_eval`` builds them (rcnn/tester.py:50-65).

    scores (R,C) float32 = ``cls_prob_reshape_output[0]``; rois (R,4) or (R,5) float32 (a leading batch
    column is dropped like rcnn/detector.py:68); bbox_deltas (R,4C) float32; im_shape = ``im_array[0].shape[-2:]``;
    scale = ``im_info[0, 2]``; thresh / max_per_image as hard-coded in tester.py:22-24; nms_thresh defaults
    to ``config.TEST.NMS``.  Returns ``boxes_this_image``: a list with ``[]`` for the background and one
    (n_j,5) float64 array ``[x1,y1,x2,y2,score]`` per foreground class; with ``with_rows`` also the roi row
    of every detection.
    """
    out = detections_batch(np.asarray(scores)[None], np.asarray(rois)[None], np.asarray(bbox_deltas)[None],
                           [im_shape], [scale], thresh, nms_thresh, max_per_image, handle, with_rows)
    return (out[0][0], out[1][0]) if with_rows else out[0]


def detections_batch(scores, rois, bbox_deltas, im_shapes, scales, thresh=0.05, nms_thresh=None, max_per_image=100,
                     handle=None, with_rows=False):
    """Extension: B images per launch.  scores (B,R,C), rois (B,R,4|5), bbox_deltas (B,R,4C); im_shapes and
    scales have one entry per image.  Returns a list of B ``boxes_this_image`` lists (and the rows)."""
    lib = _lib.load()
    h = _lib.handle_for(handle, scores)
    scores = _lib.host(scores, np.float32)
    rois = np.asarray(rois)
    if rois.ndim != 3 or rois.shape[2] not in (4, 5):
        raise ValueError("rois must be (B,R,4) or (B,R,5)")
    rois = _lib.host(rois[:, :, -4:], np.float32)
    bbox_deltas = _lib.host(bbox_deltas, np.float32)
    if scores.ndim != 3:
        raise ValueError("scores must be (B,R,C)")
    B, R, C = scores.shape
    if rois.shape[:2] != (B, R) or bbox_deltas.shape != (B, R, 4 * C):
        raise ValueError("box head: scores %s, rois %s, bbox_deltas %s do not agree"
                         % (scores.shape, rois.shape, bbox_deltas.shape))
    if len(im_shapes) != B or len(scales) != B:
        raise ValueError("box head: one im_shape and one scale per image")
    im = np.array([[float(s[0]), float(s[1]), float(sc)] for s, sc in zip(im_shapes, scales)], dtype=np.float64)
    p = _params(B, R, C, thresh, nms_thresh, max_per_image)
    dets = np.empty((B, C, R, 5), dtype=np.float64)
    rows = np.empty((B, C, R), dtype=np.int32)
    count = np.zeros((B, C, 2), dtype=np.int32)
    _lib.check(lib.mxr_box_head_host(h.ptr, ctypes.byref(p), _lib.hptr(rois), _lib.hptr(bbox_deltas), _lib.hptr(scores),
                                     _lib.hptr(im), _lib.hptr(dets), _lib.hptr(rows), _lib.hptr(count)))
    out, out_rows = [], []
    for b in range(B):
        out.append([[]] + [dets[b, j, :count[b, j, 0]].copy() for j in range(1, C)])
        out_rows.append([[]] + [rows[b, j, :count[b, j, 0]].astype(np.int64) for j in range(1, C)])
    return (out, out_rows) if with_rows else out


def detections_device(scores, rois, bbox_deltas, im, thresh=0.05, nms_thresh=None, max_per_image=100, handle=None):
    """Device-resident form: all inputs are device arrays (scores (B,R,C) f32, rois (B,R,4) f32, bbox_deltas
    (B,R,4C) f32, im (B,3) f64 = [height, width, scale]); nothing is synchronised.  Returns device tensors
    ``dets`` (B,C,R,5) f64, ``rows`` (B,C,R) i32, ``count`` (B,C,2) i32 ([after the cap, after NMS])."""
    import torch
    lib = _lib.load()
    h = _lib.handle_for(handle, scores)
    vs, vr, vd = _lib.device_view(scores, np.float32), _lib.device_view(rois, np.float32), _lib.device_view(bbox_deltas, np.float32)
    vi = _lib.device_view(im, np.float64)
    B, R, C = vs.shape
    if tuple(vr.shape) != (B, R, 4) or tuple(vd.shape) != (B, R, 4 * C) or tuple(vi.shape) != (B, 3):
        raise ValueError("box head: scores %s, rois %s, bbox_deltas %s, im %s do not agree"
                         % (vs.shape, vr.shape, vd.shape, vi.shape))
    p = _params(B, R, C, thresh, nms_thresh, max_per_image)
    dev = scores.device
    dets = torch.empty((B, C, R, 5), dtype=torch.float64, device=dev)
    rows = torch.empty((B, C, R), dtype=torch.int32, device=dev)
    count = torch.empty((B, C, 2), dtype=torch.int32, device=dev)
    _lib.check(lib.mxr_box_head(h.ptr, _lib.current_stream_ptr(scores), ctypes.byref(p), vr.ptr, vd.ptr, vs.ptr, vi.ptr,
                                dets.data_ptr(), rows.data_ptr(), count.data_ptr()))
    return dets, rows, count
