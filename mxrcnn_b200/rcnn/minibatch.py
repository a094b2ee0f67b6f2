"""Hot-path functions of the reference's rcnn/minibatch.py: ``assign_anchor`` and ``sample_rois``
(same signatures, return structures, dtypes and layouts).  ``get_minibatch`` / ``get_image_array``
(image IO and resizing) are out of scope (component 4).

Random draws stay on the host, in numpy's global RNG and in the reference's call order; the GPU
does the IoU / argmax / labelling / target arithmetic.
"""
import ctypes

import numpy as np
import numpy.random as npr

from .. import _lib
from ..helper.processing.bbox_regression import expand_bbox_regression_targets
from ..helper.processing.generate_anchor import generate_anchors
from .config import config


def _fg_bg_lists(overlaps, fg_thresh, bg_hi, bg_lo):
    """(fg_indexes, bg_indexes) as np.where would give them, computed on the GPU."""
    ov = np.ascontiguousarray(overlaps)
    if ov.dtype not in (np.float32, np.float64):
        ov = ov.astype(np.float64)
    n = ov.shape[0]
    fg = np.empty((n,), dtype=np.int32)
    bg = np.empty((n,), dtype=np.int32)
    cnt = np.zeros((2,), dtype=np.int32)
    _lib.check(_lib.load().mxr_fg_bg_lists_host(_lib.handle_for(None, overlaps).ptr, _lib.hptr(ov),
                                                1 if ov.dtype == np.float64 else 0, n, float(fg_thresh), float(bg_hi),
                                                float(bg_lo), _lib.hptr(fg), _lib.hptr(bg), _lib.hptr(cnt)))
    return fg[:cnt[0]].astype(np.int64), bg[:cnt[1]].astype(np.int64)


def sample_rois(roidb, fg_rois_per_image, rois_per_image, num_classes):
    """Random sample of fg / bg rois from a precomputed roidb record (reference: minibatch.py:151-201).
    Returns (rois, labels, bbox_targets, bbox_inside_weights, overlaps).  RNG order: fg, bg, gap."""
    labels = roidb['max_classes']
    overlaps = roidb['max_overlaps']
    rois = roidb['boxes']
    fg_indexes, bg_indexes = _fg_bg_lists(overlaps, config.TRAIN.FG_THRESH, config.TRAIN.BG_THRESH_HI,
                                          config.TRAIN.BG_THRESH_LO)
    fg_rois_per_this_image = np.minimum(fg_rois_per_image, fg_indexes.size)
    if fg_indexes.size > 0:
        fg_indexes = npr.choice(fg_indexes, size=fg_rois_per_this_image, replace=False)
    bg_rois_per_this_image = rois_per_image - fg_rois_per_this_image
    bg_rois_per_this_image = np.minimum(bg_rois_per_this_image, bg_indexes.size)
    if bg_indexes.size > 0:
        bg_indexes = npr.choice(bg_indexes, size=bg_rois_per_this_image, replace=False)
    keep_indexes = np.append(fg_indexes, bg_indexes)
    if keep_indexes.shape[0] < rois_per_image:
        gap = rois_per_image - keep_indexes.shape[0]
        gap_indexes = npr.choice(range(len(rois)), size=gap, replace=False)
        keep_indexes = np.append(keep_indexes, gap_indexes)
    keep_indexes = keep_indexes.astype(np.int64)
    labels = labels[keep_indexes]
    labels[fg_rois_per_this_image:] = 0
    overlaps = overlaps[keep_indexes]
    rois = rois[keep_indexes]
    bbox_targets, bbox_inside_weights = \
        expand_bbox_regression_targets(roidb['bbox_targets'][keep_indexes, :], num_classes)
    return rois, labels, bbox_targets, bbox_inside_weights, overlaps


def _anchor_draws(counts, G):
    """The host half of assign_anchor between its two device phases (minibatch.py:318-354): the subsampling draws
    from numpy's global RNG in the reference's order, and the example weights.  ``counts`` = (n_inside, n_fg, n_bg)
    of one image.  Returns (fg ranks to disable, bg ranks to disable, positive weight, negative weight)."""
    T = config.TRAIN
    n_inside, n_fg, n_bg = (int(c) for c in counts)
    if G > 0 and n_inside == 0:
        # the reference dies in overlaps.argmax(axis=0) on a (0, G) array (minibatch.py:298)
        raise ValueError("attempt to get argmax of an empty sequence")
    # subsample positives, then negatives: npr.choice(inds, size=k, replace=False) is
    # inds[npr.permutation(len(inds))[:k]] on the legacy RandomState, so only the ranks are needed
    num_fg = int(T.RPN_FG_FRACTION * T.RPN_BATCH_SIZE)
    fg_ranks = np.zeros((0,), dtype=np.int32)
    if n_fg > num_fg:
        fg_ranks = np.ascontiguousarray(npr.permutation(n_fg)[:n_fg - num_fg], dtype=np.int32)
    n_fg_final = n_fg - fg_ranks.size
    num_bg = T.RPN_BATCH_SIZE - n_fg_final
    bg_ranks = np.zeros((0,), dtype=np.int32)
    if n_bg > num_bg:
        bg_ranks = np.ascontiguousarray(npr.permutation(n_bg)[:n_bg - num_bg], dtype=np.int32)
    n_bg_final = n_bg - bg_ranks.size
    with np.errstate(divide='ignore', invalid='ignore'):
        if T.RPN_POSITIVE_WEIGHT < 0:
            num_examples = n_fg_final + n_bg_final
            positive_weight = float(np.float64(1.0) / np.float64(num_examples))
            negative_weight = positive_weight
        else:
            assert ((T.RPN_POSITIVE_WEIGHT > 0) & (T.RPN_POSITIVE_WEIGHT < 1))
            positive_weight = float(np.float64(T.RPN_POSITIVE_WEIGHT) / np.float64(n_fg_final))
            negative_weight = float(np.float64(1.0 - T.RPN_POSITIVE_WEIGHT) / np.float64(n_fg_final))
    return fg_ranks, bg_ranks, positive_weight, negative_weight


def assign_anchor_batch(feat_shape, gt_boxes_list, im_info_list, out, feat_stride=16, scales=(8, 16, 32),
                        ratios=(0.5, 1, 2), allowed_border=0):
    """assign_anchor for every image of a batch in two launches (extension, used by mxrcnn_b200/rcnn/loader.py).

    ``out``: dict of the four batch tensors 'label' (B, A*H*W), 'bbox_target', 'bbox_inside_weight',
    'bbox_outside_weight' (B, 4A, H, W), contiguous float32 DEVICE arrays; row i receives what
    ``assign_anchor(feat_shape, gt_boxes_list[i], im_info_list[i], ...)`` returns, and the draws from numpy's RNG are
    made image by image in list order, as the reference's loop does (rcnn/loader.py:302-311)."""
    lib, h = _lib.load(), _lib.handle_for(None, *tuple(out.values()))
    T = config.TRAIN
    B = len(gt_boxes_list)
    scales_f32 = np.array(scales, dtype=np.float32)                 # minibatch.py:243
    base_anchors = generate_anchors(base_size=16, ratios=list(ratios), scales=scales_f32)
    A = base_anchors.shape[0]
    fh, fw = int(feat_shape[-2]), int(feat_shape[-1])
    N = A * fh * fw
    gts = [np.ascontiguousarray(g, dtype=np.float32).reshape(-1, 5) if np.asarray(g).size > 0 else np.zeros((0, 5), np.float32)
           for g in gt_boxes_list]                                  # `if gt_boxes.size > 0` (:292)
    offsets = np.zeros((B + 1,), dtype=np.int32)
    offsets[1:] = np.cumsum([g.shape[0] for g in gts])
    gt_all = np.ascontiguousarray(np.concatenate(gts, axis=0)) if offsets[-1] else np.zeros((1, 5), np.float32)
    im_hw = np.ascontiguousarray([[np.asarray(i, dtype=np.float32)[0][0], np.asarray(i, dtype=np.float32)[0][1]]
                                  for i in im_info_list], dtype=np.float32)
    p = _lib.AnchorTargetParams(fh, fw, A, int(feat_stride), int(allowed_border), 0.0, 0.0, 0,
                                float(T.RPN_NEGATIVE_OVERLAP), float(T.RPN_POSITIVE_OVERLAP),
                                1 if T.RPN_CLOBBER_POSITIVES else 0)
    counts = np.zeros((B, 3), dtype=np.int32)
    _lib.check(lib.mxr_anchor_target_batch_phase1(h.ptr, ctypes.byref(p), base_anchors.ctypes.data_as(_lib.c_double_p), B,
                                                  im_hw.ctypes.data_as(_lib.c_float_p), _lib.hptr(gt_all),
                                                  offsets.ctypes.data_as(_lib.c_int_p), counts.ctypes.data_as(_lib.c_int_p)))
    ranks, n_dis, weights = [], np.zeros((B, 2), dtype=np.int32), np.zeros((B, 2), dtype=np.float64)
    for i in range(B):
        fg_ranks, bg_ranks, weights[i, 0], weights[i, 1] = _anchor_draws(counts[i], gts[i].shape[0])
        n_dis[i] = (fg_ranks.size, bg_ranks.size)
        ranks += [fg_ranks, bg_ranks]
    flat = np.ascontiguousarray(np.concatenate(ranks), dtype=np.int32) if n_dis.sum() else np.zeros((1,), np.int32)
    iw = np.ascontiguousarray(T.RPN_BBOX_INSIDE_WEIGHTS, dtype=np.float32)
    views = [_lib.device_view(out[k], np.float32) for k in ('label', 'bbox_target', 'bbox_inside_weight', 'bbox_outside_weight')]
    if int(np.prod(views[0].shape)) != B * N or any(int(np.prod(v.shape)) != 4 * B * N for v in views[1:]):
        raise ValueError("assign_anchor_batch: `out` arrays do not have %d / %d elements" % (B * N, 4 * B * N))
    _lib.check(lib.mxr_anchor_target_batch_phase2(
        h.ptr, _lib.current_stream_ptr(out['label']), flat.ctypes.data_as(_lib.c_int_p), n_dis.ctypes.data_as(_lib.c_int_p),
        weights.ctypes.data_as(_lib.c_double_p), iw.ctypes.data_as(_lib.c_float_p), views[0].ptr, views[1].ptr,
        views[2].ptr, views[3].ptr))
    return out


def assign_anchor(feat_shape, gt_boxes, im_info, feat_stride=16,
                  scales=(8, 16, 32), ratios=(0.5, 1, 2), allowed_border=0, out=None):
    """RPN labels / bbox targets / weights for one image (reference: minibatch.py:204-395).

    Returns the reference's dict: 'label' (1, A*H*W) float32 in (a,h,w) order, 'bbox_target',
    'bbox_inside_weight', 'bbox_outside_weight' (1, 4A, H, W) float32, plus 'gt_boxes' when
    config.END2END == 1.

    ``out`` (extension, used by mxrcnn_b200/rcnn/loader.py): a dict with those four keys holding contiguous float32
    DEVICE arrays of those shapes (e.g. one image's slices of the batch tensors); the results are written there on
    the arrays' current stream and never visit the host.
    """
    lib, h = _lib.load(), _lib.handle_for(None, *(tuple(out.values()) if out else ()))
    T = config.TRAIN
    info = np.asarray(im_info, dtype=np.float32)[0]
    scales_f32 = np.array(scales, dtype=np.float32)                 # minibatch.py:243
    base_anchors = generate_anchors(base_size=16, ratios=list(ratios), scales=scales_f32)
    A = base_anchors.shape[0]
    fh, fw = int(feat_shape[-2]), int(feat_shape[-1])
    gt = np.ascontiguousarray(gt_boxes, dtype=np.float32)
    G = int(gt.shape[0]) if gt.size > 0 else 0                      # `if gt_boxes.size > 0` (:292)
    p = _lib.AnchorTargetParams(fh, fw, A, int(feat_stride), int(allowed_border), float(info[0]), float(info[1]), G,
                                float(T.RPN_NEGATIVE_OVERLAP), float(T.RPN_POSITIVE_OVERLAP),
                                1 if T.RPN_CLOBBER_POSITIVES else 0)
    counts = np.zeros((3,), dtype=np.int32)
    _lib.check(lib.mxr_anchor_target_phase1(h.ptr, ctypes.byref(p), base_anchors.ctypes.data_as(_lib.c_double_p),
                                            _lib.hptr(gt) if G else None, counts.ctypes.data_as(_lib.c_int_p)))
    fg_ranks, bg_ranks, positive_weight, negative_weight = _anchor_draws(counts, G)
    iw = np.ascontiguousarray(T.RPN_BBOX_INSIDE_WEIGHTS, dtype=np.float32)
    N = A * fh * fw
    if out is not None:
        views = [_lib.device_view(out[k], np.float32) for k in ('label', 'bbox_target', 'bbox_inside_weight', 'bbox_outside_weight')]
        if int(np.prod(views[0].shape)) != N or any(int(np.prod(v.shape)) != 4 * N for v in views[1:]):
            raise ValueError("assign_anchor: `out` arrays do not have %d / %d elements" % (N, 4 * N))
        _lib.check(lib.mxr_anchor_target_phase2(
            h.ptr, _lib.current_stream_ptr(out['label']), fg_ranks.ctypes.data_as(_lib.c_int_p), int(fg_ranks.size),
            bg_ranks.ctypes.data_as(_lib.c_int_p), int(bg_ranks.size), positive_weight, negative_weight,
            iw.ctypes.data_as(_lib.c_float_p), views[0].ptr, views[1].ptr, views[2].ptr, views[3].ptr))
        res = dict(out)
        if config.END2END == 1:
            res.update({'gt_boxes': gt_boxes})
        return res
    label = np.empty((1, N), dtype=np.float32)
    tgt = np.empty((1, 4 * A, fh, fw), dtype=np.float32)
    inw = np.empty((1, 4 * A, fh, fw), dtype=np.float32)
    outw = np.empty((1, 4 * A, fh, fw), dtype=np.float32)
    _lib.check(lib.mxr_anchor_target_phase2_host(
        h.ptr, fg_ranks.ctypes.data_as(_lib.c_int_p), int(fg_ranks.size), bg_ranks.ctypes.data_as(_lib.c_int_p),
        int(bg_ranks.size), positive_weight, negative_weight, iw.ctypes.data_as(_lib.c_float_p), _lib.hptr(label),
        _lib.hptr(tgt), _lib.hptr(inw), _lib.hptr(outw)))
    out = {'label': label, 'bbox_target': tgt, 'bbox_inside_weight': inw, 'bbox_outside_weight': outw}
    if config.END2END == 1:
        out.update({'gt_boxes': gt_boxes})
    return out
