"""``proposal_target`` custom operator -- drop-in for the reference's rcnn/rpn/proposal_target.py
(same classes, argument / output names, shapes, and the module-level ``_sample_rois``)."""
import ctypes
import logging

import numpy as np
import numpy.random as npr

from ... import _lib
from ... import operator as mxop
from ...helper.processing.bbox_regression import bbox_overlaps_reduced
from ..config import config
from ..minibatch import _fg_bg_lists


def _as_numpy(x):
    if hasattr(x, "asnumpy"):
        return x.asnumpy()
    if _lib.is_device_array(x):
        return x.cpu().numpy()
    return np.asarray(x)


class ProposalTargetOperator(mxop.CustomOp):
    def __init__(self, num_classes, is_train=False):
        super(ProposalTargetOperator, self).__init__()
        self._num_classes = int(num_classes)
        self.cfg_key = 'TRAIN' if is_train else 'TEST'
        self._img_per_batch = 1

    def forward(self, is_train, req, in_data, out_data, aux):
        assert config.TRAIN.BATCH_SIZE % self._img_per_batch == 0, \
            'IMAGESPERBATCH {} must devide BATCHSIZE {}'.format(self._img_per_batch, config.TRAIN.BATCH_SIZE)
        num_images = self._img_per_batch
        assert num_images == 1, "only support signle image"
        rois_per_image = config.TRAIN.BATCH_SIZE // self._img_per_batch          # py2 integer division (quirk Q5)
        fg_rois_per_image = np.round(config.TRAIN.FG_FRACTION * rois_per_image).astype(int)
        all_rois = _as_numpy(in_data[0])
        gt_boxes = _as_numpy(in_data[1])
        gt_boxes = gt_boxes[np.where(gt_boxes[:, :5].mean(axis=1) != -1)]          # strip the -1 padding rows
        zeros = np.zeros((gt_boxes.shape[0], 1), dtype=gt_boxes.dtype)
        all_rois = np.vstack((all_rois, np.hstack((zeros, gt_boxes[:, :-1]))))     # gt boxes join the candidates
        assert np.all(all_rois[:, 0] == 0), 'Only single item batches are supported'
        labels, rois, bbox_targets, bbox_inside_weights = _sample_rois(
            all_rois, gt_boxes, fg_rois_per_image, rois_per_image, self._num_classes, self.cfg_key)
        self.assign(out_data[0], req[0], _like(out_data[0], rois))
        self.assign(out_data[1], req[1], _like(out_data[1], labels))
        self.assign(out_data[2], req[2], _like(out_data[2], bbox_targets))
        self.assign(out_data[3], req[3], _like(out_data[3], bbox_inside_weights))
        self.assign(out_data[4], req[4], _like(out_data[4], np.array(bbox_inside_weights > 0).astype(np.float32)))

    def backward(self, req, out_grad, in_data, out_data, in_grad, aux):
        self.assign(in_grad[0], req[0], 0)
        self.assign(in_grad[1], req[1], 0)


def _like(dst, src):
    """numpy result -> whatever `dst` is (device tensors get a device copy)."""
    if _lib.is_device_array(dst):
        import torch
        return torch.from_numpy(np.ascontiguousarray(src)).to(dst.device)
    return src


@mxop.register("proposal_target")
class ProposalTargetProp(mxop.CustomOpProp):
    def __init__(self, num_classes, is_train=False):
        super(ProposalTargetProp, self).__init__(need_top_grad=False)
        self._num_classes = int(num_classes)
        self._is_train = True if is_train == 'True' else False
        self.cfg_key = 'TRAIN' if self._is_train else 'TEST'
        self._img_per_batch = 1

    def list_arguments(self):
        return ['rpn_roi', 'gt_boxes']

    def list_outputs(self):
        return ['roi', 'label', 'bbox_target', 'bbox_inside_weight', 'bbox_outside_weight']

    def infer_shape(self, in_shape):
        rpn_roi_shape, gt_boxes_shape = in_shape[0], in_shape[1]
        batch_size = config.TRAIN.BATCH_SIZE // self._img_per_batch
        c4 = self._num_classes * 4
        return [rpn_roi_shape, gt_boxes_shape], [(batch_size, 5), (batch_size,), (batch_size, c4), (batch_size, c4),
                                                 (batch_size, c4)]

    def create_operator(self, ctx, shapes, dtypes):
        return ProposalTargetOperator(self._num_classes, self._is_train)

    def declare_backward_dependency(self, out_grad, in_data, out_data):
        return []


def _sample_rois(all_rois, gt_boxes, fg_rois_per_image, rois_per_image, num_classes, key):
    """Random sample of RoIs with labels and class-expanded regression targets
    (reference: proposal_target.py:135-196).  Returns (labels, rois, bbox_targets, bbox_inside_weights).
    RNG order: fg (no replacement), fg padding (with replacement, if short), bg, bg padding."""
    lib, h = _lib.load(), _lib.handle_for(None, all_rois)
    T = config.TRAIN
    rois_f32 = np.ascontiguousarray(all_rois, dtype=np.float32)
    gt_f32 = np.ascontiguousarray(gt_boxes, dtype=np.float32)
    if gt_f32.shape[0] == 0:
        raise ValueError("attempt to get argmax of an empty sequence")  # overlaps.argmax(axis=1) on (n,0), :143
    # ONE launch: IoU + argmax + max in float64 like the reference (:140-144) and the three index lists the sampler may
    # draw from (:148, :160-166); rois, gt boxes and assignments stay on the device for the gather below
    R, G = rois_f32.shape[0], gt_f32.shape[0]
    fg_buf, bg_buf, fb_buf = (np.empty((R,), dtype=np.int32) for _ in range(3))
    cnt = np.zeros((3,), dtype=np.int32)
    c4 = 4 * int(num_classes)
    _lib.check(lib.mxr_proposal_target_phase1_host(
        h.ptr, _lib.hptr(rois_f32), R, _lib.hptr(gt_f32), G, int(rois_per_image), int(num_classes), float(T.FG_THRESH),
        float(T.BG_THRESH_HI), float(T.BG_THRESH_LO), float(T.BG_THRESH_HI + 0.2), 0.0, _lib.hptr(fg_buf), _lib.hptr(bg_buf),
        _lib.hptr(fb_buf), _lib.hptr(cnt), None, None))
    fg_inds = fg_buf[:cnt[0]].astype(np.int64)
    bg_inds_ = bg_buf[:cnt[1]].astype(np.int64)
    fg_rois_per_this_image = min(fg_rois_per_image, fg_inds.size)
    if fg_inds.size > 0:
        fg_inds = npr.choice(fg_inds, size=fg_rois_per_this_image, replace=False)
    if fg_inds.size < fg_rois_per_image:
        fg_inds_ = npr.choice(fg_inds, size=fg_rois_per_image - fg_inds.size, replace=True)
        fg_inds = np.hstack((fg_inds_, fg_inds))
    if len(bg_inds_) == 0 and key == 'TRAIN':
        bg_inds = fb_buf[:cnt[2]].astype(np.int64)
    else:
        bg_inds = bg_inds_
    if len(bg_inds) == 0:
        logging.log(logging.ERROR, "currently len(bg_inds) is zero")
    bg_rois_per_this_image = rois_per_image - len(fg_inds)
    bg_rois_per_this_image = min(bg_rois_per_this_image, bg_inds.size)
    if bg_inds.size > 0:
        bg_inds = npr.choice(bg_inds, size=bg_rois_per_this_image, replace=False)
    if bg_inds.size < rois_per_image - fg_rois_per_image:
        bg_inds_ = npr.choice(bg_inds, size=rois_per_image - fg_rois_per_image - bg_inds.size, replace=True)
        bg_inds = np.hstack((bg_inds_, bg_inds))
    keep_inds = np.ascontiguousarray(np.append(fg_inds, bg_inds), dtype=np.int32)
    K = int(keep_inds.size)
    p = _lib.ProposalTargetParams(rois_f32.shape[0], gt_f32.shape[0], K, int(fg_rois_per_this_image), int(num_classes),
                                  1 if T.BBOX_NORMALIZATION_PRECOMPUTED else 0)
    for i in range(4):
        p.means[i] = float(T.BBOX_MEANS[i])
        p.stds[i] = float(T.BBOX_STDS[i])
        p.inside_weights[i] = float(T.BBOX_INSIDE_WEIGHTS[i])
    rois = np.empty((K, 5), dtype=np.float32)
    labels = np.empty((K,), dtype=np.float32)
    bbox_targets = np.empty((K, c4), dtype=np.float32)
    inside = np.empty((K, c4), dtype=np.float32)
    outside = np.empty((K, c4), dtype=np.float32)
    _lib.check(lib.mxr_proposal_target_phase2_host(
        h.ptr, ctypes.byref(p), _lib.hptr(keep_inds), _lib.hptr(rois), _lib.hptr(labels), _lib.hptr(bbox_targets),
        _lib.hptr(inside), _lib.hptr(outside)))
    return labels, rois, bbox_targets, inside
