"""``proposal`` custom operator -- drop-in for the reference's rcnn/rpn/proposal.py.

Same class names, constructor arguments (string-typed, as MXNet hands them over), method
signatures, argument / output names and shapes.  The arithmetic runs in
libmxrcnn_b200.so (decode + clip + min-size filter, radix-select top-k, striped bitmask
NMS); this file only does what the reference does on the host: config lookup, the TRAIN
mode crop size, NaN -> ValueError, and the numpy RNG draw that pads short outputs.
"""
import ctypes
import logging

import numpy as np
import numpy.random as npr

from ... import _lib
from ... import operator as mxop
from ...helper.processing.generate_anchor import generate_anchors
from ..config import config


def _as_numpy(x):
    return x.asnumpy() if hasattr(x, "asnumpy") else np.asarray(x)


def proposal_forward(cls_prob, bbox_pred, im_info, anchors, feat_stride, pre_nms_top_n, post_nms_top_n, nms_thresh,
                     min_size, train_crop=False, out_rois=None, out_scores=None, handle=None, first_stripe=0,
                     want_indices=False, rng=None):
    """Functional form used by ProposalOperator.forward (reference: rcnn/rpn/proposal.py:57-149).

    Inputs either all numpy (host path: the library stages the copies) or all device arrays.
    Batch B >= 1 is an extension; B == 1 is the reference.  Returns a dict with ``rois``
    (B*post,5), ``scores`` (B*post,1), ``nkeep`` (B,), and with ``want_indices`` also ``keep``
    and ``anchor_index`` (B, post).  Short outputs are padded with ``npr.choice`` exactly as
    proposal.py:135-139 (one draw per short image, image order).
    """
    lib = _lib.load()
    h = _lib.handle_for(handle, cls_prob)
    on_device = _lib.is_device_array(cls_prob)
    if on_device:
        vc, vb, vi = _lib.device_view(cls_prob, np.float32), _lib.device_view(bbox_pred, np.float32), \
            _lib.device_view(im_info, np.float32)
        shape_c, shape_b = vc.shape, vb.shape
    else:
        cls_prob = _lib.host(cls_prob, np.float32)
        bbox_pred = _lib.host(bbox_pred, np.float32)
        im_info = _lib.host(im_info, np.float32)
        shape_c, shape_b = cls_prob.shape, bbox_pred.shape
    anchors = np.ascontiguousarray(anchors, dtype=np.float64)
    A = anchors.shape[0]
    B, _, H, W = shape_c
    if shape_c[1] != 2 * A or tuple(shape_b) != (B, 4 * A, H, W):
        raise ValueError("proposal: cls_prob %s / bbox_pred %s do not match %d anchors" % (shape_c, shape_b, A))
    crop_h = crop_w = 0
    if train_crop:
        # proposal.py:70-71: int(im_info[0] / feat_stride) in float32 (needs im_info on the host)
        info_h = _as_numpy(im_info.cpu() if on_device and hasattr(im_info, "cpu") else im_info).astype(np.float32)
        crops = {(int(r[0] / float(feat_stride)), int(r[1] / float(feat_stride))) for r in info_h.reshape(-1, 3)}
        if len(crops) != 1:
            raise ValueError("proposal: images of one batch must share the TRAIN-mode crop")
        crop_h, crop_w = crops.pop()
        if crop_h > H or crop_w > W:
            raise ValueError("proposal: im_info implies a %dx%d map but the score map is %dx%d" % (crop_h, crop_w, H, W))
    post = int(post_nms_top_n)
    n_crop = A * (crop_h or H) * (crop_w or W)
    rows = post if post > 0 else (min(int(pre_nms_top_n), n_crop) if pre_nms_top_n > 0 else n_crop)
    p = _lib.ProposalParams(B, A, H, W, crop_h, crop_w, float(feat_stride), int(pre_nms_top_n), post,
                            float(nms_thresh), float(min_size), int(first_stripe))
    anc = anchors.ctypes.data_as(_lib.c_double_p)
    if on_device:
        import torch  # device tensors come from the harness, which already runs torch
        dev = cls_prob.device
        rois = out_rois if out_rois is not None else torch.empty((B * rows, 5), dtype=torch.float32, device=dev)
        scores = out_scores if out_scores is not None else torch.empty((B * rows, 1), dtype=torch.float32, device=dev)
        keep = torch.empty((B, rows), dtype=torch.int32, device=dev) if want_indices else None
        aidx = torch.empty((B, rows), dtype=torch.int32, device=dev) if want_indices else None
        nk = torch.empty((B, 2), dtype=torch.int32, device=dev)
        stream = _lib.current_stream_ptr(cls_prob)
        _lib.check(lib.mxr_proposal_forward(
            h.ptr, stream, ctypes.byref(p), vc.ptr, vb.ptr, vi.ptr, anc, rows, rois.data_ptr(), scores.data_ptr(),
            keep.data_ptr() if want_indices else None, aidx.data_ptr() if want_indices else None, nk.data_ptr()))
        nk_h = nk.cpu().numpy()  # the reference synchronises here too (asnumpy, proposal.py:57)
        if (nk_h[:, 1] & 1).any():
            raise ValueError("there is nan in input scores or bbox_deltas")
        for b in range(B):
            n = int(nk_h[b, 0])
            if post > 0 and n < post:
                pad = _draw_pad(n, post - n, rng)
                _lib.check(lib.mxr_proposal_pad(h.ptr, stream, b, rows, n, pad.ctypes.data_as(_lib.c_int_p), pad.size,
                                                rois.data_ptr(), scores.data_ptr()))
                if want_indices:
                    idx = torch.from_numpy(pad.astype(np.int64)).to(dev)
                    keep[b, n:post] = keep[b, idx]
                    aidx[b, n:post] = aidx[b, idx]
        out = {"rois": rois, "scores": scores, "nkeep": nk_h[:, 0].copy()}
    else:
        rois = np.empty((B * rows, 5), dtype=np.float32)
        scores = np.empty((B * rows, 1), dtype=np.float32)
        keep = np.empty((B, rows), dtype=np.int32)
        aidx = np.empty((B, rows), dtype=np.int32)
        nk = np.zeros((B, 2), dtype=np.int32)
        _lib.check(lib.mxr_proposal_forward_host(
            h.ptr, ctypes.byref(p), _lib.hptr(cls_prob), _lib.hptr(bbox_pred), _lib.hptr(im_info), anc, rows,
            _lib.hptr(rois), _lib.hptr(scores), _lib.hptr(keep), _lib.hptr(aidx), _lib.hptr(nk)))
        for b in range(B):
            n = int(nk[b, 0])
            if post > 0 and n < post:
                pad = _draw_pad(n, post - n, rng)
                r0 = b * rows
                rois[r0 + n:r0 + post] = rois[r0 + pad]
                scores[r0 + n:r0 + post] = scores[r0 + pad]
                keep[b, n:post] = keep[b, pad]
                aidx[b, n:post] = aidx[b, pad]
        out = {"rois": rois, "scores": scores, "nkeep": nk[:, 0].copy()}
    if want_indices:
        out["keep"], out["anchor_index"] = keep, aidx
    return out


def _draw_pad(nkeep, count, rng):
    """proposal.py:135-139.  ``npr.choice(keep, size=k)`` draws ``randint(0, len(keep), k)`` from the
    legacy global RandomState and indexes ``keep`` with it; row r of the output IS keep[r], so the
    draw is used as a row index.  Empty ``keep`` raises like the reference (quirk Q7)."""
    if nkeep == 0:
        logging.log(logging.ERROR, "currently len(keep) is zero")
    rs = rng if rng is not None else npr
    return np.asarray(rs.choice(np.arange(nkeep), size=count), dtype=np.int32)


class ProposalOperator(mxop.CustomOp):
    def __init__(self, feat_stride, scales, ratios, is_train=False, output_score=False):
        super(ProposalOperator, self).__init__()
        self._feat_stride = float(feat_stride)
        # kwargs arrive as strings '(8, 16, 32)' (proposal.py:23-24); tuples/lists are accepted too
        self._scales = _parse_tuple(scales)
        self._ratios = _parse_tuple(ratios).tolist()
        self._anchors = generate_anchors(base_size=self._feat_stride, scales=self._scales, ratios=self._ratios)
        self._num_anchors = self._anchors.shape[0]
        self._output_score = output_score
        self.cfg_key = "TRAIN" if is_train else "TEST"

    def forward(self, is_train, req, in_data, out_data, aux):
        cfg = config[self.cfg_key]
        ins = [x if (_lib.is_device_array(x) or isinstance(x, np.ndarray)) else _as_numpy(x) for x in in_data[:3]]
        direct = (_lib.is_device_array(ins[0]) and _lib.is_device_array(out_data[0]) and req[0] in ("write", "inplace")
                  and (not self._output_score or (len(out_data) > 1 and _lib.is_device_array(out_data[1])
                                                   and req[1] in ("write", "inplace"))))
        res = proposal_forward(
            ins[0], ins[1], ins[2], self._anchors, self._feat_stride, cfg.RPN_PRE_NMS_TOP_N, cfg.RPN_POST_NMS_TOP_N,
            cfg.RPN_NMS_THRESH, cfg.RPN_MIN_SIZE, train_crop=(self.cfg_key == "TRAIN"),
            out_rois=out_data[0] if direct else None,
            out_scores=out_data[1] if (direct and self._output_score) else None)
        if not direct:
            self.assign(out_data[0], req[0], res["rois"])
            if self._output_score:
                self.assign(out_data[1], req[1], res["scores"])

    def backward(self, req, out_grad, in_data, out_data, in_grad, aux):
        self.assign(in_grad[0], req[0], 0)
        self.assign(in_grad[1], req[1], 0)


def _parse_tuple(v):
    if isinstance(v, str):
        return np.array([float(t) for t in v.strip()[1:-1].split(",") if t.strip()], dtype=float)
    return np.asarray(v, dtype=float).ravel()


@mxop.register("proposal")
class ProposalProp(mxop.CustomOpProp):
    def __init__(self, feat_stride, scales, ratios, is_train=False, output_score=False):
        super(ProposalProp, self).__init__(need_top_grad=False)
        self._feat_stride = feat_stride
        self._scales = scales
        self._ratios = ratios
        self._is_train = True if is_train == "True" else False  # proposal.py:171 (string compare)
        self._output_score = output_score  # kept raw like proposal.py:172: any non-empty string is truthy
        self.cfg_key = "TRAIN" if self._is_train else "TEST"

    def list_arguments(self):
        return ["cls_prob", "bbox_pred", "im_info"]

    def list_outputs(self):
        return ["output", "score"] if self._output_score else ["output"]

    def infer_shape(self, in_shape):
        cls_prob_shape, bbox_pred_shape = in_shape[0], in_shape[1]
        assert cls_prob_shape[0] == bbox_pred_shape[0], "ROI number does not equal in cls and reg"
        batch_size = cls_prob_shape[0]
        if batch_size > 1:
            raise ValueError("Only single item batches are supported")
        im_info_shape = (batch_size, 3)
        output_shape = (config[self.cfg_key].RPN_POST_NMS_TOP_N, 5)
        score_shape = (config[self.cfg_key].RPN_POST_NMS_TOP_N, 1)
        if self._output_score:
            return [cls_prob_shape, bbox_pred_shape, im_info_shape], [output_shape, score_shape]
        return [cls_prob_shape, bbox_pred_shape, im_info_shape], [output_shape]

    def create_operator(self, ctx, shapes, dtypes):
        return ProposalOperator(self._feat_stride, self._scales, self._ratios, self._is_train, self._output_score)

    def declare_backward_dependency(self, out_grad, in_data, out_data):
        return []
