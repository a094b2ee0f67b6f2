"""CPU restatement (numpy) of the mx-rcnn RPN proposal / box-processing hot path.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py): the checker for the CUDA path and the
``cpu_baseline`` leg of bench.py.  Never imported by the product package.

Every function cites the reference ``file:line`` (relative to the reference root) whose
arithmetic, operation order and dtype flow it follows.  Pinned against the real reference
by ``tests/golden/*.npz`` (written by ``oracle/make_golden.py``) and, where the reference
tree is mounted, live by ``tests/test_oracle_vs_reference.py``.

Two deliberate differences from the reference, both documented in DESIGN.md:
  * sort ties: the reference calls ``argsort()`` with numpy's default (unstable) kind, so
    the order of EQUAL scores is not defined by the reference.  The restatement (and the
    CUDA path) use ``argsort(kind='stable')[::-1]`` = descending score, then descending
    index.  Identical whenever scores are distinct.
  * random draws are taken from an explicit ``RandomState`` (the reference uses the global
    one) in exactly the reference's call order, and are recorded so the CUDA path can be
    fed the same draws.
"""
import logging

import numpy as np

# ----------------------------------------------------------------------------------------
# configuration values read by the hot path (rcnn/config.py:20-70)
# ----------------------------------------------------------------------------------------


class Cfg(dict):
    __getattr__ = dict.__getitem__
    __setattr__ = dict.__setitem__


def default_config():
    """Values of rcnn/config.py:4-70 that the hot path reads."""
    c = Cfg()
    c.END2END = 0
    c.EPS = 1e-14
    c.TRAIN = Cfg(
        BATCH_SIZE=128, BATCH_IMAGES=2, FG_FRACTION=0.25, FG_THRESH=0.5, BG_THRESH_HI=0.5, BG_THRESH_LO=0.1,
        BBOX_INSIDE_WEIGHTS=np.array([1.0, 1.0, 1.0, 1.0]),
        RPN_BATCH_SIZE=256, RPN_FG_FRACTION=0.5, RPN_POSITIVE_OVERLAP=0.7, RPN_NEGATIVE_OVERLAP=0.3,
        RPN_CLOBBER_POSITIVES=False, RPN_BBOX_INSIDE_WEIGHTS=(1.0, 1.0, 1.0, 1.0), RPN_POSITIVE_WEIGHT=-1.0,
        RPN_NMS_THRESH=0.7, RPN_PRE_NMS_TOP_N=12000, RPN_POST_NMS_TOP_N=6000, RPN_MIN_SIZE=16,
        BBOX_NORMALIZATION_PRECOMPUTED=False, BBOX_MEANS=(0.0, 0.0, 0.0, 0.0), BBOX_STDS=(0.1, 0.1, 0.2, 0.2),
        BBOX_REGRESSION_THRESH=0.5)
    c.TEST = Cfg(NMS=0.3, RPN_NMS_THRESH=0.7, RPN_PRE_NMS_TOP_N=6000, RPN_POST_NMS_TOP_N=300, RPN_MIN_SIZE=16)
    return c


def _desc_order(v, tie="stable"):
    """``v.argsort()[::-1]`` (rcnn/rpn/proposal.py:123, helper/processing/nms.py:19)."""
    if tie == "stable":
        return v.argsort(kind="stable")[::-1]
    return v.argsort()[::-1]


# ----------------------------------------------------------------------------------------
# A0  helper/processing/generate_anchor.py:8-72
# ----------------------------------------------------------------------------------------
def generate_anchors(base_size=16, ratios=(0.5, 1, 2), scales=2 ** np.arange(3, 6)):
    ratios = np.asarray(ratios, dtype=np.float64)
    scales = np.asarray(scales)

    def whc(a):  # :23-32
        w = a[2] - a[0] + 1
        h = a[3] - a[1] + 1
        return w, h, a[0] + 0.5 * (w - 1), a[1] + 0.5 * (h - 1)

    def mk(ws, hs, cx, cy):  # :35-46
        ws, hs = ws[:, None], hs[:, None]
        return np.hstack((cx - 0.5 * (ws - 1), cy - 0.5 * (hs - 1), cx + 0.5 * (ws - 1), cy + 0.5 * (hs - 1)))

    base = np.array([1, 1, base_size, base_size]) - 1  # :14
    w, h, cx, cy = whc(base)  # :49-60 ratio enumeration
    size_ratios = (w * h) / ratios
    ws = np.round(np.sqrt(size_ratios))
    hs = np.round(ws * ratios)
    ratio_anchors = mk(ws, hs, cx, cy)
    rows = []
    for i in range(ratio_anchors.shape[0]):  # :16-18, scale enumeration :63-72
        w, h, cx, cy = whc(ratio_anchors[i, :])
        rows.append(mk(w * scales, h * scales, cx, cy))
    return np.vstack(rows)


# ----------------------------------------------------------------------------------------
# A2/A3/A10  helper/processing/bbox_transform.py
# ----------------------------------------------------------------------------------------
def bbox_pred(boxes, box_deltas):
    """bbox_transform.py:35-78 (is_train clamp branch :56-64 is dead code: no caller enables it)."""
    if boxes.shape[0] == 0:
        return np.zeros((0, box_deltas.shape[1]))  # :43-44
    boxes = boxes.astype(np.float64, copy=False)  # :46
    w = boxes[:, 2] - boxes[:, 0] + 1.0
    h = boxes[:, 3] - boxes[:, 1] + 1.0
    cx = boxes[:, 0] + 0.5 * (w - 1.0)
    cy = boxes[:, 1] + 0.5 * (h - 1.0)
    dx, dy, dw, dh = (box_deltas[:, k::4] for k in range(4))  # :52-55
    pcx = dx * w[:, None] + cx[:, None]  # :65-66
    pcy = dy * h[:, None] + cy[:, None]
    pw = np.exp(dw) * w[:, None]  # :67-68  exp evaluated in the dtype of the deltas (f32)
    ph = np.exp(dh) * h[:, None]
    out = np.zeros(box_deltas.shape)  # :70 float64
    out[:, 0::4] = pcx - 0.5 * (pw - 1.0)
    out[:, 1::4] = pcy - 0.5 * (ph - 1.0)
    out[:, 2::4] = pcx + 0.5 * (pw - 1.0)
    out[:, 3::4] = pcy + 0.5 * (ph - 1.0)
    return out


def clip_boxes(boxes, im_shape):
    """bbox_transform.py:81-96, in place; x uses im_shape[1], y uses im_shape[0]."""
    for k in range(4):
        lim = im_shape[1] - 1 if k % 2 == 0 else im_shape[0] - 1
        boxes[:, k::4] = np.maximum(np.minimum(boxes[:, k::4], lim), 0)
    return boxes


def clip_pad(t, pad_shape):
    """bbox_transform.py:98-111."""
    H, W = t.shape[2:]
    h, w = pad_shape
    if h < H:
        t = t[:, :, :h, :].copy()
    if w < W:
        t = t[:, :, :, :w].copy()
    return t


def bbox_transform(ex, gt):
    """bbox_transform.py:8-32 (arithmetic in the promoted dtype of the inputs)."""
    ew = ex[:, 2] - ex[:, 0] + 1.0
    eh = ex[:, 3] - ex[:, 1] + 1.0
    ecx = ex[:, 0] + 0.5 * (ew - 1.0)
    ecy = ex[:, 1] + 0.5 * (eh - 1.0)
    gw = gt[:, 2] - gt[:, 0] + 1.0
    gh = gt[:, 3] - gt[:, 1] + 1.0
    gcx = gt[:, 0] + 0.5 * (gw - 1.0)
    gcy = gt[:, 1] + 0.5 * (gh - 1.0)
    return np.vstack(((gcx - ecx) / (ew + 1e-14), (gcy - ecy) / (eh + 1e-14),
                      np.log(gw / ew), np.log(gh / eh))).transpose()


# ----------------------------------------------------------------------------------------
# A5  helper/processing/nms.py:4-38
# ----------------------------------------------------------------------------------------
def nms(dets, thresh, tie="stable", stop_after=0, audit=None):
    """Greedy NMS, arithmetic in ``dets.dtype``.  Returns a python list (quirk Q6).

    ``stop_after`` > 0 stops once that many boxes are kept (legal: keep[:post] is a prefix,
    rcnn/rpn/proposal.py:132-133) -- used only for bounded audits / CPU timing samples.
    ``audit`` (dict) receives ``pairs`` (IoU evaluations) and ``min_margin`` = min |ovr-thresh|.
    """
    x1, y1, x2, y2, sc = (dets[:, k] for k in range(5))
    areas = (x2 - x1 + 1) * (y2 - y1 + 1)  # :18
    order = _desc_order(sc, tie)  # :19
    keep = []
    pairs, margin = 0, np.inf
    while order.size > 0:  # :22
        i = order[0]
        keep.append(i)
        if stop_after and len(keep) >= stop_after:
            break
        rest = order[1:]
        xx1 = np.maximum(x1[i], x1[rest])  # :25-28
        yy1 = np.maximum(y1[i], y1[rest])
        xx2 = np.minimum(x2[i], x2[rest])
        yy2 = np.minimum(y2[i], y2[rest])
        w = np.maximum(0.0, xx2 - xx1 + 1)  # :30-31
        h = np.maximum(0.0, yy2 - yy1 + 1)
        inter = w * h
        ovr = inter / (areas[i] + areas[rest] - inter)  # :33
        if audit is not None and ovr.size:
            pairs += ovr.size
            margin = min(margin, float(np.min(np.abs(ovr - thresh))))
        order = order[np.where(ovr <= thresh)[0] + 1]  # :35-36
    if audit is not None:
        audit["pairs"] = pairs
        audit["min_margin"] = margin
    return keep


def nest(dets, thresh=0.90):
    """helper/processing/nms.py:40-70: drop box i if inter(i,j)/area(i) > thresh for any j != i."""
    x1, y1, x2, y2 = (dets[:, k] for k in range(4))
    areas = (x2 - x1 + 1) * (y2 - y1 + 1)
    n = len(dets)
    if n == 0:
        return []
    xx1 = np.maximum(x1[:, None], x1[None, :])
    yy1 = np.maximum(y1[:, None], y1[None, :])
    xx2 = np.minimum(x2[:, None], x2[None, :])
    yy2 = np.minimum(y2[:, None], y2[None, :])
    w = np.maximum(0, xx2 - xx1 + 1)
    h = np.maximum(0, yy2 - yy1 + 1)
    frac = (w * h) / areas[:, None]
    bad = frac > thresh
    np.fill_diagonal(bad, False)
    return [int(i) for i in np.where(~bad.any(axis=1))[0]]


# ----------------------------------------------------------------------------------------
# A7/A9  helper/processing/bbox_regression.py
# ----------------------------------------------------------------------------------------
def bbox_overlaps(boxes, query_boxes):
    """bbox_regression.py:11-31, vectorised with the same per-pair operation order (f64).

    Only columns 0..3 of either input are read.  ``bbox_overlaps_loop`` is the literal
    double loop, used by the tests on small cases to pin this vectorisation.
    """
    b = np.asarray(boxes, dtype=np.float64)
    q = np.asarray(query_boxes, dtype=np.float64)
    n_, k_ = b.shape[0], q.shape[0]
    if n_ == 0 or k_ == 0:
        return np.zeros((n_, k_), dtype=np.float64)
    qa = (q[:, 2] - q[:, 0] + 1) * (q[:, 3] - q[:, 1] + 1)  # :21
    ba = (b[:, 2] - b[:, 0] + 1) * (b[:, 3] - b[:, 1] + 1)  # :27
    iw = np.minimum(b[:, None, 2], q[None, :, 2]) - np.maximum(b[:, None, 0], q[None, :, 0]) + 1  # :23
    ih = np.minimum(b[:, None, 3], q[None, :, 3]) - np.maximum(b[:, None, 1], q[None, :, 1]) + 1  # :25
    ok = (iw > 0) & (ih > 0)
    inter = iw * ih
    with np.errstate(divide="ignore", invalid="ignore"):
        ov = inter / (ba[:, None] + qa[None, :] - inter)  # :28-29
    return np.where(ok, ov, 0.0)


def bbox_overlaps_loop(boxes, query_boxes):
    """Literal restatement of the double loop bbox_regression.py:18-31 (small cases only)."""
    n_, k_ = boxes.shape[0], query_boxes.shape[0]
    out = np.zeros((n_, k_), dtype=np.float64)
    for k in range(k_):
        qa = (query_boxes[k, 2] - query_boxes[k, 0] + 1) * (query_boxes[k, 3] - query_boxes[k, 1] + 1)
        for n in range(n_):
            iw = min(boxes[n, 2], query_boxes[k, 2]) - max(boxes[n, 0], query_boxes[k, 0]) + 1
            if iw > 0:
                ih = min(boxes[n, 3], query_boxes[k, 3]) - max(boxes[n, 1], query_boxes[k, 1]) + 1
                if ih > 0:
                    ba = (boxes[n, 2] - boxes[n, 0] + 1) * (boxes[n, 3] - boxes[n, 1] + 1)
                    out[n, k] = iw * ih / float(ba + qa - iw * ih)
    return out


def expand_bbox_regression_targets(data, num_classes, inside_weights=(1.0, 1.0, 1.0, 1.0)):
    """bbox_regression.py:67-85."""
    cls = data[:, 0]
    t = np.zeros((cls.size, 4 * num_classes), dtype=np.float32)
    w = np.zeros(t.shape, dtype=np.float32)
    for i in np.where(cls > 0)[0]:
        s = int(4 * cls[i])
        t[i, s:s + 4] = data[i, 1:]
        w[i, s:s + 4] = inside_weights
    return t, w


# ----------------------------------------------------------------------------------------
# A1  rcnn/rpn/proposal.py:39-149
# ----------------------------------------------------------------------------------------
def proposal_stages(cls_prob, bbox_deltas, im_info, feat_stride=16, scales=(8, 16, 32), ratios=(0.5, 1, 2),
                    train_crop=False, pre_nms_topN=6000, post_nms_topN=300, nms_thresh=0.7, min_size=16,
                    tie="stable", rng=None, nms_stop_after=0, audit=False):
    """ProposalOperator.forward, returning every intermediate stage in a dict.

    ``train_crop`` = the cfg_key == 'TRAIN' branch at proposal.py:70-71 (feature-map crop, quirk Q1).
    ``nms_stop_after``: bound the greedy loop (timing samples only; outputs identical because
    keep[:post] is a prefix).
    """
    fs = float(feat_stride)  # proposal.py:22
    base = generate_anchors(base_size=fs, scales=np.asarray(scales, dtype=np.float64), ratios=list(ratios))  # :25
    A = base.shape[0]
    scores = cls_prob[:, A:, :, :]  # :57
    if np.isnan(scores).any():
        raise ValueError("there is nan in input scores")  # :58-59
    if np.isnan(bbox_deltas).any():
        raise ValueError("there is nan in input bbox_deltas")  # :61-62
    im = im_info[0, :]  # :63
    height, width = scores.shape[-2:]
    if train_crop:
        height, width = int(im[0] / fs), int(im[1] / fs)  # :70-71
    sx = np.arange(0, width) * fs  # :77-80
    sy = np.arange(0, height) * fs
    sx, sy = np.meshgrid(sx, sy)
    shifts = np.vstack((sx.ravel(), sy.ravel(), sx.ravel(), sy.ravel())).transpose()
    K = shifts.shape[0]
    anchors = (base.reshape((1, A, 4)) + shifts.reshape((1, K, 4)).transpose((1, 0, 2))).reshape((K * A, 4))  # :88-91
    d = clip_pad(bbox_deltas, (height, width))  # :99
    d = d.transpose((0, 2, 3, 1)).reshape((-1, 4))  # :100
    s = scores.transpose((0, 2, 3, 1)).reshape((-1, 1))  # :107 (NOT cropped: quirk Q1)
    props = bbox_pred(anchors, d)  # :110
    props = clip_boxes(props, im[:2])  # :113
    ms = min_size * im[2]  # :117  (f32 scalar under numpy>=2)
    ws = props[:, 2] - props[:, 0] + 1  # :158-160
    hs = props[:, 3] - props[:, 1] + 1
    fkeep = np.where((ws >= ms) & (hs >= ms))[0]
    out = {"anchors": anchors, "decoded": props.copy(), "filter_keep": fkeep, "min_size_scaled": ms}
    if audit:
        out["margin_filter"] = float(min(np.min(np.abs(ws - ms)), np.min(np.abs(hs - ms)))) if ws.size else np.inf
    props = props[fkeep, :]
    s = s[fkeep]
    order = _desc_order(s.ravel(), tie)  # :123
    if pre_nms_topN > 0:
        order = order[:pre_nms_topN]  # :124-125
    props = props[order, :]
    s = s[order]
    dets = np.hstack((props, s))  # :131  float64 (n,5)
    aud = {} if audit else None
    keep = nms(dets, nms_thresh, tie=tie, stop_after=nms_stop_after, audit=aud)
    if audit:
        out["nms_pairs"], out["margin_nms"] = aud["pairs"], aud["min_margin"]
    out["nms_keep_full_len"] = len(keep)
    if post_nms_topN > 0:
        keep = keep[:post_nms_topN]  # :132-133
    nkeep = len(keep)
    pad_draw = None
    if len(keep) < post_nms_topN:  # :135-139
        if len(keep) == 0:
            logging.log(logging.ERROR, "currently len(keep) is zero")
        rs = rng if rng is not None else np.random
        # npr.choice(keep, size=k) == keep[randint(0, len(keep), k)] on the legacy RandomState
        pad = rs.choice(keep, size=post_nms_topN - len(keep))
        pad_draw = np.asarray(pad)
        keep = np.hstack((keep, pad))
    keep = np.asarray(keep, dtype=np.int64)
    props_k = props[keep, :]
    s_k = s[keep]
    blob = np.hstack((np.zeros((props_k.shape[0], 1), dtype=np.float32), props_k.astype(np.float32, copy=False)))  # :145-146
    out.update({
        "order": fkeep[order],  # original (h,w,a) anchor-row index of each sorted candidate
        "dets": dets, "keep": keep, "nkeep": nkeep, "pad_draw": pad_draw,
        "keep_anchor_index": fkeep[order][keep] if keep.size else np.zeros((0,), np.int64),
        "rois": blob, "scores": s_k.astype(np.float32, copy=False),
    })
    return out


# ----------------------------------------------------------------------------------------
# A6  rcnn/minibatch.py:204-395
# ----------------------------------------------------------------------------------------
def assign_anchor(feat_shape, gt_boxes, im_info, feat_stride=16, scales=(8, 16, 32), ratios=(0.5, 1, 2),
                  allowed_border=0, cfg=None, rng=None, record=None):
    """RPN anchor labels / targets / weights.  ``record`` (dict) receives intermediates."""
    cfg = cfg or default_config()
    T = cfg.TRAIN
    rs = rng if rng is not None else np.random
    im = im_info[0]  # :242
    sc = np.array(scales, dtype=np.float32)  # :243
    base = generate_anchors(base_size=16, ratios=list(ratios), scales=sc)  # :244
    A = base.shape[0]
    fh, fw = feat_shape[-2:]
    sx = np.arange(0, fw) * feat_stride  # :260-263
    sy = np.arange(0, fh) * feat_stride
    sx, sy = np.meshgrid(sx, sy)
    shifts = np.vstack((sx.ravel(), sy.ravel(), sx.ravel(), sy.ravel())).transpose()
    K = shifts.shape[0]
    allanc = (base.reshape((1, A, 4)) + shifts.reshape((1, K, 4)).transpose((1, 0, 2))).reshape((K * A, 4))  # :270-271
    total = int(K * A)
    inside = np.where((allanc[:, 0] >= -allowed_border) & (allanc[:, 1] >= -allowed_border) &
                      (allanc[:, 2] < im[1] + allowed_border) & (allanc[:, 3] < im[0] + allowed_border))[0]  # :275-278
    anc = allanc[inside, :]
    labels = np.empty((len(inside),), dtype=np.float32)  # :289-290
    labels.fill(-1)
    argmax = None
    if gt_boxes.size > 0:
        ov = bbox_overlaps(anc.astype(np.float64), gt_boxes.astype(np.float64))  # :295
        argmax = ov.argmax(axis=1)  # :296
        maxov = ov[np.arange(len(inside)), argmax]
        gt_arg = ov.argmax(axis=0)  # :298-300
        gt_max = ov[gt_arg, np.arange(ov.shape[1])]
        gt_arg = np.where(ov == gt_max)[0]  # quirk Q4: equality, all ties, incl. gt whose max is 0
        if not T.RPN_CLOBBER_POSITIVES:
            labels[maxov < T.RPN_NEGATIVE_OVERLAP] = 0  # :302-304
        labels[gt_arg] = 1  # :307
        labels[maxov >= T.RPN_POSITIVE_OVERLAP] = 1  # :310
        if T.RPN_CLOBBER_POSITIVES:
            labels[maxov < T.RPN_NEGATIVE_OVERLAP] = 0  # :312-314
        if record is not None:
            record.update(max_overlaps=maxov, argmax_overlaps=argmax, gt_max_overlaps=gt_max)
    else:
        labels[:] = 0  # :316
    if record is not None:
        record.update(inds_inside=inside, labels_presample=labels.copy())
    num_fg = int(T.RPN_FG_FRACTION * T.RPN_BATCH_SIZE)  # :319
    fg = np.where(labels == 1)[0]
    fg_disable = bg_disable = None
    if len(fg) > num_fg:  # :321-325
        fg_disable = rs.choice(fg, size=(len(fg) - num_fg), replace=False)
        labels[fg_disable] = -1
    num_bg = T.RPN_BATCH_SIZE - np.sum(labels == 1)  # :328
    bg = np.where(labels == 0)[0]
    if len(bg) > num_bg:  # :330-334
        bg_disable = rs.choice(bg, size=(len(bg) - num_bg), replace=False)
        labels[bg_disable] = -1
    if record is not None:
        record.update(fg_disable=fg_disable, bg_disable=bg_disable, n_fg_pre=len(fg), n_bg_pre=len(bg))
    tgt = np.zeros((len(inside), 4), dtype=np.float32)  # :336-338
    if gt_boxes.size > 0:
        tgt[:] = bbox_transform(anc, gt_boxes[argmax, :4]).astype(np.float32, copy=False)
    inw = np.zeros((len(inside), 4), dtype=np.float32)  # :340-341
    inw[labels == 1, :] = np.array(T.RPN_BBOX_INSIDE_WEIGHTS)
    outw = np.zeros((len(inside), 4), dtype=np.float32)  # :343-354
    if T.RPN_POSITIVE_WEIGHT < 0:
        nex = np.sum(labels >= 0)
        with np.errstate(divide="ignore"):
            pw = np.ones((1, 4)) * 1.0 / nex
            nw = np.ones((1, 4)) * 1.0 / nex
    else:
        assert (T.RPN_POSITIVE_WEIGHT > 0) & (T.RPN_POSITIVE_WEIGHT < 1)
        pw = T.RPN_POSITIVE_WEIGHT / np.sum(labels == 1)
        nw = (1.0 - T.RPN_POSITIVE_WEIGHT) / np.sum(labels == 1)
    outw[labels == 1, :] = pw
    outw[labels == 0, :] = nw

    def unmap(data, fill):  # :221-231
        if data.ndim == 1:
            r = np.empty((total,), dtype=np.float32)
            r.fill(fill)
            r[inside] = data
        else:
            r = np.empty((total,) + data.shape[1:], dtype=np.float32)
            r.fill(fill)
            r[inside, :] = data
        return r

    labels = unmap(labels, -1)  # :366-369
    tgt, inw, outw = unmap(tgt, 0), unmap(inw, 0), unmap(outw, 0)
    labels = labels.reshape((1, fh, fw, A)).transpose(0, 3, 1, 2).reshape((1, A * fh * fw))  # :381-382
    tgt = tgt.reshape((1, fh, fw, A * 4)).transpose(0, 3, 1, 2)  # :383-385
    inw = inw.reshape((1, fh, fw, A * 4)).transpose((0, 3, 1, 2))
    outw = outw.reshape((1, fh, fw, A * 4)).transpose((0, 3, 1, 2))
    label = {"label": labels, "bbox_target": tgt, "bbox_inside_weight": inw, "bbox_outside_weight": outw}
    if cfg.END2END == 1:
        label.update({"gt_boxes": gt_boxes})  # :392-393
    return label


# ----------------------------------------------------------------------------------------
# A8  rcnn/rpn/proposal_target.py:34-76,120-196
# ----------------------------------------------------------------------------------------
def _pt_compute_targets(ex_rois, gt_rois, labels, cfg):
    """proposal_target.py:120-133."""
    assert ex_rois.shape[0] == gt_rois.shape[0] and ex_rois.shape[1] == 4 and gt_rois.shape[1] == 4
    t = bbox_transform(ex_rois, gt_rois)
    if cfg.TRAIN.BBOX_NORMALIZATION_PRECOMPUTED:
        t = (t - np.array(cfg.TRAIN.BBOX_MEANS)) / np.array(cfg.TRAIN.BBOX_STDS)
    return np.hstack((labels[:, np.newaxis], t)).astype(np.float32, copy=False)


def sample_rois_e2e(all_rois, gt_boxes, fg_rois_per_image, rois_per_image, num_classes, key="TRAIN",
                    cfg=None, rng=None, record=None):
    """``_sample_rois`` of proposal_target.py:135-196.  RNG order: fg(no-repl), fg(repl, cond.),
    bg(no-repl), bg(repl, cond.)."""
    cfg = cfg or default_config()
    T = cfg.TRAIN
    rs = rng if rng is not None else np.random
    ov = bbox_overlaps(np.ascontiguousarray(all_rois[:, 1:5], dtype=np.float64),
                       np.ascontiguousarray(gt_boxes[:, :4], dtype=np.float64))  # :140-142
    assign = ov.argmax(axis=1)
    maxov = ov.max(axis=1)
    labels = gt_boxes[assign, 4]  # :145
    fg = np.where(maxov >= T.FG_THRESH)[0]  # :148
    fg_this = min(fg_rois_per_image, fg.size)  # :151
    if fg.size > 0:
        fg = rs.choice(fg, size=fg_this, replace=False)  # :153-154
    if fg.size < fg_rois_per_image:  # :155-157 (padding is PREPENDED: quirk Q3)
        fg_ = rs.choice(fg, size=fg_rois_per_image - fg.size, replace=True)
        fg = np.hstack((fg_, fg))
    bg_ = np.where((maxov < T.BG_THRESH_HI) & (maxov >= T.BG_THRESH_LO))[0]  # :160-161
    if len(bg_) == 0 and key == "TRAIN":
        bg = np.where((maxov < T.BG_THRESH_HI + 0.2) & (maxov >= 0))[0]  # :162-164
    else:
        bg = bg_
    if len(bg) == 0:
        logging.log(logging.ERROR, "currently len(bg_inds) is zero")  # :168-169
    bg_this = rois_per_image - len(fg)  # :173-174
    bg_this = min(bg_this, bg.size)
    if bg.size > 0:
        bg = rs.choice(bg, size=bg_this, replace=False)  # :176-177
    if bg.size < rois_per_image - fg_rois_per_image:  # :178-180
        bg_ = rs.choice(bg, size=rois_per_image - fg_rois_per_image - bg.size, replace=True)
        bg = np.hstack((bg_, bg))
    keep = np.append(fg, bg)  # :183
    labels = labels[keep]
    labels[fg_this:] = 0  # :187 (quirk Q3)
    rois = all_rois[keep]
    data = _pt_compute_targets(rois[:, 1:5], gt_boxes[assign[keep], :4], labels, cfg)  # :190-191
    bt, bw = expand_bbox_regression_targets(data, num_classes, T.BBOX_INSIDE_WEIGHTS)  # :193-194
    if record is not None:
        record.update(max_overlaps=maxov, gt_assignment=assign, keep_inds=np.asarray(keep, dtype=np.int64),
                      fg_this=int(fg_this))
    return labels, rois, bt, bw


def proposal_target_forward(rpn_rois, gt_boxes_padded, num_classes, cfg=None, rng=None, key="TRAIN", record=None):
    """ProposalTargetOperator.forward, proposal_target.py:34-76.  Returns the 5 outputs."""
    cfg = cfg or default_config()
    img_per_batch = 1
    assert cfg.TRAIN.BATCH_SIZE % img_per_batch == 0
    rois_per_image = cfg.TRAIN.BATCH_SIZE // img_per_batch  # :39 (py2 int division, quirk Q5)
    fg_per = np.round(cfg.TRAIN.FG_FRACTION * rois_per_image).astype(int)  # :40
    all_rois = rpn_rois
    gt = gt_boxes_padded[np.where(gt_boxes_padded[:, :5].mean(axis=1) != -1)]  # :43
    zeros = np.zeros((gt.shape[0], 1), dtype=gt.dtype)
    all_rois = np.vstack((all_rois, np.hstack((zeros, gt[:, :-1]))))  # :46-49
    assert np.all(all_rois[:, 0] == 0), "Only single item batches are supported"  # :51
    labels, rois, bt, bw = sample_rois_e2e(all_rois, gt, fg_per, rois_per_image, num_classes, key, cfg, rng, record)
    return rois, labels, bt, bw, np.array(bw > 0).astype(np.float32)  # :72-76


# ----------------------------------------------------------------------------------------
# A11  rcnn/minibatch.py:151-201
# ----------------------------------------------------------------------------------------
def sample_rois(roidb, fg_rois_per_image, rois_per_image, num_classes, cfg=None, rng=None, record=None):
    """Fast R-CNN sampling from a precomputed roidb.  RNG order: fg, bg, gap."""
    cfg = cfg or default_config()
    T = cfg.TRAIN
    rs = rng if rng is not None else np.random
    labels = roidb["max_classes"]
    overlaps = roidb["max_overlaps"]
    rois = roidb["boxes"]
    fg = np.where(overlaps >= T.FG_THRESH)[0]  # :166
    fg_this = np.minimum(fg_rois_per_image, fg.size)
    if fg.size > 0:
        fg = rs.choice(fg, size=fg_this, replace=False)  # :170-171
    bg = np.where((overlaps < T.BG_THRESH_HI) & (overlaps >= T.BG_THRESH_LO))[0]  # :174
    bg_this = np.minimum(rois_per_image - fg_this, bg.size)
    if bg.size > 0:
        bg = rs.choice(bg, size=bg_this, replace=False)  # :179-180
    keep = np.append(fg, bg)  # :183
    if keep.shape[0] < rois_per_image:  # :186-189
        gap = rois_per_image - keep.shape[0]
        keep = np.append(keep, rs.choice(range(len(rois)), size=gap, replace=False))
    keep = keep.astype(np.int64)
    labels = labels[keep]
    labels[fg_this:] = 0  # :194
    overlaps = overlaps[keep]
    rois = rois[keep]
    bt, bw = expand_bbox_regression_targets(roidb["bbox_targets"][keep, :], num_classes, T.BBOX_INSIDE_WEIGHTS)
    if record is not None:
        record.update(keep_indexes=keep, fg_this=int(fg_this))
    return rois, labels, bt, bw, overlaps


# ----------------------------------------------------------------------------------------
# N1  test-time box head: rcnn/detector.py:73-79 + the class loop of rcnn/tester.py:50-65
# ----------------------------------------------------------------------------------------
def box_head(rois, bbox_deltas, scores, im_shape, scale, thresh=0.05, nms_thresh=0.3, max_per_image=100, tie="stable"):
    """Detections of one image.  ``rois`` (R,4) f32 (batch column already dropped, detector.py:68),
    ``bbox_deltas`` (R,4C) f32, ``scores`` (R,C) f32, ``im_shape`` = (h, w) of the network input,
    ``scale`` = im_info[0, 2] (np.float32) or a python float.

    Returns ``(dets, rows)``: per class j (index 0 = background = empty) the (n_j,5) float64 array
    ``all_boxes[j][i]`` of tester.py:56,64 and the roi row each detection came from.
    """
    pred = clip_boxes(bbox_pred(rois, bbox_deltas), im_shape)  # detector.py:73-74
    C = scores.shape[1]
    dets = [np.zeros((0, 5))]
    rows = [np.zeros((0,), dtype=np.int64)]
    for j in range(1, C):  # tester.py:50-56
        idx = np.where(scores[:, j] > thresh)[0]  # f32 array vs python float: compared in float32
        cls_scores = scores[idx, j]
        cls_boxes = pred[idx, j * 4:(j + 1) * 4] / scale  # f64 / (f32 scalar | python float) -> f64
        cls_dets = np.hstack((cls_boxes, cls_scores[:, np.newaxis]))
        keep = np.asarray(nms(cls_dets, nms_thresh, tie=tie), dtype=np.int64)
        dets.append(cls_dets[keep, :])
        rows.append(idx[keep])
    if max_per_image > 0:  # tester.py:58-65
        image_scores = np.hstack([dets[j][:, -1] for j in range(1, C)])
        if len(image_scores) > max_per_image:
            image_thresh = np.sort(image_scores)[-max_per_image]
            for j in range(1, C):
                k = np.where(dets[j][:, -1] >= image_thresh)[0]
                dets[j] = dets[j][k, :]
                rows[j] = rows[j][k]
    return dets, rows


# ----------------------------------------------------------------------------------------
# N3  offline roidb users of bbox_overlaps
# ----------------------------------------------------------------------------------------
def compute_bbox_regression_targets(rois, overlaps, labels, cfg=None):
    """helper/processing/bbox_regression.py:34-64 -> (k,5) float32 [class, dx, dy, dw, dh]."""
    cfg = cfg or default_config()
    rois = rois.astype(np.float64, copy=False)  # :43
    gt_inds = np.where(overlaps == 1)[0]  # :46
    ex_inds = np.where(overlaps >= cfg.TRAIN.BBOX_REGRESSION_THRESH)[0]  # :50
    ex_gt = bbox_overlaps(rois[ex_inds, :], rois[gt_inds, :])  # :53
    gt_assignment = ex_gt.argmax(axis=1)  # :57
    gt_rois = rois[gt_inds[gt_assignment], :]
    ex_rois = rois[ex_inds, :]
    targets = np.zeros((rois.shape[0], 5), dtype=np.float32)
    targets[ex_inds, 0] = labels[ex_inds]
    targets[ex_inds, 1:] = bbox_transform(ex_rois, gt_rois)  # f64 -> f32 on assignment
    return targets


def add_bbox_regression_targets(roidb, cfg=None):
    """helper/processing/roidb.py:42-90: per-image targets, then per-class mean/std normalisation (in place)."""
    cfg = cfg or default_config()
    num_images = len(roidb)
    num_classes = roidb[0]["gt_overlaps"].shape[1]
    for r in roidb:
        r["bbox_targets"] = compute_bbox_regression_targets(r["boxes"], r["max_overlaps"], r["max_classes"], cfg)
    if cfg.TRAIN.BBOX_NORMALIZATION_PRECOMPUTED:  # :58-61
        means = np.tile(np.array(cfg.TRAIN.BBOX_MEANS), (num_classes, 1))
        stds = np.tile(np.array(cfg.TRAIN.BBOX_STDS), (num_classes, 1))
    else:  # :62-79
        class_counts = np.zeros((num_classes, 1)) + cfg.EPS
        sums = np.zeros((num_classes, 4))
        squared_sums = np.zeros((num_classes, 4))
        for im_i in range(num_images):
            t = roidb[im_i]["bbox_targets"]
            for cls in range(1, num_classes):
                idx = np.where(t[:, 0] == cls)[0]
                if idx.size > 0:
                    class_counts[cls] += idx.size
                    sums[cls, :] += t[idx, 1:].sum(axis=0)
                    squared_sums[cls, :] += (t[idx, 1:] ** 2).sum(axis=0)
        means = sums / class_counts
        stds = np.sqrt(squared_sums / class_counts - means ** 2)
    for im_i in range(num_images):  # :82-88
        t = roidb[im_i]["bbox_targets"]
        for cls in range(1, num_classes):
            idx = np.where(t[:, 0] == cls)[0]
            roidb[im_i]["bbox_targets"][idx, 1:] -= means[cls, :]
            roidb[im_i]["bbox_targets"][idx, 1:] /= stds[cls, :]
    return means.ravel(), stds.ravel()


def box_list_overlaps(boxes, gt_boxes, gt_classes, num_classes):
    """The per-image body of IMDB.create_roidb_from_box_list (helper/dataset/imdb.py:43-55): dense (n, C) float32
    with, per box, its best IoU in the column of that gt's class (the reference then wraps it in a csr_matrix)."""
    overlaps = np.zeros((boxes.shape[0], num_classes), dtype=np.float32)
    if gt_boxes is not None and gt_boxes.size > 0:
        gt_overlaps = bbox_overlaps(boxes.astype(np.float64), gt_boxes.astype(np.float64))  # :50
        argmaxes = gt_overlaps.argmax(axis=1)
        maxes = gt_overlaps.max(axis=1)
        idx = np.where(maxes > 0)[0]
        overlaps[idx, gt_classes[argmaxes[idx]]] = maxes[idx]
    return overlaps


def recall_gt_overlaps(boxes, gt_boxes):
    """Greedy one-to-one matching of helper/dataset/imdb.py:151-172: the IoU with which each gt box is covered,
    in the order the loop records them."""
    overlaps = bbox_overlaps(boxes.astype(np.float64), gt_boxes.astype(np.float64))
    out = np.zeros((gt_boxes.shape[0]))
    for j in range(gt_boxes.shape[0]):
        argmax_overlaps = overlaps.argmax(axis=0)
        max_overlaps = overlaps.max(axis=0)
        gt_ind = max_overlaps.argmax()
        box_ind = argmax_overlaps[gt_ind]
        out[j] = overlaps[box_ind, gt_ind]
        overlaps[box_ind, :] = -1
        overlaps[:, gt_ind] = -1
    return out
