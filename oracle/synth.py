"""Seeded synthetic inputs for the hot path (SURVEY.md section 8d recipes).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  bench.py and the tests build their
inputs here so that the CUDA path, the restatement and the real reference all see the
same arrays.  Everything derives from ``np.random.RandomState(seed)``.
"""
import numpy as np

# Named workloads of BASELINE.json ``configs``.
CONFIGS = {
    # C1/C2: VGG16 stride-16 map 38x63, 9 anchors, pre 6000, NMS 0.7, post 300
    "voc_test": dict(H=38, W=63, scales=(8, 16, 32), ratios=(0.5, 1, 2), feat_stride=16, im_info=(600, 1000, 1.6),
                     pre=6000, post=300, thresh=0.7, min_size=16),
    # reference TRAIN defaults (rcnn/config.py:47-50)
    "voc_train": dict(H=38, W=63, scales=(8, 16, 32), ratios=(0.5, 1, 2), feat_stride=16, im_info=(600, 1000, 1.6),
                      pre=12000, post=6000, thresh=0.7, min_size=16),
    # alternate-training generation (train_alternate.py:30-31)
    "voc_alt": dict(H=38, W=63, scales=(8, 16, 32), ratios=(0.5, 1, 2), feat_stride=16, im_info=(600, 1000, 1.6),
                    pre=-1, post=2000, thresh=0.7, min_size=16),
    # C5: WIDER-face stress
    "wider": dict(H=75, W=100, scales=(4, 8, 16, 32), ratios=(0.5, 1, 2), feat_stride=16, im_info=(1200, 1600, 1.0),
                  pre=12000, post=2000, thresh=0.7, min_size=16),
    # small case for fast tests
    "tiny": dict(H=6, W=9, scales=(8, 16, 32), ratios=(0.5, 1, 2), feat_stride=16, im_info=(96, 144, 1.0),
                 pre=200, post=30, thresh=0.7, min_size=16),
}


def proposal_inputs(seed, H=38, W=63, A=9, im_info=(600, 1000, 1.6), pad_hw=None, **_):
    """cls_prob (1,2A,H,W) f32, bbox_pred (1,4A,H,W) f32, im_info (1,3) f32.

    Scores are provably distinct: ``(permutation(N)+0.5)/N`` in (h,w,a) order.
    Deltas: dx,dy ~ N(0,0.1); dw,dh ~ N(0,0.25).  ``pad_hw=(Hp,Wp)`` embeds the map in a
    larger zero-padded one (TRAIN-mode crop / quirk Q1 tests).
    """
    rs = np.random.RandomState(seed)
    N = A * H * W
    fg = ((rs.permutation(N) + 0.5) / N).astype(np.float32).reshape(H, W, A)
    d = np.empty((N, 4), dtype=np.float32)
    d[:, 0:2] = rs.normal(0, 0.1, size=(N, 2))
    d[:, 2:4] = rs.normal(0, 0.25, size=(N, 2))
    d = d.reshape(H, W, A * 4)
    Hp, Wp = pad_hw if pad_hw else (H, W)
    cls = np.zeros((1, 2 * A, Hp, Wp), dtype=np.float32)
    box = np.zeros((1, 4 * A, Hp, Wp), dtype=np.float32)
    cls[0, A:, :H, :W] = fg.transpose(2, 0, 1)
    cls[0, :A, :H, :W] = 1.0 - fg.transpose(2, 0, 1)
    box[0, :, :H, :W] = d.transpose(2, 0, 1)
    if pad_hw:
        # give the padded region distinct (low) scores too, so quirk Q1 has something to read
        extra = Hp * Wp * A - N
        pad_scores = ((rs.permutation(extra) + 0.25) / (4.0 * (extra + N))).astype(np.float32)
        full = cls[0, A:].transpose(1, 2, 0).copy()
        mask = np.ones((Hp, Wp), dtype=bool)
        mask[:H, :W] = False
        full[mask] = pad_scores.reshape(-1, A)
        cls[0, A:] = full.transpose(2, 0, 1)
    info = np.array([im_info], dtype=np.float32)
    return cls, box, info


def proposal_batch(seeds, cfg):
    A = len(cfg["scales"]) * len(cfg["ratios"])
    parts = [proposal_inputs(s, H=cfg["H"], W=cfg["W"], A=A, im_info=cfg["im_info"]) for s in seeds]
    return (np.concatenate([p[0] for p in parts]), np.concatenate([p[1] for p in parts]),
            np.concatenate([p[2] for p in parts]))


def gt_boxes(seed, G=64, im_h=600, im_w=1000, num_classes=21):
    """(G,5) f32 [x1,y1,x2,y2,cls]: w~U(20,0.5 im_w), h~U(20,0.5 im_h), uniform position."""
    rs = np.random.RandomState(seed + 100003)
    if G == 0:
        return np.zeros((0, 5), dtype=np.float32)
    w = rs.uniform(20, 0.5 * im_w, G)
    h = rs.uniform(20, 0.5 * im_h, G)
    x1 = rs.uniform(0, im_w - 1 - w)
    y1 = rs.uniform(0, im_h - 1 - h)
    cls = rs.randint(1, num_classes, G)
    return np.stack([x1, y1, x1 + w, y1 + h, cls], axis=1).astype(np.float32)


def jittered_rois(seed, gt, R=2000, im_h=600, im_w=1000, sigma=40.0):
    """(R,5) f32 [0, gt[j,:4] + N(0,sigma)] clipped to the image with x2>=x1+1, y2>=y1+1."""
    rs = np.random.RandomState(seed + 200003)
    j = rs.randint(0, gt.shape[0], R)
    b = gt[j, :4].astype(np.float64) + rs.normal(0, sigma, size=(R, 4))
    b[:, 0] = np.clip(b[:, 0], 0, im_w - 2)
    b[:, 1] = np.clip(b[:, 1], 0, im_h - 2)
    b[:, 2] = np.clip(np.maximum(b[:, 2], b[:, 0] + 1), 0, im_w - 1)
    b[:, 3] = np.clip(np.maximum(b[:, 3], b[:, 1] + 1), 0, im_h - 1)
    return np.hstack((np.zeros((R, 1)), b)).astype(np.float32)


def pad_gt(gt, rows=100):
    """gt padded with -1 rows as the loader does (rcnn/loader.py:319, tensor_vstack)."""
    out = -np.ones((max(rows, gt.shape[0]), 5), dtype=np.float32)
    out[:gt.shape[0]] = gt
    return out


def class_dets(seed, n=300, im_h=600, im_w=1000, dtype=np.float64, clusters=12):
    """(n,5) dets for per-class NMS tests: boxes clustered around a few centres so that
    many pairs overlap; distinct f32-representable scores."""
    rs = np.random.RandomState(seed + 300003)
    c = rs.randint(0, clusters, n)
    cx = rs.uniform(50, im_w - 50, clusters)[c] + rs.normal(0, 12, n)
    cy = rs.uniform(50, im_h - 50, clusters)[c] + rs.normal(0, 12, n)
    w = rs.uniform(30, 200, clusters)[c] * rs.uniform(0.8, 1.25, n)
    h = rs.uniform(30, 200, clusters)[c] * rs.uniform(0.8, 1.25, n)
    x1 = np.clip(cx - w / 2, 0, im_w - 1)
    y1 = np.clip(cy - h / 2, 0, im_h - 1)
    x2 = np.clip(cx + w / 2, 0, im_w - 1)
    y2 = np.clip(cy + h / 2, 0, im_h - 1)
    sc = ((rs.permutation(n) + 0.5) / n).astype(np.float32)
    return np.stack([x1, y1, x2, y2, sc], axis=1).astype(dtype)


def roidb_entry(seed, R=2000, num_classes=21):
    """A precomputed roidb record as rcnn/minibatch.py:151 sample_rois consumes."""
    rs = np.random.RandomState(seed + 400003)
    gt = gt_boxes(seed, G=8)
    rois = jittered_rois(seed, gt, R=R)[:, 1:5]
    ov = rs.uniform(0, 1, R).astype(np.float32)
    cls = rs.randint(1, num_classes, R).astype(np.int64)
    tg = np.zeros((R, 5), dtype=np.float32)
    tg[:, 0] = np.where(ov >= 0.5, cls, 0)
    tg[:, 1:] = rs.normal(0, 0.5, size=(R, 4)).astype(np.float32) * (tg[:, :1] > 0)
    return {"max_classes": cls, "max_overlaps": ov, "boxes": rois.astype(np.uint16), "bbox_targets": tg}


def gt_roidb_and_box_list(seed, n_images=4, R=300, C=21, max_gt=6, im_h=375, im_w=500):
    """What IMDB.create_roidb_from_box_list consumes (helper/dataset/imdb.py:32): a ground-truth roidb in the
    layout of helper/dataset/pascal_voc.py (uint16 boxes, int32 classes, csr one-hot overlaps) and one (R,4) uint16
    proposal array per image (jittered ground truth plus boxes far from everything).  Image 1 has no ground truth."""
    import scipy.sparse
    gt_roidb, box_list = [], []
    for i in range(n_images):
        rs = np.random.RandomState(seed * 1000 + i + 900001)
        G = 0 if i == 1 else int(rs.randint(1, max_gt + 1))
        gt = gt_boxes(seed * 1000 + i, G=G, im_h=im_h, im_w=im_w, num_classes=C)
        ov = np.zeros((G, C), dtype=np.float32)
        ov[np.arange(G), gt[:, 4].astype(np.int64)] = 1.0
        gt_roidb.append({"boxes": gt[:, :4].astype(np.uint16), "gt_classes": gt[:, 4].astype(np.int32),
                         "gt_overlaps": scipy.sparse.csr_matrix(ov), "flipped": False})
        if G:
            near = jittered_rois(seed * 1000 + i, gt, R=R - R // 4, im_h=im_h, im_w=im_w, sigma=15.0)[:, 1:5]
        else:
            near = np.zeros((0, 4), dtype=np.float32)
        n_far = R - near.shape[0]
        x1 = rs.uniform(0, im_w - 40, n_far)
        y1 = rs.uniform(0, im_h - 40, n_far)
        far = np.stack([x1, y1, x1 + rs.uniform(8, 39, n_far), y1 + rs.uniform(8, 39, n_far)], axis=1)
        box_list.append(np.vstack((near, far)).astype(np.uint16))
    return gt_roidb, box_list


def box_head_inputs(seed, R=300, C=21, G=6, im_h=600, im_w=1000, scale=1.6, peaked=0.35):
    """Inputs of the test-time box head (detector.py:60-74): ``rois`` (R,4) f32 jittered around G objects,
    ``deltas`` (R,4C) f32 ~ N(0,0.1)/N(0,0.2), ``scores`` (R,C) f32 softmax rows: a fraction ``peaked`` of the rois
    is confident about its object's class, the rest is mostly background.  Scores are made distinct per class."""
    rs = np.random.RandomState(seed + 700001)
    gt = gt_boxes(seed, G=G, im_h=im_h, im_w=im_w, num_classes=C)
    rois = jittered_rois(seed, gt, R=R, im_h=im_h, im_w=im_w, sigma=25.0)[:, 1:5].astype(np.float32)
    owner = rs.randint(0, G, R) if G else np.zeros(R, dtype=np.int64)
    deltas = np.empty((R, 4 * C), dtype=np.float32)
    deltas[:, 0::4] = rs.normal(0, 0.1, (R, C))
    deltas[:, 1::4] = rs.normal(0, 0.1, (R, C))
    deltas[:, 2::4] = rs.normal(0, 0.2, (R, C))
    deltas[:, 3::4] = rs.normal(0, 0.2, (R, C))
    logits = rs.normal(0, 1.0, (R, C))
    logits[:, 0] += 3.0
    hot = rs.uniform(size=R) < peaked
    if G:
        cls = gt[owner, 4].astype(np.int64)
        logits[hot, cls[hot]] += rs.uniform(3.0, 8.0, int(hot.sum()))
    e = np.exp(logits - logits.max(axis=1, keepdims=True))
    scores = (e / e.sum(axis=1, keepdims=True)).astype(np.float32)
    return rois, deltas, scores, (im_h, im_w), np.float32(scale)
