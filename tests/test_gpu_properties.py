"""GPU: size-independent properties at BASELINE.json's full sizes (no oracle loop in the way):
greedy-NMS invariants, idempotence, permutation invariance, sampling quotas."""
import numpy as np
import pytest

from oracle import restate, synth

pytestmark = pytest.mark.gpu


def _pair_ovr(a, b):
    """ovr of helper/processing/nms.py:25-33 for every (a_i, b_j), float64, same operation order."""
    xx1 = np.maximum(a[:, None, 0], b[None, :, 0])
    yy1 = np.maximum(a[:, None, 1], b[None, :, 1])
    xx2 = np.minimum(a[:, None, 2], b[None, :, 2])
    yy2 = np.minimum(a[:, None, 3], b[None, :, 3])
    w = np.maximum(0.0, xx2 - xx1 + 1)
    h = np.maximum(0.0, yy2 - yy1 + 1)
    inter = w * h
    aa = (a[:, 2] - a[:, 0] + 1) * (a[:, 3] - a[:, 1] + 1)
    ab = (b[:, 2] - b[:, 0] + 1) * (b[:, 3] - b[:, 1] + 1)
    return inter / (aa[:, None] + ab[None, :] - inter)


@pytest.mark.parametrize("cfgname,seed", [("voc_test", 61), ("voc_train", 62), ("wider", 63)])
def test_greedy_nms_invariants_full_size(cfgname, seed):
    """On the op's own f64 dets (6000 / 12000 boxes): kept boxes never suppress each other, and every
    dropped box is suppressed by a kept box that precedes it -- together these characterise greedy NMS."""
    from mxrcnn_b200.helper.processing.nms import nms
    cfg = synth.CONFIGS[cfgname]
    A = len(cfg["scales"]) * len(cfg["ratios"])
    cls, box, info = synth.proposal_inputs(seed, H=cfg["H"], W=cfg["W"], A=A, im_info=cfg["im_info"])
    st = restate.proposal_stages(cls, box, info, cfg["feat_stride"], cfg["scales"], cfg["ratios"], False, cfg["pre"], 1,
                                 cfg["thresh"], cfg["min_size"], nms_stop_after=1)
    dets = st["dets"]                      # descending scores, distinct
    n = dets.shape[0]
    assert n == cfg["pre"]
    keep = np.asarray(nms(dets, cfg["thresh"]), dtype=np.int64)
    assert np.all(np.diff(keep) > 0)       # keep order = score order = row order here
    kept = dets[keep]
    ov = _pair_ovr(kept[:, :4], dets[:, :4])              # (K, n)
    earlier = keep[:, None] < np.arange(n)[None, :]       # kept row precedes the column
    sup = ((ov > cfg["thresh"]) & earlier).any(axis=0)
    is_kept = np.zeros(n, bool)
    is_kept[keep] = True
    assert not sup[is_kept].any(), "a kept box is suppressed by an earlier kept box"
    assert sup[~is_kept].all(), "a dropped box has no earlier kept suppressor"
    # idempotence: NMS of the survivors keeps all of them
    assert nms(kept, cfg["thresh"]) == list(range(len(keep)))
    # the op's output is the prefix of this list
    from mxrcnn_b200.helper.processing.generate_anchor import generate_anchors
    from mxrcnn_b200.rcnn.rpn.proposal import proposal_forward
    anchors = generate_anchors(float(cfg["feat_stride"]), list(cfg["ratios"]), np.array(cfg["scales"], dtype=float))
    out = proposal_forward(cls, box, info, anchors, cfg["feat_stride"], cfg["pre"], cfg["post"], cfg["thresh"],
                           cfg["min_size"], want_indices=True, rng=np.random.RandomState(seed))
    m = min(cfg["post"], len(keep))
    assert int(out["nkeep"][0]) == m and np.array_equal(out["keep"][0, :m], keep[:m])
    assert np.all(np.diff(out["scores"][:m, 0]) < 0)      # sorted by descending score


def test_nms_row_permutation_invariance():
    from mxrcnn_b200.helper.processing.nms import nms
    dets = synth.class_dets(71, n=5000, clusters=100)
    perm = np.random.RandomState(71).permutation(5000)
    a = nms(dets, 0.5)
    b = nms(dets[perm], 0.5)
    assert [int(perm[i]) for i in b] == a


def test_anchor_target_quotas_full_size():
    """C3: 21 546 anchors x 64 gt: at most 128 positives, at most 256 sampled, weights sum to 4 per map."""
    from mxrcnn_b200.rcnn.minibatch import assign_anchor
    info = np.array([[600, 1000, 1.6]], np.float32)
    for seed in (81, 82, 83):
        gt = synth.gt_boxes(seed, G=64)
        np.random.seed(seed)
        out = assign_anchor((1, 18, 38, 63), gt, info)
        lab = out["label"]
        assert (lab == 1).sum() <= 128 and (lab >= 0).sum() <= 256
        assert set(np.unique(lab)) <= {-1.0, 0.0, 1.0}
        n_ex = (lab >= 0).sum()
        np.testing.assert_allclose(out["bbox_outside_weight"].sum(), 4.0 * n_ex * np.float32(1.0 / n_ex), rtol=1e-6)
        inw = out["bbox_inside_weight"].reshape(9, 4, 38, 63)
        assert np.array_equal(inw[:, 0] == 1, lab.reshape(9, 38, 63) == 1)


@pytest.mark.parametrize("dtype,n", [(np.float32, 16384), (np.float64, 8192)])
def test_nms_invariants_at_the_largest_single_launch_size(dtype, n):
    """The longest list the fused NMS kernel takes (sort in the leader's shared memory), arithmetic in the dtype of
    dets like numpy: kept boxes do not suppress each other, every dropped box has an earlier kept suppressor."""
    from mxrcnn_b200.helper.processing.nms import nms
    dets = synth.class_dets(91, n=n, dtype=dtype, clusters=n // 40)
    dets[:, 4] = np.random.RandomState(91).permutation(n).astype(dtype) / dtype(n)   # distinct scores
    thr = 0.5
    keep = np.asarray(nms(dets, thr), dtype=np.int64)
    order = np.argsort(-dets[:, 4], kind="stable")
    pos = np.empty(n, np.int64)
    pos[order] = np.arange(n)
    assert np.all(np.diff(pos[keep]) > 0)                 # keep order = descending score
    kept = dets[keep]
    is_kept = np.zeros(n, bool)
    is_kept[keep] = True
    sup = np.zeros(n, bool)
    t = dtype(thr)
    for c0 in range(0, n, 2048):                          # (K, 2048) blocks of the IoU matrix
        ov = _pair_ovr(kept[:, :4], dets[c0:c0 + 2048, :4])
        assert ov.dtype == dtype
        earlier = pos[keep][:, None] < pos[c0:c0 + 2048][None, :]
        sup[c0:c0 + 2048] = ((ov > t) & earlier).any(axis=0)
    assert not sup[is_kept].any(), "a kept box is suppressed by an earlier kept box"
    assert sup[~is_kept].all(), "a dropped box has no earlier kept suppressor"
    assert nms(kept, thr) == list(range(len(keep)))
