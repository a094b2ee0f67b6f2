"""CPU: host-side logic of the mirrors that needs neither the GPU nor the reference tree."""
import pickle

import numpy as np

from oracle import restate, synth


def test_greedy_cover_equals_reference_matching_loop():
    """helper/dataset/imdb.py:153-170 restated two ways (column-wise here, as in the reference in the oracle)."""
    from mxrcnn_b200.helper.dataset.imdb import AREA_RANGES, greedy_cover
    rs = np.random.RandomState(3)
    for n, k in [(50, 4), (7, 7), (12, 9), (200, 1), (5, 0)]:  # the reference's loop asserts n >= k too
        gt = synth.gt_boxes(n + k, G=k) if k else np.zeros((0, 5), np.float32)
        boxes = synth.jittered_rois(n, gt, R=n)[:, 1:5] if k else rs.uniform(0, 100, (n, 4)).astype(np.float32)
        m = restate.bbox_overlaps(boxes.astype(np.float64), gt[:, :4].astype(np.float64)) if k else np.zeros((n, 0))
        want = restate.recall_gt_overlaps(boxes, gt[:, :4])
        assert np.array_equal(greedy_cover(m.copy()), want)
    assert AREA_RANGES["all"] == (0, 1e10) and AREA_RANGES["medium"] == (32 ** 2, 96 ** 2)


def test_rpn_wire_format_round_trip(tmp_path):
    """rcnn/rpn/generate.py:60-76 layout: one pickled list of (post,4) float32 arrays in original image coordinates."""
    from mxrcnn_b200.rcnn.rpn import generate
    rois = np.concatenate([synth.jittered_rois(i, synth.gt_boxes(i, G=3), R=20) for i in range(3)]).astype(np.float32)
    rois[:, 0] = np.repeat(np.arange(3), 20)
    info = np.array([[600, 1000, 1.6], [600, 800, 1.25], [480, 640, 1.0]], dtype=np.float32)
    boxes = generate.batch_to_box_list(rois, info, 20)
    assert len(boxes) == 3 and all(b.shape == (20, 4) and b.dtype == np.float32 for b in boxes)
    assert np.array_equal(boxes[1], rois[20:40, 1:] / np.float32(1.25))
    path = generate.write_rpn_file(str(tmp_path), "voc_2007_trainval", boxes)
    assert path.endswith("rpn_data/voc_2007_trainval_rpn.pkl")
    with open(path, "rb") as f:
        assert f.read(2) == b"\x80\x02"  # protocol 2: what the reference's python 2 cPickle writes and reads
    back = generate.read_rpn_file(str(tmp_path), "voc_2007_trainval")
    assert len(back) == 3 and all(np.array_equal(a, b) for a, b in zip(back, boxes))
    with open(path, "rb") as f:
        again = pickle.load(f)  # pascal_voc.py:199-200
    assert np.array_equal(again[2], boxes[2])


def test_box_head_oracle_cap_and_threshold_semantics():
    """tester.py:51 compares float32 scores with a python float in float32; tester.py:60-65 keeps ties at the cap."""
    rois, deltas, scores, shape, scale = synth.box_head_inputs(5, R=60, C=4, G=3)
    scores = scores.copy()
    scores[0, 1] = np.float32(0.05)               # equal to the threshold in float32: not a detection
    scores[1, 1] = np.nextafter(np.float32(0.05), np.float32(1))
    dets, rows = restate.box_head(rois, deltas, scores, shape, scale, nms_thresh=1.0, max_per_image=0)  # no suppression
    assert 0 not in rows[1] and 1 in rows[1]
    dets, rows = restate.box_head(rois, deltas, scores, shape, scale, max_per_image=0)
    total = sum(len(d) for d in dets)
    capped, _ = restate.box_head(rois, deltas, scores, shape, scale, max_per_image=max(1, total // 2))
    kept = np.concatenate([d[:, 4] for d in capped[1:]])
    assert len(kept) >= total // 2 and kept.min() >= np.sort(np.concatenate([d[:, 4] for d in dets[1:]]))[-max(1, total // 2)]


def test_anchor_draws_consume_the_rng_like_the_reference_choice_calls():
    """minibatch._anchor_draws hands the device RANKS: npr.choice(inds, size=k, replace=False) of the reference
    (rcnn/minibatch.py:320-334) equals inds[npr.permutation(len(inds))[:k]] on the legacy global RandomState, draw for
    draw.  Checked here for the quotas of the default configuration, RNG state included."""
    from mxrcnn_b200.rcnn import minibatch
    from mxrcnn_b200.rcnn.config import config
    T = config.TRAIN
    num_fg = int(T.RPN_FG_FRACTION * T.RPN_BATCH_SIZE)
    for seed, (n_inside, n_fg, n_bg) in enumerate([(8151, 300, 7000), (8151, 20, 8000), (500, 129, 100), (40, 0, 40)]):
        fg_inds = np.arange(n_fg) * 3 + 1                 # any index arrays of those lengths
        bg_inds = np.arange(n_bg) * 2 + 5
        np.random.seed(seed)
        want_fg = np.random.choice(fg_inds, size=n_fg - num_fg, replace=False) if n_fg > num_fg else np.zeros((0,), int)
        kept_fg = n_fg - want_fg.size
        num_bg = T.RPN_BATCH_SIZE - kept_fg
        want_bg = np.random.choice(bg_inds, size=n_bg - num_bg, replace=False) if n_bg > num_bg else np.zeros((0,), int)
        state = np.random.get_state()[1].copy()
        np.random.seed(seed)
        fg_r, bg_r, pw, nw = minibatch._anchor_draws((n_inside, n_fg, n_bg), G=3)
        assert np.array_equal(np.random.get_state()[1], state)
        assert np.array_equal(fg_inds[fg_r], want_fg) and np.array_equal(bg_inds[bg_r], want_bg)
        n_examples = kept_fg + (n_bg - want_bg.size)
        assert pw == nw == 1.0 / n_examples
    import pytest
    with pytest.raises(ValueError):
        minibatch._anchor_draws((0, 0, 0), G=2)          # the reference dies in argmax of an empty array
