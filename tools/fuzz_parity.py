"""Randomised parity sweep (not part of the test suite: minutes of oracle time): the proposal op, nms / nms_batched, assign_anchor batches
and recall matching on random sizes, dtypes, thresholds and degenerate mixtures, each compared with oracle/restate.py.
Usage: python tools/fuzz_parity.py [seconds per family, default 60] [proposal,nms,assign_anchor,recall,nest]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import restate, synth  # noqa: E402


def fuzz_proposal(budget):
    from mxrcnn_b200.helper.processing.generate_anchor import generate_anchors
    from mxrcnn_b200.rcnn.rpn.proposal import proposal_forward
    rs = np.random.RandomState(99)
    t0, n_cases = time.time(), 0
    while time.time() - t0 < budget:
        H, W = int(rs.randint(3, 50)), int(rs.randint(3, 70))
        scales = [(8, 16, 32), (4, 8, 16, 32), (16,)][rs.randint(3)]
        ratios = [(0.5, 1, 2), (1,)][rs.randint(2)]
        A = len(scales) * len(ratios)
        N = A * H * W
        pre = int(rs.choice([-1, 40, 300, 1000, 6000, 12000]))
        post = int(rs.choice([1, 30, 300, 2000]))
        thresh = float(rs.choice([0.3, 0.5, 0.7, 0.9]))
        min_size = int(rs.choice([0, 16, 48]))
        info_t = (H * 16 - int(rs.randint(0, 16)), W * 16 - int(rs.randint(0, 16)), float(rs.choice([1.0, 1.6, 0.6])))
        seed = int(rs.randint(1 << 30))
        cls, box, info = synth.proposal_inputs(seed, H=H, W=W, A=A, im_info=info_t)
        kind = rs.randint(4)
        fg = cls[0, A:].astype(np.float64)
        if kind == 1:    # few distinct values: long tie runs, crowded buckets
            fg = np.round(fg * rs.choice([6, 40, 300])) / rs.choice([6, 40, 300])
        elif kind == 2:  # all scores inside a sliver of the range
            fg = rs.choice([0.2, 0.7, 0.999]) + fg * rs.choice([1e-3, 3e-4, 0.05])
        elif kind == 3:  # wild deltas
            box = (box * np.float32(rs.choice([3.0, 8.0]))).astype(np.float32)
        fg = np.clip(fg, 0, 1).astype(np.float32)
        cls[0, A:] = fg
        cls[0, :A] = 1 - fg
        anchors = generate_anchors(16.0, list(ratios), np.array(scales, dtype=float))
        try:
            st = restate.proposal_stages(cls, box, info, 16, scales, ratios, False, pre, post, thresh, min_size,
                                         rng=np.random.RandomState(seed), nms_stop_after=post)
        except ValueError:  # nothing passes the min-size filter: the reference dies in npr.choice on an empty list
            try:
                proposal_forward(cls, box, info, anchors, 16, pre, post, thresh, min_size, rng=np.random.RandomState(seed))
            except ValueError:
                n_cases += 1
                continue
            raise AssertionError(("proposal: the oracle raised, the op did not", H, W, A, min_size, seed))
        out = proposal_forward(cls, box, info, anchors, 16, pre, post, thresh, min_size, want_indices=True,
                               rng=np.random.RandomState(seed))
        n = int(out["nkeep"][0])
        tag = ("proposal", H, W, A, pre, post, thresh, min_size, kind, seed)
        assert n == st["nkeep"], tag
        assert np.array_equal(out["keep"][0, :n], st["keep"][:n]), tag
        assert np.array_equal(out["anchor_index"][0, :n], st["keep_anchor_index"][:n]), tag
        assert np.array_equal(out["scores"][:n], st["scores"][:n]), tag
        np.testing.assert_allclose(out["rois"][:n], st["rois"][:n], rtol=1e-5, atol=3e-4 * (8.0 if kind == 3 else 1.0))
        n_cases += 1
    return n_cases


def fuzz_nms(budget):
    from mxrcnn_b200.helper.processing.nms import nms, nms_batched
    rs = np.random.RandomState(1234)
    t0, n_cases = time.time(), 0
    while time.time() - t0 < budget:
        dt = np.float32 if rs.rand() < 0.5 else np.float64
        n = int(rs.choice([1, 2, 31, 32, 33, 100, 511, 512, 513, 700, 1024, 1500, 2500]))
        thr = float(rs.choice([0.3, 0.5, 0.7, 1.0 / 3.0, 0.25, 0.9, 0.05]))
        kind = rs.randint(4)
        if kind == 0:
            d = synth.class_dets(int(rs.randint(1 << 30)), n=n, dtype=dt, clusters=max(2, n // int(rs.choice([10, 40, 100]))))
        elif kind == 1:  # integer grid: exact ties with the threshold
            x1 = rs.randint(0, 30, n) * 4.0
            y1 = rs.randint(0, 30, n) * 4.0
            w = rs.choice([8, 12, 16, 24, 32], n)
            h = rs.choice([8, 12, 16, 24, 32], n)
            d = np.stack([x1, y1, x1 + w - 1, y1 + h - 1, rs.rand(n)], axis=1).astype(dt)
        elif kind == 2:  # tied scores, duplicates, inverted and sub-pixel boxes mixed in
            d = synth.class_dets(int(rs.randint(1 << 30)), n=n, dtype=dt, clusters=max(2, n // 20))
            d[:, 4] = np.round(d[:, 4] * 8) / 8
            m = rs.rand(n) < 0.1
            d[m, 2] = d[m, 0] - rs.rand(int(m.sum())).astype(dt) * 3
            m = rs.rand(n) < 0.1
            d[m, 2:4] = d[m, 0:2] + rs.rand(int(m.sum()), 2).astype(dt) * 0.5
            if n > 4:
                d[n // 2:n // 2 + n // 4] = d[:n // 4]
        else:  # large offsets / fractional coordinates
            d = synth.class_dets(int(rs.randint(1 << 30)), n=n, dtype=dt, clusters=max(2, n // 30))
            d[:, :4] = d[:, :4] * dt(rs.choice([1.0 / 3.0, 7.0, 1e-2])) + dt(rs.choice([0.1, 1e4, -5e3, 3e5]))
        want = [int(i) for i in restate.nms(d, thr)]
        got = nms(d, thr)
        assert got == want, ("nms", dt, n, thr, kind)
        if n_cases % 5 == 0:
            parts = [d[: n // 3], d[n // 3: n // 3], d[n // 3:]]
            gb = nms_batched(parts, thr)
            assert gb == [[int(i) for i in restate.nms(p, thr)] if p.shape[0] else [] for p in parts], ("batched", dt, n, thr)
        n_cases += 1
    return n_cases


def fuzz_nest(budget):
    from mxrcnn_b200.helper.processing.nms import nest
    rs = np.random.RandomState(11)
    t0, n_cases = time.time(), 0
    with np.errstate(divide="ignore", invalid="ignore"):
        while time.time() - t0 < budget:
            dt = np.float32 if rs.rand() < 0.5 else np.float64
            n = int(rs.choice([1, 2, 33, 255, 256, 257, 600]))
            d = synth.class_dets(int(rs.randint(1 << 30)), n=n, dtype=dt, clusters=max(2, n // 15))
            if rs.rand() < 0.4 and n > 8:
                d[n // 2:n // 2 + n // 8] = d[:n // 8]                      # duplicates
                m = rs.rand(n) < 0.05
                d[m, 2] = d[m, 0] - dt(rs.choice([1.0, 5.0]))               # zero / negative areas
            th = float(rs.choice([0.9, 0.5, 0.99, 0.0, -0.25, 1.0]))
            assert nest(d, th) == restate.nest(d, th), ("nest", dt, n, th)
            n_cases += 1
    return n_cases


def fuzz_anchor(budget):
    from mxrcnn_b200.rcnn.loader import anchor_labels
    rs = np.random.RandomState(77)
    cfg = restate.default_config()
    t0, n_cases = time.time(), 0
    while time.time() - t0 < budget:
        fh, fw = int(rs.randint(6, 40)), int(rs.randint(6, 64))
        feat = (1, 18, fh, fw)
        B = int(rs.randint(1, 5))
        im_h, im_w = fh * 16 - int(rs.randint(0, 16)), fw * 16 - int(rs.randint(0, 16))
        gts, infos = [], []
        for _ in range(B):
            G = int(rs.choice([0, 1, 2, 7, 20, 64, 65, 130]))
            gts.append(synth.gt_boxes(int(rs.randint(1 << 30)), G=G, im_h=im_h, im_w=im_w) if G else np.zeros((0, 5), np.float32))
            infos.append(np.array([[im_h, im_w, 1.0]], np.float32))
        seed = int(rs.randint(1 << 30))
        try:
            np.random.seed(seed)
            want = [restate.assign_anchor(feat, gts[i], infos[i], rng=np.random, cfg=cfg) for i in range(B)]
        except ValueError:
            continue  # no anchor inside the image: the reference raises
        np.random.seed(seed)
        dev = anchor_labels(feat, gts, infos, device="cuda:0")
        for i in range(B):
            assert np.array_equal(dev[0][i].cpu().numpy(), want[i]["label"][0]), ("label", fh, fw, i)
            assert np.array_equal(dev[2][i].cpu().numpy(), want[i]["bbox_inside_weight"][0])
            assert np.array_equal(dev[3][i].cpu().numpy(), want[i]["bbox_outside_weight"][0])
            np.testing.assert_allclose(dev[1][i].cpu().numpy(), want[i]["bbox_target"][0], rtol=2e-6, atol=1e-7)
        n_cases += 1
    return n_cases


def fuzz_recall(budget):
    from mxrcnn_b200.helper.dataset.imdb import recall_cover
    rs = np.random.RandomState(5)
    t0, n_cases = time.time(), 0
    while time.time() - t0 < budget:
        k = int(rs.choice([1, 2, 5, 33, 64, 200]))
        n = k + int(rs.choice([0, 1, 50, 900]))
        gt = synth.gt_boxes(int(rs.randint(1 << 30)), G=k)[:, :4].astype(np.float64)
        boxes = synth.jittered_rois(int(rs.randint(1 << 30)), np.hstack([gt, np.ones((k, 1))]).astype(np.float32), R=n)[:, 1:5]
        boxes = boxes.astype(np.float64)
        if rs.rand() < 0.5 and n > 3:
            boxes[n // 2:] = boxes[: n - n // 2]
        assert np.array_equal(recall_cover(boxes, gt), restate.recall_gt_overlaps(boxes, gt)), ("recall", n, k)
        n_cases += 1
    return n_cases


if __name__ == "__main__":
    budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
    only = sys.argv[2].split(",") if len(sys.argv) > 2 else None
    for name, fn in (("proposal", fuzz_proposal), ("nms", fuzz_nms), ("assign_anchor batches", fuzz_anchor),
                     ("recall matching", fuzz_recall), ("nest", fuzz_nest)):
        if only is None or name.split()[0] in only:
            print("%s: %d random cases equal to the oracle" % (name, fn(budget)), flush=True)
