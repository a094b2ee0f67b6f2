"""CPU, build container only: oracle/restate.py against the real reference imported live.

Skipped wherever /root/reference is not mounted (e.g. the GPU box); the committed
tests/golden fixtures cover the same comparison there.
"""
import numpy as np
import pytest

from oracle import ref_shim, restate, synth

pytestmark = pytest.mark.skipif(not ref_shim.available(), reason="reference tree not mounted")


@pytest.fixture(scope="module")
def R():
    return ref_shim.ref()


@pytest.mark.parametrize("seed", [10, 11])
def test_nms_live(R, seed):
    for dt, th in [(np.float64, 0.3), (np.float32, 0.7)]:
        dets = synth.class_dets(seed, n=400, dtype=dt)
        assert restate.nms(dets, th) == [int(i) for i in R.nms(dets, th)]


def test_overlaps_live(R):
    gt = synth.gt_boxes(3, G=5).astype(np.float64)
    rois = synth.jittered_rois(3, gt.astype(np.float32), R=120)[:, 1:5].astype(np.float64)
    want = R.bbox_overlaps(rois, gt)
    assert np.array_equal(restate.bbox_overlaps(rois, gt), want)
    assert np.array_equal(restate.bbox_overlaps_loop(rois, gt), want)


@pytest.mark.parametrize("seed", [20, 21])
def test_assign_anchor_live(R, seed):
    gt = synth.gt_boxes(seed, G=5, im_h=300, im_w=500)
    info = np.array([[300, 500, 0.8]], np.float32)
    np.random.seed(seed)
    want = R.assign_anchor((1, 18, 19, 31), gt, info)
    got = restate.assign_anchor((1, 18, 19, 31), gt, info, rng=np.random.RandomState(seed))
    for k in want:
        assert np.array_equal(want[k], got[k]), k


def test_sample_rois_live(R):
    gt = synth.gt_boxes(5, G=4)
    rois = np.vstack((synth.jittered_rois(5, gt, R=500), np.hstack((np.zeros((4, 1), np.float32), gt[:, :4]))))
    R.config.TRAIN.BG_THRESH_LO = 0.0
    np.random.seed(5)
    want = R._sample_rois(rois, gt, 32, 128, 21, "TRAIN")
    R.config.TRAIN.BG_THRESH_LO = 0.1
    cfg = restate.default_config()
    cfg.TRAIN.BG_THRESH_LO = 0.0
    got = restate.sample_rois_e2e(rois, gt, 32, 128, 21, "TRAIN", cfg=cfg, rng=np.random.RandomState(5))
    for a, b in zip(want, got):
        assert np.array_equal(a, b)


@pytest.mark.parametrize("seed,kw", [(30, {}), (31, dict(R=120, G=3, peaked=0.8)), (32, dict(C=2, R=500, G=20))])
def test_box_head_live(R, seed, kw):
    """The reference's own Detector.im_detect + pred_eval (driven as in oracle/make_golden.py) on two images."""
    from oracle.make_golden import run_reference_box_head
    cases = [synth.box_head_inputs(seed, **kw), synth.box_head_inputs(seed + 100, **kw)]
    all_boxes = run_reference_box_head(R, cases, nms_thresh=0.3)
    for i, case in enumerate(cases):
        mine, _ = restate.box_head(*case, thresh=0.05, nms_thresh=0.3, max_per_image=100)
        for j in range(1, case[2].shape[1]):
            assert np.array_equal(np.asarray(all_boxes[j][i]).reshape(-1, 5), mine[j]), (i, j)


def test_rpn_wire_format_live(R, tmp_path):
    """rcnn/rpn/generate.py:43-77 run for real (fake detector / iterator / imdb) next to the mirror's writer: the two
    files hold the same list, and each side reads the other's file."""
    import pickle
    from mxrcnn_b200.rcnn.rpn import generate as mine
    rs = np.random.RandomState(3)
    rois, infos = [], []
    for i in range(3):
        b = synth.jittered_rois(i, synth.gt_boxes(i, G=4), R=50)
        rois.append(b * np.float32(1.6))
        infos.append(np.array([[600, 1000, 1.6 if i != 1 else 1.25]], dtype=np.float32))

    class Det(object):
        calls = 0

        def im_detect(self, im, im_info):
            self.calls += 1
            return rois[self.calls - 1], rs.uniform(size=(50, 1)).astype(np.float32)

    class Batch(object):
        def __init__(self, info):
            self.data = {"data": np.zeros((1, 3, 4, 4), np.float32), "im_info": info}

    class Data(list):
        shuffle = False

    class Imdb(object):
        num_images, name = 3, "synthetic"
        root_path = str(tmp_path / "ref")
    (tmp_path / "ref").mkdir()
    (tmp_path / "mine").mkdir()
    want = R.generate_detections(Det(), Data(Batch(i) for i in infos), Imdb(), vis=False)
    got = mine.batch_to_box_list(np.concatenate(rois), np.concatenate(infos), 50)
    assert len(got) == len(want) == 3
    for a, b in zip(got, want):
        assert a.dtype == b.dtype == np.float32 and np.array_equal(a, b)
    mine.write_rpn_file(str(tmp_path / "mine"), "synthetic", got)
    theirs = mine.read_rpn_file(str(tmp_path / "ref"), "synthetic")           # mirror reads the reference's file
    with open(mine.rpn_file_path(str(tmp_path / "mine"), "synthetic"), "rb") as f:  # what pascal_voc.py:199-200 does
        ours = pickle.load(f)
    for a, b, c in zip(theirs, ours, want):
        assert np.array_equal(a, c) and np.array_equal(b, c)
