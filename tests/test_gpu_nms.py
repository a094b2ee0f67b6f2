"""GPU parity: nms(dets, thresh) through the C ABI -- keep lists must be bit-exact."""
import numpy as np
import pytest

from oracle import restate, synth
from tests.util import golden

pytestmark = pytest.mark.gpu


def gpu_nms(dets, thresh, **kw):
    from mxrcnn_b200.helper.processing.nms import nms
    return nms(dets, thresh, **kw)


def test_nms_vs_reference_fixture():
    h = golden("helpers")
    for i in range(6):
        seed, n, size, th = h["nms%d_spec" % i]
        dets = synth.class_dets(int(seed), n=int(n), dtype=np.float64 if size == 8 else np.float32)
        keep = gpu_nms(dets, float(th))
        assert isinstance(keep, list) and all(isinstance(k, int) for k in keep)
        assert np.array_equal(np.asarray(keep, dtype=np.int64), h["nms%d_keep" % i]), i
    assert gpu_nms(np.zeros((0, 5)), 0.3) == []


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("n,thresh", [(1, 0.5), (63, 0.3), (64, 0.3), (65, 0.7), (1000, 0.5), (5000, 0.3), (9000, 0.7)])
def test_nms_vs_oracle(dtype, n, thresh):
    dets = synth.class_dets(n * 7 + 1, n=n, dtype=dtype, clusters=max(2, n // 40))
    assert gpu_nms(dets, thresh) == [int(i) for i in restate.nms(dets, thresh)]


def test_nms_stage_level_on_oracle_dets():
    """Stage-level parity (unconditional): the oracle's own f64 (n,5) dets of the proposal op
    (rcnn/rpn/proposal.py:131) fed to the CUDA NMS give the oracle's keep list, all 6000 boxes."""
    cfg = synth.CONFIGS["voc_test"]
    cls, box, info = synth.proposal_inputs(3, H=cfg["H"], W=cfg["W"], A=9, im_info=cfg["im_info"])
    st = restate.proposal_stages(cls, box, info, 16, cfg["scales"], cfg["ratios"], False, cfg["pre"], 0, cfg["thresh"],
                                 cfg["min_size"])
    assert st["dets"].dtype == np.float64 and st["dets"].shape == (6000, 5)
    keep = gpu_nms(st["dets"], cfg["thresh"])
    assert keep == [int(i) for i in st["keep"]]
    assert gpu_nms(st["dets"], cfg["thresh"], max_keep=300) == keep[:300]


def test_nms_ties_and_degenerate_boxes():
    rs = np.random.RandomState(5)
    dets = synth.class_dets(9, n=500)
    dets[:, 4] = np.round(dets[:, 4] * 10) / 10          # many equal scores
    assert gpu_nms(dets, 0.4) == [int(i) for i in restate.nms(dets, 0.4)]
    d32 = dets.astype(np.float32)
    assert gpu_nms(d32, 0.4) == [int(i) for i in restate.nms(d32, 0.4)]
    same = np.tile(np.array([[10., 10., 50., 50., 0.5]]), (130, 1))   # identical boxes: only one survives
    assert gpu_nms(same, 0.5) == [int(i) for i in restate.nms(same, 0.5)] == [129]
    # zero-area / inverted boxes follow the arithmetic, whatever it gives
    weird = np.array([[5, 5, 5, 5, .9], [5, 5, 4, 4, .8], [0, 0, 100, 100, .7], [5, 5, 5, 5, .6], [50, 50, 40, 60, .5]],
                     dtype=np.float64)
    for th in (0.0, 0.3, 1.0):
        assert gpu_nms(weird, th) == [int(i) for i in restate.nms(weird, th)]
    # negative scores and -0.0/+0.0 compare like numpy
    dets = synth.class_dets(12, n=200)
    dets[:, 4] = rs.permutation(200) - 100.0
    dets[7, 4], dets[9, 4] = 0.0, -0.0
    assert gpu_nms(dets, 0.5) == [int(i) for i in restate.nms(dets, 0.5)]


def test_nms_device_array_and_nan():
    import torch
    dets = synth.class_dets(77, n=700)
    t = torch.from_numpy(dets).cuda()
    assert gpu_nms(t, 0.3) == [int(i) for i in restate.nms(dets, 0.3)]
    bad = dets.copy()
    bad[3, 1] = np.nan
    with pytest.raises(ValueError):
        gpu_nms(bad, 0.3)


def test_nms_batched_per_class_loop():
    """rcnn/tester.py:50-56: one nms per class; the batched entry point must equal the loop."""
    from mxrcnn_b200.helper.processing.nms import nms_batched
    for dt in (np.float64, np.float32):
        dets = [synth.class_dets(200 + c, n=n, dtype=dt) for c, n in enumerate([300, 0, 17, 64, 1, 257, 300, 129])]
        got = nms_batched(dets, 0.3)
        want = [[int(i) for i in restate.nms(d, 0.3)] if d.shape[0] else [] for d in dets]
        assert got == want
    assert nms_batched([], 0.3) == []


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_nms_cluster_and_striped_paths_agree(dtype, monkeypatch):
    from mxrcnn_b200 import _lib
    monkeypatch.setenv("MXRCNN_B200_NO_CLUSTER", "1")
    striped = _lib.Handle(0)
    monkeypatch.delenv("MXRCNN_B200_NO_CLUSTER")
    for n, th in [(300, 0.3), (1025, 0.5), (4000, 0.7), (8192, 0.4)]:
        dets = synth.class_dets(n + 3, n=n, dtype=dtype, clusters=max(2, n // 50))
        a = gpu_nms(dets, th)
        b = gpu_nms(dets, th, handle=striped)
        assert a == b == [int(i) for i in restate.nms(dets, th)]
        assert gpu_nms(dets, th, max_keep=50) == a[:50]
    striped.close()


def test_nest_vs_reference_fixture_and_oracle():
    from mxrcnn_b200.helper.processing.nms import nest
    h = golden("helpers")
    dets = synth.class_dets(40, n=60, dtype=np.float32)
    assert nest(dets, 0.9) == [int(i) for i in h["nest_keep"]]
    for n, dt, th in [(1, np.float64, 0.9), (130, np.float64, 0.5), (700, np.float32, 0.9), (300, np.float64, 0.99)]:
        d = synth.class_dets(90 + n, n=n, dtype=dt, clusters=max(2, n // 30))
        assert nest(d, th) == restate.nest(d, th)
    assert nest(np.zeros((0, 5)), 0.9) == []
    # inputs for which "no intersection" does not mean "keep": zero / negative areas, thresholds at and below zero
    with np.errstate(divide="ignore", invalid="ignore"):
        for dt in (np.float32, np.float64):
            d = synth.class_dets(17, n=260, dtype=dt, clusters=9)
            d[5, 2], d[9, 3] = d[5, 0] - 1, d[9, 1] - 1          # area 0
            d[20, 2] = d[20, 0] - 7                              # negative area
            d[40:50] = d[30:40]                                  # duplicates: contained in each other
            for th in (0.9, 0.0, -0.5, 1.0):
                assert nest(d, th) == restate.nest(d, th), (dt, th)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_nms_fused_and_first_generation_cluster_kernels_agree(dtype, monkeypatch):
    """nms_fused.cuh (float32 screen + exact test in the dtype of dets) against nms_cluster.cuh (every pair exact)."""
    from mxrcnn_b200 import _lib
    monkeypatch.setenv("MXRCNN_B200_V1", "1")
    first = _lib.Handle(0)
    monkeypatch.delenv("MXRCNN_B200_V1")
    sized = {}
    for c in (2, 4, 8, 16):
        monkeypatch.setenv("MXRCNN_B200_CLUSTER", str(c))
        sized[c] = _lib.Handle(0)
    monkeypatch.delenv("MXRCNN_B200_CLUSTER")
    for n, th in [(2, 0.5), (33, 0.3), (511, 0.5), (512, 0.5), (513, 0.5), (1024, 0.7), (1500, 0.3), (6000, 0.7)]:
        dets = synth.class_dets(n + 11, n=n, dtype=dtype, clusters=max(2, n // 50))
        want = [int(i) for i in restate.nms(dets, th)]
        assert gpu_nms(dets, th) == want
        assert gpu_nms(dets, th, handle=first) == want
        for c, h in sized.items():
            assert gpu_nms(dets, th, handle=h) == want, (n, c)
            assert gpu_nms(dets, th, max_keep=37, handle=h) == want[:37], (n, c)
    first.close()
    for h in sized.values():
        h.close()


def _grid_dets(seed, n, dtype, shift=0.0, step=1.0):
    """Boxes on an integer grid with few distinct sizes: many pairs whose IoU equals a simple fraction exactly."""
    rs = np.random.RandomState(seed)
    x1 = rs.randint(0, 40, n) * step + shift
    y1 = rs.randint(0, 40, n) * step + shift
    w = rs.choice([10, 20, 30, 40], n) * step
    h = rs.choice([10, 20, 30, 40], n) * step
    d = np.stack([x1, y1, x1 + w - 1, y1 + h - 1, rs.rand(n)], axis=1)
    return d.astype(dtype)


@pytest.mark.parametrize("dtype", [np.float32, np.float64])
def test_nms_screen_agrees_with_exact_arithmetic_on_adversarial_dets(dtype):
    """Inputs chosen against the float32 screen of nms_fused.cuh: IoU exactly on the threshold, coordinates so large
    that float32 cannot hold them (every pair goes to the exact test), tiny fractional sizes, thresholds 0 and 1."""
    for th in (0.5, 1.0 / 3.0, 0.25, 0.2):
        for seed in range(3):
            dets = _grid_dets(seed, 700, dtype)
            assert gpu_nms(dets, th) == [int(i) for i in restate.nms(dets, th)], (th, seed)
    far = _grid_dets(5, 600, dtype, shift=3.0e6 if dtype == np.float64 else 6.0e5)
    assert gpu_nms(far, 0.5) == [int(i) for i in restate.nms(far, 0.5)]
    huge = _grid_dets(6, 300, dtype, shift=1.0e12, step=1.0e9)
    assert gpu_nms(huge, 0.5) == [int(i) for i in restate.nms(huge, 0.5)]
    small = _grid_dets(7, 600, dtype, shift=0.125, step=1.0 / 64.0)
    assert gpu_nms(small, 0.5) == [int(i) for i in restate.nms(small, 0.5)]
    frac = synth.class_dets(31, n=900, dtype=dtype, clusters=12)
    frac[:, :4] = frac[:, :4] * dtype(1.0 / 3.0) + dtype(1000.1)
    for th in (0.0, 0.3, 0.999, 1.0, 1.5, -0.5):
        assert gpu_nms(frac, th) == [int(i) for i in restate.nms(frac, th)], th
