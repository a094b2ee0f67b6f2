"""GPU parity: the CUDA proposal operator (through the C ABI) against the fixtures produced by
the real reference and against oracle/restate.py on the same seeded inputs."""
import numpy as np
import pytest

from oracle import restate, synth
from tests.util import golden, golden_names, proposal_case

pytestmark = pytest.mark.gpu

RTOL = 1e-5  # decoded boxes: exp() is evaluated by CUDA's expf, numpy's differs in the last ulp
# 1e-5 relative is the north_star's bar.  A clipped coordinate is exactly 0 or exactly the limit on both sides; an unclipped
# one differs by |d exp| * width <= 2 ulp(f32) * width, i.e. <= 3e-4 px for the widest anchor: ATOL covers the coordinates
# that land within a few pixels of 0, where a relative bar on x1 would be a bar of 1e-9 px on the box width.
ATOL = 3e-4


def run(cls, box, info, cfg, train=False, **kw):
    from mxrcnn_b200.helper.processing.generate_anchor import generate_anchors
    from mxrcnn_b200.rcnn.rpn.proposal import proposal_forward
    anchors = generate_anchors(float(cfg["feat_stride"]), list(cfg["ratios"]), np.array(cfg["scales"], dtype=float))
    return proposal_forward(cls, box, info, anchors, cfg["feat_stride"], cfg["pre"], cfg["post"], cfg["thresh"],
                            cfg["min_size"], train_crop=train, want_indices=True, **kw)


def check_against(st, out, post):
    n = int(out["nkeep"][0])
    assert n == st["nkeep"]
    assert np.array_equal(out["keep"][0, :n], st["keep"][:n])
    assert np.array_equal(out["anchor_index"][0, :n], st["keep_anchor_index"][:n])
    assert np.array_equal(out["scores"][:n], st["scores"][:n])          # scores are gathered: bit-exact
    np.testing.assert_allclose(out["rois"][:n], st["rois"][:n], rtol=RTOL, atol=ATOL)
    assert np.all(out["rois"][:, 0] == 0)


@pytest.mark.parametrize("name", golden_names("proposal_tiny") + golden_names("proposal_voc_test")
                         + golden_names("proposal_voc_alt") + golden_names("proposal_wider")
                         + golden_names("proposal_voc_train"))
def test_proposal_vs_reference_fixture(name):
    g = golden(name)
    cls, box, info, cfg, train = proposal_case(g)
    out = run(cls, box, info, cfg, train, rng=np.random.RandomState(int(g["seed"])))
    post = cfg["post"]
    assert out["rois"].shape == (post, 5) and out["rois"].dtype == np.float32
    assert out["scores"].shape == (post, 1)
    n = int(g["nkeep"])
    assert int(out["nkeep"][0]) == n
    assert np.array_equal(out["keep"][0, :n], g["keep"][:n])
    assert np.array_equal(out["anchor_index"][0, :n], g["keep_anchor_index"][:n])
    # full outputs incl. the rows padded from the RNG (same draws: same seed, same call order)
    assert np.array_equal(out["scores"], g["scores"])
    np.testing.assert_allclose(out["rois"], g["rois"], rtol=RTOL, atol=ATOL)


@pytest.mark.parametrize("seed", [7, 8])
def test_proposal_padding_draws(seed):
    g = golden("proposal_pad_s%d" % seed)
    cfg = dict(synth.CONFIGS["tiny"], post=int(g["post"]), thresh=float(g["thresh"]))
    cls, box, info = synth.proposal_inputs(seed, H=cfg["H"], W=cfg["W"], A=9, im_info=cfg["im_info"])
    out = run(cls, box, info, cfg, rng=np.random.RandomState(seed))
    assert int(out["nkeep"][0]) == int(g["nkeep"]) < cfg["post"]
    assert np.array_equal(out["scores"], g["scores"])
    np.testing.assert_allclose(out["rois"], g["rois"], rtol=RTOL, atol=ATOL)


@pytest.mark.parametrize("cfgname,seed", [("tiny", 11), ("tiny", 12), ("voc_test", 13), ("voc_train", 14)])
def test_proposal_vs_oracle(cfgname, seed):
    cfg = synth.CONFIGS[cfgname]
    A = len(cfg["scales"]) * len(cfg["ratios"])
    cls, box, info = synth.proposal_inputs(seed, H=cfg["H"], W=cfg["W"], A=A, im_info=cfg["im_info"])
    st = restate.proposal_stages(cls, box, info, cfg["feat_stride"], cfg["scales"], cfg["ratios"], False, cfg["pre"],
                                 cfg["post"], cfg["thresh"], cfg["min_size"], rng=np.random.RandomState(seed))
    out = run(cls, box, info, cfg, rng=np.random.RandomState(seed))
    check_against(st, out, cfg["post"])


def test_proposal_score_ties_follow_documented_rule():
    # equal scores: descending score, then descending row index (argsort(kind='stable')[::-1]);
    # the reference's default-kind argsort leaves this order undefined
    cfg = synth.CONFIGS["tiny"]
    cls, box, info = synth.proposal_inputs(21, H=cfg["H"], W=cfg["W"], A=9, im_info=cfg["im_info"])
    q = np.round(cls[0, 9:] * 12) / 12.0          # 13 distinct values over 486 anchors
    cls[0, 9:] = q.astype(np.float32)
    st = restate.proposal_stages(cls, box, info, 16, cfg["scales"], cfg["ratios"], False, 150, cfg["post"], 0.7, 16)
    out = run(cls, box, info, dict(cfg, pre=150))
    check_against(st, out, cfg["post"])


def test_proposal_nan_raises_value_error():
    cfg = synth.CONFIGS["tiny"]
    cls, box, info = synth.proposal_inputs(0, H=cfg["H"], W=cfg["W"], A=9, im_info=cfg["im_info"])
    bad = cls.copy()
    bad[0, 12, 2, 3] = np.nan
    with pytest.raises(ValueError):
        run(bad, box, info, cfg)
    bad = box.copy()
    bad[0, 5, 1, 1] = np.nan
    with pytest.raises(ValueError):
        run(cls, bad, info, cfg)
    ok = cls.copy()
    ok[0, 0, 0, 0] = np.nan  # background half is never read (proposal.py:57)
    run(ok, box, info, cfg)


def test_proposal_operator_dropin_numpy():
    """The reference's own calling convention: Prop built from string kwargs, forward(in_data, out_data)."""
    from mxrcnn_b200 import operator as mxop
    from mxrcnn_b200.config import config
    import mxrcnn_b200.rcnn.rpn.proposal  # noqa: F401  (registers 'proposal')
    g = golden("proposal_voc_test_s0")
    cls, box, info, cfg, _ = proposal_case(g)
    config.TEST.RPN_PRE_NMS_TOP_N, config.TEST.RPN_POST_NMS_TOP_N = cfg["pre"], cfg["post"]
    prop, op = mxop.create("proposal", feat_stride=16, scales=(8, 16, 32), ratios=(0.5, 1, 2), output_score=True)
    assert prop.list_arguments() == ["cls_prob", "bbox_pred", "im_info"]
    assert prop.list_outputs() == ["output", "score"]
    in_shapes, out_shapes = prop.infer_shape([cls.shape, box.shape, info.shape])
    assert out_shapes == [(300, 5), (300, 1)]
    with pytest.raises(ValueError):
        prop.infer_shape([(2,) + cls.shape[1:], (2,) + box.shape[1:], (2, 3)])
    outs = [np.zeros(s, np.float32) for s in out_shapes]
    op.forward(False, ["write", "write"], [cls, box, info], outs, [])
    assert np.array_equal(outs[1], g["scores"])
    np.testing.assert_allclose(outs[0], g["rois"], rtol=RTOL, atol=ATOL)
    grads = [np.ones_like(cls), np.ones_like(box)]
    op.backward(["write", "write"], [], [cls, box, info], outs, grads, [])
    assert not grads[0].any() and not grads[1].any()


def test_proposal_device_arrays_and_batch():
    """Device-pointer entry point (torch tensors via data_ptr/DLPack), batch of 8 = C2; every image
    must equal its own single-image host-path result."""
    import torch
    cfg = synth.CONFIGS["voc_test"]
    seeds = list(range(30, 38))
    cls, box, info = synth.proposal_batch(seeds, cfg)
    dev = torch.device("cuda:0")
    out = run(torch.from_numpy(cls).to(dev), torch.from_numpy(box).to(dev), torch.from_numpy(info).to(dev), cfg)
    rois = out["rois"].cpu().numpy().reshape(8, cfg["post"], 5)
    scores = out["scores"].cpu().numpy().reshape(8, cfg["post"])
    keep = out["keep"].cpu().numpy()
    for i, s in enumerate(seeds):
        one = run(cls[i:i + 1], box[i:i + 1], info[i:i + 1], cfg)
        assert np.array_equal(one["keep"][0], keep[i])
        assert np.array_equal(one["scores"][:, 0], scores[i])
        assert np.array_equal(one["rois"][:, 1:], rois[i][:, 1:])
        assert np.all(rois[i][:, 0] == i)
        if i < 2:
            st = restate.proposal_stages(cls[i:i + 1], box[i:i + 1], info[i:i + 1], 16, cfg["scales"], cfg["ratios"],
                                         False, cfg["pre"], cfg["post"], cfg["thresh"], cfg["min_size"],
                                         nms_stop_after=cfg["post"])
            assert np.array_equal(st["keep"], keep[i])


def test_proposal_first_stripe_width_does_not_change_results():
    cfg = synth.CONFIGS["voc_test"]
    cls, box, info = synth.proposal_inputs(40, H=cfg["H"], W=cfg["W"], A=9, im_info=cfg["im_info"])
    base = run(cls, box, info, cfg)
    for fs in (64, 256, 4096):
        o = run(cls, box, info, cfg, first_stripe=fs)
        assert np.array_equal(o["keep"], base["keep"]) and np.array_equal(o["rois"], base["rois"])


@pytest.mark.parametrize("cfgname,seed", [("tiny", 51), ("voc_test", 52), ("voc_train", 53), ("voc_alt", 54), ("wider", 55)])
def test_fused_cluster_path_equals_multi_kernel_path(cfgname, seed, monkeypatch):
    """The op has two implementations: the fused cluster kernel (default) and the multi-kernel path
    (MXRCNN_B200_NO_CLUSTER=1; also the fallback for very large maps).  Same results, bit for bit."""
    from mxrcnn_b200 import _lib
    cfg = synth.CONFIGS[cfgname]
    A = len(cfg["scales"]) * len(cfg["ratios"])
    cls, box, info = synth.proposal_inputs(seed, H=cfg["H"], W=cfg["W"], A=A, im_info=cfg["im_info"])
    fused = run(cls, box, info, cfg, rng=np.random.RandomState(seed))
    monkeypatch.setenv("MXRCNN_B200_NO_CLUSTER", "1")
    h = _lib.Handle(0)
    monkeypatch.delenv("MXRCNN_B200_NO_CLUSTER")
    multi = run(cls, box, info, cfg, rng=np.random.RandomState(seed), handle=h)
    for k in ("rois", "scores", "keep", "anchor_index", "nkeep"):
        assert np.array_equal(fused[k], multi[k]), k
    h.close()


@pytest.mark.parametrize("cfgname,seed", [("tiny", 61), ("voc_test", 62), ("voc_train", 63), ("wider", 65)])
def test_cluster_sizes_agree(cfgname, seed, monkeypatch):
    """The fused kernel runs on clusters of 2 ... 16 CTAs (even sizes), chosen per call from the batch size
    (MXRCNN_B200_CLUSTER pins one): same kernel, same results bit for bit."""
    from mxrcnn_b200 import _lib
    cfg = synth.CONFIGS[cfgname]
    A = len(cfg["scales"]) * len(cfg["ratios"])
    cls, box, info = synth.proposal_inputs(seed, H=cfg["H"], W=cfg["W"], A=A, im_info=cfg["im_info"])
    auto = run(cls, box, info, cfg, rng=np.random.RandomState(seed))
    for size in ("2", "4", "6", "8", "12", "16"):
        monkeypatch.setenv("MXRCNN_B200_CLUSTER", size)
        h = _lib.Handle(0)
        monkeypatch.delenv("MXRCNN_B200_CLUSTER")
        pinned = run(cls, box, info, cfg, rng=np.random.RandomState(seed), handle=h)
        for k in ("rois", "scores", "keep", "anchor_index", "nkeep"):
            assert np.array_equal(auto[k], pinned[k]), (size, k)
        h.close()


@pytest.mark.parametrize("B", [6, 20, 45, 80])
def test_batch_size_picks_cluster_size_without_changing_results(B):
    """6 / 20 / 45 / 80 images per call land on 16- / 4- / 2-CTA clusters (several waves at 80); every image must
    equal its own single-image result."""
    cfg = synth.CONFIGS["voc_test"]
    seeds = [200 + (i % 5) for i in range(B)]
    cls, box, info = synth.proposal_batch(seeds, cfg)
    out = run(cls, box, info, cfg, rng=np.random.RandomState(0))
    ones = {}
    for i in range(5):
        ones[200 + i] = run(cls[i:i + 1], box[i:i + 1], info[i:i + 1], cfg, rng=np.random.RandomState(0))
    post = cfg["post"]
    for i, s in enumerate(seeds):
        assert np.array_equal(out["keep"][i], ones[s]["keep"][0]), i
        assert np.array_equal(out["rois"][i * post:(i + 1) * post, 1:], ones[s]["rois"][:, 1:]), i
        assert np.array_equal(out["scores"][i * post:(i + 1) * post], ones[s]["scores"]), i


def test_two_handles_share_kernel_attributes():
    """The dynamic shared-memory opt-in belongs to (kernel, device): a second handle that needs less must
    not lower it under the first one (regression: 'invalid argument' at launch)."""
    from mxrcnn_b200 import _lib
    big, small = synth.CONFIGS["voc_test"], synth.CONFIGS["tiny"]
    a, b = _lib.Handle(0), _lib.Handle(0)
    ins_big = synth.proposal_inputs(71, H=big["H"], W=big["W"], A=9, im_info=big["im_info"])
    ins_small = synth.proposal_inputs(72, H=small["H"], W=small["W"], A=9, im_info=small["im_info"])
    first = run(*ins_big, big, handle=a, rng=np.random.RandomState(1))
    run(*ins_small, small, handle=b, rng=np.random.RandomState(1))
    again = run(*ins_big, big, handle=a, rng=np.random.RandomState(1))
    assert np.array_equal(first["rois"], again["rois"]) and np.array_equal(first["keep"], again["keep"])
    a.close()
    b.close()


@pytest.mark.parametrize("cfgname,lo,width", [("tiny", 0.7, 5e-4), ("voc_test", 0.9990, 9e-4), ("voc_test", 0.25, 0.5)])
def test_head_selection_when_scores_crowd_few_bins(cfgname, lo, width):
    """The head is normally cut at a linear score-bin boundary found in one pass; when the scores crowd into one
    bin (first two cases: all of them inside 1/1024) the kernel falls back to the exact radix select.  Same results
    either way: equal to the oracle."""
    cfg = synth.CONFIGS[cfgname]
    A = len(cfg["scales"]) * len(cfg["ratios"])
    cls, box, info = synth.proposal_inputs(81, H=cfg["H"], W=cfg["W"], A=A, im_info=cfg["im_info"])
    fg = (lo + cls[0, A:].astype(np.float64) * width).astype(np.float32)  # same order, squeezed into [lo, lo + width)
    cls[0, A:] = fg
    cls[0, :A] = 1 - fg
    out = run(cls, box, info, cfg, rng=np.random.RandomState(0))
    st = restate.proposal_stages(cls, box, info, cfg["feat_stride"], cfg["scales"], cfg["ratios"], False, cfg["pre"],
                                 cfg["post"], cfg["thresh"], cfg["min_size"], nms_stop_after=cfg["post"])
    check_against(st, out, cfg["post"])


@pytest.mark.parametrize("cfgname,seed", [("tiny", 91), ("voc_test", 92), ("voc_train", 93), ("voc_alt", 94), ("wider", 95)])
def test_first_generation_fused_kernel_agrees(cfgname, seed, monkeypatch):
    """MXRCNN_B200_V1=1 selects the first-generation fused kernel (sorted head, float64 pair tests everywhere);
    the default kernel (bucketed list, float32 screen in front of the float64 test) must give the same bits."""
    from mxrcnn_b200 import _lib
    cfg = synth.CONFIGS[cfgname]
    A = len(cfg["scales"]) * len(cfg["ratios"])
    cls, box, info = synth.proposal_inputs(seed, H=cfg["H"], W=cfg["W"], A=A, im_info=cfg["im_info"])
    new = run(cls, box, info, cfg, rng=np.random.RandomState(seed))
    monkeypatch.setenv("MXRCNN_B200_V1", "1")
    h = _lib.Handle(0)
    monkeypatch.delenv("MXRCNN_B200_V1")
    old = run(cls, box, info, cfg, rng=np.random.RandomState(seed), handle=h)
    for k in ("rois", "scores", "keep", "anchor_index", "nkeep"):
        assert np.array_equal(new[k], old[k]), k
    h.close()


@pytest.mark.parametrize("name", golden_names("proposal_tiny") + golden_names("proposal_voc_test")
                         + golden_names("proposal_voc_train") + golden_names("proposal_wider"))
def test_stage_filter_count_matches_reference(name):
    """proposal.py:113-120: the number of boxes that pass _filter_boxes, counted on the device by the fused kernel,
    equals the fixture's n_filter (real reference); the pre-NMS list length follows proposal.py:123-127."""
    from mxrcnn_b200 import _lib
    g = golden(name)
    cls, box, info, cfg, train = proposal_case(g)
    h = _lib.Handle(0)
    run(cls, box, info, cfg, train, rng=np.random.RandomState(int(g["seed"])), handle=h)
    c = h.proposal_counts(1)[0]
    h.close()
    assert int(c[0]) == int(g["n_filter"])
    assert int(c[1]) == (min(cfg["pre"], int(g["n_filter"])) if cfg["pre"] > 0 else int(g["n_filter"]))
    assert int(c[2]) >= int(g["nkeep"]) or int(c[2]) == cfg["post"]


@pytest.mark.parametrize("cfgname,seed,train,pad", [("tiny", 101, False, None), ("tiny", 102, True, (8, 11)),
                                                    ("voc_test", 103, False, None), ("wider", 104, False, None)])
def test_stage_decode_clip_and_filter_all_anchors(cfgname, seed, train, pad):
    """mxr_decode_clip: EVERY anchor's decoded + clipped box (proposal.py:110-113) within 1e-5 relative of the oracle's
    float64 boxes, and the min-size decision (proposal.py:117, 156-161) of EVERY anchor equal to the oracle's
    filter_keep -- the stage the fused kernel otherwise shows only through the rows that survive NMS."""
    import ctypes
    import torch
    from mxrcnn_b200 import _lib
    from mxrcnn_b200.helper.processing.generate_anchor import generate_anchors
    cfg = synth.CONFIGS[cfgname]
    A = len(cfg["scales"]) * len(cfg["ratios"])
    cls, box, info = synth.proposal_inputs(seed, H=cfg["H"], W=cfg["W"], A=A, im_info=cfg["im_info"], pad_hw=pad)
    st = restate.proposal_stages(cls, box, info, cfg["feat_stride"], cfg["scales"], cfg["ratios"], train, cfg["pre"],
                                 cfg["post"], cfg["thresh"], cfg["min_size"], nms_stop_after=cfg["post"])
    Hm, Wm = box.shape[2], box.shape[3]
    crop_h = int(info[0, 0] / np.float32(cfg["feat_stride"])) if train else 0
    crop_w = int(info[0, 1] / np.float32(cfg["feat_stride"])) if train else 0
    n = A * (crop_h or Hm) * (crop_w or Wm)
    assert st["decoded"].shape == (n, 4)
    anchors = generate_anchors(float(cfg["feat_stride"]), list(cfg["ratios"]), np.array(cfg["scales"], dtype=float))
    p = _lib.ProposalParams(1, A, Hm, Wm, crop_h, crop_w, float(cfg["feat_stride"]), cfg["pre"], cfg["post"],
                            cfg["thresh"], float(cfg["min_size"]), 0)
    dev = torch.device("cuda:0")
    box_d, info_d = torch.from_numpy(box).to(dev), torch.from_numpy(info).to(dev)
    out_b = torch.zeros((n, 4), dtype=torch.float64, device=dev)
    out_v = torch.zeros((n,), dtype=torch.int32, device=dev)
    h = _lib.Handle(0)
    _lib.check(_lib.load().mxr_decode_clip(h.ptr, int(torch.cuda.current_stream().cuda_stream), ctypes.byref(p),
                                           box_d.data_ptr(), info_d.data_ptr(), anchors.ctypes.data_as(_lib.c_double_p),
                                           out_b.data_ptr(), out_v.data_ptr()))
    torch.cuda.synchronize()
    h.close()
    np.testing.assert_allclose(out_b.cpu().numpy(), st["decoded"], rtol=RTOL, atol=ATOL)
    valid = np.zeros(n, dtype=bool)
    valid[st["filter_keep"]] = True
    assert np.array_equal(out_v.cpu().numpy().astype(bool), valid)


@pytest.mark.parametrize("cfgname,seed", [("tiny", 111), ("voc_test", 112), ("voc_alt", 113)])
def test_stage_full_sorted_order(cfgname, seed):
    """proposal.py:123-127: the COMPLETE descending order of the pre-NMS list, not only the few hundred rows a 0.7 NMS
    keeps.  With an NMS threshold just below 1 only exact duplicates (boxes clipped onto the same rectangle) are
    suppressed and post = pre lets everything else through, so the op's anchor-index output is the sorted candidate
    list minus those duplicates: it must equal the oracle's, row for row, and cover (nearly) the whole list."""
    cfg = dict(synth.CONFIGS[cfgname])
    A = len(cfg["scales"]) * len(cfg["ratios"])
    cls, box, info = synth.proposal_inputs(seed, H=cfg["H"], W=cfg["W"], A=A, im_info=cfg["im_info"])
    n_valid = A * cfg["H"] * cfg["W"]
    k = min(cfg["pre"], n_valid) if cfg["pre"] > 0 else n_valid
    cfg.update(thresh=0.9999999, post=k)
    st = restate.proposal_stages(cls, box, info, cfg["feat_stride"], cfg["scales"], cfg["ratios"], False, cfg["pre"],
                                 cfg["post"], cfg["thresh"], cfg["min_size"], rng=np.random.RandomState(seed))
    out = run(cls, box, info, cfg, rng=np.random.RandomState(seed))
    n = st["nkeep"]
    assert int(out["nkeep"][0]) == n
    if cfgname != "tiny":
        assert n >= 0.98 * len(st["order"])  # the order of (nearly) all candidates is what is being compared
    assert np.array_equal(out["anchor_index"][0, :n], st["keep_anchor_index"][:n])
    assert np.array_equal(out["keep"][0, :n], st["keep"][:n])
    assert np.array_equal(np.asarray(st["order"])[st["keep"][:n]], st["keep_anchor_index"][:n])


@pytest.mark.parametrize("cfgname,seed,ties", [("tiny", 121, False), ("voc_test", 122, False), ("voc_test", 123, True),
                                               ("voc_train", 124, False), ("voc_alt", 125, False)])
def test_stage_topk_sorted_is_the_oracles_order(cfgname, seed, ties):
    """mxr_topk_sorted: the whole pre-NMS list -- every candidate, duplicates included -- in processing order
    equals the oracle's ``order`` (proposal.py:113-127 with the documented tie rule), row for row."""
    import ctypes
    import torch
    from mxrcnn_b200 import _lib
    from mxrcnn_b200.helper.processing.generate_anchor import generate_anchors
    cfg = synth.CONFIGS[cfgname]
    A = len(cfg["scales"]) * len(cfg["ratios"])
    cls, box, info = synth.proposal_inputs(seed, H=cfg["H"], W=cfg["W"], A=A, im_info=cfg["im_info"])
    if ties:  # few distinct scores: the order inside a run of equal scores is the rule's, not the sort's
        q = (np.round(cls[0, A:].astype(np.float64) * 500) / 500).astype(np.float32)
        cls[0, A:] = q
        cls[0, :A] = 1 - q
    st = restate.proposal_stages(cls, box, info, cfg["feat_stride"], cfg["scales"], cfg["ratios"], False, cfg["pre"], 1,
                                 cfg["thresh"], cfg["min_size"], nms_stop_after=1)
    want = np.asarray(st["order"])
    n_all = A * cfg["H"] * cfg["W"]
    cap = min(cfg["pre"], n_all) if cfg["pre"] > 0 else n_all
    anchors = generate_anchors(float(cfg["feat_stride"]), list(cfg["ratios"]), np.array(cfg["scales"], dtype=float))
    p = _lib.ProposalParams(1, A, cfg["H"], cfg["W"], 0, 0, float(cfg["feat_stride"]), cfg["pre"], cfg["post"],
                            cfg["thresh"], float(cfg["min_size"]), 0)
    dev = torch.device("cuda:0")
    cls_d, box_d, info_d = (torch.from_numpy(x).to(dev) for x in (cls, box, info))
    order = torch.full((cap,), -7, dtype=torch.int32, device=dev)
    cnt = torch.zeros((1, 2), dtype=torch.int32, device=dev)
    h = _lib.Handle(0)
    _lib.check(_lib.load().mxr_topk_sorted(h.ptr, int(torch.cuda.current_stream().cuda_stream), ctypes.byref(p),
                                           cls_d.data_ptr(), box_d.data_ptr(), info_d.data_ptr(),
                                           anchors.ctypes.data_as(_lib.c_double_p), cap, order.data_ptr(), cnt.data_ptr()))
    torch.cuda.synchronize()
    h.close()
    assert int(cnt[0, 0]) == want.size and int(cnt[0, 1]) == 0
    assert np.array_equal(order.cpu().numpy()[:want.size], want)


def _adversarial(kind, seed, cfgname):
    cfg = dict(synth.CONFIGS[cfgname])
    A = len(cfg["scales"]) * len(cfg["ratios"])
    cls, box, info = synth.proposal_inputs(seed, H=cfg["H"], W=cfg["W"], A=A, im_info=cfg["im_info"])
    rs = np.random.RandomState(seed + 7)
    if kind == "wild_deltas":      # boxes far outside the image, clipped onto its border: exact duplicates, huge exp()
        box = (box * 10).astype(np.float32)
    elif kind == "overflow":       # exp() overflows to inf for some anchors, deltas of 1e30 for others
        flat = box.reshape(-1)
        flat[rs.choice(flat.size, 200, replace=False)] = np.float32(95.0)
        flat[rs.choice(flat.size, 50, replace=False)] = np.float32(1e30)
        flat[rs.choice(flat.size, 50, replace=False)] = np.float32(-1e30)
    elif kind == "zero_deltas":    # boxes are the anchors themselves: integer coordinates, sides EQUAL to the threshold
        box[:] = 0
        cfg["min_size"] = 184      # scale 1.0 below: anchors of side 184 sit exactly on `ws >= min_size` (proposal.py:117)
        info = np.array([[cfg["im_info"][0], cfg["im_info"][1], 1.0]], dtype=np.float32)
    elif kind == "tiny_image":     # everything is clipped into a 20x30 window: degenerate and identical boxes
        info = np.array([[20, 30, 1.0]], dtype=np.float32)
        cfg["min_size"] = 0
    elif kind == "thr_one":
        cfg["thresh"] = 1.0        # outside (0,1): the float32 screen is off, every pair takes the float64 test
    elif kind == "thr_zero":
        cfg["thresh"] = 0.0
    return cls, box, info, cfg


@pytest.mark.parametrize("kind", ["wild_deltas", "overflow", "zero_deltas", "tiny_image", "thr_one", "thr_zero"])
def test_screens_agree_with_float64_on_adversarial_inputs(kind, monkeypatch):
    """The fused kernel decides the min-size filter and the IoU test on float32 images of the boxes inside rigorous
    error bounds and hands everything else to the reference's float64 arithmetic.  Inputs built to sit ON the
    thresholds, to overflow, or to collapse onto identical boxes must give exactly what the float64-only kernel
    (MXRCNN_B200_V1) and the oracle give."""
    from mxrcnn_b200 import _lib
    cfgname = "tiny" if kind in ("thr_one", "thr_zero") else "voc_test"
    cls, box, info, cfg = _adversarial(kind, 300, cfgname)
    new = run(cls, box, info, cfg, rng=np.random.RandomState(1))
    monkeypatch.setenv("MXRCNN_B200_V1", "1")
    h = _lib.Handle(0)
    monkeypatch.delenv("MXRCNN_B200_V1")
    old = run(cls, box, info, cfg, rng=np.random.RandomState(1), handle=h)
    h.close()
    for k in ("rois", "scores", "keep", "anchor_index", "nkeep"):
        assert np.array_equal(new[k], old[k]), (kind, k)
    if cfgname == "tiny" or kind in ("zero_deltas", "tiny_image"):
        with np.errstate(all="ignore"):
            st = restate.proposal_stages(cls, box, info, cfg["feat_stride"], cfg["scales"], cfg["ratios"], False, cfg["pre"],
                                         cfg["post"], cfg["thresh"], cfg["min_size"], rng=np.random.RandomState(1))
        n = st["nkeep"]
        assert int(new["nkeep"][0]) == n
        assert np.array_equal(new["anchor_index"][0, :n], st["keep_anchor_index"][:n])
