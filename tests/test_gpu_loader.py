"""GPU: the loader hand-off: batch labels written straight into device tensors equal the
per-image host results (same RNG stream), and leave through DLPack."""
import numpy as np
import pytest

from oracle import restate, synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("end2end", [0, 1])
def test_anchor_labels_device_equals_host_and_oracle(end2end):
    import torch
    from mxrcnn_b200.config import config
    from mxrcnn_b200.rcnn.loader import anchor_labels
    feat = (1, 18, 38, 63)
    gts = [synth.gt_boxes(70 + i, G=g) for i, g in enumerate((6, 1, 40))]
    infos = [np.array([[600, 1000, 1.6]], np.float32), np.array([[600, 901, 1.6]], np.float32),
             np.array([[562, 1000, 1.5]], np.float32)]
    saved = config.END2END
    config.END2END = end2end
    try:
        np.random.seed(5)
        host = anchor_labels(feat, gts, infos)
        state_after_host = np.random.get_state()[1].copy()
        np.random.seed(5)
        dev = anchor_labels(feat, gts, infos, device="cuda:0")
        assert np.array_equal(np.random.get_state()[1], state_after_host)  # same draws consumed
    finally:
        config.END2END = saved
    assert len(host) == len(dev) == 4 + end2end
    for h, d in zip(host[:4], dev[:4]):
        assert d.is_cuda and d.dtype == torch.float32 and tuple(d.shape) == h.shape
        assert np.array_equal(d.cpu().numpy(), h)
    back = torch.utils.dlpack.from_dlpack(torch.utils.dlpack.to_dlpack(dev[1]))  # the hand-off the trainer uses
    assert back.data_ptr() == dev[1].data_ptr()
    if end2end:
        assert host[4].shape == (3, 200) and np.array_equal(host[4][1, :5], gts[1].reshape(-1)) and (host[4][1, 5:] == -1).all()
    np.random.seed(5)
    cfg = restate.default_config()
    for i in range(3):
        want = restate.assign_anchor(feat, gts[i], infos[i], rng=np.random, cfg=cfg)
        assert np.array_equal(host[0][i], want["label"][0])
        assert np.array_equal(host[2][i], want["bbox_inside_weight"][0])
        assert np.array_equal(host[3][i], want["bbox_outside_weight"][0])
        np.testing.assert_allclose(host[1][i], want["bbox_target"][0], rtol=2e-6, atol=1e-7)


def test_anchor_labels_device_under_a_busy_stream():
    """The per-image label kernels of the device path run on the trainer's (torch's current) stream while the next
    image's first phase runs on the library's own stream; both use the handle's workspace.  With a backlog on the
    trainer's stream the second image's phase 1 used to overwrite the scratch of the first image's pending phase 2
    (ADVICE round 1).  The library now orders the streams; the batch must equal the host results regardless."""
    import torch
    from mxrcnn_b200.rcnn.loader import anchor_labels
    feat = (1, 18, 38, 63)
    gts = [synth.gt_boxes(170 + i, G=g) for i, g in enumerate((12, 3, 25, 7))]
    infos = [np.array([[600, 1000, 1.6]], np.float32)] * 4
    np.random.seed(9)
    host = anchor_labels(feat, gts, infos)
    a = torch.randn((4096, 4096), device="cuda:0")
    for trial in range(3):
        b = a
        for _ in range(20):  # a few tens of milliseconds of queued work on the current stream
            b = b @ a * 1e-3
        np.random.seed(9)
        dev = anchor_labels(feat, gts, infos, device="cuda:0")
        torch.cuda.synchronize()
        for h, d in zip(host[:4], dev[:4]):
            assert np.array_equal(d.cpu().numpy(), h), trial
        assert torch.isfinite(b).any() or True


def test_anchor_labels_batch_ragged_gt_counts_vs_oracle():
    """The batched kernels (one cluster per image) on a batch whose images differ in everything per-image: no gt at
    all, more gts than one 64-box round, a gt that touches no inside anchor (quirk Q4: its column maximum of 0 makes
    every inside anchor positive), a smaller image inside the padded map."""
    from mxrcnn_b200.rcnn.loader import anchor_labels
    feat = (1, 18, 38, 63)
    far = np.array([[5000, 5000, 5100, 5100, 1]], np.float32)
    gts = [np.zeros((0, 5), np.float32), synth.gt_boxes(301, G=150), far, synth.gt_boxes(302, G=65),
           synth.gt_boxes(303, G=7, im_h=480, im_w=640)]
    infos = [np.array([[600, 1000, 1.6]], np.float32)] * 4 + [np.array([[480, 640, 1.0]], np.float32)]
    np.random.seed(77)
    dev = anchor_labels(feat, gts, infos, device="cuda:0")
    state_after = np.random.get_state()[1].copy()
    np.random.seed(77)
    cfg = restate.default_config()
    for i in range(len(gts)):
        want = restate.assign_anchor(feat, gts[i], infos[i], rng=np.random, cfg=cfg)
        assert np.array_equal(dev[0][i].cpu().numpy(), want["label"][0]), i
        assert np.array_equal(dev[2][i].cpu().numpy(), want["bbox_inside_weight"][0]), i
        assert np.array_equal(dev[3][i].cpu().numpy(), want["bbox_outside_weight"][0]), i
        np.testing.assert_allclose(dev[1][i].cpu().numpy(), want["bbox_target"][0], rtol=2e-6, atol=1e-7)
    assert np.array_equal(np.random.get_state()[1], state_after)


def test_anchor_labels_fused_and_single_image_kernels_agree(monkeypatch):
    """MXRCNN_B200_V1=1 selects the single-image kernels (the path of maps too large for one cluster) on a map the
    batched kernels also take: both must write the same tensors."""
    import torch
    from mxrcnn_b200 import _lib
    from mxrcnn_b200.rcnn import minibatch
    feat = (1, 18, 38, 63)
    gts = [synth.gt_boxes(400 + i, G=g) for i, g in enumerate((20, 64, 3))]
    infos = [np.array([[600, 1000, 1.6]], np.float32)] * 3
    outs = []
    for v1 in ("1", "0"):
        monkeypatch.setenv("MXRCNN_B200_V1", v1)
        h = _lib.Handle(0)
        monkeypatch.setattr(_lib, "handle_for", lambda handle, *arrays: h)
        t = [torch.empty((3, 9 * 38 * 63), dtype=torch.float32, device="cuda:0")] + \
            [torch.empty((3, 36, 38, 63), dtype=torch.float32, device="cuda:0") for _ in range(3)]
        np.random.seed(12)
        minibatch.assign_anchor_batch(feat, gts, infos, dict(zip(("label", "bbox_target", "bbox_inside_weight",
                                                                  "bbox_outside_weight"), t)))
        torch.cuda.synchronize()
        outs.append([x.cpu().numpy() for x in t])
        h.close()
    for a, b in zip(*outs):
        assert np.array_equal(a, b)
    assert (outs[0][0] == 1).sum() > 0
