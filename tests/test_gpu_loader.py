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
