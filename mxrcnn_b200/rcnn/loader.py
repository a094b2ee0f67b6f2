"""Loader hand-off: the label half of ``AnchorLoader.get_batch`` (reference:
rcnn/loader.py:293-329).

The reference runs ``assign_anchor`` per image on the host, stacks the numpy results with ``tensor_vstack`` and
uploads them with ``mx.nd.array``.  ``anchor_labels`` writes every image's labels / targets / weights straight into
its row of four device-resident batch tensors; they reach the trainer zero-copy through DLPack
(``mx.nd.from_dlpack(torch.utils.dlpack.to_dlpack(t))`` or ``__dlpack__``).  Image decoding, ``get_minibatch`` and the
iterator plumbing stay the reference's.
"""
import numpy as np

from . import minibatch
from .config import config

LABEL_KEYS = ('label', 'bbox_target', 'bbox_inside_weight', 'bbox_outside_weight')


def anchor_labels(feat_shape, gt_boxes_list, im_info_list, feat_stride=16, anchor_scales=(8, 16, 32),
                  anchor_ratios=(0.5, 1, 2), allowed_border=0, device=None):
    """Labels of one batch: image i owns ``gt_boxes_list[i]`` (G_i,5) and ``im_info_list[i]`` (1,3); all images
    share ``feat_shape`` because the loader pads the data first (loader.py:283-286).

    Returns ``[label (B, A*H*W), bbox_target, bbox_inside_weight, bbox_outside_weight (B, 4A, H, W)]`` -- the order of
    ``self.label`` at loader.py:321-324 -- as float32 torch tensors on ``device`` (e.g. ``"cuda:0"``), or as numpy
    arrays when ``device`` is None.  With ``config.END2END == 1`` a fifth entry holds the gt boxes flattened per image
    and padded with -1 (loader.py:316-318,328), on the host: it is an input of the graph's proposal_target op.
    The RNG draws of assign_anchor happen image by image in list order, exactly as in the reference loop.
    """
    B = len(gt_boxes_list)
    assert B == len(im_info_list) and B > 0
    fh, fw = int(feat_shape[-2]), int(feat_shape[-1])
    A = len(anchor_scales) * len(anchor_ratios)
    N = A * fh * fw
    if device is None:
        per_image = [minibatch.assign_anchor(feat_shape, gt, info, feat_stride, anchor_scales, anchor_ratios, allowed_border)
                     for gt, info in zip(gt_boxes_list, im_info_list)]
        out = [np.concatenate([p[k] for p in per_image], axis=0) for k in LABEL_KEYS]
    else:
        import torch  # device tensors and streams come from the harness / trainer side
        out = [torch.empty((B, N), dtype=torch.float32, device=device)] + \
              [torch.empty((B, 4 * A, fh, fw), dtype=torch.float32, device=device) for _ in range(3)]
        # the whole batch in two launches (one cluster per image) and two synchronisations
        minibatch.assign_anchor_batch(feat_shape, gt_boxes_list, im_info_list, dict(zip(LABEL_KEYS, out)), feat_stride,
                                      anchor_scales, anchor_ratios, allowed_border)
    if config.END2END == 1:
        flat = [np.asarray(gt, dtype=np.float32).reshape(1, -1) for gt in gt_boxes_list]
        width = max(f.shape[1] for f in flat)
        gts = -np.ones((B, width), dtype=np.float32)
        for i, f in enumerate(flat):
            gts[i, :f.shape[1]] = f
        out.append(gts)
    return out
