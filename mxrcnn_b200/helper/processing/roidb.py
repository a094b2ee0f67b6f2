"""``add_bbox_regression_targets`` of the reference's helper/processing/roidb.py: the
roidb-wide user of ``bbox_overlaps``.  ``prepare_roidb`` is dataset IO (image paths, cv2.imread) and stays the
reference's."""
import numpy as np

from ...config import config
from .bbox_regression import compute_bbox_regression_targets


def _class_rows(targets, num_classes):
    """rows of each foreground class inside one image's (k,5) target array"""
    return [None] + [np.flatnonzero(targets[:, 0] == c) for c in range(1, num_classes)]


def add_bbox_regression_targets(roidb):
    """Adds ``roidb[i]['bbox_targets']`` ((k,5) float32, GPU: IoU + argmax + transform per image) and normalises
    them per class in place; returns ``(means.ravel(), stds.ravel())`` (reference: roidb.py:42-90).

    The moments are accumulated on the host in float64, image by image and class by class with the same numpy
    reductions as the reference, so given identical targets the means / stds are bit-identical.
    """
    print('add bounding box regression targets')
    assert len(roidb) > 0
    assert 'max_classes' in roidb[0]
    num_classes = roidb[0]['gt_overlaps'].shape[1]
    rows = []
    for entry in roidb:
        entry['bbox_targets'] = compute_bbox_regression_targets(entry['boxes'], entry['max_overlaps'], entry['max_classes'])
        rows.append(_class_rows(entry['bbox_targets'], num_classes))
    if config.TRAIN.BBOX_NORMALIZATION_PRECOMPUTED:  # fixed values instead of empirical ones (:58-61)
        means = np.tile(np.array(config.TRAIN.BBOX_MEANS), (num_classes, 1))
        stds = np.tile(np.array(config.TRAIN.BBOX_STDS), (num_classes, 1))
    else:
        count = np.zeros((num_classes, 1)) + config.EPS
        s1 = np.zeros((num_classes, 4))
        s2 = np.zeros((num_classes, 4))
        for entry, per_class in zip(roidb, rows):
            t = entry['bbox_targets']
            for c in range(1, num_classes):
                r = per_class[c]
                if r.size > 0:
                    count[c] += r.size
                    s1[c, :] += t[r, 1:].sum(axis=0)
                    s2[c, :] += (t[r, 1:] ** 2).sum(axis=0)
        means = s1 / count
        stds = np.sqrt(s2 / count - means ** 2)  # var(x) = E(x^2) - E(x)^2
    for entry, per_class in zip(roidb, rows):
        for c in range(1, num_classes):
            r = per_class[c]
            entry['bbox_targets'][r, 1:] -= means[c, :]
            entry['bbox_targets'][r, 1:] /= stds[c, :]
    return means.ravel(), stds.ravel()
