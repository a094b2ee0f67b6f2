"""GPU-backed versions of the two ``IMDB`` methods that consume ``bbox_overlaps`` (reference:
helper/dataset/imdb.py:32-64 and :109-188).

Dataset readers (pascal_voc.py ...) subclass the reference's IMDB.  To move their IoU work to the GPU either put
``RoidbOverlapMixin`` in front of it (``class PascalVOC(RoidbOverlapMixin, IMDB)``) or assign the two functions onto
the class.  Paths, flipping, merging and evaluation IO are out of scope and stay the reference's.
"""
import numpy as np
import scipy.sparse

from ..processing.bbox_regression import bbox_overlaps, bbox_overlaps_reduced

# name -> [min area, max area] of a ground-truth box, as tabulated in imdb.py:122-125
AREA_RANGES = {'all': (0, 1e10), 'small': (0, 32 ** 2), 'medium': (32 ** 2, 96 ** 2), 'large': (96 ** 2, 1e10),
               '96-128': (96 ** 2, 128 ** 2), '128-256': (128 ** 2, 256 ** 2), '256-512': (256 ** 2, 512 ** 2),
               '512-inf': (512 ** 2, 1e10)}


def class_overlaps(boxes, gt_boxes, gt_classes, num_classes):
    """Dense (n, C) float32: each proposal's best IoU, stored in the column of the class of the gt box it overlaps
    most (imdb.py:45-55).  Only the row argmax / row max leave the GPU, not the (n,k) matrix."""
    out = np.zeros((boxes.shape[0], num_classes), dtype=np.float32)
    if boxes.shape[0] and gt_boxes is not None and gt_boxes.size > 0:
        best, best_iou, _ = bbox_overlaps_reduced(boxes.astype(np.float64), gt_boxes.astype(np.float64))
        hit = np.flatnonzero(best_iou > 0)
        out[hit, np.asarray(gt_classes)[best[hit]]] = best_iou[hit]
    return out


def greedy_cover(overlaps):
    """One-to-one matching of imdb.py:153-170 on an (n,k) IoU matrix (modified in place): k rounds, each takes the gt
    box with the best remaining cover, records that IoU and retires the proposal and the gt box."""
    k = overlaps.shape[1]
    covered = np.zeros(k)
    for j in range(k):
        g = int(overlaps.max(axis=0).argmax())
        b = int(overlaps[:, g].argmax())
        covered[j] = overlaps[b, g]
        assert covered[j] >= 0
        overlaps[b, :] = -1
        overlaps[:, g] = -1
    return covered


def recall_cover(boxes, gt_boxes):
    """``greedy_cover(bbox_overlaps(boxes, gt_boxes))`` in one launch, without the (n,k) matrix ever leaving the GPU
    (or existing: every gt column keeps its best live proposal; imdb.py:149-170)."""
    from ... import _lib
    b = np.ascontiguousarray(np.asarray(boxes, dtype=np.float64)[:, :4])
    q = np.ascontiguousarray(np.asarray(gt_boxes, dtype=np.float64)[:, :4])
    covered = np.zeros((q.shape[0],), dtype=np.float64)
    if q.shape[0]:
        _lib.check(_lib.load().mxr_recall_cover_host(_lib.handle_for(None).ptr, b.ctypes.data_as(_lib.c_double_p), b.shape[0],
                                                     q.ctypes.data_as(_lib.c_double_p), q.shape[0],
                                                     covered.ctypes.data_as(_lib.c_double_p)))
        assert (covered >= 0).all()   # imdb.py:165
    return covered


class RoidbOverlapMixin(object):
    """Needs ``self.num_images`` and ``self.num_classes`` like the reference's IMDB."""

    def create_roidb_from_box_list(self, box_list, gt_roidb):
        """Same contract as imdb.py:32-64: one roidb entry per image with csr ``gt_overlaps``."""
        assert len(box_list) == self.num_images, 'number of boxes matrix must match number of images'
        roidb = []
        for i, boxes in enumerate(box_list):
            gt = gt_roidb[i] if gt_roidb is not None else None
            dense = class_overlaps(boxes, gt['boxes'] if gt else None, gt['gt_classes'] if gt else None, self.num_classes)
            roidb.append({'boxes': boxes, 'gt_classes': np.zeros((boxes.shape[0],), dtype=np.int32),
                          'gt_overlaps': scipy.sparse.csr_matrix(dense), 'flipped': False})
        return roidb

    def evaluate_recall(self, roidb, candidate_boxes=None, thresholds=None, area='all', limit=None):
        """Same report as imdb.py:109-188 (printed line for line).  Extension: the numbers are also returned as
        ``{'ar', 'recalls', 'thresholds', 'gt_overlaps'}`` -- the reference returns None."""
        assert area in AREA_RANGES, 'unknown area range: {}'.format(area)
        lo, hi = AREA_RANGES[area]
        found, num_pos = [], 0
        for i in range(self.num_images):
            entry = roidb[i]
            full = entry['gt_overlaps'].toarray().max(axis=1) == 1  # excludes crowd annotations
            gt = entry['boxes'][np.flatnonzero((entry['gt_classes'] > 0) & full), :]
            gt_area = (gt[:, 2] - gt[:, 0] + 1) * (gt[:, 3] - gt[:, 1] + 1)
            gt = gt[np.flatnonzero((gt_area >= lo) & (gt_area <= hi)), :]
            num_pos += gt.shape[0]
            boxes = entry['boxes'][np.flatnonzero(entry['gt_classes'] == 0), :] if candidate_boxes is None \
                else candidate_boxes[i]
            if boxes.shape[0] == 0:
                continue
            if limit is not None:
                boxes = boxes[:limit, :]
            found.append(recall_cover(boxes, gt))
        gt_overlaps = np.sort(np.concatenate(found)) if found else np.zeros(0)
        if thresholds is None:
            thresholds = np.arange(0.5, 0.95 + 1e-5, 0.05)
        recalls = np.zeros_like(thresholds)
        for i, t in enumerate(thresholds):
            recalls[i] = (gt_overlaps >= t).sum() / float(num_pos)
        ar = recalls.mean()
        print('average recall: {:.3f}'.format(ar))
        for t, r in zip(thresholds, recalls):
            print('recall @{:.2f}: {:.3f}'.format(t, r))
        return {'ar': ar, 'recalls': recalls, 'thresholds': thresholds, 'gt_overlaps': gt_overlaps}


class IMDB(RoidbOverlapMixin):
    """Bare container with the reference constructor's fields (imdb.py:13-21), enough to use the two methods."""

    def __init__(self, name):
        self.name = name
        self.classes = []
        self.num_classes = 0
        self.image_set_index = []
        self.num_images = 0
        self.config = {}
