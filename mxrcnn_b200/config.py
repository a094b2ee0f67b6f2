"""Hot-path configuration: same keys, defaults and mutability as the reference's global
``config`` (rcnn/config.py:4-70).  Scripts mutate it exactly as they mutate the reference's
(e.g. tools/test_rpn.py:14-16 sets TEST.RPN_PRE_NMS_TOP_N)."""
import numpy as np


class AttrDict(dict):
    """easydict.EasyDict stand-in (attribute access on a dict)."""

    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError:
            raise AttributeError(k)

    def __setattr__(self, k, v):
        self[k] = v


config = AttrDict()
config.EPS = 1e-14
config.PIXEL_MEANS = np.array([[[123.68, 116.779, 103.939]]])
config.SCALES = (600,)
config.MAX_SIZE = 1000
config.USE_GPU_NMS = True  # rcnn/config.py:13 -- never read by the reference; here NMS always runs on the GPU
config.GPU_ID = 0
config.END2END = 0

config.TRAIN = AttrDict()
config.TRAIN.FINETUNE = False
config.TRAIN.BATCH_SIZE = 128
config.TRAIN.HAS_RPN = False
config.TRAIN.ASPECT_GROUPING = True
config.TRAIN.BATCH_IMAGES = 2
config.TRAIN.FG_FRACTION = 0.25
config.TRAIN.FG_THRESH = 0.5
config.TRAIN.BG_THRESH_HI = 0.5
config.TRAIN.BG_THRESH_LO = 0.1
config.TRAIN.BBOX_REGRESSION_THRESH = 0.5
config.TRAIN.BBOX_INSIDE_WEIGHTS = np.array([1.0, 1.0, 1.0, 1.0])
config.TRAIN.RPN_BATCH_SIZE = 256
config.TRAIN.RPN_FG_FRACTION = 0.5
config.TRAIN.RPN_POSITIVE_OVERLAP = 0.7
config.TRAIN.RPN_NEGATIVE_OVERLAP = 0.3
config.TRAIN.RPN_CLOBBER_POSITIVES = False
config.TRAIN.RPN_BBOX_INSIDE_WEIGHTS = (1.0, 1.0, 1.0, 1.0)
config.TRAIN.RPN_POSITIVE_WEIGHT = -1.0
config.TRAIN.RPN_NMS_THRESH = 0.7
config.TRAIN.RPN_PRE_NMS_TOP_N = 12000
config.TRAIN.RPN_POST_NMS_TOP_N = 6000
config.TRAIN.RPN_MIN_SIZE = 16
config.TRAIN.BBOX_NORMALIZATION_PRECOMPUTED = False
config.TRAIN.BBOX_MEANS = (0.0, 0.0, 0.0, 0.0)
config.TRAIN.BBOX_STDS = (0.1, 0.1, 0.2, 0.2)
config.TRAIN.BBOX_MEANS_INV = (0.0, 0.0, 0.0, 0.0)
config.TRAIN.BBOX_STDS_INV = (10.0, 10.0, 5.0, 5.0)
config.TRAIN.IMS_PER_BATCH = 1

config.TEST = AttrDict()
config.TEST.HAS_RPN = False
config.TEST.BATCH_IMAGES = 1
config.TEST.NMS = 0.3
config.TEST.DEDUP_BOXES = 1. / 16.
config.TEST.RPN_NMS_THRESH = 0.7
config.TEST.RPN_PRE_NMS_TOP_N = 6000
config.TEST.RPN_POST_NMS_TOP_N = 300
config.TEST.RPN_MIN_SIZE = 16
