"""helper/processing/generate_anchor.py of the reference, same signature."""
import ctypes

import numpy as np

from ... import _lib


def generate_anchors(base_size=16, ratios=[0.5, 1, 2], scales=2 ** np.arange(3, 6)):
    """(len(ratios)*len(scales), 4) float64 anchors around a (0,0,base-1,base-1) window, ratio-major
    (reference: helper/processing/generate_anchor.py:8-72)."""
    r = np.ascontiguousarray(ratios, dtype=np.float64).ravel()
    s = np.ascontiguousarray(scales, dtype=np.float64).ravel()
    out = np.empty((r.size * s.size, 4), dtype=np.float64)
    _lib.check(_lib.load().mxr_generate_anchors(
        float(base_size), r.ctypes.data_as(_lib.c_double_p), int(r.size), s.ctypes.data_as(_lib.c_double_p),
        int(s.size), out.ctypes.data_as(_lib.c_double_p)))
    return out
