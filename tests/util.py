"""Shared helpers for the tests: golden loading and input reconstruction from seeds."""
import glob
import os

import numpy as np

from oracle import synth

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"), allow_pickle=False)


def golden_names(prefix):
    return sorted(os.path.basename(p)[:-4] for p in glob.glob(os.path.join(GOLDEN, prefix + "*.npz")))


def proposal_case(g):
    """(cls, box, info, cfg, train) for a tests/golden/proposal_<config>_s<seed>[_train].npz fixture."""
    cfg = synth.CONFIGS[str(g["config"])]
    A = len(cfg["scales"]) * len(cfg["ratios"])
    pad = tuple(int(x) for x in g["pad_hw"])
    pad = pad if pad != (0, 0) else None
    cls, box, info = synth.proposal_inputs(int(g["seed"]), H=cfg["H"], W=cfg["W"], A=A, im_info=cfg["im_info"], pad_hw=pad)
    return cls, box, info, cfg, bool(g["train"])


def box_head_case(g):
    """(inputs of synth.box_head_inputs, nms threshold, per-class dets, per-class roi rows) of a box_head_* fixture."""
    import ast
    kw = ast.literal_eval(str(g["kw"]))
    case = synth.box_head_inputs(int(g["seed"]), **kw)
    off = g["offsets"]
    dets = [g["dets"][off[j]:off[j + 1]] for j in range(len(off) - 1)]
    rows = [g["rows"][off[j]:off[j + 1]] for j in range(len(off) - 1)]
    return case, float(g["nms_thresh"]), dets, rows
