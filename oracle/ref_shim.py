"""Import the REAL reference (walkoncross/mx-rcnn) hot-path modules without copying them.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Works only where the reference tree
is mounted (``$MXRCNN_REF``, default ``/root/reference``) -- i.e. in the build container.
It is used by ``oracle/make_golden.py`` to produce ``tests/golden/*.npz`` and by the
``not gpu`` tests (when the tree exists) to re-validate ``oracle/restate.py`` live.

The reference is Python 2 on top of an un-vendored MXNet.  The shim
  * installs stub modules ``easydict`` (attribute dict) and ``mxnet`` /
    ``mxnet.operator`` (``CustomOp.assign``: ``dst[:] = src``, ``CustomOpProp``,
    ``register``) -- MXNet only transports arrays on this path;
  * installs a meta-path finder for the ``helper`` and ``rcnn`` packages that reads each
    source file from the reference tree and, IN MEMORY ONLY, rewrites
      - ``print X``  ->  ``print(X)`` (balanced across continuation lines),
      - the implicit relative import ``from bbox_transform import`` (bbox_regression.py:8),
      - the two py2 integer divisions ``BATCH_SIZE / self._img_per_batch``
        (proposal_target.py:39,104) to ``//``;
  * aliases ``np.float = float`` (removed in numpy>=1.24; used at bbox_transform.py:46,
    bbox_regression.py:20,43, minibatch.py:295, proposal_target.py:141-142) and
    ``builtins.xrange = range`` (generate_anchor.py:18).
No reference source text is written to disk or into this repository.
"""
import builtins
import importlib
import importlib.abc
import importlib.util
import os
import re
import sys
import types

import numpy as np

REF_ROOT = os.environ.get("MXRCNN_REF", "/root/reference")


def available():
    return os.path.isfile(os.path.join(REF_ROOT, "rcnn", "rpn", "proposal.py"))


def _rewrite_prints(src):
    """py2 ``print expr`` statements -> ``print(expr)``; keeps line count stable."""
    lines = src.split("\n")
    out = []
    i = 0
    while i < len(lines):
        line = lines[i]
        m = re.match(r"^(\s*)print\s+(.*)$", line)
        if m and not re.match(r"^\s*print\s*\(.*\)\s*$", line):
            indent, expr = m.group(1), m.group(2)
            depth = expr.count("(") + expr.count("[") - expr.count(")") - expr.count("]")
            extra = []
            while depth > 0 and i + 1 < len(lines):
                i += 1
                extra.append(lines[i])
                depth += lines[i].count("(") + lines[i].count("[") - lines[i].count(")") - lines[i].count("]")
            if extra:
                out.append(indent + "print(" + expr)
                out.extend(extra[:-1])
                out.append(extra[-1] + ")")
            else:
                out.append(indent + "print(" + expr + ")")
        else:
            out.append(line)
        i += 1
    return "\n".join(out)


def _rewrite(src, relpath):
    src = _rewrite_prints(src)
    src = src.replace("from bbox_transform import", "from helper.processing.bbox_transform import")
    src = src.replace("from bbox_regression import", "from helper.processing.bbox_regression import")  # roidb.py:10
    src = src.replace("areas.has_key(area)", "(area in areas)")  # helper/dataset/imdb.py:124 (py2 dict method)
    src = src.replace("config.TRAIN.BATCH_SIZE / self._img_per_batch",
                      "config.TRAIN.BATCH_SIZE // self._img_per_batch")
    src = src.replace("config.TRAIN.BATCH_SIZE / config.TRAIN.BATCH_IMAGES",
                      "config.TRAIN.BATCH_SIZE // config.TRAIN.BATCH_IMAGES")
    return src


class _RefLoader(importlib.abc.Loader):
    def __init__(self, path, is_pkg):
        self.path, self.is_pkg = path, is_pkg

    def create_module(self, spec):
        return None

    def exec_module(self, module):
        if self.path is None:  # namespace-like package without __init__
            return
        with open(self.path, "r") as f:
            src = f.read()
        code = compile(_rewrite(src, self.path), self.path, "exec")
        exec(code, module.__dict__)


class _RefFinder(importlib.abc.MetaPathFinder):
    ROOTS = ("helper", "rcnn")

    def find_spec(self, fullname, path=None, target=None):
        if fullname.split(".")[0] not in self.ROOTS:
            return None
        rel = os.path.join(REF_ROOT, *fullname.split("."))
        if os.path.isdir(rel):
            init = os.path.join(rel, "__init__.py")
            loader = _RefLoader(init if os.path.isfile(init) else None, True)
            spec = importlib.util.spec_from_loader(fullname, loader, is_package=True)
            spec.submodule_search_locations = [rel]
            return spec
        if os.path.isfile(rel + ".py"):
            return importlib.util.spec_from_loader(fullname, _RefLoader(rel + ".py", False))
        return None


class _EasyDict(dict):
    def __getattr__(self, k):
        try:
            return self[k]
        except KeyError:
            raise AttributeError(k)

    def __setattr__(self, k, v):
        self[k] = v


def _install_stubs():
    if "easydict" not in sys.modules:
        m = types.ModuleType("easydict")
        m.EasyDict = _EasyDict
        sys.modules["easydict"] = m
    if "mxnet" not in sys.modules:
        mx = types.ModuleType("mxnet")
        op = types.ModuleType("mxnet.operator")

        class CustomOp(object):
            def assign(self, dst, req, src):
                if req == "null":
                    return
                if req == "add":
                    dst[:] += src
                else:
                    dst[:] = src

        class CustomOpProp(object):
            def __init__(self, need_top_grad=True):
                self.need_top_grad_ = need_top_grad

        _registry = {}

        def register(name):
            def deco(cls):
                _registry[name] = cls
                return cls
            return deco

        op.CustomOp, op.CustomOpProp, op.register, op._registry = CustomOp, CustomOpProp, register, _registry
        mx.operator = op
        # what rcnn/detector.py touches: mx.cpu(), mx.nd.array / mx.nd.zeros (float32 like MXNet's default dtype)
        nd = types.ModuleType("mxnet.ndarray")
        nd.array = lambda a, ctx=None: NDArrayStub(np.asarray(a, dtype=np.float32))
        nd.zeros = lambda shape, ctx=None: NDArrayStub(np.zeros(shape, dtype=np.float32))
        mx.nd = nd
        mx.cpu = lambda *a: "cpu"
        sys.modules["mxnet.ndarray"] = nd
        sys.modules["mxnet"] = mx
        sys.modules["mxnet.operator"] = op


_installed = False


def install():
    """Idempotent. Returns True when the reference tree is importable."""
    global _installed
    if not available():
        return False
    if _installed:
        return True
    if not hasattr(np, "float"):
        np.float = float  # noqa: NPY001 -- reference uses the removed alias
    if not hasattr(builtins, "xrange"):
        builtins.xrange = range
    _install_stubs()
    if "cPickle" not in sys.modules:  # rcnn/tester.py:4 (py2 module name)
        import pickle
        sys.modules["cPickle"] = pickle
    sys.meta_path.insert(0, _RefFinder())
    _installed = True
    return True


class NDArrayStub(object):
    """Minimal stand-in for mx.nd.NDArray: ``asnumpy()`` and slice assignment."""

    def __init__(self, arr):
        self.arr = np.array(arr)

    def asnumpy(self):
        return self.arr.copy()

    def __setitem__(self, k, v):
        self.arr[k] = v

    def __getitem__(self, k):
        return self.arr[k]

    @property
    def shape(self):
        return self.arr.shape


def ref():
    """Namespace with the reference's hot-path callables (the real code)."""
    if not install():
        raise RuntimeError("reference tree not found at %s" % REF_ROOT)
    ns = types.SimpleNamespace()
    ns.config = importlib.import_module("rcnn.config").config
    bt = importlib.import_module("helper.processing.bbox_transform")
    ns.bbox_pred, ns.clip_boxes, ns.clip_pad, ns.bbox_transform = bt.bbox_pred, bt.clip_boxes, bt.clip_pad, bt.bbox_transform
    nm = importlib.import_module("helper.processing.nms")
    ns.nms, ns.nest = nm.nms, nm.nest
    br = importlib.import_module("helper.processing.bbox_regression")
    ns.bbox_overlaps = br.bbox_overlaps
    ns.expand_bbox_regression_targets = br.expand_bbox_regression_targets
    ns.compute_bbox_regression_targets = br.compute_bbox_regression_targets
    ns.generate_anchors = importlib.import_module("helper.processing.generate_anchor").generate_anchors
    prop = importlib.import_module("rcnn.rpn.proposal")
    ns.ProposalOperator, ns.ProposalProp = prop.ProposalOperator, prop.ProposalProp
    pt = importlib.import_module("rcnn.rpn.proposal_target")
    ns.ProposalTargetOperator, ns.ProposalTargetProp = pt.ProposalTargetOperator, pt.ProposalTargetProp
    ns._sample_rois = pt._sample_rois
    mb = importlib.import_module("rcnn.minibatch")
    ns.assign_anchor, ns.sample_rois = mb.assign_anchor, mb.sample_rois
    ns.add_bbox_regression_targets = importlib.import_module("helper.processing.roidb").add_bbox_regression_targets
    ns.IMDB = importlib.import_module("helper.dataset.imdb").IMDB
    ns.generate_detections = importlib.import_module("rcnn.rpn.generate").generate_detections
    ns.Detector = importlib.import_module("rcnn.detector").Detector
    ns.pred_eval = importlib.import_module("rcnn.tester").pred_eval
    return ns
