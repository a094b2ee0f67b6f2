"""CPU: the C-ABI library loads and exports every symbol include/mxrcnn_b200.h declares."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    from mxrcnn_b200 import build, _lib
    build.build()
    return _lib.load()


def declared_symbols():
    text = open(os.path.join(ROOT, "include", "mxrcnn_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(mxr_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound(lib):
    from mxrcnn_b200 import _lib
    names = declared_symbols()
    assert len(names) >= 10
    for n in names:
        assert hasattr(lib, n), "declared in the header but not exported: %s" % n
        assert n in _lib.SIGNATURES, "no ctypes signature for %s" % n
    for n in _lib.SIGNATURES:
        assert n in names, "bound but not declared in include/mxrcnn_b200.h: %s" % n


def test_generate_anchors_host_table(lib):
    # host-side constant table (no GPU involved): reference helper/processing/generate_anchor.py:8
    from mxrcnn_b200.helper.processing.generate_anchor import generate_anchors
    from oracle import restate
    from tests.util import golden
    h = golden("helpers")
    got = generate_anchors(base_size=16.0, ratios=[0.5, 1, 2], scales=np.array([8, 16, 32], dtype=float))
    assert got.dtype == np.float64 and np.array_equal(got, h["anchors_voc"])
    assert np.array_equal(generate_anchors(16, [0.5, 1, 2], np.array([4, 8, 16, 32.])), h["anchors_wider"])
    for ratios, scales in [([0.5, 1, 2], [2, 4]), ([0.25, 0.5, 1, 2, 4], [1, 3, 7]), ([1], [8])]:
        assert np.array_equal(generate_anchors(16, ratios, np.array(scales, dtype=float)),
                              restate.generate_anchors(16, ratios, np.array(scales, dtype=float)))
    with pytest.raises(ValueError):
        generate_anchors(16, list(np.linspace(0.1, 3, 9)), np.arange(1, 9, dtype=float))  # 72 > 64 anchors


def test_missing_library_fails_loudly(monkeypatch):
    from mxrcnn_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", "/nonexistent/libmxrcnn_b200.so")
    with pytest.raises(_lib.MxrError):
        _lib.load()


def test_error_status_maps_to_exception(lib):
    from mxrcnn_b200 import _lib
    st = lib.mxr_generate_anchors(16.0, None, 0, None, 0, None)
    assert st == -1
    with pytest.raises(ValueError):
        _lib.check(st)
