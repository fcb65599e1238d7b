"""CPU: the C-ABI library builds for sm_100a, loads, and exports every symbol include/isb.h declares
(no compute calls -- there is no GPU here)."""
import ctypes
import os
import re

import pytest

from conftest import ROOT


@pytest.fixture(scope="module")
def lib():
    import __graft_entry__ as g
    g.build()
    from isb import _lib
    return _lib


def header_symbols():
    text = open(os.path.join(ROOT, "include", "isb.h")).read()
    return sorted(set(re.findall(r"ISB_API[^;(]*?\b(isb_\w+)\s*\(", text)))


def test_header_declares_entry_points():
    syms = header_symbols()
    assert "isb_yolact_detect" in syms and "isb_create" in syms and len(syms) >= 15


def test_library_exports_every_declared_symbol(lib):
    cdll = lib.load()
    for name in header_symbols():
        assert hasattr(cdll, name), f"libisb.so does not export {name}"
    assert sorted(lib.SYMBOLS) == header_symbols(), "python binding list and header disagree"
    assert cdll.isb_version() == 100


def test_every_entry_point_cites_the_reference():
    text = open(os.path.join(ROOT, "include", "isb.h")).read()
    assert text.count("utils_bbox.py:") >= 8


def test_struct_layout_matches_header(lib):
    # field order / sizes of the ctypes mirrors (LP64): catches drift between isb.h and _lib.py
    assert ctypes.sizeof(lib.Limits) == 32
    assert ctypes.sizeof(lib.YolactIn) == 8 * 4 + 5 * 8 + 4 * 4 + 8   # + conf_is_logits, padded to 8
    assert ctypes.sizeof(lib.YolactOut) == 7 * 8
    assert ctypes.sizeof(lib.YolactMapOut) == 5 * 8


def test_no_cpu_fallback(lib):
    import torch
    from isb.yolact import BBoxUtility
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        BBoxUtility().decode_boxes(torch.zeros(4, 4), torch.zeros(4, 4))


def test_product_does_not_import_oracle():
    pkg = os.path.join(ROOT, "instancesegmentation-_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h")):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle", src, re.M), f"{f} imports the oracle"
