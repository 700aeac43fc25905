"""CPU-only checks of the boundary: the library loads, exports every symbol include/ppx.h
declares, and refuses to run without a GPU (no CPU fallback)."""
import ctypes
import os
import re

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared():
    src = open(os.path.join(ROOT, "include", "ppx.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(ppx_[a-z0-9_]+)\s*\(", src)))


def test_header_symbols_exported():
    import pathplanning_b200.lib as L
    lib = ctypes.CDLL(L.LIB_PATH)
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/ppx.h but not exported by libppx.so"
    assert set(L.EXPORTED) == set(names), set(L.EXPORTED) ^ set(names)


def test_struct_layouts_match_header():
    import pathplanning_b200 as pp
    import pathplanning_b200.lib as L
    assert pp.RS_EDGE.itemsize == 72 and pp.CSTEER_EDGE.itemsize == 64 and pp.LINE_EDGE.itemsize == 48
    assert ctypes.sizeof(L.MapDescC) == 16 + 8 * 10
    assert ctypes.sizeof(L.RSCostParamsC) == 24


def test_no_cpu_fallback():
    import torch
    import pathplanning_b200 as pp
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(pp.PPXError):
        pp.Context(0)


def test_product_does_not_touch_oracle():
    """The product path must never import / link / execute anything under oracle/."""
    for tree in ("pathplanning_b200", "include", "dropin"):
        for dirpath, dirs, files in os.walk(os.path.join(ROOT, tree)):
            dirs[:] = [d for d in dirs if d != "_build"]
            for f in files:
                if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")) or f == "Makefile":
                    text = open(os.path.join(dirpath, f), errors="ignore").read()
                    assert "refapi" not in text and "pp_oracle" not in text and "libpp_ref" not in text, f
                    assert not re.search(r"[\"'/]oracle/", text), f
