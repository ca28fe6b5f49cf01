"""CPU checks of the C-ABI boundary: the library loads, exports exactly what
include/vio360.h declares, and refuses to run without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from conftest import ROOT, have_gpu, pkg


def _declared():
    txt = open(os.path.join(ROOT, "include", "vio360.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(vio360_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol():
    b = pkg("bindings")
    lib = C.CDLL(b.LIB_PATH)
    names = _declared()
    assert len(names) >= 30
    for n in names:
        assert hasattr(lib, n), "missing export " + n
    assert sorted(b.EXPORTS) == names


def test_no_cpu_fallback():
    if have_gpu():
        pytest.skip("GPU present")
    b = pkg("bindings")
    with pytest.raises(b.Vio360Error) as e:
        b.Context(0)
    assert "no CPU fallback" in str(e.value) or "CUDA" in str(e.value)


def test_product_code_never_touches_the_oracle():
    """oracle/ is test infrastructure: nothing under the package or include/ may name it."""
    bad = []
    for base in (os.path.join(ROOT, "360_visual_inertial_odometry_b200"), os.path.join(ROOT, "include")):
        for dp, _, fns in os.walk(base):
            for fn in fns:
                if fn.endswith((".py", ".cu", ".cuh", ".h", ".cpp", ".hpp")):
                    s = open(os.path.join(dp, fn), errors="ignore").read()
                    if re.search(r"oracle[/_.]|ba_ref|fe_ref|import cv2|cv2\.", s):
                        bad.append(os.path.join(dp, fn))
    assert not bad, bad
