import importlib
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = "360_visual_inertial_odometry_b200"
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pkg(mod=None):
    return importlib.import_module(PKG if mod is None else PKG + "." + mod)


def have_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def rel_err(a, b):
    """max-norm relative error: max|a-b| / max(max|b|, tiny)"""
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    if a.size == 0:
        return 0.0
    return float(np.max(np.abs(a - b)) / max(float(np.max(np.abs(b))), 1e-300))


@pytest.fixture(scope="session")
def synthetic():
    return pkg("synthetic")


@pytest.fixture(scope="session")
def bindings():
    return pkg("bindings")


@pytest.fixture(scope="session")
def ctx(bindings):
    c = bindings.Context(0)
    yield c
    c.close()


@pytest.fixture(scope="session")
def emu_lib():
    """host-compiled twin of the device arithmetic (tests/emu)."""
    import ctypes
    src = os.path.join(ROOT, "tests", "emu", "ba_math_emu.cpp")
    so = os.path.join(ROOT, "tests", "emu", "libba_math_emu.so")
    hdr = os.path.join(ROOT, PKG, "csrc", "ba_math.cuh")
    if not os.path.exists(so) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-x", "c++", src, "-o", so,
                               "-Wno-unknown-pragmas"])
    lib = ctypes.CDLL(so)
    lib.emu_evaluate.restype = ctypes.c_double
    return lib
