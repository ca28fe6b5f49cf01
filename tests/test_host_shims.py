"""The C++ host shims (package host/) driven the way the reference's callers drive the
classes they replace: see tests/host/test_host_shims.cpp."""
import os
import subprocess

import pytest

from conftest import ROOT, have_gpu

BIN = os.path.join(ROOT, "tests", "host", "test_host_shims")


def _ensure_built():
    if not os.path.exists(BIN):
        subprocess.check_call(["make", "-C", ROOT, "hosttest"], stdout=subprocess.DEVNULL)


def test_host_shims_fail_loudly_without_gpu():
    if have_gpu():
        pytest.skip("GPU present")
    _ensure_built()
    p = subprocess.run([BIN], capture_output=True, text=True)
    assert p.returncode == 2 and "no CPU fallback" in p.stdout


@pytest.mark.gpu
def test_host_shims_on_gpu():
    _ensure_built()
    p = subprocess.run([BIN], capture_output=True, text=True, timeout=300)
    print(p.stdout, p.stderr)
    assert p.returncode == 0 and "HOST_SHIM_TEST PASS" in p.stdout
