"""Shared fixtures.  `-m "not gpu"` runs here on the CPU; `-m gpu` are the parity tests proper and run on a B200."""
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
sys.path.insert(0, os.path.join(ROOT, "cuda-bezier-triangle-raytracer_b200"))
sys.path.insert(0, os.path.join(ROOT, "tests"))


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def oracle_built():
    """plain-C oracle + (when /root/reference is mounted) the compiled reference; prebuilt files are kept otherwise"""
    subprocess.check_call(["make", "-C", os.path.join(ROOT, "oracle")], stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    import bzref
    if not bzref.available():
        pytest.skip("oracle/_ref/libbzref.so absent and /root/reference not mounted")
    return True


@pytest.fixture(scope="session")
def shim():
    """product device arithmetic + cull hierarchy compiled for the CPU (tests/host_shim.cpp) -- test infrastructure"""
    import common
    return common.load_shim()
