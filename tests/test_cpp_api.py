"""The host C++ API (host/bz_mesh.h, bz_bezier_mesh.h, bz_bezier_lens.h) used the way a reference caller would."""
import os
import subprocess

import pytest

import common

PKG = os.path.join(common.ROOT, "cuda-bezier-triangle-raytracer_b200")


def build(tmp_path):
    exe = str(tmp_path / "cpp_api_check")
    cuda = os.environ.get("CUDA_HOME", "/usr/local/cuda")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-I" + os.path.join(PKG, "host"), "-I" + os.path.join(common.ROOT, "include"), "-I" + os.path.join(cuda, "include"),
                           os.path.join(common.ROOT, "tests", "cpp_api_check.cpp"), "-L" + PKG, "-lbz_b200", "-L" + os.path.join(cuda, "lib64"), "-lcudart",
                           "-Wl,-rpath," + PKG, "-Wl,-rpath," + os.path.join(cuda, "lib64"), "-o", exe])
    return exe


def test_cpp_api_without_device(tmp_path):
    import bezier_b200 as bz
    if bz.device_count() > 0:
        pytest.skip("a CUDA device is present")
    out = subprocess.check_output([build(tmp_path)]).decode()
    assert out.startswith("ok cpu") and "no usable CUDA device" in out


@pytest.mark.gpu
def test_cpp_api_on_device(tmp_path):
    out = subprocess.check_output([build(tmp_path), "gpu"]).decode()
    assert out.strip().splitlines()[-1] == "ok gpu"      # (NCCL prints its version banner on stdout when the communicator is made)
