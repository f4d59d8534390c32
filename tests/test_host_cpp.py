"""The C++ host mirror (include/slamgpu_fastslam.hpp) compiles with plain g++ against the C ABI and behaves: on a
CPU-only box it throws (the product has no CPU path), on a GPU box it runs a FastSLAM2 cycle."""
import os
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
EXE = os.path.join(ROOT, "tests", "host_cpp", "host_smoke")


def build():
    import slam_b200
    slam_b200.build()
    lib = os.path.join(ROOT, "slam_b200")
    subprocess.check_call(["g++", "-std=c++14", "-O1", "-I", os.path.join(ROOT, "include"), "-o", EXE,
                           os.path.join(ROOT, "tests", "host_cpp", "host_smoke.cpp"), "-L", lib, "-lslamgpu",
                           "-Wl,-rpath," + lib])


def run():
    return subprocess.run([EXE], capture_output=True, text=True, timeout=120)


def test_host_cpp_throws_without_device():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    build()
    r = run()
    assert r.returncode == 3 and "no CUDA device" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
def test_host_cpp_cycle_on_gpu():
    build()
    r = run()
    assert r.returncode == 0 and "HOST_CPP_OK" in r.stdout, r.stdout + r.stderr
