"""CPU: the C-ABI library builds for sm_100a, loads, exports every symbol include/slamgpu.h declares, and refuses to
compute without a CUDA device (there is no CPU path in the product)."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    import slam_b200
    slam_b200.build()
    from slam_b200 import _lib
    return _lib.load()


def header_symbols():
    text = open(os.path.join(ROOT, "include", "slamgpu.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(slamgpu_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported(lib):
    from slam_b200 import _lib
    syms = header_symbols()
    assert len(syms) >= 25
    for s in syms:
        assert hasattr(lib, s), "missing export " + s
    assert sorted(_lib.EXPORTS) == syms  # the ctypes table and the header agree


def test_version_and_error_text(lib):
    assert b"sm_100a" in lib.slamgpu_version()
    assert isinstance(lib.slamgpu_last_error(), bytes)


def test_product_does_not_link_the_oracle():
    import subprocess
    from slam_b200 import _lib
    out = subprocess.run(["ldd", _lib.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in out and "slamref" not in out
    for root, _, files in os.walk(os.path.join(ROOT, "slam_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                src = open(os.path.join(root, f)).read()
                assert "pyoracle" not in src and "liboracle" not in src and "slam_oracle.h" not in src, f


def test_no_device_fails_loudly(lib):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    import slam_b200
    with pytest.raises(slam_b200.SlamGpuError) as e:
        slam_b200.FastSLAM(slam_b200.FASTSLAM2, 100, 8)
    assert e.value.code == -2
    with pytest.raises(slam_b200.SlamGpuError):
        slam_b200.compute_jacobians([0, 0, 0, 0.01, 0, 0, 0.0003, 1, 1, 0.1, 0, 0, 0.1], 1)


def test_bad_config_is_rejected(lib):
    from slam_b200 import _lib
    h = C.c_void_p()
    cfg = _lib.Config()
    cfg.n_particles, cfg.nf_cap, cfg.method = 0, 8, 2
    assert lib.slamgpu_create(C.byref(cfg), C.byref(h)) == -1
    cfg.n_particles, cfg.method = 100, 7
    assert lib.slamgpu_create(C.byref(cfg), C.byref(h)) == -1
    assert b"method" in lib.slamgpu_last_error()
