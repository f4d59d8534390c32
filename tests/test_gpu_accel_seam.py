"""GPU: the literal accelerator seam, compiled and run.

oracle/_ref/slam-backend-ref-accel is the REFERENCE's own program — core.cpp, the FastSLAM classes, the wrappers and
SLAMBackendApplication, all unmodified — built with the reference's own switch -DJACOBIAN_ACCELERATOR=1
(src/backend/CMakeLists.txt:45-47), so that computeJacobians takes its accelerator branch (core.cpp:586-664) through the
reference's own AcceleratorHandler class (AcceleratorHandler.h:10-33).  The only replaced part is the class's
implementation: getMemoryPointer / setN / start / isDone run on slam_b200/libslamgpu.so
(oracle/ref_accel/AcceleratorHandler_slamgpu.cpp: start() -> slamgpu_compute_jacobians) instead of /dev/mem and the HLS
driver.  Every Jacobian of the run — FastSLAM2's proposal, likelihood and feature update call computeJacobians 5-6 times
per particle and update — is then computed on the B200, inside the reference's own filter.

Held against tests/golden/host_traces.npz, the trace of the same program WITHOUT the switch (CPU Jacobians): the true
path and observation counts are bit-identical, and the estimate follows to 1e-4 until the first knife-edge resampling
index differs (the GPU's sqrt / atan2 / division are correctly rounded or within an ulp of libm, so this is far into
the run)."""
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
EXE = os.path.join(ROOT, "oracle", "_ref", "slam-backend-ref-accel")


@pytest.mark.parametrize("method,follow", [("FASTSLAM2", 64), ("FASTSLAM1", 400)])
def test_reference_program_with_gpu_jacobians(tmp_path, method, follow):
    if not os.path.exists(EXE):
        pytest.skip("oracle/_ref/slam-backend-ref-accel not built (needs /root/reference at build time)")
    g = np.load(os.path.join(HERE, "golden", "host_traces.npz"))[method]
    trace = tmp_path / "trace.txt"
    env = dict(os.environ, SLAMREF_TRACE=str(trace), SLAMREF_ACCEL_STATS="1")
    r = subprocess.run([EXE, "-m", "ring_small.mat", "-method", method], cwd=os.path.join(HERE, "data"), env=env,
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "records through slamgpu" in r.stderr
    n_records = int(r.stderr.split("AcceleratorHandler:")[1].split()[0])
    assert n_records > 10000, r.stderr[-500:]                      # every computeJacobians call went to the device
    a = np.loadtxt(str(trace))
    assert a.shape == g.shape
    assert np.array_equal(a[:, 0], g[:, 0]) and np.array_equal(a[:, 7], g[:, 7])
    assert np.array_equal(a[:, 1:4], g[:, 1:4]), "true vehicle path"
    scale = max(1.0, np.abs(g[:, 4:6]).max())
    err = np.maximum(np.abs(a[:, 4:6] - g[:, 4:6]).max(axis=1) / scale,
                     np.abs((a[:, 6] - g[:, 6] + np.pi) % (2 * np.pi) - np.pi))
    bad = np.nonzero(err > 1e-4)[0]
    followed = int(bad[0]) if bad.size else len(err)
    print("%s through the accelerator seam: %d records on the B200, trace within 1e-4 of the CPU reference for %d of %d "
          "iterations" % (method, n_records, followed, len(err)))
    assert followed >= follow, (followed, err[:followed + 1].max())
    # the filter keeps tracking after the realisations part ways
    assert np.abs(a[:, 4:6] - a[:, 1:3]).max() < 1.0
