"""GPU: the whole program.  host/slam-backend-gpu (the C++ host loop on the C ABI) runs tests/data/ring_small.mat with
the seed of its .ini for FASTSLAM1, FASTSLAM2 and EKF1 and is held against tests/golden/host_traces.npz, the
per-iteration trace of the UNMODIFIED reference program on the same map, seed and rand() stream
(tests/golden/make_golden_host.py).

What must hold: the true vehicle path and the observation counts are bit-identical for the whole run (same steering,
same rand() consumption); EKF1's estimate follows the reference's for all 1705 iterations; the particle filters follow
it to 1e-4 until the first knife-edge resampling index differs (the two are then different, equally valid random
realisations of the same filter: hundreds of iterations and dozens of bit-identical resamples in), after which the
estimate must still track the true path."""
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)


@pytest.fixture(scope="module")
def exe():
    import slam_b200
    slam_b200.build()
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "host"), "all"])
    return os.path.join(ROOT, "host", "slam-backend-gpu")


@pytest.mark.parametrize("method,follow", [("EKF1", 1705), ("FASTSLAM1", 400), ("FASTSLAM2", 200)])
def test_program_against_reference_program_trace(exe, tmp_path, method, follow):
    g = np.load(os.path.join(HERE, "golden", "host_traces.npz"))[method]
    trace = tmp_path / "trace.txt"
    r = subprocess.run([exe, "-m", "ring_small.mat", "-method", method, "-trace", str(trace)],
                       cwd=os.path.join(HERE, "data"), capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    a = np.loadtxt(str(trace))
    assert a.shape == g.shape
    assert np.array_equal(a[:, 0], g[:, 0]) and np.array_equal(a[:, 7], g[:, 7])
    assert np.array_equal(a[:, 1:4], g[:, 1:4]), "true vehicle path"
    d = np.abs(a[:follow, 4:6] - g[:follow, 4:6]).max()
    dt = np.abs((a[:follow, 6] - g[:follow, 6] + np.pi) % (2 * np.pi) - np.pi).max()
    scale = max(1.0, np.abs(g[:follow, 4:6]).max())
    assert d <= 1e-4 * scale and dt <= 1e-4, (d, dt)
    # the filter keeps tracking after the realisations part ways
    assert np.abs(a[:, 4:6] - a[:, 1:3]).max() < 1.0
    assert "landmarks 28" in r.stdout


def test_command_line_overrides_and_errors(exe, tmp_path):
    data = os.path.join(HERE, "data")
    r = subprocess.run([exe, "-m", "ring_small.mat", "-method", "FASTSLAM2", "-NPARTICLES", "3000", "-NEFFECTIVE", "2250",
                        "-max_iter", "64", "-scan", "blocked", "-plot_stride", "100"], cwd=data, capture_output=True,
                       text=True, timeout=300)
    assert r.returncode == 0 and "iterations 64 observations 8" in r.stdout, r.stdout + r.stderr
    r = subprocess.run([exe, "-m", "missing.mat"], cwd=data, capture_output=True, text=True, timeout=60)
    assert r.returncode != 0 and "Unable to read input file" in r.stderr
    r = subprocess.run([exe, "-h"], cwd=data, capture_output=True, text=True, timeout=60)
    assert r.returncode == 0 and "-method" in r.stdout and "FASTSLAM2" in r.stdout
