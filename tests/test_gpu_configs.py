"""GPU: BASELINE.json configs[0..2] end to end.  host/slam-backend-gpu runs the reference's own example maps with the
reference's own configuration values (rebuilt from tests/golden/config_traces.npz — the GPU box has no /root/reference)
and is held against the trace of the UNMODIFIED reference program on the same inputs and rand() stream
(tests/golden/make_golden_configs.py):

    cfg0  FASTSLAM1  example_webmap    cfg1  FASTSLAM2  example_loop1    cfg2  EKF1  example_loop902 (117 landmarks)

Bars: iteration count, true vehicle path and observation counts bit-identical over the whole run (same steering, same
rand() consumption, same data association sizes); the estimate within 1e-4 (relative to the map scale; 1e-4 rad) of the
reference's for the first `follow` iterations — the whole run for EKF1; for the particle filters until the first
knife-edge resampling index falls the other way (see tests/test_gpu_host_app.py) — and tracking the true path to the
end like the reference does."""
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
EVERY = 8


@pytest.fixture(scope="module")
def exe():
    import slam_b200
    slam_b200.build()
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "host"), "all"])
    return os.path.join(ROOT, "host", "slam-backend-gpu")


def write_inputs(g, tag, d):
    with open(os.path.join(d, "map.mat"), "w") as f:
        for name in ("lm", "wp"):
            a = g["%s_%s" % (tag, name)]
            f.write("%s 2 %d\n" % (name, a.shape[0]))
            for x, y in a:
                f.write("%s %s\n" % (x, y))
    with open(os.path.join(d, "map.ini"), "w") as f:
        f.write("\n".join(str(s) for s in g[tag + "_ini"]) + "\n")


# follow: iterations over which the estimate must equal the reference's to 1e-4 (position relative to max(1, |pose|),
# heading absolute).  Measured on B200 (tools/config_follow.py): EKF1 all 34418; FASTSLAM1 688 (85 observation
# updates); FASTSLAM2 136 (16 updates; the same with -math reference: a 100-particle FastSLAM2 amplifies a 1e-7
# weight difference into a different resampling index within a few updates).  Asserted with margin.
# (Round 2: the fast arithmetic changed in its last bits — bearing innovation without an arctangent — and the FASTSLAM2
# realisation now parts from the reference's at a different update; the floor asserted for it is a few updates, the
# measured count is printed, and tests/test_gpu_host_app.py::test_tracking_quality_over_seeds holds the quality of the
# estimate against the reference over 12 seeds.)
@pytest.mark.parametrize("tag,follow", [("cfg2", None), ("cfg0", 600), ("cfg1", 40)])
def test_config_against_reference_program(exe, tmp_path, tag, follow):
    g = np.load(os.path.join(HERE, "golden", "config_traces.npz"))
    method = str(g[tag + "_method"])
    write_inputs(g, tag, str(tmp_path))
    trace = tmp_path / "trace.txt"
    r = subprocess.run([exe, "-m", "map.mat", "-method", method, "-trace", str(trace)], cwd=str(tmp_path),
                       capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    a = np.loadtxt(str(trace))
    assert a.shape[0] == int(g[tag + "_iterations"])
    ref = g[tag + "_trace"]
    a8 = a[(a[:, 0].astype(int) % EVERY) == 0]
    assert a8.shape == ref.shape
    assert np.array_equal(a8[:, 0], ref[:, 0]) and np.array_equal(a8[:, 7], ref[:, 7]), "iterations / observation counts"
    assert np.array_equal(a8[:, 1:4], ref[:, 1:4]), "true vehicle path"
    assert np.array_equal(a[-1, 1:4], g[tag + "_last"][1:4])
    from gpu_util import followed_iterations
    nrow, err = followed_iterations(a8, ref)
    print("%s (%s): follows the reference program to 1e-4 for %d of %d iterations (max %.1e)"
          % (tag, method, nrow * EVERY, ref.shape[0] * EVERY, err))
    assert nrow >= (ref.shape[0] if follow is None else follow // EVERY), (nrow * EVERY, err)
    # To the end of the run the estimate tracks the true path like the reference's own estimate does.  After the
    # realisations part this is a statement about a random process, not about bits: the RMS error over the run is held
    # to 2x the reference's + 0.5 m and the worst error to 3x + 1 m (the reference program itself, over 12 seeds of
    # cfg 0: RMS 0.21 .. 1.07 m, worst error 0.47 .. 1.47 m; tools/cfg_track.py prints the same for this program).
    e_ref = np.hypot(ref[:, 4] - ref[:, 1], ref[:, 5] - ref[:, 2])
    e = np.hypot(a8[:, 4] - a8[:, 1], a8[:, 5] - a8[:, 2])
    rms_ref, rms = np.sqrt(np.mean(e_ref ** 2)), np.sqrt(np.mean(e ** 2))
    print("%s: RMS position error %.3f m (reference %.3f), worst %.3f m (reference %.3f)"
          % (tag, rms, rms_ref, e.max(), e_ref.max()))
    assert rms <= 2.0 * rms_ref + 0.5 and e.max() <= 3.0 * e_ref.max() + 1.0, (rms, rms_ref, e.max(), e_ref.max())
