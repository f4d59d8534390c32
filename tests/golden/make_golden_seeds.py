"""Generates tests/golden/host_seed_stats.npz: how well the UNMODIFIED reference program (oracle/_ref/slam-backend-ref)
tracks the true path on tests/data/ring_small.mat over a set of rand() seeds, per method: RMS and maximum position
error of the estimate and the final error.  A 100-particle filter turns a 1e-7 weight difference into a different
resampling index within a few updates, so two correct implementations part ways as random realisations; what must
agree is the QUALITY of the estimate over seeds (tests/test_gpu_host_app.py::test_tracking_quality_over_seeds).
Run in the build container only:  python tests/golden/make_golden_seeds.py"""
import os
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
DATA = os.path.join(ROOT, "tests", "data")
SEEDS = list(range(101, 113))


def stats(trace):
    e = np.hypot(trace[:, 4] - trace[:, 1], trace[:, 5] - trace[:, 2])
    return np.array([np.sqrt(np.mean(e ** 2)), e.max(), e[-1]])


def main():
    exe = os.path.join(ROOT, "oracle", "_ref", "slam-backend-ref")
    out = {"seeds": np.array(SEEDS)}
    for method in ("FASTSLAM1", "FASTSLAM2"):
        rows = []
        for seed in SEEDS:
            with tempfile.NamedTemporaryFile(suffix=".txt") as t:
                env = dict(os.environ, SLAMREF_TRACE=t.name)
                subprocess.run([exe, "-m", "ring_small.mat", "-method", method, "-SWITCH_SEED_RANDOM", str(seed)],
                               cwd=DATA, env=env, check=True, stdout=subprocess.DEVNULL)
                rows.append(stats(np.loadtxt(t.name)))
        out[method] = np.array(rows)
        print(method, "rms mean %.4f  max %.4f" % (out[method][:, 0].mean(), out[method][:, 1].max()), out[method][:, 0].round(3))
    np.savez_compressed(os.path.join(HERE, "host_seed_stats.npz"), **out)


if __name__ == "__main__":
    main()
