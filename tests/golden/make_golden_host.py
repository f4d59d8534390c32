"""Generates tests/golden/host_traces.npz: the per-iteration true / estimated pose of the UNMODIFIED reference program
(oracle/_ref/slam-backend-ref = the reference's wrappers + SLAMBackendApplication + slam-backend.cpp with
oracle/ref_app/TracePlot.cpp in place of the zmq NetworkPlot) on tests/data/ring_small.mat for the three methods.
Run in the build container only:  python tests/golden/make_golden_host.py"""
import os
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
DATA = os.path.join(ROOT, "tests", "data")


def main():
    exe = os.path.join(ROOT, "oracle", "_ref", "slam-backend-ref")
    out = {}
    for method in ("FASTSLAM1", "FASTSLAM2", "EKF1"):
        with tempfile.NamedTemporaryFile(suffix=".txt") as t:
            env = dict(os.environ, SLAMREF_TRACE=t.name)
            subprocess.run([exe, "-m", "ring_small.mat", "-method", method], cwd=DATA, env=env, check=True,
                           stdout=subprocess.DEVNULL)
            out[method] = np.loadtxt(t.name)
        print(method, out[method].shape)
    np.savez_compressed(os.path.join(HERE, "host_traces.npz"), **out)


if __name__ == "__main__":
    main()
