"""Generates tests/golden/*.npz from the UNMODIFIED reference (oracle/_ref/libslamref.so, built in
place from /root/reference by oracle/Makefile).  Run in the build container only:

    python tests/golden/make_golden.py

The vectors travel to the GPU box, the reference does not.
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from oracle.pyoracle import Oracle, Ref  # noqa: E402
from replay import make_records, run_ref, save_golden  # noqa: E402
from scenario import RAND_MAX  # noqa: E402


def main():
    o, r = Oracle(), Ref()
    # drives: FastSLAM1 / FastSLAM2, 100 particles (the reference's NPARTICLES default)
    for method, seed in ((1, 11), (2, 12)):
        P, cap = 100, 32
        recs = make_records(o, seed, method, P, n_observe=10)
        outs = run_ref(r, recs, method, P, cap)
        save_golden(os.path.join(HERE, "drive_fs%d_p100.npz" % method), recs, outs,
                    dict(method=method, P=P, nf_cap=cap, seed=seed))
    # computeJacobians known-answer records (core.cpp:586-664 layout)
    rng = np.random.default_rng(5)
    recs_in, recs_out = [], []
    for t in range(64):
        n = 4
        xv = rng.normal(0, 5, 3)
        xv[2] = rng.uniform(-3.2, 3.2)
        rec = [*xv, 0.01, 0, 0, 0.0003046]
        for i in range(n):
            xf = xv[:2] + rng.normal(0, 15, 2)
            A = rng.normal(0, 0.3, (2, 2))
            Pf = A @ A.T + 0.01 * np.eye(2)
            rec += [*xf, Pf[0, 0], Pf[1, 0], Pf[0, 1], Pf[1, 1]]
        rec = np.array(rec, np.float32)
        recs_in.append(rec)
        recs_out.append(r.compute_jacobians(rec, n))
    np.savez_compressed(os.path.join(HERE, "jacobians_kat.npz"), rec=np.stack(recs_in), out=np.stack(recs_out), n=4)
    # stratifiedResample on reference-valid N
    res = {}
    for N in (100, 1000, 5000):
        assert r.stratified_count(N) == N
        for trial in range(2):
            w = rng.random(N).astype(np.float32) ** (4 if trial else 1)
            ints = rng.integers(0, RAND_MAX, size=N).astype(np.int32)
            r.set_stream(ints)
            keep, neff = r.stratified_resample(w)
            res["w_%d_%d" % (N, trial)] = w
            res["ints_%d_%d" % (N, trial)] = ints
            res["keep_%d_%d" % (N, trial)] = keep
            res["neff_%d_%d" % (N, trial)] = neff
    np.savez_compressed(os.path.join(HERE, "resample_kat.npz"), **res)
    print("golden vectors written to", HERE)


if __name__ == "__main__":
    main()
