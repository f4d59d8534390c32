"""CPU: the C restatement (oracle/slam_oracle.c) against golden vectors produced by the unmodified
reference (tests/golden/make_golden.py).  This pins the oracle; it runs everywhere (no reference needed)."""
import os

import numpy as np
import pytest

from replay import compare_states, load_golden, run_oracle

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

# Continuous state: the oracle is bit-exact with the reference except FastSLAM2::gaussEvaluate, whose
# Jacobi-SVD solve (fastslam2.cpp:132) is restated as a triangular solve (~4e-6 relative per call).
TOL = 1e-4


def test_jacobians_kat_bit_exact(oracle):
    z = np.load(os.path.join(G, "jacobians_kat.npz"))
    n = int(z["n"])
    for rec, out in zip(z["rec"], z["out"]):
        got = oracle.compute_jacobians(rec, n)
        assert np.array_equal(got, out)


@pytest.mark.parametrize("N", [100, 1000, 5000])
def test_resample_kat_bit_exact(oracle, N):
    z = np.load(os.path.join(G, "resample_kat.npz"))
    for trial in range(2):
        w = z["w_%d_%d" % (N, trial)]
        unif = z["ints_%d_%d" % (N, trial)] / float(2147483647)
        wn, keep, neff = oracle.resample_indices(w, unif, scan_mode=0)
        assert np.array_equal(keep, z["keep_%d_%d" % (N, trial)])
        assert neff == z["neff_%d_%d" % (N, trial)]


@pytest.mark.parametrize("method", [1, 2])
def test_drive_matches_reference_golden(oracle, method):
    recs, outs, meta = load_golden(os.path.join(G, "drive_fs%d_p100.npz" % method))
    got = run_oracle(oracle, recs, method, int(meta["P"]), int(meta["nf_cap"]), scan_mode=0)
    n_upd = 0
    for i, (a, b) in enumerate(zip(got, outs)):
        compare_states(a, b, TOL, "record %d (%s)" % (i, recs[i]["kind"]))
        if recs[i]["kind"] == "update":
            n_upd += 1
    assert n_upd == 10
