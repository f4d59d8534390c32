"""Generates tests/golden/ekf_drive.npz from the UNMODIFIED reference EKFSLAM (oracle/_ref/libslamref.so through
oracle/ref_shim_ekf.cpp).  Run in the build container only:  python tests/golden/make_golden_ekf.py

The drive (tests/ekf_scenario.py) free-runs through EKFSLAM::sim; for every observe tick and every 16th control tick
the state before and after the call is stored, together with the association the call made, so that the tests can
replay single calls from the reference's own state (per-call parity) as well as the whole drive."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

from ekf_scenario import GATE_AUGMENT, GATE_REJECT, QE, RE, SIGMA_T, WHEELBASE, ekf_records  # noqa: E402
from oracle.pyekf import EkfRef  # noqa: E402

SEED, N_LM, N_TICKS = 3, 16, 1760


def main():
    lm, recs = ekf_records(seed=SEED, n_landmarks=N_LM, n_ticks=N_TICKS)
    r = EkfRef(WHEELBASE, SIGMA_T, GATE_REJECT, GATE_AUGMENT)
    out = {"seed": SEED, "n_landmarks": N_LM, "n_ticks": N_TICKS}
    ticks = []
    for i, rc in enumerate(recs):
        keep = rc["observe"] or i % 16 == 0
        if keep:
            x0, P0 = r.get()
        idf, nzf, nzn = r.sim(rc["V"], rc["G"], QE, rc["dt"], rc["phi"], rc["observe"], rc["z"], rc["idz"], None, RE, RE)
        if keep:
            x1, P1 = r.get()
            ticks.append(i)
            out["x0_%d" % i], out["P0_%d" % i], out["x1_%d" % i], out["P1_%d" % i] = x0, P0, x1, P1
            out["idf_%d" % i] = idf
            out["nz_%d" % i] = np.array([nzf, nzn], np.int32)
    out["ticks"] = np.array(ticks, np.int32)
    xf, Pf = r.get()
    out["x_final"], out["P_final"] = xf, Pf
    np.savez_compressed(os.path.join(HERE, "ekf_drive.npz"), **out)
    print("ekf_drive.npz: %d stored calls, final n = %d" % (len(ticks), len(xf)))


if __name__ == "__main__":
    main()
