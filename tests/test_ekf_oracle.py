"""CPU: the EKF-SLAM restatement (oracle/ekf_oracle.c) pinned against the UNMODIFIED reference EKFSLAM — through the
golden drive the reference produced (tests/golden/ekf_drive.npz, always) and live against oracle/_ref/libslamref.so
(when it is present: the build container)."""
import os

import numpy as np
import pytest

from ekf_scenario import GATE_AUGMENT, GATE_REJECT, QE, RE, SIGMA_T, WHEELBASE, block_relerr, ekf_records

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def golden():
    g = np.load(os.path.join(HERE, "golden", "ekf_drive.npz"))
    lm, recs = ekf_records(seed=int(g["seed"]), n_landmarks=int(g["n_landmarks"]), n_ticks=int(g["n_ticks"]))
    return g, recs


def new_oracle(**kw):
    import subprocess
    subprocess.check_call(["make", "-s", "-C", os.path.join(HERE, "..", "oracle"), "liboracle.so"])
    from oracle.pyekf import EkfOracle
    return EkfOracle(WHEELBASE, SIGMA_T, GATE_REJECT, GATE_AUGMENT, **kw)


def test_oracle_replays_every_stored_reference_call(golden):
    """Each stored call starts from the reference's own state: association bit-exact, state within 1e-6."""
    g, recs = golden
    o = new_oracle()
    worst = {}
    for i in g["ticks"]:
        rc = recs[int(i)]
        o.set(g["x0_%d" % i], g["P0_%d" % i])
        idf, nzf, nzn = o.sim(rc["V"], rc["G"], QE, rc["dt"], rc["phi"], rc["observe"], rc["z"], rc["idz"], None, RE, RE)
        assert np.array_equal(idf, g["idf_%d" % i]) and [nzf, nzn] == list(g["nz_%d" % i]), i
        x, P = o.get()
        for k, v in block_relerr(x, P, g["x1_%d" % i], g["P1_%d" % i]).items():
            worst[k] = max(worst.get(k, 0.0), float(v))
    assert max(worst.values()) <= 1e-6, worst


def test_oracle_free_run_reaches_the_reference_final_state(golden):
    """The whole 1760-tick drive without re-synchronisation (loop closure included): same map size, final state close."""
    g, recs = golden
    o = new_oracle()
    for rc in recs:
        o.sim(rc["V"], rc["G"], QE, rc["dt"], rc["phi"], rc["observe"], rc["z"], rc["idz"], None, RE, RE)
    x, P = o.get()
    assert len(x) == len(g["x_final"])
    e = block_relerr(x, P, g["x_final"], g["P_final"])
    assert max(e.values()) <= 1e-3, e


def test_oracle_vs_live_reference_known_association_and_gates():
    """Live against the linked reference: known association (the by-value table quirk of ekfslamwrapper.cpp:80-83:
    every sighting augments) and the nis / nd gate quantities of single pairs."""
    from oracle.pyekf import EkfRef
    if not EkfRef.available():
        pytest.skip("oracle/_ref/libslamref.so not built here")
    lm, recs = ekf_records(seed=9, n_landmarks=10, n_ticks=200)
    for known in (True, False):
        o = new_oracle(association_known=known)
        r = EkfRef(WHEELBASE, SIGMA_T, GATE_REJECT, GATE_AUGMENT, association_known=known)
        table = -np.ones(len(lm), np.int32)
        for i, rc in enumerate(recs):
            a = r.sim(rc["V"], rc["G"], QE, rc["dt"], rc["phi"], rc["observe"], rc["z"], rc["idz"], table.copy(), RE, RE)
            b = o.sim(rc["V"], rc["G"], QE, rc["dt"], rc["phi"], rc["observe"], rc["z"], rc["idz"], table.copy(), RE, RE)
            assert np.array_equal(a[0], b[0]) and a[1:] == b[1:], (known, i)
            xr, Pr = r.get()
            xo, Po = o.get()
            assert len(xr) == len(xo)
            assert max(block_relerr(xo, Po, xr, Pr).values()) <= 1e-4, (known, i)
            o.set(xr, Pr)
            if len(xr) > 3 and rc["observe"] and len(rc["z"]):
                for j in range(min(3, (len(xr) - 3) // 2)):
                    na, nb = r.association(rc["z"][0], RE, j), o.association(rc["z"][0], RE, j)
                    assert np.allclose(na, nb, rtol=1e-5, atol=1e-6), (na, nb)
