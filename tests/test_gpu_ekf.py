"""GPU: the EKF-SLAM entry point (slamgpu_ekf_*, K9 kernels) through the C ABI against
  * the golden drive produced by the UNMODIFIED reference (tests/golden/ekf_drive.npz): every stored EKFSLAM::sim call
    replayed from the reference's own state — association bit-exact, x / P within 1e-4 norm-relative;
  * the CPU oracle (oracle/ekf_oracle.c) on the free-running drive, loop closure included.
"""
import os

import numpy as np
import pytest

from ekf_scenario import GATE_AUGMENT, GATE_REJECT, QE, RE, SIGMA_T, WHEELBASE, block_relerr, ekf_records

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))
TOL = 1e-4


@pytest.fixture(scope="module")
def sb():
    import slam_b200
    slam_b200.build()
    return slam_b200


@pytest.fixture(scope="module")
def golden():
    g = np.load(os.path.join(HERE, "golden", "ekf_drive.npz"))
    lm, recs = ekf_records(seed=int(g["seed"]), n_landmarks=int(g["n_landmarks"]), n_ticks=int(g["n_ticks"]))
    return g, recs


def test_every_stored_reference_call(sb, golden):
    g, recs = golden
    f = sb.EKFSLAM(32, WHEELBASE, SIGMA_T, GATE_REJECT, GATE_AUGMENT)
    worst = {}
    n_assoc = 0
    for i in g["ticks"]:
        rc = recs[int(i)]
        f.set(g["x0_%d" % i], g["P0_%d" % i])
        idf, nzf, nzn = f.sim(rc["V"], rc["G"], QE, rc["dt"], rc["phi"], rc["observe"], rc["z"], rc["idz"], None, RE, RE)
        assert np.array_equal(idf, g["idf_%d" % i]) and [nzf, nzn] == list(g["nz_%d" % i]), i
        n_assoc += nzf
        x, P = f.get()
        assert len(x) == len(g["x1_%d" % i])
        for k, v in block_relerr(x, P, g["x1_%d" % i], g["P1_%d" % i]).items():
            worst[k] = max(worst.get(k, 0.0), float(v))
    assert n_assoc > 300 and f.launches > 0
    assert max(worst.values()) <= TOL, worst


def test_free_running_drive_against_oracle_and_reference_final_state(sb, golden):
    """1760 ticks without re-synchronisation: every association equals the oracle's, the map ends at the reference's
    size and the final state is within 1e-3 of the reference's (accumulated rounding of a 220-update chain)."""
    from oracle.pyekf import EkfOracle
    g, recs = golden
    f = sb.EKFSLAM(32, WHEELBASE, SIGMA_T, GATE_REJECT, GATE_AUGMENT)
    o = EkfOracle(WHEELBASE, SIGMA_T, GATE_REJECT, GATE_AUGMENT)
    for i, rc in enumerate(recs):
        a = f.sim(rc["V"], rc["G"], QE, rc["dt"], rc["phi"], rc["observe"], rc["z"], rc["idz"], None, RE, RE)
        b = o.sim(rc["V"], rc["G"], QE, rc["dt"], rc["phi"], rc["observe"], rc["z"], rc["idz"], None, RE, RE)
        assert np.array_equal(a[0], b[0]) and a[1:] == b[1:], i
    x, P = f.get()
    xo, Po = o.get()
    assert len(x) == len(xo) == len(g["x_final"])
    assert max(block_relerr(x, P, xo, Po).values()) <= 1e-3
    assert max(block_relerr(x, P, g["x_final"], g["P_final"]).values()) <= 1e-3
    xv, Pvv = f.pose()
    assert np.array_equal(xv, x[:3]) and np.array_equal(Pvv, P[:3, :3])


@pytest.mark.parametrize("heading,batch", [(True, True), (False, True), (True, False)])
def test_known_association_and_switches(sb, heading, batch):
    """SWITCH_ASSOCIATION_KNOWN with the wrapper's by-value table (never updated: every sighting augments,
    ekfslamwrapper.cpp:80-83), a table that does know some landmarks, and the heading / batch-update switches."""
    from oracle.pyekf import EkfOracle
    lm, recs = ekf_records(seed=9, n_landmarks=10, n_ticks=160)
    f = sb.EKFSLAM(64, WHEELBASE, SIGMA_T, GATE_REJECT, GATE_AUGMENT, use_heading=heading, batch_update=batch,
                   association_known=True)
    o = EkfOracle(WHEELBASE, SIGMA_T, GATE_REJECT, GATE_AUGMENT, use_heading=heading, batch_update=batch,
                  association_known=True)
    table = -np.ones(len(lm), np.int32)
    first_seen = {}
    for i, rc in enumerate(recs):
        if i >= 80:  # second half: a table that maps every landmark to the slot of its first sighting
            for k, v in first_seen.items():
                table[k] = v
        n_before = o.n
        a = f.sim(rc["V"], rc["G"], QE, rc["dt"], rc["phi"], rc["observe"], rc["z"], rc["idz"], table.copy(), RE, RE)
        b = o.sim(rc["V"], rc["G"], QE, rc["dt"], rc["phi"], rc["observe"], rc["z"], rc["idz"], table.copy(), RE, RE)
        assert np.array_equal(a[0], b[0]) and a[1:] == b[1:], i
        if rc["observe"] and i < 80:
            new = [int(k) for k in rc["idz"] if table[k] == -1]
            for j, k in enumerate(new):
                first_seen.setdefault(k, (n_before - 3) // 2 + j)
        x, P = f.get()
        xo, Po = o.get()
        assert len(x) == len(xo)
        assert max(block_relerr(x, P, xo, Po).values()) <= TOL, i
        o.set(x, P)
    assert f.n > 3


def test_capacity_and_argument_errors(sb):
    f = sb.EKFSLAM(2, WHEELBASE, SIGMA_T, GATE_REJECT, GATE_AUGMENT, m_cap=4)
    z = np.array([[5.0, 0.1], [6.0, -0.4], [7.0, 1.0]], np.float32)
    with pytest.raises(sb.SlamGpuError) as e:  # three new landmarks, room for two
        f.sim(1.0, 0.0, QE, 0.025, 0.0, True, z, None, None, RE, RE)
    assert e.value.code == -5
    with pytest.raises(sb.SlamGpuError):       # more observations than m_cap
        f.sim(1.0, 0.0, QE, 0.025, 0.0, True, np.zeros((5, 2), np.float32) + 3, None, None, RE, RE)
    idf, nzf, nzn = f.sim(1.0, 0.0, QE, 0.025, 0.0, True, z[:2], None, None, RE, RE)
    assert (nzf, nzn) == (0, 2) and f.n == 7
