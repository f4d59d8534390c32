"""CPU: the product's per-particle arithmetic (slam_b200/csrc/slam_math.cuh, the functions the CUDA kernels call)
compiled for the host by tests/host_emul and diffed against the oracle.  This catches logic errors in the device
math without a GPU; the kernels themselves are checked on the B200 by the -m gpu tests."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

from replay import compare_states, make_records, snapshot
from scenario import DT, Q_DEFAULT, R_DEFAULT, SIGMA_PHI, WHEELBASE, synthetic_state

HERE = os.path.dirname(os.path.abspath(__file__))
f32p = np.ctypeslib.ndpointer(np.float32, flags="C_CONTIGUOUS")
i32p = np.ctypeslib.ndpointer(np.int32, flags="C_CONTIGUOUS")


@pytest.fixture(scope="module")
def emul():
    d = os.path.join(HERE, "host_emul")
    so = os.path.join(d, "libhostemul.so")
    src = os.path.join(d, "host_emul.cpp")
    hdr = os.path.join(HERE, "..", "slam_b200", "csrc", "slam_math.cuh")
    if (not os.path.exists(so)) or os.path.getmtime(so) < max(os.path.getmtime(src), os.path.getmtime(hdr)):
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-msse2", "-mfpmath=sse", "-ffp-contract=off", "-fPIC",
                               "-shared", "-o", so, src, "-lm"])
    L = C.CDLL(so)
    L.he_trig_offset.restype = C.c_float
    L.he_trig_offset.argtypes = [C.c_float]
    L.he_compute_jacobians.argtypes = [f32p, C.c_int, f32p]
    L.he_predict.argtypes = [C.c_int, C.c_long, f32p, f32p, C.c_float, C.c_float, f32p, C.c_float, C.c_float,
                             C.c_float, C.c_float, C.c_int, C.c_void_p]
    L.he_fs2_update_particles.restype = C.c_int
    L.he_fs2_update_particles.argtypes = [C.c_long, C.c_int, f32p, f32p, f32p, f32p, f32p, f32p, i32p, C.c_int, f32p,
                                          C.c_int, f32p, f32p]
    L.he_fs1_update_particles.restype = C.c_int
    L.he_fs1_update_particles.argtypes = [C.c_long, C.c_int, f32p, f32p, f32p, f32p, f32p, f32p, i32p, C.c_int, f32p,
                                          C.c_int, f32p]
    L.he_predict_fast.argtypes = [C.c_int, C.c_long, f32p, f32p, C.c_float, C.c_float, f32p, C.c_float, C.c_float,
                                  C.c_float, C.c_float, C.c_int]
    L.he_fs2_update_particles_fast.restype = C.c_int
    L.he_fs2_update_particles_fast.argtypes = [C.c_long, C.c_int, f32p, f32p, f32p, f32p, f32p, f32p, i32p, C.c_int,
                                               f32p, f32p]
    return L


def _f(a):
    return np.ascontiguousarray(a, np.float32)


def emul_update(L, method, s, zf, idf, zn, gauss, R=R_DEFAULT):
    zf = _f(zf).reshape(-1, 2)
    zn = _f(zn).reshape(-1, 2)
    K, Kn = zf.shape[0], zn.shape[0]
    zf_ = zf if K else np.zeros((1, 2), np.float32)
    zn_ = zn if Kn else np.zeros((1, 2), np.float32)
    idf_ = np.ascontiguousarray(idf, np.int32) if K else np.zeros(1, np.int32)
    R = _f(R).ravel()
    if method == 2:
        s.Nf = L.he_fs2_update_particles(s.P, s.Nf, s.xv, s.Pv, s.w, s.lm_mean, s.lm_cov, zf_, idf_, K, zn_, Kn, R,
                                         _f(gauss))
    else:
        s.Nf = L.he_fs1_update_particles(s.P, s.Nf, s.xv, s.Pv, s.w, s.lm_mean, s.lm_cov, zf_, idf_, K, zn_, Kn, R)


def test_trig_offset_matches_oracle_bitwise(emul, oracle):
    rng = np.random.default_rng(0)
    vals = np.concatenate([rng.uniform(-20, 20, 5000), [np.pi, -np.pi, 2 * np.pi, -2 * np.pi, 0.0, 3.1415925, 3.1415927,
                                                        6.2831850, 6.2831855, -6.2831855]]).astype(np.float32)
    for a in vals:
        assert emul.he_trig_offset(float(a)) == oracle.trig_offset(a), a


def test_jacobians_kat(emul):
    z = np.load(os.path.join(HERE, "golden", "jacobians_kat.npz"))
    n = int(z["n"])
    for rec, out in zip(z["rec"], z["out"]):
        got = np.zeros(16 * n, np.float32)
        emul.he_compute_jacobians(_f(rec), n, got)
        np.testing.assert_allclose(got, out, rtol=2e-6, atol=1e-7)


@pytest.mark.parametrize("method,heading,noise", [(1, True, True), (2, True, False), (2, False, True)])
def test_drive_matches_oracle(emul, oracle, method, heading, noise):
    """Same records through both; the oracle's resample keeps the two particle sets aligned."""
    from oracle.pyoracle import State
    P, cap = 64, 32
    recs = make_records(oracle, 40 + method, method, P, n_observe=8, use_heading=heading, predict_noise=noise)
    a, b = State(P, cap), State(P, cap)
    Q = _f(Q_DEFAULT).ravel()
    for i, r in enumerate(recs):
        if r["kind"] == "predict":
            g = None if r["gauss"] is None else _f(r["gauss"])
            oracle.predict(method, a, r["V"], r["G"], Q_DEFAULT, r["dt"], r["phi_true"], WHEELBASE, SIGMA_PHI,
                           r["use_heading"], r["gauss"])
            emul.he_predict(method, P, b.xv, b.Pv, r["V"], r["G"], Q, r["dt"], r["phi_true"], WHEELBASE, SIGMA_PHI,
                            int(r["use_heading"]), None if g is None else g.ctypes.data_as(C.c_void_p))
        else:
            oracle.update_particles(method, a, r["zf"], r["idf"], r["zn"], R_DEFAULT, r["gauss"])
            emul_update(emul, method, b, r["zf"], r["idf"], r["zn"], r["gauss"])
            compare_states(snapshot(b), snapshot(a), 1e-4, "record %d before resample" % i, tol_w=1e-4)
            ka, _, _ = oracle.resample(a, r["unif"], r["n_effective"], r["do_resample"])
            kb, _, _ = oracle.resample(b, r["unif"], r["n_effective"], r["do_resample"])
            assert np.array_equal(ka, kb)
        compare_states(snapshot(b), snapshot(a), 1e-4, "record %d (%s)" % (i, r["kind"]), tol_w=1e-4)


def test_synthetic_cfg4_shape_update(emul, oracle):
    """One FastSLAM2 update on the cfg-4 style synthetic state (32 observed landmarks), small P."""
    from oracle.pyoracle import State
    P, Nf, K = 256, 64, 32
    d = synthetic_state(P, Nf, K, seed=3)
    rng = np.random.default_rng(9)
    gauss = rng.standard_normal((3, P)).astype(np.float32)
    a, b = State(P, Nf), State(P, Nf)
    for s in (a, b):
        s.xv[:], s.Pv[:], s.w[:], s.lm_mean[:], s.lm_cov[:] = d["xv"], d["Pv"], d["w"], d["lm_mean"], d["lm_cov"]
        s.Nf = Nf
    oracle.update_particles(2, a, d["zf"], d["idf"], np.zeros((0, 2), np.float32), d["R"], gauss)
    emul_update(emul, 2, b, d["zf"], d["idf"], np.zeros((0, 2), np.float32), gauss, d["R"])
    wa, wb = a.w / a.w.sum(), b.w / b.w.sum()
    errs = compare_states(snapshot(b), snapshot(a), 1e-4, "cfg4-shape update", tol_w=1e9)
    assert np.abs(wa - wb).max() <= 1e-4 * wa.max(), errs


@pytest.mark.parametrize("method,heading", [(2, True), (2, False), (1, True)])
def test_fast_predict_ticks_match_oracle(emul, oracle, method, heading):
    """SLAMGPU_MATH_FAST predict (predict_tick_fast: sparse Gv / Joseph forms, host-side control terms) against the
    oracle's dense reference-order predict over 5 cycles of 8 noise-free ticks, steering through both signs."""
    from oracle.pyoracle import State
    P = 128
    rng = np.random.default_rng(21)
    a, b = State(P, 1), State(P, 1)
    a.xv[:] = rng.normal(0, 1, (3, P)).astype(np.float32)
    a.xv[2] = rng.uniform(-3.1, 3.1, P).astype(np.float32)
    b.xv[:] = a.xv
    Q = _f(Q_DEFAULT).ravel()
    phi = 0.3
    for tick in range(40):
        V, G = 3.0 + 0.1 * tick, 0.5 * np.sin(0.3 * tick)
        phi = float(np.float32(phi + V * DT * np.sin(G) / WHEELBASE))
        oracle.predict(method, a, V, G, Q_DEFAULT, DT, phi, WHEELBASE, SIGMA_PHI, heading, None)
        emul.he_predict_fast(method, P, b.xv, b.Pv, V, G, Q, DT, phi, WHEELBASE, SIGMA_PHI, int(heading))
        compare_states(snapshot(b), snapshot(a), 1e-4, "fast predict tick %d" % tick, tol_w=1e-4)


def test_fast_fs2_update_matches_oracle(emul, oracle):
    """SLAMGPU_MATH_FAST FastSLAM2 update (information-form proposal, shared factors, symmetric-part S) on the cfg-4
    style synthetic state against the oracle."""
    from oracle.pyoracle import State
    P, Nf, K = 256, 64, 32
    d = synthetic_state(P, Nf, K, seed=5)
    rng = np.random.default_rng(10)
    gauss = rng.standard_normal((3, P)).astype(np.float32)
    a, b = State(P, Nf), State(P, Nf)
    for s_ in (a, b):
        s_.xv[:], s_.Pv[:], s_.w[:], s_.lm_mean[:], s_.lm_cov[:] = d["xv"], d["Pv"], d["w"], d["lm_mean"], d["lm_cov"]
        s_.Nf = Nf
    oracle.update_particles(2, a, d["zf"], d["idf"], np.zeros((0, 2), np.float32), d["R"], gauss)
    emul.he_fs2_update_particles_fast(P, Nf, b.xv, b.Pv, b.w, b.lm_mean, b.lm_cov, _f(d["zf"]),
                                      np.ascontiguousarray(d["idf"], np.int32), K, _f(d["R"]).ravel(), gauss)
    wa, wb = a.w / a.w.sum(), b.w / b.w.sum()
    errs = compare_states(snapshot(b), snapshot(a), 1e-4, "fast cfg4-shape update", tol_w=1e9)
    assert np.abs(wa - wb).max() <= 2e-4 * wa.max(), errs
