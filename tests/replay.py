"""Record / replay of a short FastSLAM drive so that the reference, the oracle and the CUDA path
consume IDENTICAL inputs (controls, observations, pre-drawn Gaussians and resample uniforms).

A record list is plain numpy data (picklable into .npz); each backend replays it and returns the
state after every record.  The reference derives its noise from an injected rand() stream; the
records carry both the raw ints (for the reference) and the Gaussians / uniforms they turn into
(for the oracle and the CUDA path) - derived with the reference's own Box-Muller restated in the
oracle (validated bit-exact in test_oracle_vs_ref.py).
"""
import json
import os

import numpy as np

from scenario import DT, Q_DEFAULT, R_DEFAULT, RAND_MAX, SIGMA_PHI, WHEELBASE, Drive


CONTEXT = ""  # set by tests: which configuration the comparisons that follow belong to (parity log only)


def make_records(oracle, seed, method, P, n_observe=10, ticks_per_observe=8, use_heading=True,
                 predict_noise=None, n_landmarks=30, max_obs=6, n_effective=None):
    """Inputs only.  method 1/2.  predict_noise default: on for FastSLAM1 (fastslam1wrapper.cpp:20), off for 2."""
    if predict_noise is None:
        predict_noise = method == 1
    if n_effective is None:
        n_effective = int(0.75 * P)
    d = Drive(seed, n_landmarks=n_landmarks, max_obs=max_obs)
    recs = []
    for step in range(n_observe):
        for t in range(ticks_per_observe):
            V, G = d.control()
            rec = dict(kind="predict", V=float(V), G=float(G), dt=DT, phi_true=float(np.float32(d.xtrue[2])),
                       use_heading=bool(use_heading), ints=None, gauss=None)
            if predict_noise:
                ints = d.ints(3 * P)  # randn(2,1) consumes 3 rand() per particle (core.cpp:386)
                rec["ints"] = ints
                rec["gauss"] = np.stack([oracle.randn(ints[3 * p:3 * p + 3], 2) for p in range(P)], 1)
            recs.append(rec)
        zf, idf, zn = d.observe(max_new=2 if step % 2 == 0 else 1)
        if step == 3:  # exercise the "no known landmark, new ones only" branch (fastslam2.cpp:36-42)
            pass
        K, Kn = zf.shape[0], zn.shape[0]
        draws = method == 2 and (K > 0 or Kn > 0)
        gints = d.ints(4 * P) if draws else np.zeros(0, np.int32)
        uints = d.rng.integers(0, RAND_MAX, size=P).astype(np.int32)
        gauss = (np.stack([oracle.randn(gints[4 * p:4 * p + 4], 3) for p in range(P)], 1) if draws
                 else np.zeros((3, P), np.float32))
        recs.append(dict(kind="update", zf=zf, idf=idf, zn=zn, ints=np.concatenate([gints, uints]), gauss=gauss,
                         unif=uints / float(RAND_MAX), n_effective=n_effective, do_resample=True))
    return recs


def snapshot(s):
    return dict(xv=s.xv.copy(), Pv=s.Pv.copy(), w=s.w.copy(), lm_mean=s.lm_mean[:s.Nf].copy(),
                lm_cov=s.lm_cov[:s.Nf].copy(), Nf=s.Nf)


def run_ref(ref, recs, method, P, nf_cap):
    from oracle.pyoracle import State
    s = State(P, nf_cap)
    h = ref.create(s)
    outs = []
    for r in recs:
        if r["kind"] == "predict":
            ref.set_stream(r["ints"] if r["ints"] is not None else np.zeros(0, np.int32))
            ref.predict(method, h, r["V"], r["G"], Q_DEFAULT, r["dt"], r["phi_true"], WHEELBASE, SIGMA_PHI,
                        r["use_heading"], r["ints"] is not None)
        else:
            ref.set_stream(r["ints"])
            ref.update(method, h, r["zf"], r["idf"], r["zn"], R_DEFAULT, r["n_effective"], r["do_resample"])
        assert ref.underflow() == 0, "rand() stream was mis-sized"
        ref.download(h, s)
        outs.append(snapshot(s))
    ref.destroy(h)
    return outs


def run_oracle(oracle, recs, method, P, nf_cap, scan_mode=0):
    from oracle.pyoracle import State
    s = State(P, nf_cap)
    outs = []
    for r in recs:
        if r["kind"] == "predict":
            oracle.predict(method, s, r["V"], r["G"], Q_DEFAULT, r["dt"], r["phi_true"], WHEELBASE, SIGMA_PHI,
                           r["use_heading"], r["gauss"])
            outs.append(snapshot(s))
        else:
            oracle.update_particles(method, s, r["zf"], r["idf"], r["zn"], R_DEFAULT, r["gauss"])
            keep, neff, did = oracle.resample(s, r["unif"], r["n_effective"], r["do_resample"], scan_mode)
            o = snapshot(s)
            o.update(keep=keep, neff=neff, did=did)
            outs.append(o)
    return outs


def save_golden(path, recs, outs, meta):
    flat = {}
    for i, (r, o) in enumerate(zip(recs, outs)):
        for k, v in r.items():
            if v is None:
                continue
            flat["r%d_%s" % (i, k)] = np.asarray(v)
        for k, v in o.items():
            flat["o%d_%s" % (i, k)] = np.asarray(v)
    flat["n"] = np.asarray(len(recs))
    for k, v in meta.items():
        flat["meta_" + k] = np.asarray(v)
    np.savez_compressed(path, **flat)


def load_golden(path):
    z = np.load(path, allow_pickle=False)
    n = int(z["n"])
    recs, outs = [], []
    for i in range(n):
        r, o = {}, {}
        for k in z.files:
            if k.startswith("r%d_" % i):
                r[k[len("r%d_" % i):]] = z[k]
            elif k.startswith("o%d_" % i):
                o[k[len("o%d_" % i):]] = z[k]
        r["kind"] = str(r["kind"])
        for k in ("V", "G", "dt", "phi_true"):
            if k in r:
                r[k] = float(r[k])
        for k in ("use_heading", "do_resample"):
            if k in r:
                r[k] = bool(r[k])
        if "n_effective" in r:
            r["n_effective"] = int(r["n_effective"])
        r.setdefault("ints", None)
        r.setdefault("gauss", None)
        o["Nf"] = int(o["Nf"])
        recs.append(r)
        outs.append(o)
    meta = {k[5:]: z[k] for k in z.files if k.startswith("meta_")}
    return recs, outs, meta


def compare_states(a, b, tol, what="", tol_w=None):
    """a (candidate) vs b (reference) per SURVEY 8c: norm-relative per object, heading modulo 2*pi."""
    from scenario import angle_diff, per_object_relerr
    assert a["Nf"] == b["Nf"], what
    errs = {}
    # poses: one object per particle = (x, y); heading separately, modulo 2*pi
    errs["xy"] = per_object_relerr(a["xv"][:2], b["xv"][:2], 0)
    errs["theta"] = float(angle_diff(a["xv"][2], b["xv"][2]).max())
    errs["Pv"] = per_object_relerr(a["Pv"], b["Pv"], 0)
    errs["w"] = per_object_relerr(a["w"], b["w"], None) if a["w"].size else 0.0
    if a["Nf"]:
        errs["lm_mean"] = per_object_relerr(a["lm_mean"], b["lm_mean"], 2)
        errs["lm_cov"] = per_object_relerr(a["lm_cov"], b["lm_cov"], 2)
    log = os.environ.get("SLAM_PARITY_LOG")
    if log:  # achieved errors, one line per comparison (how the per-mode bounds in the tests were set)
        with open(log, "a") as fh:
            fh.write(json.dumps({"context": CONTEXT, "what": what, "errs": errs}) + "\n")
    bad = {k: v for k, v in errs.items() if not v <= (tol_w if (k == "w" and tol_w) else tol)}
    assert not bad, "%s: %s (tol %g)" % (what, bad, tol)
    return errs
