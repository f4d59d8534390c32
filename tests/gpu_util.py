"""Helpers for the -m gpu parity tests: move an oracle State to/from the CUDA backend."""
import numpy as np

import slam_b200
from scenario import SIGMA_PHI, WHEELBASE


def make_backend(method, P, nf_cap, use_heading=True, n_effective=None, scan_mode=slam_b200.SCAN_REFERENCE,
                 resample=True, spare_rows=0, math_mode=slam_b200.MATH_FAST):
    return slam_b200.FastSLAM(method, P, nf_cap, wheel_base=WHEELBASE, sigma_phi=SIGMA_PHI, n_effective=n_effective,
                              use_heading=use_heading, resample=resample, scan_mode=scan_mode, spare_rows=spare_rows,
                              math_mode=math_mode)


def upload_state(f, s):
    f.upload(s.xv, s.Pv, s.w, s.lm_mean[:s.Nf] if s.Nf else None, s.lm_cov[:s.Nf] if s.Nf else None, s.Nf)


def upload_snapshot(f, o):
    Nf = int(o["Nf"])
    f.upload(o["xv"], o["Pv"], o["w"], o["lm_mean"] if Nf else None, o["lm_cov"] if Nf else None, Nf)


def gpu_snapshot(f):
    d = f.download()
    if "lm_mean" not in d or d["Nf"] == 0:
        d["lm_mean"] = np.zeros((0, f.P, 2), np.float32)
        d["lm_cov"] = np.zeros((0, f.P, 4), np.float32)
    return d


def followed_iterations(a, g, tol=1e-4):
    """Rows over which trace `a` stays within tol of trace `g` (columns 4:6 estimate x, y relative to max(1, |xy|), column
    6 heading, absolute, modulo 2 pi), counted from the first row up to the first violation."""
    scale = np.maximum(1.0, np.abs(g[:, 4:6]).max(axis=1))
    err = np.maximum(np.abs(a[:, 4:6] - g[:, 4:6]).max(axis=1) / scale,
                     np.abs((a[:, 6] - g[:, 6] + np.pi) % (2 * np.pi) - np.pi))
    bad = np.nonzero(~(err <= tol))[0]
    n = int(bad[0]) if bad.size else len(err)
    return n, (float(err[:n].max()) if n else 0.0)
