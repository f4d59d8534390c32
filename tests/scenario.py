"""Seeded scenario generators shared by the parity tests and the golden-vector script.

Nothing here touches the GPU or the reference; it only produces inputs (numpy) in the flat layout
of include/slamgpu.h.  Names follow the reference: zf/idf = observations of known landmarks and
their landmark slots, zn = observations of new landmarks, R/Q = observation/control noise.
"""
import numpy as np

RAND_MAX = 2147483647

Q_DEFAULT = np.array([[0.3 ** 2, 0], [0, (3.0 * np.pi / 180) ** 2]], np.float32)   # core.cpp:977-995 defaults
R_DEFAULT = np.array([[0.1 ** 2, 0], [0, (1.0 * np.pi / 180) ** 2]], np.float32)
WHEELBASE = 4.0
SIGMA_PHI = float(np.float32(1.0 * np.pi / 180))  # sigmaT default (core.cpp:999)
DT = 0.025


def relerr(a, b):
    """Norm-relative error of one object: ||a-b||inf / max(||b||inf, tiny)."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    if a.size == 0:
        return 0.0
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-30))


def per_object_relerr(a, b, obj_axes):
    """Max over objects of the norm-relative error; obj_axes = axes that span ONE object."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    if a.size == 0:
        return 0.0
    num = np.abs(a - b).max(axis=obj_axes)
    den = np.maximum(np.abs(b).max(axis=obj_axes), 1e-30)
    return float((num / den).max())


def angle_diff(a, b):
    d = np.asarray(a, np.float64) - np.asarray(b, np.float64)
    return np.abs((d + np.pi) % (2 * np.pi) - np.pi)


class Drive:
    """A small deterministic drive through a random landmark field, producing per-tick inputs.

    Known data association as in the reference wrappers (fastslam2wrapper.cpp:84-88): landmark slot ==
    order of first sighting, shared by every particle.
    """

    def __init__(self, seed, n_landmarks=30, extent=30.0, max_obs=6):
        self.rng = np.random.default_rng(seed)
        self.lm = self.rng.uniform(-extent, extent, (n_landmarks, 2)).astype(np.float32)
        self.xtrue = np.zeros(3, np.float64)
        self.table = -np.ones(n_landmarks, np.int64)  # dataAssociationTable, core.cpp:91-120
        self.Nf = 0
        self.max_obs = max_obs
        self.step = 0

    def control(self):
        V = 3.0
        G = 0.08 * np.sin(0.7 * self.step)
        x = self.xtrue
        x[0] += V * DT * np.cos(G + x[2])
        x[1] += V * DT * np.sin(G + x[2])
        x[2] += V * DT * np.sin(G) / WHEELBASE
        return np.float32(V), np.float32(G)

    def observe(self, max_new=2):
        """Returns zf (K,2), idf (K,), zn (Kn,2) like dataAssociationKnown."""
        self.step += 1
        d = self.lm - self.xtrue[:2].astype(np.float32)
        order = np.argsort(np.hypot(d[:, 0], d[:, 1]))[: self.max_obs]
        zf, idf, zn = [], [], []
        new = 0
        for i in order:
            rng_, brg = np.hypot(*d[i]), np.arctan2(d[i, 1], d[i, 0]) - self.xtrue[2]
            z = [rng_ + self.rng.normal(0, 0.1), brg + self.rng.normal(0, np.pi / 180)]
            if self.table[i] >= 0:
                zf.append(z)
                idf.append(self.table[i])
            elif new < max_new:
                zn.append(z)
                self.table[i] = self.Nf + new
                new += 1
        self.Nf += new
        return (np.array(zf, np.float32).reshape(-1, 2), np.array(idf, np.int32),
                np.array(zn, np.float32).reshape(-1, 2))

    def ints(self, n):
        return self.rng.integers(1, RAND_MAX, size=n).astype(np.int32)


def synthetic_state(P, Nf, K, seed=1, extent=100.0, nf_cap=None):
    """cfg-4 style synthetic state, vectorised numpy.  Returns dict of flat arrays + obs.

    Map: Nf landmarks uniform in [-extent, extent]^2; true pose (0,0,0); particle poses = true pose +
    N(0, diag(0.05^2, 0.05^2, (0.5deg)^2)); Pv = a fixed SPD matrix typical of 8 predict ticks;
    landmark means = truth + N(0, 0.2^2 I); cov = Gz R Gz^T of addFeature (core.cpp:498-500).
    """
    rng = np.random.default_rng(seed)
    nf_cap = nf_cap or Nf
    lm = rng.uniform(-extent, extent, (Nf, 2)).astype(np.float32)
    xv = np.zeros((3, P), np.float32)
    xv[0] = rng.normal(0, 0.05, P)
    xv[1] = rng.normal(0, 0.05, P)
    xv[2] = rng.normal(0, 0.5 * np.pi / 180, P)
    A = np.array([[0.02, 0.001, 0.0], [0.001, 0.03, 0.002], [0.0, 0.002, 0.004]], np.float64)
    Pv0 = (A @ A.T + 1e-6 * np.eye(3)).astype(np.float32)
    Pv = np.repeat(Pv0.reshape(9, 1), P, axis=1).astype(np.float32)
    Pv *= (1.0 + 0.1 * rng.random((1, P), np.float32))
    w = np.full(P, np.float32(1.0) / np.float32(P), np.float32)
    lm_mean = np.empty((nf_cap, P, 2), np.float32)
    lm_cov = np.empty((nf_cap, P, 4), np.float32)
    R = R_DEFAULT
    for l in range(Nf):
        lm_mean[l] = lm[l] + rng.normal(0, 0.2, (P, 2)).astype(np.float32)
        r = float(np.hypot(*lm[l]))
        b = float(np.arctan2(lm[l, 1], lm[l, 0]))
        Gz = np.array([[np.cos(b), -r * np.sin(b)], [np.sin(b), r * np.cos(b)]])
        C = Gz @ R.astype(np.float64) @ Gz.T
        lm_cov[l] = np.array([C[0, 0], C[1, 0], C[0, 1], C[1, 1]], np.float32)
    lm_mean[Nf:] = 0
    lm_cov[Nf:] = 0
    order = np.argsort(np.hypot(lm[:, 0], lm[:, 1]))[:K]
    idf = order.astype(np.int32)
    zf = np.stack([np.hypot(lm[order, 0], lm[order, 1]) + rng.normal(0, 0.1, K),
                   np.arctan2(lm[order, 1], lm[order, 0]) + rng.normal(0, np.pi / 180, K)], 1).astype(np.float32)
    return dict(xv=xv, Pv=Pv, w=w, lm_mean=lm_mean, lm_cov=lm_cov, zf=zf, idf=idf, lm=lm, R=R)
