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


from slam_b200.synthetic import SYNTH_SIGMA_B, SYNTH_SIGMA_R, synthetic_state  # noqa: E402,F401  (shared with bench.py)
