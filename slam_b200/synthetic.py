"""Synthetic FastSLAM2 workloads (BASELINE.json configs 4 and 5): inputs only, numpy, no device code.

`synthetic_state` builds a particle set in the flat layout of include/slamgpu.h; `SyntheticDrive` produces the
per-step host inputs of the reference's wrapper loop (controls for 8 predict ticks per observation,
fastslam2wrapper.cpp:51-69, then the K nearest landmarks as known-association observations).
"""
import numpy as np

RAND_MAX = 2147483647
WHEELBASE = 4.0                                  # core.cpp:977-995 defaults
DT_CONTROLS = 0.025
TICKS_PER_OBSERVE = 8                            # DT_OBSERVE / DT_CONTROLS = 0.2 / 0.025 in the shipped .ini files
Q_DEFAULT = np.array([[0.3 ** 2, 0], [0, (3.0 * np.pi / 180) ** 2]], np.float32)
SIGMA_PHI = float(np.float32(1.0 * np.pi / 180))

# Observation noise of the synthetic K = 32 workload.  With the reference defaults (0.1 m, 1 deg) every
# observation multiplies the fp32 weight by ~17 on average (1/(2*pi*sqrt(det 2R)) * e^-1), so the reference's own
# unnormalised weight product overflows to inf at K = 32 (measured on the oracle: 33 of 256 particles).  The synthetic
# workload therefore uses the noise level at which that factor is O(1); the arithmetic per update is unchanged.
SYNTH_SIGMA_R = 0.5
SYNTH_SIGMA_B = 5.0 * np.pi / 180


def synthetic_state(P, Nf, K, seed=1, extent=100.0, nf_cap=None, sigma_r=SYNTH_SIGMA_R, sigma_b=SYNTH_SIGMA_B,
                    particle_spread=0.1, map_seed=None):
    """Returns dict of flat arrays + one observation set.

    Map: Nf landmarks uniform in [-extent, extent]^2; true pose (0,0,0); particle poses = true pose +
    N(0, diag(0.05^2, 0.05^2, (0.5deg)^2)); Pv = a fixed SPD matrix typical of 8 predict ticks;
    landmark cov = Gz R Gz^T of addFeature (core.cpp:498-500), landmark means = truth + N(0, cov) common to
    all particles + particle_spread * N(0, cov) per particle.

    map_seed: when given, the map and the common map error come from their own generator, so that shards built with
    different `seed`s (different particles per rank) still share ONE world and hence one observation sequence.
    """
    rng = np.random.default_rng(seed)
    rng_map = rng if map_seed is None else np.random.default_rng(map_seed)
    nf_cap = nf_cap or Nf
    lm = rng_map.uniform(-extent, extent, (Nf, 2)).astype(np.float32)
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
    R = np.array([[sigma_r ** 2, 0], [0, sigma_b ** 2]], np.float32)
    for l in range(Nf):
        r = float(np.hypot(*lm[l]))
        b = float(np.arctan2(lm[l, 1], lm[l, 0]))
        Gz = np.array([[np.cos(b), -r * np.sin(b)], [np.sin(b), r * np.cos(b)]])
        C = Gz @ R.astype(np.float64) @ Gz.T
        L = np.linalg.cholesky(C + 1e-12 * np.eye(2))
        # map error mostly common to all particles (it is correlated through the shared trajectory), plus a
        # smaller per-particle part: keeps the weights of a K = 32 update from collapsing onto a few particles
        common = rng_map.standard_normal(2) @ L.T
        lm_mean[l] = lm[l] + (common + particle_spread * (rng.standard_normal((P, 2)) @ L.T)).astype(np.float32)
        lm_cov[l] = np.array([C[0, 0], C[1, 0], C[0, 1], C[1, 1]], np.float32)
    lm_mean[Nf:] = 0
    lm_cov[Nf:] = 0
    order = np.argsort(np.hypot(lm[:, 0], lm[:, 1]))[:K]
    idf = order.astype(np.int32)
    zf = np.stack([np.hypot(lm[order, 0], lm[order, 1]) + rng.normal(0, sigma_r, K),
                   np.arctan2(lm[order, 1], lm[order, 0]) + rng.normal(0, sigma_b, K)], 1).astype(np.float32)
    return dict(xv=xv, Pv=Pv, w=w, lm_mean=lm_mean, lm_cov=lm_cov, zf=zf, idf=idf, lm=lm, R=R)


class SyntheticDrive:
    """True vehicle + observations for the synthetic workload: constant speed, gentle steering, the K landmarks
    nearest the true pose observed each step with known association (all already in the map: Kn = 0)."""

    def __init__(self, lm, K, R, seed=12345, V=3.0):
        self.lm = np.asarray(lm, np.float64)
        self.K = K
        self.sig = np.sqrt(np.diag(np.asarray(R, np.float64)))
        self.rng = np.random.default_rng(seed)
        self.x = np.zeros(3)
        self.V = V
        self.tick = 0

    def controls(self):
        """One control tick: returns (V, G, dt, phi_true) after moving the true vehicle (core.cpp:35-39)."""
        G = 0.05 * np.sin(0.01 * self.tick)
        self.tick += 1
        x = self.x
        x[0] += self.V * DT_CONTROLS * np.cos(G + x[2])
        x[1] += self.V * DT_CONTROLS * np.sin(G + x[2])
        x[2] += self.V * DT_CONTROLS * np.sin(G) / WHEELBASE
        x[2] = (x[2] + np.pi) % (2 * np.pi) - np.pi
        return float(self.V), float(G), DT_CONTROLS, float(np.float32(x[2]))

    def observe(self):
        """zf [K][2], idf [K] for the K nearest landmarks (getObservations + addObservationNoise, core.cpp:185-273,438-449)."""
        d = self.lm - self.x[:2]
        rng_ = np.hypot(d[:, 0], d[:, 1])
        order = np.argsort(rng_)[: self.K]
        brg = np.arctan2(d[order, 1], d[order, 0]) - self.x[2]
        brg = (brg + np.pi) % (2 * np.pi) - np.pi
        zf = np.stack([rng_[order] + self.rng.normal(0, self.sig[0], self.K),
                       brg + self.rng.normal(0, self.sig[1], self.K)], 1).astype(np.float32)
        return zf, order.astype(np.int32)
