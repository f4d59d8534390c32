"""Seeded EKF-SLAM drive shared by the EKF parity tests and the golden-vector script: per-tick inputs of
EKFSLAM::sim (ekfslam.cpp:18-43) as wrappers/ekfslamwrapper.cpp:47-83 produces them (noisy controls every
DT_CONTROLS, range/bearing of the landmarks within MAX_RANGE every DT_OBSERVE, a noisy compass).  numpy only."""
import numpy as np

DT = 0.025
WHEELBASE = 1.0                       # example_loop902.ini
SIGMA_V, SIGMA_G = 0.3, 3.0 * np.pi / 180
SIGMA_R, SIGMA_B = 0.1, 1.0 * np.pi / 180
SIGMA_T = float(np.float32(1.0 * np.pi / 180))
GATE_REJECT, GATE_AUGMENT = 4.0, 25.0
MAX_RANGE = 10.0
QE = np.array([[SIGMA_V ** 2, 0], [0, SIGMA_G ** 2]], np.float32)
RE = np.array([[SIGMA_R ** 2, 0], [0, SIGMA_B ** 2]], np.float32)


def ekf_records(seed=3, n_landmarks=24, radius=12.0, n_ticks=480, observe_every=8, max_range=MAX_RANGE):
    """Vehicle on a circle of the given radius through a ring of landmarks; returns (landmarks, records)."""
    rng = np.random.default_rng(seed)
    ang = np.sort(rng.uniform(0, 2 * np.pi, n_landmarks))
    rad = radius + rng.uniform(-4.0, 4.0, n_landmarks)
    lm = np.stack([rad * np.cos(ang), rad * np.sin(ang) + radius], 1)   # circle centred at (0, radius)
    xt = np.zeros(3)
    V = 2.0
    G = float(np.arctan(WHEELBASE / radius))   # steady left turn of that radius
    recs = []
    for t in range(n_ticks):
        xt[0] += V * DT * np.cos(G + xt[2])
        xt[1] += V * DT * np.sin(G + xt[2])
        xt[2] = (xt[2] + V * DT * np.sin(G) / WHEELBASE + np.pi) % (2 * np.pi) - np.pi
        Vn = float(np.float32(V + rng.normal(0, SIGMA_V)))
        Gn = float(np.float32(G + rng.normal(0, SIGMA_G)))
        phi = float(np.float32(xt[2] + SIGMA_T * rng.random()))
        rec = dict(V=Vn, G=Gn, dt=DT, phi=phi, observe=False, z=np.zeros((0, 2), np.float32),
                   idz=np.zeros(0, np.int32))
        if (t + 1) % observe_every == 0:
            d = lm - xt[:2]
            r = np.hypot(d[:, 0], d[:, 1])
            b = np.arctan2(d[:, 1], d[:, 0]) - xt[2]
            vis = np.nonzero((r < max_range))[0]
            z = np.stack([r[vis] + rng.normal(0, SIGMA_R, len(vis)),
                          (b[vis] + rng.normal(0, SIGMA_B, len(vis)) + np.pi) % (2 * np.pi) - np.pi], 1)
            rec.update(observe=True, z=z.astype(np.float32), idz=vis.astype(np.int32))
        recs.append(rec)
    return lm.astype(np.float32), recs


def block_relerr(x, P, xr, Pr):
    """SURVEY 8c tolerance form for the joint state: norm-relative error of the pose, of each landmark 2-vector
    (heading modulo 2 pi) and of P as one object."""
    x, P, xr, Pr = (np.asarray(a, np.float64) for a in (x, P, xr, Pr))
    e = {}
    e["xy"] = np.abs(x[:2] - xr[:2]).max() / max(np.abs(xr[:2]).max(), 1.0)
    e["theta"] = abs((x[2] - xr[2] + np.pi) % (2 * np.pi) - np.pi)
    if len(x) > 3:
        a, b = x[3:].reshape(-1, 2), xr[3:].reshape(-1, 2)
        e["lm"] = (np.abs(a - b).max(1) / np.maximum(np.abs(b).max(1), 1e-30)).max()
    e["P"] = np.abs(P - Pr).max() / max(np.abs(Pr).max(), 1e-30)
    return e
