"""Secondary line of SURVEY.md 8(d): FastSLAM1 on the bench's synthetic inputs (2^20 particles x 1024 landmarks x 32
observations per step, 8 predict ticks with per-particle control noise per update, one GPU).  Prints one JSON line.
Noise is staged in HBM once ([2][P] Gaussians for the predict ticks, P uniforms for the resample) and reused every
step: a timing run, not a filter run.      python tools/bench_fs1.py [--steps 20 --warmup 3]"""
import argparse
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from slam_b200.synthetic import (Q_DEFAULT, SIGMA_PHI, TICKS_PER_OBSERVE, WHEELBASE, SyntheticDrive,  # noqa: E402
                                 synthetic_state)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--particles", type=int, default=1 << 20)
    ap.add_argument("--landmarks", type=int, default=1024)
    ap.add_argument("--obs", type=int, default=32)
    a = ap.parse_args()
    import torch

    import slam_b200
    slam_b200.build()
    P, Nf, K = a.particles, a.landmarks, a.obs
    tile = min(4096, P)
    d = synthetic_state(tile, Nf, K, seed=1, map_seed=1)
    f = slam_b200.FastSLAM(slam_b200.FASTSLAM1, P, Nf, wheel_base=WHEELBASE, sigma_phi=SIGMA_PHI,
                           n_effective=int(0.75 * P), use_heading=True, scan_mode=slam_b200.SCAN_BLOCKED, device=0)
    f.upload_tiled(d["xv"], d["Pv"], np.full(tile, np.float32(1.0 / P), np.float32), d["lm_mean"], d["lm_cov"], Nf)
    stream = torch.cuda.Stream()
    f.set_stream(stream.cuda_stream)
    rng = np.random.default_rng(5)
    f.stage_noise(rng.standard_normal((2, P), dtype=np.float32), rng.random(P))
    drive = SyntheticDrive(d["lm"], K, d["R"], seed=12345)
    plan = []
    for _ in range(a.warmup + a.steps):
        ticks = [drive.controls() for _ in range(TICKS_PER_OBSERVE)]
        plan.append((ticks, drive.observe()))
    resamples = 0

    def one_step(k):
        ticks, (zf, idf) = plan[k]
        for (V, G, dt, phi) in ticks:
            f.predict(V, G, Q_DEFAULT, dt, phi)
        neff, did = f.update(zf, idf, np.zeros((0, 2), np.float32), d["R"])
        f.estimate()
        return int(did)

    for k in range(a.warmup):
        one_step(k)
    f.timing_enable(True)
    f.timing_reset()
    l0 = f.launches
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    with torch.cuda.stream(stream):
        e0.record(stream)
        for k in range(a.steps):
            resamples += one_step(a.warmup + k)
        e1.record(stream)
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.steps
    fam = {k: v["ms"] / a.steps for k, v in f.timing_read().items()}
    events = f.timing_read()["update"]["events"]
    peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"] if os.path.exists(
        os.path.join(ROOT, "MEASURED_PEAKS.json")) else 6552.0
    upd_ms = f.timing_read()["update"]["ms"] / max(1, events)
    alg = P * (48 * K + 16)  # 48 B per particle-landmark; xv (12 B) and w (4 B) per particle
    print(json.dumps({"metric": "particle*landmark updates/s (FastSLAM1, synthetic map)", "value": P * K / (ms * 1e-3),
                      "unit": "updates/s", "n_gpus": 1, "steps": a.steps, "warmup": a.warmup, "ms_per_step": ms,
                      "dtype": "f32", "data": "synthetic",
                      "config": {"workload": "FastSLAM1 synthetic: %d particles x %d landmarks x %d observations/step, "
                                             "8 noisy predict ticks per update" % (P, Nf, K),
                                 "resample": "gather ran on %d of %d steps" % (resamples, a.steps)},
                      "gpu_launches": int(f.launches - l0), "family_ms_per_step": fam,
                      "roofline": {"bound": "hbm", "kernel": "k_fs1_update_fast", "achieved": alg / (upd_ms * 1e-3) / 1e9,
                                   "peak": peak, "unit": "GB/s", "frac": alg / (upd_ms * 1e-3) / 1e9 / peak,
                                   "ms_per_launch": upd_ms}}))


if __name__ == "__main__":
    main()
