"""Worker for tests/test_gpu_sharded.py (run under torchrun, one rank per GPU, world 2 / 4 / 8): the particle set is
block-sharded over the ranks, every rank runs the same predict/update cycle through the C ABI, and rank 0 checks the
gathered result against the ORACLE run on the whole set in one piece: keep[] bit-exact (global indices), weights and
nEff bit-exact after normalisation, poses / landmarks within 1e-4, and the migrated particles a bit-exact copy of their
sources.  Phases 2-4: fused re-columning across ranks, the device generator, and the sequential-merge chain."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)

import slam_b200  # noqa: E402
from oracle.pyoracle import Oracle, State  # noqa: E402
from replay import compare_states, snapshot  # noqa: E402
from slam_b200.synthetic import Q_DEFAULT, SIGMA_PHI, WHEELBASE, SyntheticDrive, synthetic_state  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    Pl, Nf, K = 2048, 20, 8          # per-rank particles: a multiple of the 1024 scan block
    Pg = Pl * world
    d = synthetic_state(Pg, Nf, K, seed=21, extent=40.0)
    rng = np.random.default_rng(5)
    f = slam_b200.FastSLAM(slam_b200.FASTSLAM2, Pl, Nf, wheel_base=WHEELBASE, sigma_phi=SIGMA_PHI,
                           n_effective=Pg + 1, use_heading=True, scan_mode=slam_b200.SCAN_BLOCKED, device=local,
                           spare_rows=6)
    idbuf = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        idbuf.copy_(torch.frombuffer(bytearray(slam_b200.FastSLAM.comm_unique_id()), dtype=torch.uint8))
    dist.broadcast(idbuf, 0)
    f.comm_init(rank, world, idbuf.cpu().numpy().tobytes())
    sl = slice(rank * Pl, (rank + 1) * Pl)
    c = np.ascontiguousarray
    f.upload(c(d["xv"][:, sl]), c(d["Pv"][:, sl]), c(d["w"][sl]), c(d["lm_mean"][:, sl]), c(d["lm_cov"][:, sl]), Nf)
    s = State(Pg, Nf)
    s.xv[:], s.Pv[:], s.w[:], s.lm_mean[:], s.lm_cov[:], s.Nf = d["xv"], d["Pv"], d["w"], d["lm_mean"], d["lm_cov"], Nf
    o = Oracle()
    drive = SyntheticDrive(d["lm"], K, d["R"], seed=3)
    ok = True
    for cycle in range(3):
        for _ in range(8):
            V, G, dt, phi = drive.controls()
            f.predict(V, G, Q_DEFAULT, dt, phi)
            o.predict(2, s, V, G, Q_DEFAULT, dt, phi, WHEELBASE, SIGMA_PHI, True, None)
        zf, idf = drive.observe()
        gauss = rng.standard_normal((3, Pg)).astype(np.float32)
        unif = rng.random(Pg)
        f.update_particles(zf, idf, np.zeros((0, 2), np.float32), d["R"], c(gauss[:, sl]))
        o.update_particles(2, s, zf, idf, np.zeros((0, 2), np.float32), d["R"], gauss)
        g = f.download()
        parts = [None] * world
        dist.all_gather_object(parts, g)
        full = {k: np.concatenate([p[k] for p in parts], axis=(0 if k == "w" else 1)) for k in
                ("xv", "Pv", "w", "lm_mean", "lm_cov")}
        full["Nf"] = Nf
        if rank == 0:
            # (unnormalised FastSLAM2 weights: measured 2.0e-3 at 4 ranks' data, see tests/test_gpu_parity.py TOL_W_BY)
            compare_states(full, snapshot(s), 1e-4, "cycle %d update" % cycle, tol_w=4e-3)
        s.w[:] = full["w"]   # resample the device's weights on both sides
        pre = {k: v.copy() for k, v in full.items() if k != "Nf"}
        keep, neff, did = f.resample(c(unif[sl]), want_keep=True)
        ex = f.estimate()
        keep_o, neff_o, did_o = o.resample(s, unif, Pg + 1, True, scan_mode=1)
        g = f.download()
        dist.all_gather_object(parts, dict(g, keep=keep))
        if rank == 0:
            keep_all = np.concatenate([p["keep"] for p in parts])
            assert np.array_equal(keep_all, keep_o), "keep[] differs from the one-piece oracle"
            assert neff == neff_o and did and did_o
            for k in ("xv", "Pv", "lm_mean", "lm_cov"):
                got = np.concatenate([p[k] for p in parts], axis=1)
                assert np.array_equal(got, np.take(pre[k], keep_all, axis=1)), "migrated " + k
                getattr(s, k)[..., :] = got if k in ("xv", "Pv") else getattr(s, k)
            s.lm_mean[:Nf] = np.concatenate([p["lm_mean"] for p in parts], axis=1)
            s.lm_cov[:Nf] = np.concatenate([p["lm_cov"] for p in parts], axis=1)
            wall = np.concatenate([p["w"] for p in parts])
            assert np.array_equal(wall, s.w)
            eo = o.estimate(s)
            assert abs(ex[0] - eo[0]) < 1e-9 and abs(ex[1] - eo[1]) < 1e-9 and ex[2] == eo[2], (ex, eo)
        # keep the oracle's continuous state identical to the devices' for the next cycle
        st = [None]
        if rank == 0:
            st = [(s.xv.copy(), s.Pv.copy(), s.lm_mean.copy(), s.lm_cov.copy())]
        dist.broadcast_object_list(st, 0)
        s.xv[:], s.Pv[:], s.lm_mean[:], s.lm_cov[:] = st[0]
    # ---- phase 2: observed rows re-columned INSIDE the update kernel, ancestors on peer devices ----------------
    # No map download between a resample and the next update, so the sharded handles take the fused path
    # (kernels.cuh K3 REMAP, sources read through the peer mappings).  Rank 0 also drives a ONE-device handle holding
    # all Pg particles with a single spare row (separate gather kernels, no fusion): per-particle arithmetic does not
    # depend on the sharding and the blocked resample is bit-identical for any device count, so after 4 cycles the
    # two must agree bit for bit.
    f.close()
    f = slam_b200.FastSLAM(slam_b200.FASTSLAM2, Pl, Nf, wheel_base=WHEELBASE, sigma_phi=SIGMA_PHI,
                           n_effective=Pg + 1, use_heading=True, scan_mode=slam_b200.SCAN_BLOCKED, device=local)
    dist.broadcast(idbuf.zero_() if rank else idbuf.copy_(torch.frombuffer(
        bytearray(slam_b200.FastSLAM.comm_unique_id()), dtype=torch.uint8)), 0)
    f.comm_init(rank, world, idbuf.cpu().numpy().tobytes())
    f.upload(c(d["xv"][:, sl]), c(d["Pv"][:, sl]), c(d["w"][sl]), c(d["lm_mean"][:, sl]), c(d["lm_cov"][:, sl]), Nf)
    f1 = None
    if rank == 0:
        f1 = slam_b200.FastSLAM(slam_b200.FASTSLAM2, Pg, Nf, wheel_base=WHEELBASE, sigma_phi=SIGMA_PHI,
                                n_effective=Pg + 1, use_heading=True, scan_mode=slam_b200.SCAN_BLOCKED, device=local,
                                spare_rows=1)
        f1.upload(d["xv"], d["Pv"], d["w"], d["lm_mean"], d["lm_cov"], Nf)
    for cycle in range(4):
        for _ in range(2):
            V, G, dt, phi = drive.controls()
            f.predict(V, G, Q_DEFAULT, dt, phi)
            if f1:
                f1.predict(V, G, Q_DEFAULT, dt, phi)
        zf, idf = drive.observe()
        gauss = rng.standard_normal((3, Pg)).astype(np.float32)
        unif = rng.random(Pg)
        f.update_particles(zf, idf, np.zeros((0, 2), np.float32), d["R"], c(gauss[:, sl]))
        keep, neff, did = f.resample(c(unif[sl]), want_keep=True)
        assert did
        if f1:
            f1.update_particles(zf, idf, np.zeros((0, 2), np.float32), d["R"], gauss)
            keep1, neff1, did1 = f1.resample(unif, want_keep=True)
            assert did1 and neff1 == neff and np.array_equal(keep1[sl], keep), "phase 2 cycle %d keep[]" % cycle
    assert f.stats()["rows_fused"] >= 3 * (K - 1), f.stats()
    g = f.download()
    dist.all_gather_object(parts, g)
    if rank == 0:
        g1 = f1.download()
        assert f1.stats()["rows_fused"] == 0
        for k in ("xv", "Pv", "w", "lm_mean", "lm_cov"):
            got = np.concatenate([p[k] for p in parts], axis=(0 if k == "w" else 1))
            assert np.array_equal(got, g1[k]), "phase 2 (fused, sharded) vs one device: " + k
        assert np.isfinite(g1["lm_mean"]).all()
    # ---- phase 3: throughput mode.  The device generator is keyed by the GLOBAL particle index, so the sharded run and
    # the one-device run draw the same noise: through slamgpu_update (no host noise at all) they stay bit-identical.
    seed = 0xC0FFEE1234
    f.seed(seed)
    if f1:
        f1.seed(seed)
    for cycle in range(3):
        for _ in range(2):
            V, G, dt, phi = drive.controls()
            f.predict(V, G, Q_DEFAULT, dt, phi)
            if f1:
                f1.predict(V, G, Q_DEFAULT, dt, phi)
        ep = f.estimate()  # after predict ticks: computed on request (a launch + one exchange of the rank partials)
        if f1:
            e1 = f1.estimate()
            assert abs(ep[0] - e1[0]) < 1e-9 and abs(ep[1] - e1[1]) < 1e-9 and ep[2] == e1[2], (ep, e1)
        zf, idf = drive.observe()
        neff, did = f.update(zf, idf, np.zeros((0, 2), np.float32), d["R"])
        keep = f.debug_read(2)
        est = f.estimate()  # after an update: the value the resample's tail already produced
        if f1:
            neff1, did1 = f1.update(zf, idf, np.zeros((0, 2), np.float32), d["R"])
            assert (neff1, did1) == (neff, did) and np.array_equal(f1.debug_read(2)[sl], keep), "phase 3 cycle %d" % cycle
            e1 = f1.estimate()
            assert abs(est[0] - e1[0]) < 1e-9 and abs(est[1] - e1[1]) < 1e-9 and est[2] == e1[2], (est, e1)
    g = f.download()
    dist.all_gather_object(parts, g)
    if rank == 0:
        g1 = f1.download()
        for k in ("xv", "Pv", "w", "lm_mean", "lm_cov"):
            got = np.concatenate([p[k] for p in parts], axis=(0 if k == "w" else 1))
            assert np.array_equal(got, g1[k]), "phase 3 (device generator, sharded) vs one device: " + k
        f1.close()
    f.close()
    # ---- phase 4: a search anomaly on any rank sends EVERY rank into the exact sequential merge, the ranks taking
    # turns in rank order (capi.cu keep_fallback); forced here.  Result: the one-piece oracle's keep[], and the gather
    # redone with it.
    os.environ["SLAMGPU_FORCE_ANOMALY"] = "1"
    f = slam_b200.FastSLAM(slam_b200.FASTSLAM2, Pl, Nf, wheel_base=WHEELBASE, sigma_phi=SIGMA_PHI,
                           n_effective=Pg + 1, use_heading=True, scan_mode=slam_b200.SCAN_BLOCKED, device=local)
    del os.environ["SLAMGPU_FORCE_ANOMALY"]
    dist.broadcast(idbuf.zero_() if rank else idbuf.copy_(torch.frombuffer(
        bytearray(slam_b200.FastSLAM.comm_unique_id()), dtype=torch.uint8)), 0)
    f.comm_init(rank, world, idbuf.cpu().numpy().tobytes())
    w4 = (np.random.default_rng(8).random(Pg) ** 6).astype(np.float32) + np.float32(1e-30)
    u4 = np.random.default_rng(9).random(Pg)
    f.upload(c(d["xv"][:, sl]), c(d["Pv"][:, sl]), c(w4[sl]), c(d["lm_mean"][:, sl]), c(d["lm_cov"][:, sl]), Nf)
    keep, neff, did = f.resample(c(u4[sl]), want_keep=True)
    assert did and f.stats()["keep_fallbacks"] == 1
    g = f.download(landmarks=False)
    dist.all_gather_object(parts, dict(g, keep=keep))
    if rank == 0:
        wn, keep_o, neff_o = o.resample_indices(w4, u4, scan_mode=1)
        keep_all = np.concatenate([p["keep"] for p in parts])
        assert neff == neff_o and np.array_equal(keep_all, keep_o), "phase 4: sequential-merge chain"
        got = np.concatenate([p["xv"] for p in parts], axis=1)
        assert np.array_equal(got, d["xv"][:, np.where(keep_o < 0, Pg - 1, keep_o)])
        print("SHARDED_OK world=%d" % world)
    f.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
