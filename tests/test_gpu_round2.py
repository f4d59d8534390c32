"""GPU (B200): the paths added in round 2, through the C ABI, against the oracle.

* the benched configuration itself (BASELINE.json configs[3]: 2^20 particles x 1024 landmarks x 32 observations,
  slamgpu_update with host noise arriving behind the kernel, fused re-columning, ~50 % resampling) over 12 cycles:
  keep[] / nEff / decision / normalised weights bit-exact for the WHOLE set every cycle, and the continuous state of
  three 2048-particle slices within 1e-4 of the oracle run along their ancestor lineages;
* throughput mode (device Philox generator) against its numpy twin (oracle/philox.py);
* rand() ints instead of doubles for the resample uniforms (same bits);
* the exact sequential merge behind a forced search anomaly; the cooperative weights kernel against its split form;
* the pose estimate produced by the resample's tail; the in-kernel noise wait with a busy copy engine.
"""
import ctypes as C
import os

import numpy as np
import pytest

from replay import compare_states, snapshot
from scenario import RAND_MAX, per_object_relerr

pytestmark = pytest.mark.gpu

TOL = 1e-4


@pytest.fixture(scope="module")
def sb():
    import slam_b200
    slam_b200.build()
    return slam_b200


def _state(oracle_mod, d, P, Nf):
    s = oracle_mod.State(P, Nf)
    s.xv[:], s.Pv[:], s.w[:], s.lm_mean[:], s.lm_cov[:], s.Nf = d["xv"], d["Pv"], d["w"], d["lm_mean"], d["lm_cov"], Nf
    return s


def _mk(sb, P, Nf, method=2, n_effective=None, env=None, **kw):
    from slam_b200.synthetic import SIGMA_PHI, WHEELBASE
    old = {}
    for k, v in (env or {}).items():
        old[k] = os.environ.get(k)
        os.environ[k] = v
    try:
        return sb.FastSLAM(method, P, Nf, wheel_base=WHEELBASE, sigma_phi=SIGMA_PHI, n_effective=n_effective,
                           use_heading=True, scan_mode=sb.SCAN_BLOCKED, **kw)
    finally:
        for k, v in old.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v


def _pinned(L, shape, dtype):
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    p = L.slamgpu_host_alloc(n)
    assert p
    return np.frombuffer((C.c_char * n).from_address(p), dtype=dtype).reshape(shape), p


def test_cfg4_shape_through_the_benched_path(sb, oracle):
    """2^20 x 1024 x 32 through slamgpu_update exactly as bench.py drives it (pinned host noise going up behind the
    kernels in chunks, resample uniforms alternately as doubles and as rand() ints, REMAP re-columning, deferred map
    resampling), 12 cycles.
    Every cycle, whole set: the oracle's blocked resample of the weights the device held gives the same keep[], nEff,
    decision and normalised weights, bit for bit, over all 2^20 particles.
    Three slices of 2048 particles, followed along their ANCESTOR lineages (slot j of a slice descends from
    keep_t[...] of earlier epochs; the run is made twice with identical inputs, the first pass yields the lineages, the
    second reads the poses of the lineage slots after every update): per cycle the oracle's update of the same
    particles with the same noise columns gives the same unnormalised weights (norm-relative, bound asserted) and the
    same poses / pose covariances (1e-4), after which the oracle continues from the device's poses — the pose sample
    and the weights amplify differences (an absolute 1e-6 m on a pose a few centimetres from the origin is already
    1e-5 relative), the landmark filters do not, so the 1024 landmark rows run free along the lineages for all 12 cycles
    and are compared at the end (1e-4)."""
    import oracle.pyoracle as po
    from slam_b200 import _lib
    from slam_b200.synthetic import Q_DEFAULT, SIGMA_PHI, WHEELBASE, SyntheticDrive, synthetic_state
    L = _lib.load()
    P, Nf, K, TILE, CYCLES, SL = 1 << 20, 1024, 32, 4096, 12, 2048
    d = synthetic_state(TILE, Nf, K, seed=1, map_seed=1)
    n_eff = int(0.75 * P)
    drive = SyntheticDrive(d["lm"], K, d["R"], seed=12345)
    rng = np.random.default_rng(7)
    inputs = []
    for cyc in range(CYCLES):
        ticks = [drive.controls() for _ in range(8)]
        zf, idf = drive.observe()
        inputs.append(dict(ticks=ticks, zf=np.ascontiguousarray(zf, np.float32), idf=np.ascontiguousarray(idf, np.int32),
                           gauss=rng.standard_normal((3, P), dtype=np.float32),
                           ints=rng.integers(1, RAND_MAX, size=P).astype(np.int32)))
    g_pin, g_ptr = _pinned(L, (3, P), np.float32)
    u_pin, u_ptr = _pinned(L, (P,), np.float64)
    i_pin, i_ptr = _pinned(L, (P,), np.int32)
    R = np.ascontiguousarray(d["R"], np.float32).ravel()
    Q = np.ascontiguousarray(Q_DEFAULT, np.float32).ravel()

    def run(read_at=None):
        f = _mk(sb, P, Nf, n_effective=n_eff)
        f.upload_tiled(d["xv"], d["Pv"], np.full(TILE, np.float32(1.0 / P), np.float32), d["lm_mean"], d["lm_cov"], Nf)
        f.debug_capture(True)
        rec = []
        for cyc, inp in enumerate(inputs):
            for (V, G, dt, phi) in inp["ticks"]:
                f.predict_raw(V, G, Q.ctypes.data, dt, phi)
            g_pin[:] = inp["gauss"]
            if cyc % 2:
                i_pin[:] = inp["ints"]
                neff, did = f.update_raw(inp["zf"].ctypes.data, inp["idf"].ctypes.data, K, R.ctypes.data, g_ptr, i_ptr,
                                         unif_ints=True)
            else:
                u_pin[:] = inp["ints"] / float(RAND_MAX)
                neff, did = f.update_raw(inp["zf"].ctypes.data, inp["idf"].ctypes.data, K, R.ctypes.data, g_ptr, u_ptr)
            r = dict(wb=f.debug_read(0), wn=f.debug_read(1), keep=f.debug_read(2), did=did, neff=neff)
            if read_at is not None:
                r["poses"] = f.download_poses_at(read_at[cyc + 1])
            rec.append(r)
        return f, rec

    f, rec = run()
    for cyc, (inp, r) in enumerate(zip(inputs, rec)):
        wn_o, keep_o, neff_o = oracle.resample_indices(r["wb"], inp["ints"] / float(RAND_MAX), scan_mode=1)
        assert np.array_equal(r["wn"], wn_o), "cycle %d: normalised weights" % cyc
        assert r["neff"] == neff_o and r["did"] == bool(neff_o < n_eff), "cycle %d: nEff / decision" % cyc
        assert np.array_equal(r["keep"], keep_o), "cycle %d: keep[] (%d differ)" % (cyc, int((r["keep"] != keep_o).sum()))
    st = f.stats()
    assert st["rows_fused"] > 0 and st["resamples"] >= 3, st          # the fused re-columning and the gather did run
    assert any(not r["did"] for r in rec)                                # ... and so did the no-resample branch
    f.close()
    # the slices and their lineages
    starts = [0, P // 2 - SL // 2, P - SL]
    A = np.concatenate([np.arange(p0, p0 + SL) for p0 in starts]).astype(np.int64)
    lineage = [None] * (CYCLES + 1)
    lineage[CYCLES] = A
    for t in range(CYCLES - 1, -1, -1):
        if rec[t]["did"]:
            k = rec[t]["keep"][lineage[t + 1]].astype(np.int64)
            lineage[t] = np.where(k < 0, P - 1, k)
        else:
            lineage[t] = lineage[t + 1]
    f, rec2 = run(read_at=lineage)
    for r, r2 in zip(rec, rec2):                                         # the run is deterministic
        assert r["did"] == r2["did"] and np.array_equal(r["keep"], r2["keep"]) and np.array_equal(r["wb"], r2["wb"])
    got = [f.download_slice(p0, SL) for p0 in starts]
    est = f.estimate()
    n = A.size
    a0 = lineage[0] % TILE
    s = po.State(n, Nf)
    s.xv[:], s.Pv[:] = d["xv"][:, a0], d["Pv"][:, a0]
    s.lm_mean[:], s.lm_cov[:], s.Nf = d["lm_mean"][:, a0], d["lm_cov"][:, a0], Nf
    s.w[:] = np.float32(1.0 / P)
    worst_w, worst_pose = 0.0, {}
    none = np.zeros((0, 2), np.float32)
    for t in range(CYCLES):
        inp, r = inputs[t], rec2[t]
        for (V, G, dt, phi) in inp["ticks"]:
            oracle.predict(2, s, V, G, Q_DEFAULT, dt, phi, WHEELBASE, SIGMA_PHI, True, None)
        oracle.update_particles(2, s, inp["zf"], inp["idf"], none, d["R"], np.ascontiguousarray(inp["gauss"][:, lineage[t]]))
        # unnormalised weights of the lineage particles against what the device held (norm-relative to the largest)
        worst_w = max(worst_w, per_object_relerr(r["wb"][lineage[t]], s.w, None))
        s.w[:] = np.float32(1.0 / P) if r["did"] else r["wn"][lineage[t]]
        gp = dict(r["poses"], lm_mean=s.lm_mean, lm_cov=s.lm_cov, Nf=Nf)
        assert np.array_equal(gp["w"], s.w), "cycle %d: weights after the resample" % t
        errs = compare_states(gp, snapshot(s), TOL, "cycle %d: poses of the lineage slots" % t)
        for k in ("xy", "theta", "Pv"):
            worst_pose[k] = max(worst_pose.get(k, 0.0), errs[k])
        s.xv[:], s.Pv[:] = r["poses"]["xv"], r["poses"]["Pv"]
    # 32 likelihood factors and the prior / proposal densities (see tests/test_gpu_parity.py): measured 2.5e-3 on B200
    assert worst_w <= 4e-3, worst_w
    full = {k: np.concatenate([g[k] for g in got], axis=(0 if k == "w" else 1)) for k in
            ("xv", "Pv", "w", "lm_mean", "lm_cov")}
    full["Nf"] = Nf
    errs = compare_states(full, snapshot(s), TOL, "cfg 4 slices after %d cycles" % CYCLES)
    assert np.array_equal(full["w"], s.w)
    assert np.isfinite(est).all()
    print("cfg4-shape parity: %d cycles, %d resamples, rows fused %d, worst unnormalised weight error %.2e, per-cycle "
          "pose errors %s, final state %s" % (CYCLES, st["resamples"], st["rows_fused"], worst_w,
                                               {k: "%.1e" % v for k, v in worst_pose.items()},
                                               {k: "%.1e" % v for k, v in errs.items()}))
    f.close()
    for p in (g_ptr, u_ptr, i_ptr):
        L.slamgpu_host_free(p)


def _small(seed=4, P=5000, Nf=12, K=6):
    from slam_b200.synthetic import synthetic_state
    return synthetic_state(P, Nf, K, seed=seed, extent=30.0), P, Nf, K


def test_device_rng_equals_its_numpy_twin(sb, oracle):
    """slamgpu_seed: the pose-sample normals, the resample uniforms and the control noise the device generates are the
    ones oracle/philox.py derives (Philox4x32-10 words -> rand()-like ints -> the reference's own arithmetic).
    keep[] / nEff: bit-exact (the uniforms are exact); continuous state: 1e-5 (device logf / sinf against libm)."""
    from oracle import philox
    from slam_b200.synthetic import Q_DEFAULT, SyntheticDrive
    d, P, Nf, K = _small()
    seed = 0x1234ABCD5678EF01
    slots = np.arange(P)
    fa, fb = _mk(sb, P, Nf, n_effective=P + 1), _mk(sb, P, Nf, n_effective=P + 1)
    fa.seed(seed)
    for f in (fa, fb):
        f.upload(d["xv"], d["Pv"], d["w"], d["lm_mean"], d["lm_cov"], Nf)
    drive = SyntheticDrive(d["lm"], K, d["R"], seed=3)
    none = np.zeros((0, 2), np.float32)
    for cyc in range(3):
        for _ in range(8):  # (fewer ticks leave Pv, zeroed by the previous pose sample, numerically singular)
            V, G, dt, phi = drive.controls()
            fa.predict(V, G, Q_DEFAULT, dt, phi)
            fb.predict(V, G, Q_DEFAULT, dt, phi)
        zf, idf = drive.observe()
        fa.update_particles(zf, idf, none, d["R"])                                # gauss = None: generated on the device
        fb.update_particles(zf, idf, none, d["R"], philox.pose_gauss(oracle, seed, slots, cyc))
        ga, gb = fa.download(), fb.download()
        # (unnormalised weights: the prior / proposal densities turn an ulp of the pose sample into ~1e-4, see
        # tests/test_gpu_parity.py; measured 2.7e-4 here)
        compare_states(ga, gb, 1e-5, "rng cycle %d" % cyc, tol_w=1e-3)
        fb.upload(w=ga["w"])                                                       # identical weights into both resamples
        ka, na, da = fa.resample(None, want_keep=True)
        kb, nb, db = fb.resample(philox.resample_unif(seed, slots, cyc), want_keep=True)
        assert na == nb and da and db and np.array_equal(ka, kb), "rng cycle %d resample" % cyc
        wn, keep_o, neff_o = oracle.resample_indices(ga["w"], philox.resample_unif(seed, slots, cyc), scan_mode=1)
        assert np.array_equal(ka, keep_o) and na == neff_o
    # control noise (FastSLAM1 draws it on every tick, fastslam1wrapper.cpp:20)
    f1, f2 = _mk(sb, P, Nf, method=1), _mk(sb, P, Nf, method=1)
    f1.seed(seed, predict_noise=True)
    for f in (f1, f2):
        f.upload(d["xv"], d["Pv"], d["w"], d["lm_mean"], d["lm_cov"], Nf)
    for tick in range(3):
        V, G, dt, phi = drive.controls()
        f1.predict(V, G, Q_DEFAULT, dt, phi)
        f2.predict(V, G, Q_DEFAULT, dt, phi, gauss=philox.control_gauss(oracle, seed, slots, tick))
    compare_states(f1.download(), f2.download(), 1e-5, "rng control noise")
    for f in (fa, fb, f1, f2):
        f.close()


def test_update_with_rand_ints_equals_doubles(sb):
    """slamgpu_update_i32 forms rand() / double(RAND_MAX) on the device: the same bits as passing the doubles."""
    from slam_b200.synthetic import Q_DEFAULT, SyntheticDrive
    d, P, Nf, K = _small(seed=9, P=70000)
    fa, fb = _mk(sb, P, Nf, n_effective=P + 1), _mk(sb, P, Nf, n_effective=P + 1)
    for f in (fa, fb):
        f.upload(d["xv"], d["Pv"], d["w"], d["lm_mean"], d["lm_cov"], Nf)
    drive = SyntheticDrive(d["lm"], K, d["R"], seed=3)
    rng = np.random.default_rng(1)
    none = np.zeros((0, 2), np.float32)
    dids = []
    for cyc in range(4):
        for _ in range(8):
            V, G, dt, phi = drive.controls()
            fa.predict(V, G, Q_DEFAULT, dt, phi)
            fb.predict(V, G, Q_DEFAULT, dt, phi)
        zf, idf = drive.observe()
        gauss = rng.standard_normal((3, P), dtype=np.float32)
        ints = rng.integers(0, RAND_MAX, size=P, endpoint=True).astype(np.int32)
        ints[:3] = [0, RAND_MAX, 1]
        na, da = fa.update(zf, idf, none, d["R"], gauss, ints / float(RAND_MAX))
        nb, db = fb.update_i32(zf, idf, none, d["R"], gauss, ints)
        assert na == nb and da == db
        assert np.array_equal(fa.debug_read(2), fb.debug_read(2))
        dids.append(da)
    ga, gb = fa.download(), fb.download()
    for k in ("xv", "Pv", "w", "lm_mean", "lm_cov"):
        assert np.array_equal(ga[k], gb[k]), k
    assert any(dids)
    fa.close()
    fb.close()


@pytest.mark.parametrize("N", [1000, 1024 * 37 + 5, 1 << 20])
def test_forced_anomaly_takes_the_sequential_merge(sb, oracle, N):
    """SLAMGPU_FORCE_ANOMALY makes every search report an unproven slot: keep[] then comes from the exact sequential
    merge (core.cpp:800-806) over the final CDF, the speculative pose gather is redone with it, and the result is the
    oracle's."""
    rng = np.random.default_rng(N)
    f = _mk(sb, N, 4, n_effective=N + 1, env={"SLAMGPU_FORCE_ANOMALY": "1"})
    w = (rng.random(N) ** 4).astype(np.float32) + np.float32(1e-30)
    xv = rng.standard_normal((3, N)).astype(np.float32)
    unif = rng.integers(0, RAND_MAX, size=N) / float(RAND_MAX)
    f.upload(xv=xv, w=w)
    keep, neff, did = f.resample(unif, want_keep=True)
    wn, keep_o, neff_o = oracle.resample_indices(w, unif, scan_mode=1)
    assert did and neff == neff_o and np.array_equal(keep, keep_o)
    assert f.stats()["keep_fallbacks"] == 1
    g = f.download(landmarks=False)
    assert np.array_equal(g["xv"], xv[:, np.where(keep_o < 0, N - 1, keep_o)])
    assert np.all(g["w"] == np.float32(1.0 / N))
    ex = f.estimate()
    assert abs(ex[0] - g["xv"][0].astype(np.float64).mean()) < 1e-9 and ex[2] == g["xv"][2, 0]
    f.close()


@pytest.mark.parametrize("N", [1, 1000, 1024 * 37 + 5, 1 << 20])
def test_cooperative_weights_kernel_equals_split_launches(sb, oracle, N):
    """k_weights_fused (one cooperative launch, two grid-wide barriers) and the same phases as three launches
    (SLAMGPU_WEIGHTS_SPLIT) share their device functions: identical normalised weights, nEff, keep[] — and both are
    the oracle's."""
    rng = np.random.default_rng(N + 1)
    fa = _mk(sb, N, 4, n_effective=0)
    fb = _mk(sb, N, 4, n_effective=0, env={"SLAMGPU_WEIGHTS_SPLIT": "1"})
    for trial in range(3):
        w = (rng.random(N) ** (1 + 4 * trial)).astype(np.float32) + np.float32(1e-30)
        unif = rng.integers(0, RAND_MAX, size=N) / float(RAND_MAX)
        out = []
        for f in (fa, fb):
            f.upload(w=w)
            keep, neff, did = f.resample(unif, want_keep=True)
            out.append((keep, neff, f.download(landmarks=False)["w"]))
        wn, keep_o, neff_o = oracle.resample_indices(w, unif, scan_mode=1)
        for keep, neff, wgot in out:
            assert neff == neff_o and np.array_equal(wgot, wn) and np.array_equal(keep, keep_o), (N, trial)
    fa.close()
    fb.close()


def test_estimate_comes_with_the_update(sb, oracle):
    """The resample's tail computes the pose estimate on the buffer the device-side decision selects; slamgpu_estimate
    right after slamgpu_update returns it without another launch.  Checked in both branches against
    computeEstimatedPosition (ParticleSLAMWrapper.cpp:56-77) of the downloaded state."""
    import oracle.pyoracle as po
    from slam_b200.synthetic import Q_DEFAULT, SyntheticDrive
    d, P, Nf, K = _small(seed=11, P=30000)
    rng = np.random.default_rng(2)
    none = np.zeros((0, 2), np.float32)
    for n_eff in (0, P + 1):  # never / always resample
        f = _mk(sb, P, Nf, n_effective=n_eff)
        f.upload(d["xv"], d["Pv"], d["w"], d["lm_mean"], d["lm_cov"], Nf)
        drive = SyntheticDrive(d["lm"], K, d["R"], seed=3)
        for cyc in range(2):
            for _ in range(8):
                V, G, dt, phi = drive.controls()
                f.predict(V, G, Q_DEFAULT, dt, phi)
            zf, idf = drive.observe()
            neff, did = f.update(zf, idf, none, d["R"], rng.standard_normal((3, P), dtype=np.float32), rng.random(P))
            assert did == (n_eff > 0)
            n0 = f.launches
            ex = f.estimate()
            assert f.launches == n0, "the estimate after an update must not launch"
            g = f.download(landmarks=False)
            s = po.State(P, 1)
            s.xv[:], s.w[:] = g["xv"], g["w"]
            eo = oracle.estimate(s)
            assert abs(ex[0] - eo[0]) < 1e-9 and abs(ex[1] - eo[1]) < 1e-9 and ex[2] == eo[2], (ex, eo)
        V, G, dt, phi = drive.controls()
        f.predict(V, G, Q_DEFAULT, dt, phi)          # moves the poses: the cached value is dropped
        ex = f.estimate()
        g = f.download(landmarks=False)
        s = po.State(P, 1)
        s.xv[:], s.w[:] = g["xv"], g["w"]
        eo = oracle.estimate(s)
        assert abs(ex[0] - eo[0]) < 1e-9 and ex[2] == eo[2]
        f.close()


def test_noise_arrives_behind_a_busy_copy_engine(sb):
    """The update kernel's warps wait for per-chunk arrival flags the COPY ENGINE writes (4-byte copies from a pinned
    constant, stream-ordered behind each chunk) — no kernel has to find an SM while every SM is full of waiting CTAs.
    A large unrelated host-to-device copy is put in front; the device noise buffer is poisoned with NaN first: the
    result must be finite and bit-identical to uploading the noise before the kernel."""
    import torch
    from slam_b200 import _lib
    from slam_b200.synthetic import Q_DEFAULT, SyntheticDrive
    L = _lib.load()
    d, P, Nf, K = _small(seed=13, P=1 << 19, Nf=8, K=4)
    fa, fb = _mk(sb, P, Nf), _mk(sb, P, Nf)
    for f in (fa, fb):
        f.upload(d["xv"], d["Pv"], d["w"], d["lm_mean"], d["lm_cov"], Nf)
    drive = SyntheticDrive(d["lm"], K, d["R"], seed=3)
    rng = np.random.default_rng(3)
    g_pin, g_ptr = _pinned(L, (3, P), np.float32)
    u_pin, u_ptr = _pinned(L, (P,), np.float64)
    big_h = torch.empty(1 << 28, dtype=torch.uint8).pin_memory()      # 256 MiB, ~5 ms of PCIe
    big_d = torch.empty(1 << 28, dtype=torch.uint8, device="cuda")
    side = torch.cuda.Stream()
    R = np.ascontiguousarray(d["R"], np.float32).ravel()
    for cyc in range(3):
        for _ in range(8):
            V, G, dt, phi = drive.controls()
            fa.predict(V, G, Q_DEFAULT, dt, phi)
            fb.predict(V, G, Q_DEFAULT, dt, phi)
        zf, idf = drive.observe()
        zf, idf = np.ascontiguousarray(zf, np.float32), np.ascontiguousarray(idf, np.int32)
        g_pin[:] = rng.standard_normal((3, P), dtype=np.float32)
        u_pin[:] = rng.random(P)
        fa.stage_noise(np.full((3, P), np.nan, np.float32), None)        # poison: a stale read would surface as NaN
        with torch.cuda.stream(side):
            for _ in range(4):
                big_d.copy_(big_h, non_blocking=True)
        na, da = fa.update_raw(zf.ctypes.data, idf.ctypes.data, K, R.ctypes.data, g_ptr, u_ptr)
        fb.update_particles(zf, idf, np.zeros((0, 2), np.float32), d["R"], g_pin.copy())
        kb, nb, db = fb.resample(u_pin.copy())
        assert na == nb and da == db
    side.synchronize()
    ga, gb = fa.download(), fb.download()
    for k in ("xv", "Pv", "w", "lm_mean", "lm_cov"):
        assert np.isfinite(ga[k]).all(), k
        assert np.array_equal(ga[k], gb[k]), k
    fa.close()
    fb.close()
    L.slamgpu_host_free(g_ptr)
    L.slamgpu_host_free(u_ptr)


def test_comm_init_rejects_more_than_2_30_particles(sb):
    """keep[], the keep log and the ancestor maps are int32 GLOBAL indices and the upper scan level holds 1024 x 1024
    block totals: P x world > 2^30 must be refused, not corrupted (the check precedes any communicator work)."""
    P = (1 << 27) + 1024
    f = sb.FastSLAM(sb.FASTSLAM2, P, 1, scan_mode=sb.SCAN_BLOCKED, spare_rows=1, log_epochs=4)
    with pytest.raises(sb.SlamGpuError) as e:
        f.comm_init(0, 8, bytes(128))
    assert e.value.code == -1 and "2^30" in str(e.value)
    f.close()


@pytest.mark.parametrize("scan", ["blocked", "reference"])
def test_zero_weight_sum_goes_on_unweighted(sb, scan):
    """Every likelihood underflowed (BASELINE.json config 0 does that at a loop closure with the shipped seed: 100
    particles, 1-degree bearing sigma): the reference divides by the zero sum (core.cpp:727-729) and is left with NaN
    weights and a NaN nEff for good.  By default the library sets w = 1/N, reports nEff = N, skips that resample and
    counts the event; keep_nan_weights = 1 asks for the reference's behaviour."""
    N = 3000
    mode = sb.SCAN_BLOCKED if scan == "blocked" else sb.SCAN_REFERENCE
    unif = np.random.default_rng(0).random(N)
    f = sb.FastSLAM(sb.FASTSLAM1, N, 4, n_effective=int(0.75 * N), scan_mode=mode)
    f.upload(w=np.zeros(N, np.float32))
    keep, neff, did = f.resample(unif, want_keep=True)
    assert not did and neff == pytest.approx(N, rel=1e-5)
    assert np.all(f.download(landmarks=False)["w"] == np.float32(1.0 / N))
    assert f.stats()["weight_resets"] == 1
    f.upload(w=np.full(N, 2.0, np.float32))                       # and an ordinary sum afterwards
    keep, neff, did = f.resample(unif, want_keep=True)
    assert neff == pytest.approx(N, rel=1e-5) and f.stats()["weight_resets"] == 1
    f.close()
    f = sb.FastSLAM(sb.FASTSLAM1, N, 4, n_effective=int(0.75 * N), scan_mode=mode, keep_nan_weights=True)
    f.upload(w=np.zeros(N, np.float32))
    keep, neff, did = f.resample(unif, want_keep=True)
    assert not did and np.isnan(neff) and np.isnan(f.download(landmarks=False)["w"]).all()
    f.close()
