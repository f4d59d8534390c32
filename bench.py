#!/usr/bin/env python
"""bench.py — FastSLAM2 synthetic workload (BASELINE.json config 4) on N B200s of one node.

A "step" is one observation cycle of the reference's wrapper loop (wrappers/fastslam2wrapper.cpp:51-117):
TICKS_PER_OBSERVE (8) x FastSLAM2::predict, one FastSLAM2::update (per-particle proposal / weight / landmark EKF for
K observed landmarks, then resampleParticles: weight normalisation, nEff, CDF, keep[] and — when nEff < NEFFECTIVE —
the gather of poses and whole landmark maps), and one pose estimate.  metric = particle*landmark updates per second
= P_total * K / step time.

    python bench.py [--gpus N] [--steps K] [--warmup W]            this repo (CUDA, through the C ABI)
    python bench.py --impl reference ...                           the reference's own CPU code on the host cores

N > 1 is launched by torchrun, one rank per GPU; particles are block-sharded (weak scaling: --particles per GPU).
Prints ONE JSON line on rank 0.
"""
import argparse
import ctypes as C
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

from slam_b200.synthetic import (Q_DEFAULT, SIGMA_PHI, TICKS_PER_OBSERVE, WHEELBASE, SyntheticDrive,  # noqa: E402
                                 synthetic_state)

METRIC = "particle*landmark updates/s (FastSLAM2, synthetic map)"
UNIT = "updates/s"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--particles", type=int, default=1 << 20, help="particles PER GPU")
    ap.add_argument("--landmarks", type=int, default=1024)
    ap.add_argument("--obs", type=int, default=32)
    ap.add_argument("--tile", type=int, default=4096, help="distinct synthetic particles generated on the host")
    ap.add_argument("--cpu-seconds", type=float, default=12.0, help="budget of the cpu_baseline leg")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--same-tile", action="store_true", help="experiment: every rank starts from the same particle tile")
    ap.add_argument("--ref-particles", type=int, default=1000, help="particles per reference instance (core.cpp:762 valid)")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks and throttle reasons during the timed region (B200_PROFILING.md recipe)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
                for n, v in zip(names, r[5:9]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def pinned_array(L, shape, dtype):
    n = int(np.prod(shape)) * np.dtype(dtype).itemsize
    p = L.slamgpu_host_alloc(n)
    if not p:
        raise RuntimeError("slamgpu_host_alloc failed")
    buf = (C.c_char * n).from_address(p)
    return np.frombuffer(buf, dtype=dtype).reshape(shape), p


# ------------------------------------------------------------------------------------------------------------
# the CPU arms
# ------------------------------------------------------------------------------------------------------------
def _cpu_state(P, Nf, K, seed):
    from oracle.pyoracle import State
    d = synthetic_state(P, Nf, K, seed=seed)
    s = State(P, Nf)
    s.xv[:], s.Pv[:], s.w[:], s.lm_mean[:], s.lm_cov[:], s.Nf = d["xv"], d["Pv"], d["w"], d["lm_mean"], d["lm_cov"], Nf
    return d, s


def _ref_worker(args):
    """One independent instance of the UNMODIFIED reference (oracle/_ref) running `steps` observation cycles."""
    P, Nf, K, warmup, steps, seed = args
    from oracle.pyoracle import Ref
    r = Ref()
    d, s = _cpu_state(P, Nf, K, seed)
    h = r.create(s)
    drive = SyntheticDrive(d["lm"], K, d["R"], seed=seed)
    r.lib.srand(seed)  # rand() falls back to the shim's deterministic generator: the reference draws its own noise
    times = []
    for it in range(warmup + steps):
        t0 = time.perf_counter()
        for _ in range(TICKS_PER_OBSERVE):
            V, G, dt, phi = drive.controls()
            r.predict(2, h, V, G, Q_DEFAULT, dt, phi, WHEELBASE, SIGMA_PHI, True, False)
        zf, idf = drive.observe()
        r.update(2, h, zf, idf, np.zeros((0, 2), np.float32), d["R"], int(0.75 * P), True)
        times.append(time.perf_counter() - t0)
    r.destroy(h)
    return times[warmup:]


def _port_worker(args):
    """The oracle's C restatement, same cycle (used where oracle/_ref was never built)."""
    P, Nf, K, warmup, steps, seed = args
    from oracle.pyoracle import Oracle
    o = Oracle()
    d, s = _cpu_state(P, Nf, K, seed)
    drive = SyntheticDrive(d["lm"], K, d["R"], seed=seed)
    rng = np.random.default_rng(seed)
    times = []
    for it in range(warmup + steps):
        gauss = rng.standard_normal((3, P)).astype(np.float32)
        unif = rng.random(P)
        t0 = time.perf_counter()
        for _ in range(TICKS_PER_OBSERVE):
            V, G, dt, phi = drive.controls()
            o.predict(2, s, V, G, Q_DEFAULT, dt, phi, WHEELBASE, SIGMA_PHI, True, None)
        zf, idf = drive.observe()
        o.update_particles(2, s, zf, idf, np.zeros((0, 2), np.float32), d["R"], gauss)
        o.resample(s, unif, int(0.75 * P), True, scan_mode=0)
        times.append(time.perf_counter() - t0)
    return times[warmup:]


def cpu_arm(P, Nf, K, warmup, steps, cores):
    """Runs `cores` independent instances (particles are independent between resamples; each instance resamples its
    own P particles).  Returns (updates/s, ms per step, kind, sample text)."""
    import multiprocessing as mp
    from oracle.pyoracle import Ref, build
    build(ref=os.path.isdir("/root/reference"))
    kind = "reference" if Ref.available() else "port"
    worker = _ref_worker if kind == "reference" else _port_worker
    jobs = [(P, Nf, K, warmup, steps, 100 + i) for i in range(cores)]
    if cores == 1:
        res = [worker(jobs[0])]
    else:
        with mp.get_context("spawn").Pool(cores) as pool:
            res = pool.map(worker, jobs)
    step_s = float(np.mean([np.mean(t) for t in res]))  # instances run concurrently
    value = cores * P * K / step_s
    sample = ("%d instance(s) x %d particles x %d landmarks, %d observed/step, %d timed cycles each "
              "(8 predict + update + resampleParticles)" % (cores, P, Nf, K, steps))
    return value, step_s * 1e3, kind, sample


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    steps = max(1, min(a.steps, 6))   # bounded: each cycle deep-copies every particle's 1024-landmark map
    warmup = max(1, min(a.warmup, 1))
    value, ms, kind, sample = cpu_arm(a.ref_particles, a.landmarks, a.obs, warmup, steps, cores)
    line = {"impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": a.gpus, "steps": steps,
            "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32", "data": "synthetic",
            "config": {"workload": "FastSLAM2 synthetic: %d particles/instance x %d landmarks x %d observations/step"
                                   % (a.ref_particles, a.landmarks, a.obs),
                       "note": "the unmodified reference cannot run 2^20 particles (core.cpp:762 assert, O(N^2) "
                               "cumulativeSum, ~200 GB of Eigen heap objects): timed at a reference-valid particle "
                               "count per instance, one instance per host core"},
            "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------------------
# the CUDA arm
# ------------------------------------------------------------------------------------------------------------
def run_ours(a):
    import torch
    import torch.distributed as dist

    import slam_b200
    from slam_b200 import _lib

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != a.gpus:
        if world == 1 and a.gpus > 1:
            raise SystemExit("--gpus %d needs torchrun (one rank per GPU)" % a.gpus)
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    slam_b200.build()
    L = _lib.load()

    P, Nf, K = a.particles, a.landmarks, a.obs
    Pg = P * world
    tile = min(a.tile, P)
    d = synthetic_state(tile, Nf, K, seed=1 if a.same_tile else 1 + rank, map_seed=1)  # one world, different particles per rank
    f = slam_b200.FastSLAM(slam_b200.FASTSLAM2, P, Nf, wheel_base=WHEELBASE, sigma_phi=SIGMA_PHI,
                           n_effective=int(0.75 * Pg), use_heading=True, scan_mode=slam_b200.SCAN_BLOCKED,
                           device=local)
    if world > 1:
        idbuf = torch.zeros(128, dtype=torch.uint8, device="cuda")
        if rank == 0:
            idbuf.copy_(torch.frombuffer(bytearray(slam_b200.FastSLAM.comm_unique_id()), dtype=torch.uint8))
        dist.broadcast(idbuf, 0)
        f.comm_init(rank, world, bytes(idbuf.cpu().numpy().tobytes()))
    f.upload_tiled(d["xv"], d["Pv"], np.full(tile, np.float32(1.0 / Pg), np.float32), d["lm_mean"], d["lm_cov"], Nf)
    stream = torch.cuda.Stream()
    f.set_stream(stream.cuda_stream)
    R = np.ascontiguousarray(d["R"], np.float32).ravel()
    Q = np.ascontiguousarray(Q_DEFAULT, np.float32).ravel()

    # host noise pool in pinned memory (the reference draws 4 rand() per particle for the pose sample and 1 per
    # particle for the resample, every update): NPOOL sets, rotated
    NPOOL = 4
    rng = np.random.default_rng(777 + rank)
    pool = []
    for i in range(NPOOL):
        g, gp = pinned_array(L, (3, P), np.float32)
        u, up = pinned_array(L, (P,), np.float64)
        g[:] = rng.standard_normal((3, P), dtype=np.float32)
        u[:] = rng.random(P)
        pool.append((g, gp, u, up))
    # The controls and observation sets of every step are host INPUTS: the synthetic drive is run ahead of the timed
    # regions (its numpy argsort / trig per observe would otherwise sit on the critical path between two steps and be
    # measured as GPU idle time), and each step's zf / idf sit in their own pinned slots.
    drive = SyntheticDrive(d["lm"], K, d["R"], seed=12345)  # same observations on every rank
    n_plan = 2 * a.warmup + 2 * a.steps + 4
    zall, zall_p = pinned_array(L, (n_plan, K, 2), np.float32)
    iall, iall_p = pinned_array(L, (n_plan, K), np.int32)
    plan = []
    for sidx in range(n_plan):
        ticks = [drive.controls() for _ in range(TICKS_PER_OBSERVE)]
        zall[sidx], iall[sidx] = drive.observe()
        plan.append(ticks)
    Qp = Q.ctypes.data
    Rp = R.ctypes.data
    cursor = {"i": 0}

    state = {"resamples": 0, "steps": 0}

    def one_step(i, resident):
        k = cursor["i"]
        cursor["i"] = k + 1
        for (V, G, dt, phi) in plan[k]:
            f.predict_raw(V, G, Qp, dt, phi)
        zp = zall_p + k * K * 2 * 4
        ip = iall_p + k * K * 4
        g, gp, u, up = pool[i % NPOOL]
        if resident:
            neff, did = f.update_raw(zp, ip, K, Rp, None, None)
        else:
            neff, did = f.update_raw(zp, ip, K, Rp, gp, up)
        est = f.estimate()
        state["resamples"] += int(did)
        state["steps"] += 1
        return neff, did, est

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---- device-resident leg ("value"): noise staged in HBM before the timed region -------------------------
    f.stage_noise(pool[0][0], pool[0][2])
    for i in range(a.warmup):
        one_step(i, True)
    f.timing_enable(True)
    f.timing_reset()
    state.update(resamples=0, steps=0)
    launches0 = f.launches
    stats0 = f.stats()
    clocks = ClockSampler(local)
    if rank == 0:
        clocks.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    with torch.cuda.stream(stream):
        e0.record(stream)
        for i in range(a.steps):
            one_step(i, True)
        e1.record(stream)
    barrier()
    dev_ms = max_over_ranks(e0.elapsed_time(e1))
    clk = clocks.stop() if rank == 0 else None
    launches = f.launches - launches0
    stats1 = f.stats()
    rows_moved = stats1["rows_moved"] - stats0["rows_moved"]
    n_resamples = stats1["resamples"] - stats0["resamples"]
    rows_fused = stats1["rows_fused"] - stats0["rows_fused"]
    remote = stats1["remote_sources"] - stats0["remote_sources"]
    remote_frac = max_over_ranks(remote / max(1, n_resamples * P)) if world > 1 else 0.0
    fam = f.timing_read()
    f.timing_enable(False)
    if os.environ.get("SLAM_BENCH_RANK_REPORT"):  # per-rank family times (the JSON line carries rank 0's)
        print("[rank %d] ms/step by family: %s  remote sources/resample: %.3f"
              % (rank, {k: round(v["ms"] / a.steps, 4) for k, v in fam.items()}, remote / max(1, n_resamples * P)),
              file=sys.stderr, flush=True)
    resample_rate = state["resamples"] / max(1, state["steps"])

    # ---- end-to-end leg: host buffers in, result out, every step ---------------------------------------------
    for i in range(min(a.warmup, 2)):
        one_step(i, False)
    barrier()
    t0 = time.perf_counter()
    for i in range(a.steps):
        one_step(i, False)
    barrier()
    e2e_s = max_over_ranks(time.perf_counter() - t0)
    h2d = 3 * P * 4 + P * 8 + K * 12
    d2h = 32 + 40

    ms_per_step = dev_ms / a.steps
    value = Pg * K / (ms_per_step * 1e-3)
    e2e_value = Pg * K / (e2e_s / a.steps)

    # ---- roofline of the dominant kernel family (device time from CUDA events inside the timed region) --------
    peak, peak_src = peaks()
    # algorithmic bytes per launch: update 48 B per particle-landmark + 116 B per particle;
    # predict 96 B per particle per flush of queued ticks; state resampling 104 B per particle per resample plus 48 B
    # per particle per landmark row actually re-columned (the reference's eager copy would be 2*Nf*24 B per particle)
    fam_ms = {k: v["ms"] for k, v in fam.items()}
    alg_total = {"update": P * (48 * K + 116) * fam["update"]["events"],
                 "predict": P * 96 * fam["predict"]["events"],
                 "gather": P * (104 * n_resamples + 48 * (rows_moved - rows_fused))}
    top = max(("update", "gather", "predict", "weights"), key=lambda k: fam_ms[k])
    roof = {}
    for k in ("update", "gather", "predict"):
        ev = fam[k]["events"]
        if ev and fam_ms[k] > 0:
            gbs = alg_total[k] / (fam_ms[k] * 1e-3) / 1e9
            roof[k] = {"achieved": gbs, "frac": gbs / peak, "ms_per_event": fam_ms[k] / ev, "events": ev,
                       "launches": fam[k]["launches"], "alg_bytes_per_event": alg_total[k] / ev}
    kern = {"update": "k_fs2_update_fast<REMAP> (cp.async landmark ring)", "gather": "k_gather_pose + k_chain + k_materialise_rows",
            "predict": "k_predict_batch_fast",
            "weights": "scan/search kernels"}[top]
    traffic = None
    tp = os.path.join(ROOT, "profiles", "r01_fs2_update_traffic.json")
    if top == "update" and (P, Nf, K) == (1 << 20, 1024, 32) and os.path.exists(tp):
        try:  # dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full capture
            traffic = json.load(open(tp))["dram_bytes_per_launch"]
        except Exception:
            traffic = None
    roofline = {"bound": "hbm", "kernel": kern, "achieved": roof.get(top, {}).get("achieved"), "peak": peak,
                "unit": "GB/s", "frac": roof.get(top, {}).get("frac"), "traffic": traffic, "peak_source": peak_src,
                "families": roof, "family_ms_total": fam_ms}

    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        v, ms, kind, sample = cpu_arm(a.ref_particles, Nf, K, 1, 3, 1)
        cpu = {"value": v, "unit": UNIT, "cores": 1, "kind": kind, "sample": sample, "ms_per_step": ms}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic",
                "config": {"workload": "FastSLAM2 synthetic: %d particles/GPU x %d landmarks x %d observations/step, "
                                       "8 predict ticks per update (BASELINE.json config 4 shape)" % (P, Nf, K),
                           "particles_total": Pg, "parallelism": "particles block-sharded over %d GPU(s)" % world,
                           "resample": "NEFFECTIVE = 0.75 N; gather ran on %.0f%% of timed steps" % (100 * resample_rate),
                           "l2": "inputs larger than L2 (the update alone streams %.1f GB per step)"
                                 % (P * (48 * K + 116) / 1e9),
                           "map_resampling": "deferred: %d landmark rows re-columned over %d resamples (eager copy: %d); "
                                             "%d of them inside the update kernel itself (fused, no extra pass)"
                                             % (rows_moved, n_resamples, n_resamples * Nf, rows_fused),
                           "migration": "%.1f%% of resampled particles came from another GPU (worst rank)"
                                        % (100 * remote_frac),
                           "predict": "the 8 ticks per step are queued and applied in one pass",
                           "noise": "pre-drawn on the host, %d sets rotated" % NPOOL},
                "clocks": clk, "gpu_launches": int(launches),
                "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "ms_per_step": e2e_s / a.steps * 1e3},
                "roofline": roofline}
        if cpu:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    f.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
