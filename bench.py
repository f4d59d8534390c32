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

from slam_b200.synthetic import (Q_DEFAULT, RAND_MAX, SIGMA_PHI, TICKS_PER_OBSERVE, WHEELBASE,  # noqa: E402
                                 SyntheticDrive, synthetic_state)

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
    ap.add_argument("--burn-in", type=int, default=100, help="untimed steps before the timed (steady-state) leg")
    ap.add_argument("--no-parity", action="store_true", help="skip the untimed oracle parity phase")
    ap.add_argument("--parity-landmarks", type=int, default=40, help="landmarks per particle in the parity phase")
    ap.add_argument("--no-config5", action="store_true", help="N > 1: skip the secondary 4096-landmark leg")
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
                                          "--format=csv,noheader,nounits", "-lms", "200"], stdout=subprocess.PIPE,
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
class Ctx:
    """torch.distributed plumbing shared by the legs."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        if self.world > 1:
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.local))

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, x):
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device="cuda")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def new_handle(self, slam_b200, P, Nf, n_effective, **kw):
        f = slam_b200.FastSLAM(slam_b200.FASTSLAM2, P, Nf, wheel_base=WHEELBASE, sigma_phi=SIGMA_PHI,
                               n_effective=n_effective, use_heading=True, scan_mode=slam_b200.SCAN_BLOCKED,
                               device=self.local, **kw)
        if self.world > 1:
            torch = self.torch
            idbuf = torch.zeros(128, dtype=torch.uint8, device="cuda")
            if self.rank == 0:
                idbuf.copy_(torch.frombuffer(bytearray(slam_b200.FastSLAM.comm_unique_id()), dtype=torch.uint8))
            self.dist.broadcast(idbuf, 0)
            f.comm_init(self.rank, self.world, bytes(idbuf.cpu().numpy().tobytes()))
        return f


def measure(cx, a, L, slam_b200, P, Nf, K, steps, warmup, burn_in, full):
    """One workload on this rank's GPU.  Legs (each: `warmup` untimed steps, then `steps` timed):
       cold     throughput mode from the balanced initial population (what a 20-step run from scratch sees)
       [burn_in untimed steps: lineages coalesce, shards drift apart, the re-column backlog builds up]
       value    throughput mode, steady state, CUDA events on the library's stream          -> the headline
       long     the value leg again over max(200, 10 K) steps (low-variance steady state)   -> `steady_long`
       profiled the same again with one CUDA-event pair around every kernel family (the events cost host time between
                the launches, which is why the headline leg runs without them)            -> the roofline
       staged   one host-drawn noise set staged in HBM and reused (round 1's `value` leg, for continuity)
       e2e      host noise in (pinned: 12 B normals + 4 B rand() int per particle), result out, wall clock
       e2e_rng  throughput mode through the same public call, wall clock (only the observation set goes in)
    full == False: cold + value only (the secondary configuration)."""
    torch = cx.torch
    world, rank = cx.world, cx.rank
    Pg = P * world
    tile = min(a.tile, P)
    d = synthetic_state(tile, Nf, K, seed=1 if a.same_tile else 1 + rank, map_seed=1)  # one world, different particles per rank
    f = cx.new_handle(slam_b200, P, Nf, int(0.75 * Pg))
    f.upload_tiled(d["xv"], d["Pv"], np.full(tile, np.float32(1.0 / Pg), np.float32), d["lm_mean"], d["lm_cov"], Nf)
    stream = torch.cuda.Stream()
    f.set_stream(stream.cuda_stream)
    R = np.ascontiguousarray(d["R"], np.float32).ravel()
    Q = np.ascontiguousarray(Q_DEFAULT, np.float32).ravel()
    pinned = []

    def pin(shape, dtype):
        arr, ptr = pinned_array(L, shape, dtype)
        pinned.append(ptr)
        return arr, ptr

    # host noise pool in pinned memory (the reference draws 4 rand() per particle for the pose sample and 1 per
    # particle for the resample, every update): NPOOL sets, rotated
    NPOOL = 4 if full else 0
    rng = np.random.default_rng(777 + rank)
    pool = []
    for i in range(NPOOL):
        g, gp = pin((3, P), np.float32)
        u, up = pin((P,), np.int32)
        g[:] = rng.standard_normal((3, P), dtype=np.float32)
        u[:] = rng.integers(1, RAND_MAX, size=P, dtype=np.int32)
        pool.append((g, gp, u, up))
    # The controls and observation sets of every step are host INPUTS: the synthetic drive is run ahead of the timed
    # regions (its numpy argsort / trig per observe would otherwise sit on the critical path between two steps and be
    # measured as GPU idle time), and each step's zf / idf sit in their own pinned slots.
    drive = SyntheticDrive(d["lm"], K, d["R"], seed=12345)  # same observations on every rank
    n_legs = 7 if full else 3
    n_plan = n_legs * (warmup + steps) + burn_in + 4 + (warmup + max(200, 10 * steps) if full else 0)
    zall, zall_p = pin((n_plan, K, 2), np.float32)
    iall, iall_p = pin((n_plan, K), np.int32)
    plan = []
    for sidx in range(n_plan):
        ticks = [drive.controls() for _ in range(TICKS_PER_OBSERVE)]
        zall[sidx], iall[sidx] = drive.observe()
        plan.append(ticks)
    Qp, Rp = Q.ctypes.data, R.ctypes.data
    cursor = {"i": 0}
    state = {"resamples": 0, "steps": 0}

    def one_step(i, mode):
        k = cursor["i"]
        cursor["i"] = k + 1
        for (V, G, dt, phi) in plan[k]:
            f.predict_raw(V, G, Qp, dt, phi)
        zp = zall_p + k * K * 2 * 4
        ip = iall_p + k * K * 4
        if mode == "host":
            g, gp, u, up = pool[i % NPOOL]
            neff, did = f.update_raw(zp, ip, K, Rp, gp, up, unif_ints=True)
        else:  # "rng" (device generator) or "staged": no noise pointers
            neff, did = f.update_raw(zp, ip, K, Rp, None, None)
        est = f.estimate()
        state["resamples"] += int(did)
        state["steps"] += 1
        return neff, did, est

    def device_leg(mode, timing, steps=steps):
        for i in range(warmup):
            one_step(i, mode)
        if timing:
            f.timing_enable(True)
            f.timing_reset()
        state.update(resamples=0, steps=0)
        snap = dict(launches=f.launches, stats=f.stats())
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        trace = [] if os.environ.get("SLAM_BENCH_TRACE") else None
        cx.barrier()
        with torch.cuda.stream(stream):
            e0.record(stream)
            for i in range(steps):
                neff, did, _ = one_step(i, mode)
                if trace is not None:  # per-step device time and migration (diagnostics; perturbs the leg slightly)
                    ev = torch.cuda.Event(enable_timing=True)
                    ev.record(stream)
                    trace.append((ev, did, f.stats()["remote_sources"], f.stats()["rows_moved"]))
            e1.record(stream)
        cx.barrier()
        if trace:
            prev, pr, pm, rows = e0, snap["stats"]["remote_sources"], snap["stats"]["rows_moved"], []
            for ev, did, rs, mv in trace:
                rows.append("%.3f%s r%.2f m%d" % (prev.elapsed_time(ev), "*" if did else " ", (rs - pr) / P, mv - pm))
                prev, pr, pm = ev, rs, mv
            print("[rank %d] %s leg, ms per step (* = resampled, r = remote fraction, m = rows re-columned): %s"
                  % (rank, mode, " | ".join(rows)), file=sys.stderr, flush=True)
        ms = cx.max_over_ranks(e0.elapsed_time(e1)) / steps
        out = dict(ms_per_step=ms, value=Pg * K / (ms * 1e-3), launches=f.launches - snap["launches"],
                   resample_rate=state["resamples"] / max(1, state["steps"]))
        st1 = f.stats()
        out["stats"] = {k: st1[k] - snap["stats"][k] for k in st1}
        if timing:
            out["fam"] = f.timing_read()
            f.timing_enable(False)
        return out

    def wall_leg(mode):
        for i in range(min(warmup, 2)):
            one_step(i, mode)
        cx.barrier()
        t0 = time.perf_counter()
        for i in range(steps):
            one_step(i, mode)
        cx.barrier()
        s = cx.max_over_ranks(time.perf_counter() - t0) / steps
        return dict(ms_per_step=s * 1e3, value=Pg * K / s)

    res = {"P": P, "Nf": Nf, "K": K, "Pg": Pg, "npool": NPOOL}
    f.seed(20261104)                       # throughput mode: the device draws what the reference draws inside update()
    # The clock sampler (nvidia-smi, polling) is started well BEFORE the timed legs and left running through them:
    # its start-up initialises NVML on every GPU of the node, which stalls running CUDA work for milliseconds (measured:
    # +0.9 ms per step over a 20-step leg at 8 GPUs when it was started right in front of the leg).
    clocks = ClockSampler(cx.local)
    if rank == 0:
        clocks.start()
        time.sleep(1.5)
    cx.barrier()
    res["cold"] = device_leg("rng", False)
    for i in range(burn_in):
        one_step(i, "rng")
    res["value"] = device_leg("rng", False)
    res["profiled"] = device_leg("rng", True)
    if full:
        # Migration between shards comes in bursts (a shard of sibling particles loses or gains weight as a whole), so
        # K = 20 steps are a noisy sample of the steady state at N > 1: the same leg over 10 K (>= 200) steps, beside it
        res["long"] = device_leg("rng", False, steps=max(200, 10 * steps))
    res["clocks"] = clocks.stop() if rank == 0 else None
    v = res["profiled"]
    remote = v["stats"]["remote_sources"] / max(1, v["stats"]["resamples"] * P)
    res["remote_frac"] = cx.max_over_ranks(remote) if world > 1 else 0.0
    if os.environ.get("SLAM_BENCH_RANK_REPORT"):  # per-rank family times (the JSON line carries rank 0's)
        print("[rank %d] Nf=%d ms/step by family: %s  remote sources/resample: %.3f"
              % (rank, Nf, {k: round(x["ms"] / steps, 4) for k, x in v["fam"].items()}, remote), file=sys.stderr, flush=True)
    if full:
        f.stage_noise(pool[0][0], pool[0][2] / float(RAND_MAX))
        res["staged"] = device_leg("staged", False)
        f.seed(20261104)
        res["e2e"] = wall_leg("host")
        res["e2e"].update(h2d_bytes_per_step=3 * P * 4 + P * 4 + K * 12, d2h_bytes_per_step=32 + 48)
        res["e2e_rng"] = wall_leg("rng")
        res["e2e_rng"].update(h2d_bytes_per_step=K * 12, d2h_bytes_per_step=32 + 48)
    f.close()
    for ptr in pinned:
        L.slamgpu_host_free(ptr)
    return res


def roofline_of(res, peak, peak_src):
    """Roofline of the dominant kernel family: algorithmic bytes (update 48 B per particle-landmark + 116 B
    per particle; predict 96 B per particle per flush of queued ticks; state resampling 104 B per particle
    per resample plus 48 B per particle per landmark row re-columned outside the update kernel) over the family's
    device time, CUDA events inside the timed region."""
    P, K, Nf = res["P"], res["K"], res["Nf"]
    v = res["profiled"]
    fam, st = v["fam"], v["stats"]
    fam_ms = {k: x["ms"] for k, x in fam.items()}
    alg_total = {"update": P * (48 * K + 116) * fam["update"]["events"],
                 "predict": P * 96 * fam["predict"]["events"],
                 "gather": P * (104 * st["resamples"] + 48 * (st["rows_moved"] - st["rows_fused"]))}
    top = max(("update", "gather", "predict", "weights"), key=lambda k: fam_ms[k])
    roof = {}
    for k in ("update", "gather", "predict"):
        ev = fam[k]["events"]
        if ev and fam_ms[k] > 0:
            gbs = alg_total[k] / (fam_ms[k] * 1e-3) / 1e9
            roof[k] = {"achieved": gbs, "frac": gbs / peak, "ms_per_event": fam_ms[k] / ev, "events": ev,
                       "launches": fam[k]["launches"], "alg_bytes_per_event": alg_total[k] / ev}
    kern = {"update": "k_fs2_update_fast<REMAP> (cp.async landmark ring)",
            "gather": "k_gather_pose + k_chain + k_materialise_rows", "predict": "k_predict_batch_fast",
            "weights": "k_weights_fused + k_search_keep"}[top]
    traffic, traffic_src = None, None
    tp = os.path.join(ROOT, "profiles", "r02_fs2_update_traffic.json")
    if top == "update" and (P, Nf, K) == (1 << 20, 1024, 32) and os.path.exists(tp):
        try:  # dram__bytes_read.sum + dram__bytes_write.sum per launch from the committed ncu --set full capture
            t = json.load(open(tp))
            traffic, traffic_src = t["dram_bytes_per_launch"], "ncu capture profiles/" + t.get("source", "r02_fs2_update_traffic.json")
        except Exception:
            traffic = None
    return {"bound": "hbm", "kernel": kern, "achieved": roof.get(top, {}).get("achieved"), "peak": peak,
            "unit": "GB/s", "frac": roof.get(top, {}).get("frac"), "traffic": traffic, "traffic_source": traffic_src,
            "peak_source": peak_src, "families": roof, "family_ms_total": fam_ms,
            "ms_per_step_with_events": v["ms_per_step"],
            "family_ms_per_step": {k: x["ms"] / max(1, res["value_steps"]) for k, x in fam.items()}}


def parity_phase(cx, slam_b200, Nf=40):
    """Untimed: N ranks x 4096 particles x 40 landmarks x 8 observations through slamgpu_update (fused re-columning,
    in-kernel exchanges, migration over peer memory), weights skewed so that a large part of every resample's sources
    sits on another rank, compared on rank 0 with the ORACLE run in one piece: keep[], nEff, the decision and the
    normalised weights bit for bit every cycle, the state after the last cycle within 1e-4.  The oracle is the checker
    here, nothing timed touches it."""
    import sys as _sys
    _sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle.pyoracle import Oracle, State
    from replay import compare_states, snapshot
    dist, world, rank = cx.dist, cx.world, cx.rank
    Pl, K, CYCLES = 4096, 8, 4
    Pg = Pl * world
    d = synthetic_state(Pg, Nf, K, seed=21, extent=40.0)
    n_eff = int(0.75 * Pg)
    f = cx.new_handle(slam_b200, Pl, Nf, n_eff)
    sl = slice(rank * Pl, (rank + 1) * Pl)
    c = np.ascontiguousarray
    skew = np.exp(3.0 * np.arange(Pg) / Pg).astype(np.float32)
    skew /= skew.sum(dtype=np.float64).astype(np.float32)
    f.upload(c(d["xv"][:, sl]), c(d["Pv"][:, sl]), c(skew[sl]), c(d["lm_mean"][:, sl]), c(d["lm_cov"][:, sl]), Nf)
    f.debug_capture(True)
    drive = SyntheticDrive(d["lm"], K, d["R"], seed=3)
    rng = np.random.default_rng(5)
    none = np.zeros((0, 2), np.float32)
    o = Oracle() if rank == 0 else None
    s = None
    if rank == 0:
        s = State(Pg, Nf)
        s.xv[:], s.Pv[:], s.w[:], s.lm_mean[:], s.lm_cov[:], s.Nf = d["xv"], d["Pv"], skew, d["lm_mean"], d["lm_cov"], Nf
    out = {"ranks": world, "particles": Pg, "landmarks": Nf, "observations": K, "cycles": CYCLES, "keep_bit_exact": True,
           "weights_bit_exact": True, "neff_bit_exact": True, "decision_equal": True, "max_rel_weight": 0.0,
           "poses_within_1e-4": True, "resamples": 0}

    def gather(x):
        if world == 1:
            return [x]
        parts = [None] * world
        dist.all_gather_object(parts, x)
        return parts

    for cyc in range(CYCLES):
        if cyc == 2:  # skew the weights again (weights only: the landmark rows stay as the resamples left them)
            f.upload(w=c(skew[sl]))
            if rank == 0:
                s.w[:] = skew
        ticks = [drive.controls() for _ in range(TICKS_PER_OBSERVE)]  # fewer ticks leave Pv (zeroed by the pose sample) near rank 2
        zf, idf = drive.observe()
        gauss = rng.standard_normal((3, Pg)).astype(np.float32)
        ints = rng.integers(1, RAND_MAX, size=Pg).astype(np.int32)
        for (V, G, dt, phi) in ticks:
            f.predict(V, G, Q_DEFAULT, dt, phi)
        if cyc % 2:
            neff, did = f.update_i32(zf, idf, none, d["R"], c(gauss[:, sl]), c(ints[sl]))
        else:
            neff, did = f.update(zf, idf, none, d["R"], c(gauss[:, sl]), c(ints[sl] / float(RAND_MAX)))
        parts = gather(dict(wb=f.debug_read(0), wn=f.debug_read(1), keep=f.debug_read(2), neff=neff, did=did))
        if rank == 0:
            for (V, G, dt, phi) in ticks:
                o.predict(2, s, V, G, Q_DEFAULT, dt, phi, WHEELBASE, SIGMA_PHI, True, None)
            o.update_particles(2, s, zf, idf, none, d["R"], gauss)
            wb = np.concatenate([p["wb"] for p in parts])
            out["max_rel_weight"] = max(out["max_rel_weight"],
                                        float(np.abs(wb.astype(np.float64) - s.w).max() / max(float(s.w.max()), 1e-300)))
            wn_o, keep_o, neff_o = o.resample_indices(wb, ints / float(RAND_MAX), scan_mode=1)
            did_o = bool(neff_o < n_eff)
            out["keep_bit_exact"] &= bool(np.array_equal(np.concatenate([p["keep"] for p in parts]), keep_o))
            out["weights_bit_exact"] &= bool(np.array_equal(np.concatenate([p["wn"] for p in parts]), wn_o))
            out["neff_bit_exact"] &= all(p["neff"] == float(neff_o) for p in parts)
            out["decision_equal"] &= all(p["did"] == did_o for p in parts)
            s.w[:] = wn_o
            if did_o:
                o.resample_gather(s, keep_o)
                out["resamples"] += 1
        # poses after the cycle (no landmark rows are touched: the next update still re-columns inside the kernel)
        pp = gather(f.download(landmarks=False))
        if rank == 0:
            gp = {k: np.concatenate([p[k] for p in pp], axis=(0 if k == "w" else 1)) for k in ("xv", "Pv", "w")}
            gp.update(lm_mean=s.lm_mean[:Nf], lm_cov=s.lm_cov[:Nf], Nf=Nf)
            try:
                errs = compare_states(gp, snapshot(s), 1e-4, "parity phase cycle %d poses" % cyc, tol_w=1e-4)
                out["max_rel_pose"] = max(out.get("max_rel_pose", 0.0), errs["xy"], errs["Pv"], errs["theta"])
            except AssertionError as e:
                out["state_error"] = str(e)[:300]
                out["poses_within_1e-4"] = False
            # the pose sample and the weights amplify differences, the landmark filters do not: the oracle continues
            # from the device's poses, its landmark rows run free for all cycles and are compared at the end
            s.xv[:], s.Pv[:] = gp["xv"], gp["Pv"]
    st = f.stats()
    remote = cx.max_over_ranks(st["remote_sources"] / max(1, st["resamples"] * Pl)) if world > 1 else 0.0
    fused = st["rows_fused"]
    parts = gather(f.download())
    ok = True
    if rank == 0:
        full = {k: np.concatenate([p[k] for p in parts], axis=(0 if k == "w" else 1)) for k in
                ("xv", "Pv", "w", "lm_mean", "lm_cov")}
        full["Nf"] = Nf
        try:
            errs = compare_states(full, snapshot(s), 1e-4, "parity phase", tol_w=1e-4)
            out["max_rel_state"] = max(v for k, v in errs.items() if k != "w")
            out["state_within_1e-4"] = True
        except AssertionError as e:
            out["state_within_1e-4"] = False
            out["state_error"] = str(e)[:300]
        out["remote_fraction"] = remote
        out["rows_fused"] = int(fused)
        ok = all(out[k] for k in ("keep_bit_exact", "weights_bit_exact", "neff_bit_exact", "decision_equal",
                                  "poses_within_1e-4", "state_within_1e-4")) and out["resamples"] >= 2
        out["ok"] = bool(ok)
    f.close()
    return out, ok


def cpu_model():
    try:
        for line in open("/proc/cpuinfo"):
            if line.startswith("model name"):
                return line.split(":", 1)[1].strip()
    except Exception:
        pass
    return "unknown"


def run_ours(a):
    cx = Ctx()
    import slam_b200
    from slam_b200 import _lib
    world, rank = cx.world, cx.rank
    if world != a.gpus and world == 1 and a.gpus > 1:
        raise SystemExit("--gpus %d needs torchrun (one rank per GPU)" % a.gpus)
    slam_b200.build()
    L = _lib.load()
    P, Nf, K = a.particles, a.landmarks, a.obs
    res = measure(cx, a, L, slam_b200, P, Nf, K, a.steps, a.warmup, a.burn_in, True)
    res["value_steps"] = a.steps
    peak, peak_src = peaks()
    roofline = roofline_of(res, peak, peak_src)
    c5 = None
    if world > 1 and not a.no_config5 and Nf != 4096:
        # BASELINE.json configs[4] as named: 2^20 particles per GPU x 4096 landmarks (103 GB of map per GPU), sharded
        s5 = max(3, min(a.steps, 10))
        r5 = measure(cx, a, L, slam_b200, P, 4096, K, s5, 3, 0, False)
        r5["value_steps"] = s5
        rf5 = roofline_of(r5, peak, peak_src)
        c5 = {"workload": "FastSLAM2 synthetic: %d particles/GPU x 4096 landmarks x %d observations/step (BASELINE.json "
                          "config 5 shape), %d GPUs" % (P, K, world), "steps": s5, "warmup": 3,
              "value": r5["value"]["value"], "ms_per_step": r5["value"]["ms_per_step"],
              "value_cold": r5["cold"]["value"], "map_bytes_per_gpu": P * (4096 + 64) * 24,
              "rows_recolumned": r5["value"]["stats"]["rows_moved"], "resamples": r5["value"]["stats"]["resamples"],
              "update_frac_of_hbm": rf5["families"].get("update", {}).get("frac"),
              "family_ms_per_step": rf5["family_ms_per_step"], "remote_fraction": r5["remote_frac"]}
    parity, parity_ok = (None, True) if a.no_parity else parity_phase(cx, slam_b200, a.parity_landmarks)

    cpu = None
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        sizes = []
        for Pc, st in ((100, 4), (1000, 3), (12000, 1)):
            v, ms, kind, sample = cpu_arm(Pc, Nf, K, 1, st, 1)
            sizes.append({"particles": Pc, "value": v, "ms_per_step": ms, "steps": st})
        mid = sizes[1]
        cpu = {"value": mid["value"], "unit": UNIT, "cores": 1, "kind": kind, "ms_per_step": mid["ms_per_step"],
               "sample": "1 instance x 1000 particles x %d landmarks, %d observed/step, 3 timed cycles (8 predict + update "
                         "+ resampleParticles); other sizes in `sizes`" % (Nf, K),
               "cpu_model": cpu_model(), "host_cores": os.cpu_count(), "sizes": sizes}

    if rank == 0:
        v, st = res["value"], res["value"]["stats"]
        line = {"metric": METRIC, "value": v["value"], "unit": UNIT, "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
                "ms_per_step": v["ms_per_step"], "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic",
                "value_cold": res["cold"]["value"], "ms_per_step_cold": res["cold"]["ms_per_step"],
                "value_staged_noise": res["staged"]["value"], "ms_per_step_staged_noise": res["staged"]["ms_per_step"],
                "steady_long": {"steps": max(200, 10 * a.steps), "value": res["long"]["value"],
                                "ms_per_step": res["long"]["ms_per_step"],
                                "resample_rate": res["long"]["resample_rate"],
                                "remote_fraction_rank0": res["long"]["stats"]["remote_sources"]
                                / max(1, res["long"]["stats"]["resamples"] * P)},
                "config": {"workload": "FastSLAM2 synthetic: %d particles/GPU x %d landmarks x %d observations/step, "
                                       "8 predict ticks per update (BASELINE.json config 4 shape)" % (P, Nf, K),
                           "particles_total": res["Pg"], "parallelism": "particles block-sharded over %d GPU(s)" % world,
                           "burn_in_steps": a.burn_in,
                           "steady_state": "`value` is timed after %d untimed burn-in steps (lineages coalesced, shards "
                                           "drifted, re-column backlog built up); `value_cold` is the same leg from the "
                                           "balanced initial population" % a.burn_in,
                           "resample": "NEFFECTIVE = 0.75 N; gather ran on %.0f%% of timed steps" % (100 * v["resample_rate"]),
                           "l2": "inputs larger than L2 (the update alone streams %.1f GB per step)"
                                 % (P * (48 * K + 116) / 1e9),
                           "map_resampling": "deferred: %d landmark rows re-columned over %d resamples (eager copy: %d); "
                                             "%d of them inside the update kernel itself (fused, no extra pass)"
                                             % (st["rows_moved"], st["resamples"], st["resamples"] * Nf, st["rows_fused"]),
                           "migration": "%.1f%% of resampled particles came from another GPU (worst rank)"
                                        % (100 * res["remote_frac"]),
                           "predict": "the 8 ticks per step are queued and applied in one pass",
                           "noise": "`value` / `value_cold` / `e2e_device_rng`: throughput mode — drawn on the device inside "
                                    "update like the reference draws it (Philox4x32-10, fresh every step); "
                                    "`value_staged_noise`: one host-drawn set staged in HBM and reused (round 1's leg); "
                                    "`e2e`: pre-drawn on the host, %d pinned sets rotated, 12 B normals + 4 B rand() int "
                                    "per particle uploaded every step" % res["npool"]},
                "clocks": res["clocks"], "gpu_launches": int(v["launches"]),
                "gpu_launches_per_step": v["launches"] / a.steps,
                "e2e": {"value": res["e2e"]["value"], "unit": UNIT,
                        "h2d_bytes_per_step": res["e2e"]["h2d_bytes_per_step"],
                        "d2h_bytes_per_step": res["e2e"]["d2h_bytes_per_step"], "ms_per_step": res["e2e"]["ms_per_step"],
                        "noise": "host (parity mode)"},
                "e2e_device_rng": {"value": res["e2e_rng"]["value"], "unit": UNIT,
                                   "h2d_bytes_per_step": res["e2e_rng"]["h2d_bytes_per_step"],
                                   "d2h_bytes_per_step": res["e2e_rng"]["d2h_bytes_per_step"],
                                   "ms_per_step": res["e2e_rng"]["ms_per_step"], "noise": "device generator"},
                "roofline": roofline}
        if c5:
            line["config5"] = c5
        if parity:
            line["parity"] = parity
        if cpu:
            line["cpu_baseline"] = cpu
        print(json.dumps(line), flush=True)
    if world > 1:
        cx.dist.destroy_process_group()
    if not parity_ok:
        raise SystemExit("bench.py: the parity phase FAILED (see `parity` in the JSON line)")


if __name__ == "__main__":
    args = parse()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)
