"""CPU: the bench.py contract.  (1) The reference arm (`--impl reference`) runs here — it times the CPU implementation
(oracle/_ref when the reference was compiled in this container, else the oracle's C restatement) — and must print ONE
JSON line with the agreed keys.  (2) The committed line of the CUDA arm (profiles/, measured on B200) must carry every
key the contract names, with a roofline fraction that is achieved / peak and an end-to-end leg that moved bytes.
(3) Without a GPU the CUDA arm must fail loudly: no JSON line, no CPU fallback."""
import glob
import json
import os
import subprocess
import sys

import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
BASE_KEYS = ["metric", "value", "unit", "n_gpus", "steps", "warmup", "ms_per_step", "higher_is_better", "scaling",
             "vs_baseline", "dtype", "data", "config"]


def json_lines(text):
    return [json.loads(l) for l in text.splitlines() if l.startswith("{")]


def test_reference_arm_prints_one_contract_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1", "--warmup", "1",
                        "--ref-particles", "100", "--landmarks", "64", "--obs", "8"], capture_output=True, text=True,
                       timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = json_lines(r.stdout)
    assert len(lines) == 1
    d = lines[0]
    for k in BASE_KEYS + ["impl", "cpu_baseline", "e2e"]:
        assert k in d, k
    assert d["impl"] == "reference" and d["higher_is_better"] is True and d["value"] > 0
    assert d["cpu_baseline"]["kind"] in ("reference", "port") and d["cpu_baseline"]["cores"] >= 1
    assert d["cpu_baseline"]["value"] == d["value"] and d["cpu_baseline"]["sample"]
    assert d["e2e"] == {"value": d["value"], "unit": d["unit"], "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}
    assert "workload" in d["config"]


def test_reference_arm_other_ranks_stay_silent():
    env = dict(os.environ, RANK="1", WORLD_SIZE="2", LOCAL_RANK="1")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                        "--warmup", "1"], capture_output=True, text=True, timeout=300, cwd=ROOT, env=env)
    assert r.returncode == 0 and json_lines(r.stdout) == []


def test_committed_cuda_line_carries_the_contract():
    """The newest committed 1-GPU line of the CUDA arm (profiles/r02_bench_1gpu.json, measured on B200)."""
    f = os.path.join(ROOT, "profiles", "r02_bench_1gpu.json")
    assert os.path.exists(f)
    d = json_lines(open(f).read())[-1]
    for k in BASE_KEYS + ["clocks", "gpu_launches", "e2e", "roofline", "cpu_baseline", "value_cold", "e2e_device_rng",
                          "parity"]:
        assert k in d, (f, k)
    assert d["unit"] == "updates/s" and d["n_gpus"] == 1 and d["scaling"] == "weak" and d["vs_baseline"] is None
    assert d["dtype"] == "f32" and d["data"] == "synthetic" and "workload" in d["config"] and d["warmup"] >= 3
    assert d["config"]["burn_in_steps"] >= 100          # `value` is the steady-state figure, `value_cold` the other one
    assert d["gpu_launches"] > 0
    e = d["e2e"]
    assert e["unit"] == d["unit"] and e["h2d_bytes_per_step"] > 16e6 and e["d2h_bytes_per_step"] > 0
    # host noise goes up every step: the end-to-end leg cannot be ahead of the device-resident legs (each leg times its
    # own K steps of the run, and K = 20 steps differ by how many of them resample)
    best = max(d["value"], d["value_cold"], d["value_staged_noise"], d.get("steady_long", {}).get("value", 0))
    assert 0 < e["value"] <= best * 1.05
    assert d["e2e_device_rng"]["h2d_bytes_per_step"] < 4096
    rf = d["roofline"]
    assert rf["bound"] == "hbm" and rf["unit"] == "GB/s" and abs(rf["frac"] - rf["achieved"] / rf["peak"]) < 1e-9
    assert rf["traffic"] is None or (rf["traffic"] > 0 and rf["traffic_source"])
    assert 0.6 <= rf["frac"] <= 1.0  # north_star: >= 60 % of the measured HBM peak on one B200
    c = d["cpu_baseline"]
    assert c["kind"] in ("reference", "port") and c["cores"] >= 1 and c["value"] > 0 and c["sample"] and c["cpu_model"]
    assert [s["particles"] for s in c["sizes"]] == [100, 1000, 12000]
    assert not set(d["clocks"]["reasons"]) & {"hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown"}
    p = d["parity"]
    assert p["ok"] and p["keep_bit_exact"] and p["weights_bit_exact"] and p["neff_bit_exact"] and p["decision_equal"]
    assert p["max_rel_state"] <= 1e-4 and p["max_rel_pose"] <= 1e-4


def test_committed_multi_gpu_lines_carry_parity():
    """Every committed N > 1 line carries the oracle parity phase run at that rank count."""
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r02_bench_[2-8]gpu.json")))
    assert any("config5" in json_lines(open(f).read())[-1] for f in files)
    for f in files:
        d = json_lines(open(f).read())[-1]
        p = d["parity"]
        assert p["ranks"] == d["n_gpus"] and p["ok"] and p["keep_bit_exact"] and p["remote_fraction"] > 0.3, f
        if "config5" in d:  # the 4096-landmark leg (run at 8 ranks; --no-config5 at 2 and 4)
            assert d["config5"]["map_bytes_per_gpu"] > 100e9, f


def test_cuda_arm_fails_loudly_without_a_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--steps", "1", "--warmup", "1", "--no-cpu-baseline"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode != 0 and json_lines(r.stdout) == []
