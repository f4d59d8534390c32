"""Developer tool: after tools/run_final_1gpu.sh came back in gpurun_out/, copy the evidence into profiles/, rebuild
the ncu summaries, the traffic file and the parity-error summary, and refresh the numbers DESIGN.md / README.md quote."""
import collections
import json
import os
import re
import shutil
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
OUT, PROF = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")


def last_json(path):
    return json.loads(open(path).read().splitlines()[-1])


def main():
    for f in ("r02_bench_1gpu.json", "r02_bench_1gpu_200steps.json", "r02_bench_reference_arm.json",
              "r02_bench_nf4096_1gpu.json", "r02_fs1_secondary.json", "r02_launches.csv", "r02_gpu_tests.log"):
        if os.path.exists(os.path.join(OUT, f)):  # (a partial re-run, tools/run_profile_1gpu.sh, brings only some of them)
            shutil.copy(os.path.join(OUT, f), os.path.join(PROF, f))
    # ncu summaries
    txt = subprocess.run(["python", os.path.join(ROOT, "tools", "ncu_summary.py"), os.path.join(OUT, "r02_prof.ncu-rep")],
                         capture_output=True, text=True).stdout
    blocks = [b for b in txt.split("\n\n") if b.strip()]
    seen, fs2, other, traffic = collections.Counter(), [], [], []
    mul = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}
    for b in blocks:
        first = b.splitlines()[0]
        name = first.split(None, 2)[2] if len(first.split(None, 2)) > 2 else first
        short = re.sub(r"\(.*", "", name).strip()
        seen[short] += 1
        if "k_fs2_update_fast" in short:
            def val(metric):
                m = re.search(metric + r"\s+(\S+)\s+(\S+)", b)
                return m.group(1), float(m.group(2))
            ru, rd = val("dram__bytes_read.sum")
            wu, wr = val("dram__bytes_write.sum")
            traffic.append(dict(kernel=short, dram_bytes=rd * mul[ru] + wr * mul[wu], us=val("gpu__time_duration.sum")[1],
                                issue_pct=val("smsp__issue_active.avg.pct_of_peak_sustained_active")[1],
                                warp_inst=val("smsp__inst_executed.sum")[1]))
        if seen[short] <= 2:
            (fs2 if "k_fs2_update_fast" in short else other).append(b)
    hdr = ("ncu --set full --clock-control none inside steady-state steps of\n`python bench.py --steps 6 --warmup 3 --burn-in 30 "
           "--no-cpu-baseline --no-parity` (throughput mode: Philox in the kernel), round 2 final code.\nSummary by "
           "tools/ncu_summary.py (first two captured launches of each kernel / instantiation).\n\n")
    open(os.path.join(PROF, "r02_fs2_update_fast.txt"), "w").write(hdr + "\n\n".join(fs2) + "\n")
    open(os.path.join(PROF, "r02_other_kernels.txt"), "w").write(hdr + "\n\n".join(other) + "\n")
    avg = sum(t["dram_bytes"] for t in traffic) / max(1, len(traffic))
    json.dump({"kernel": "k_fs2_update_fast (instantiations as captured)", "dram_bytes_per_launch": avg,
               "launches_captured": len(traffic), "per_launch": traffic, "algorithmic_bytes": 1732247552,
               "source": "r02_fs2_update_fast.txt (ncu --set full of the final code)"},
              open(os.path.join(PROF, "r02_fs2_update_traffic.json"), "w"), indent=1)
    # parity errors
    mx, n = collections.defaultdict(lambda: collections.defaultdict(float)), collections.Counter()
    if not os.path.exists(os.path.join(OUT, "r02_parity_errors.log")):
        for t in traffic:
            print(t)
        print(dict(seen))
        return
    for line in open(os.path.join(OUT, "r02_parity_errors.log")):
        d = json.loads(line)
        key = d["context"] or d["what"].split(":")[0][:40]
        n[key] += 1
        for k, v in d["errs"].items():
            mx[key][k] = max(mx[key][k], v)
    with open(os.path.join(PROF, "r02_parity_errors_summary.txt"), "w") as f:
        f.write("Achieved errors of the CUDA path against the oracle / the reference's golden vectors on B200 (round 2, final "
                "code):\nmaximum over every compare_states() call of `pytest -m gpu` (SLAM_PARITY_LOG), norm-relative per object "
                "as in\nSURVEY 8c (xy, Pv, lm_mean, lm_cov), radians (theta), norm-relative to the largest weight (w).\n"
                "math=0: SLAMGPU_MATH_FAST, math=1: SLAMGPU_MATH_REFERENCE_ORDER.  The last context label also covers the\n"
                "comparisons of later tests that set none.\n\n")
        for k, v in mx.items():
            f.write("%-52s n=%-4d %s\n" % (k, n[k], "  ".join("%s %.1e" % (a, b) for a, b in v.items())))
    for t in traffic:
        print(t)
    print(dict(seen))


if __name__ == "__main__":
    main()
