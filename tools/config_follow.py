"""Prints how far host/slam-backend-gpu follows the reference program's trace on BASELINE configs[0..2]
(tests/golden/config_traces.npz): first observation-tick row where the estimate differs by more than 1e-4."""
import os, subprocess, sys, tempfile, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_gpu_configs import write_inputs
g = np.load(os.path.join(ROOT, "tests", "golden", "config_traces.npz"))
exe = os.path.join(ROOT, "host", "slam-backend-gpu")
MATH = sys.argv[1] if len(sys.argv) > 1 else "fast"
for tag in ("cfg0", "cfg1", "cfg2"):
    with tempfile.TemporaryDirectory() as d:
        write_inputs(g, tag, d)
        t0 = time.time()
        r = subprocess.run([exe, "-m", "map.mat", "-method", str(g[tag + "_method"]), "-trace", "t.txt", "-math", MATH], cwd=d,
                           capture_output=True, text=True)
        dt = time.time() - t0
        a = np.loadtxt(os.path.join(d, "t.txt"))
    ref = g[tag + "_trace"]
    a8 = a[(a[:, 0].astype(int) % 8) == 0]
    n = min(len(a8), len(ref))
    scale = max(1.0, np.abs(ref[:, 1:3]).max())
    bad = np.nonzero((np.abs(a8[:n, 4:6] - ref[:n, 4:6]).max(axis=1) > 1e-4 * scale))[0]
    dpos = np.abs(a8[:n, 4:6] - ref[:n, 4:6]).max(axis=1)
    dth = np.abs((a8[:n, 6] - ref[:n, 6] + np.pi) % (2 * np.pi) - np.pi)
    pn = np.maximum(1.0, np.abs(ref[:n, 4:6]).max(axis=1))
    for name, bad_rows in (("pos>1e-4*|pose|", np.nonzero(dpos > 1e-4 * pn)[0]), ("theta>1e-4", np.nonzero(dth > 1e-4)[0]),
                           ("pos>1e-3", np.nonzero(dpos > 1e-3)[0]), ("theta>1e-3", np.nonzero(dth > 1e-3)[0])):
        print("   ", tag, name, "first at iteration", int(ref[bad_rows[0], 0]) if len(bad_rows) else None)
    print("    first rows dpos", np.array2string(dpos[:40], precision=6, max_line_width=200))
    print("    first rows dth ", np.array2string(dth[:40], precision=6, max_line_width=200))
    print(tag, str(g[tag + "_method"]), "rc", r.returncode, "%.1fs" % dt, "rows", a.shape[0], int(g[tag + "_iterations"]),
          "true path equal", np.array_equal(a8[:n, 1:4], ref[:n, 1:4]), "nobs equal", np.array_equal(a8[:n, 7], ref[:n, 7]),
          "first divergence at iteration", (int(ref[bad[0], 0]) if len(bad) else None),
          "max est err", np.abs(a8[:, 4:6] - a8[:, 1:3]).max(), "ref est err", np.abs(ref[:, 4:6] - ref[:, 1:3]).max(),
          r.stdout.strip().splitlines()[-1][:120] if r.stdout.strip() else "")
