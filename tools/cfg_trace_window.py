"""Developer tool (GPU box): estimate error of host/slam-backend-gpu on a config around an iteration window."""
import os, subprocess, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_gpu_configs import write_inputs
tag, lo, hi = sys.argv[1], int(sys.argv[2]), int(sys.argv[3])
extra = sys.argv[4:]
g = np.load(os.path.join(ROOT, "tests", "golden", "config_traces.npz"))
d = tempfile.mkdtemp()
write_inputs(g, tag, d)
r = subprocess.run([os.path.join(ROOT, "host", "slam-backend-gpu"), "-m", "map.mat", "-method", str(g[tag + "_method"]),
                    "-trace", os.path.join(d, "t.txt"), "-max_iter", str(hi)] + extra, cwd=d, capture_output=True, text=True)
print(r.stdout[-600:], r.stderr[-600:])
a = np.loadtxt(os.path.join(d, "t.txt"))
for row in a[lo:hi:8]:
    e = np.hypot(row[4] - row[1], row[5] - row[2])
    print("it %5d true (%.2f %.2f %.3f) est (%.2f %.2f %.3f) err %.3f nobs %d" % (row[0], row[1], row[2], row[3], row[4], row[5], row[6], e, row[7]))
