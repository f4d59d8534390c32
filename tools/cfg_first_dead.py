"""Developer tool (GPU box): first iteration at which the estimate's heading becomes exactly 0 (dead weights)."""
import os, subprocess, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_gpu_configs import write_inputs
tag = sys.argv[1]
extra = sys.argv[2:]
g = np.load(os.path.join(ROOT, "tests", "golden", "config_traces.npz"))
d = tempfile.mkdtemp()
write_inputs(g, tag, d)
r = subprocess.run([os.path.join(ROOT, "host", "slam-backend-gpu"), "-m", "map.mat", "-method", str(g[tag + "_method"]),
                    "-trace", os.path.join(d, "t.txt")] + extra, cwd=d, capture_output=True, text=True)
a = np.loadtxt(os.path.join(d, "t.txt"))
dead = np.nonzero(a[:, 6] == 0.0)[0]
dead = dead[dead > 50]
print("first dead row", dead[:5], "of", len(a))
if dead.size:
    i0 = dead[0]
    for row in a[max(0, i0 - 40):i0 + 10]:
        e = np.hypot(row[4] - row[1], row[5] - row[2])
        print("it %5d true (%.2f %.2f %.3f) est (%.2f %.2f %.4f) err %.3f nobs %d" % (row[0], row[1], row[2], row[3], row[4], row[5], row[6], e, row[7]))
