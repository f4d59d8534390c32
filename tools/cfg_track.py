"""Developer tool (GPU box): runs host/slam-backend-gpu on one of BASELINE.json's configs rebuilt from
tests/golden/config_traces.npz and prints how well the estimate tracks the true path: python tools/cfg_track.py cfg0"""
import os, subprocess, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_gpu_configs import write_inputs
tag = sys.argv[1] if len(sys.argv) > 1 else "cfg0"
extra = sys.argv[2:]
g = np.load(os.path.join(ROOT, "tests", "golden", "config_traces.npz"))
d = tempfile.mkdtemp()
write_inputs(g, tag, d)
subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "host"), "all"])
r = subprocess.run([os.path.join(ROOT, "host", "slam-backend-gpu"), "-m", "map.mat", "-method", str(g[tag + "_method"]),
                    "-trace", os.path.join(d, "t.txt")] + extra, cwd=d, capture_output=True, text=True)
a = np.loadtxt(os.path.join(d, "t.txt"))
e = np.hypot(a[:, 4] - a[:, 1], a[:, 5] - a[:, 2])
bad = np.nonzero(e > 3.0)[0]
print(tag, os.environ.get("SLAMGPU_NO_COARSE"), os.environ.get("SLAMGPU_REMAP_MIN_ROWS"), extra,
      "rms %.3f max %.3f final %.3f first>3m at %s" % (np.sqrt((e ** 2).mean()), e.max(), e[-1], bad[0] if bad.size else None))
