"""Developer tool (GPU box): nEff / decision per update of host/slam-backend-gpu on a config (SLAM_HOST_DIAG)."""
import os, subprocess, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_gpu_configs import write_inputs
tag, max_iter = sys.argv[1], sys.argv[2]
g = np.load(os.path.join(ROOT, "tests", "golden", "config_traces.npz"))
d = tempfile.mkdtemp()
write_inputs(g, tag, d)
subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "host"), "all"])
r = subprocess.run([os.path.join(ROOT, "host", "slam-backend-gpu"), "-m", "map.mat", "-method", str(g[tag + "_method"]),
                    "-max_iter", max_iter] + sys.argv[3:], cwd=d, env=dict(os.environ, SLAM_HOST_DIAG="1"),
                   capture_output=True, text=True)
lines = [l for l in r.stderr.splitlines() if l.startswith("[diag]")]
print("\n".join(lines[-40:]))
