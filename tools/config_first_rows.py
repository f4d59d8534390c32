"""Debug aid: first rows of host/slam-backend-gpu's trace on the fixture configs (see config_follow.py)."""
import os, subprocess, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests"))
from test_gpu_configs import write_inputs
g = np.load(os.path.join(ROOT, "tests", "golden", "config_traces.npz"))
exe = os.path.join(ROOT, "host", "slam-backend-gpu")
for tag in sys.argv[1:] or ("cfg0", "cfg2"):
    with tempfile.TemporaryDirectory() as d:
        write_inputs(g, tag, d)
        subprocess.run([exe, "-m", "map.mat", "-method", str(g[tag + "_method"]), "-trace", "t.txt", "-max_iter", "40"],
                       cwd=d, capture_output=True, text=True)
        print(tag)
        print(open(os.path.join(d, "t.txt")).read())
