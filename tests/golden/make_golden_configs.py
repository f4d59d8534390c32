"""Generates tests/golden/config_traces.npz: BASELINE.json configs[0..2] — the UNMODIFIED reference program
(oracle/_ref/slam-backend-ref, see make_golden_host.py) run on the reference's own example maps with their own
configuration files:

    cfg0  FASTSLAM1 on data/example_webmap.mat    (100 particles, known association)
    cfg1  FASTSLAM2 on data/example_loop1.mat     (100 particles)
    cfg2  EKF1      on data/example_loop902.mat   (117 landmarks in one joint state)

The fixture holds the INPUTS as vectors (landmark / waypoint coordinates and the KEY = value items of the .ini) and
the OUTPUT trace at every observation tick (each 8th iteration; the complete traces are 65 k lines), so that the GPU
test can rebuild the input files on the GPU box, where /root/reference does not exist.
Run in the build container only:  python tests/golden/make_golden_configs.py   (the EKF run takes ~3.5 min of CPU)"""
import os
import subprocess
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF_DATA = "/root/reference/data"
CONFIGS = [("cfg0", "example_webmap", "FASTSLAM1"), ("cfg1", "example_loop1", "FASTSLAM2"),
           ("cfg2", "example_loop902", "EKF1")]
EVERY = 8
SEED = 20240607  # the example .ini files say SWITCH_SEED_RANDOM = 0, i.e. srand(time(NULL)) (slamwrapper.cpp:48-52): a fixed
                 # seed is passed on the command line ("-KEY value" overrides any key) and recorded in the fixture's items


def read_map(path):
    """The numbers of a .mat file as text tokens (kept as text: the program parses them with strtof)."""
    blocks = {}
    with open(path) as f:
        lines = [l.split() for l in f if l.strip() and not l.lstrip().startswith("#")]
    i = 0
    while i < len(lines):
        name, rows, cols = lines[i][0], int(float(lines[i][1])), int(float(lines[i][2]))
        assert rows == 2
        blocks[name] = np.array([[t[0], t[1]] for t in lines[i + 1:i + 1 + cols]])
        i += 1 + cols
    return blocks["lm"], blocks["wp"]


def read_ini(path):
    items = []
    with open(path) as f:
        for l in f:
            l = l.strip()
            if not l or l[0] in "#:" or "=" not in l:
                continue
            k, v = l.split("=")[0:2]
            items.append("%s = %s" % (k.strip(), v.strip()))
    return np.array(items)


def main():
    exe = os.path.join(ROOT, "oracle", "_ref", "slam-backend-ref")
    out = {}
    for tag, name, method in CONFIGS:
        lm, wp = read_map(os.path.join(REF_DATA, name + ".mat"))
        with tempfile.NamedTemporaryFile(suffix=".txt") as t:
            env = dict(os.environ, SLAMREF_TRACE=t.name)
            subprocess.run([exe, "-m", name + ".mat", "-method", method, "-SWITCH_SEED_RANDOM", str(SEED)],
                           cwd=REF_DATA, env=env, check=True,
                           stdout=subprocess.DEVNULL)
            tr = np.loadtxt(t.name)
        out[tag + "_method"] = np.array(method)
        out[tag + "_lm"] = lm
        out[tag + "_wp"] = wp
        items = [i for i in read_ini(os.path.join(REF_DATA, name + ".ini")) if not i.startswith("SWITCH_SEED_RANDOM")]
        out[tag + "_ini"] = np.array(items + ["SWITCH_SEED_RANDOM = %d" % SEED])
        out[tag + "_iterations"] = np.array(tr.shape[0])
        out[tag + "_last"] = tr[-1]
        out[tag + "_trace"] = tr[(tr[:, 0].astype(int) % EVERY) == 0]
        print(tag, method, name, tr.shape, "->", out[tag + "_trace"].shape)
    np.savez_compressed(os.path.join(HERE, "config_traces.npz"), **out)


if __name__ == "__main__":
    main()
