"""Writes tests/data/ring_small.mat / .ini: a small world in the reference's text formats (core.cpp:855-962 map file,
utils.cpp:504-566 ini) for the host-loop parity tests.  Our own data — the reference's shipped maps do not travel."""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def main():
    rng = np.random.default_rng(7)
    n_lm, n_wp, radius = 28, 12, 14.0
    ang = np.sort(rng.uniform(0, 2 * np.pi, n_lm))
    rad = radius + rng.uniform(-5, 5, n_lm)
    lm = np.stack([rad * np.cos(ang), rad * np.sin(ang)], 1)
    wa = np.linspace(0, 2 * np.pi, n_wp, endpoint=False)
    wp = np.stack([radius * np.cos(wa) * 0.9, radius * np.sin(wa) * 0.9], 1)
    with open(os.path.join(HERE, "ring_small.mat"), "w") as f:
        f.write("# landmarks and waypoints of a ring world (tests/data/make_map.py)\n")
        f.write("lm 2 %d\n" % n_lm)
        for x, y in lm:
            f.write("%.6f %.6f\n" % (x, y))
        f.write("\nwp 2 %d\n" % n_wp)
        for x, y in wp:
            f.write("%.6f %.6f\n" % (x, y))
    with open(os.path.join(HERE, "ring_small.ini"), "w") as f:
        f.write("# ini of the ring world: keys as in the reference's data/*.ini\n"
                "Vtrue = 2.0\nMAXG = 0.6\nRATEG = 0.5\nWHEELBASE = 2\nDT_CONTROLS = 0.025\n"
                ": a colon starts a comment too\n"
                "sigmaV = 0.2\nsigmaG = 0.035\nMAX_RANGE = 12.0\nDT_OBSERVE = 0.2\n"
                "sigmaR = 0.1\nsigmaB = 0.017453292519943\nsigmaT = 0.017453292519943\n"
                "GATE_REJECT = 4.0\nGATE_AUGMENT = 25.0\nAT_WAYPOINT = 1.0\nNUMBER_LOOPS = 1\n"
                "NPARTICLES = 100\nNEFFECTIVE = 75\n"
                "SWITCH_CONTROL_NOISE = 1\nSWITCH_SENSOR_NOISE = 1\nSWITCH_INFLATE_NOISE = 0\nSWITCH_PREDICT_NOISE = 0\n"
                "SWITCH_SAMPLE_PROPOSAL = 1\nSWITCH_HEADING_KNOWN = 1\nSWITCH_RESAMPLE = 1\nSWITCH_PROFILE = 0\n"
                "SWITCH_SEED_RANDOM = 4242\nSWITCH_ASSOCIATION_KNOWN = 0\nSWITCH_BATCH_UPDATE = 1\nSWITCH_USE_IEKF = 0\n")


if __name__ == "__main__":
    main()
