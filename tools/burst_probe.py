"""Developer tool (2+ GPUs, under torchrun): what a migration burst costs.  Every slot of rank 0 is made to descend from a
particle of rank 1 (rank 0's weights ~ 0), then one FastSLAM2 update runs on the freshly resampled set: the K observed
rows of every migrated particle have to come over NVLink.  Prints the device time of the update / gather families per
rank.  SLAMGPU_REMAP_MIN_ROWS=1000 makes the library re-column the rows in a separate pass instead of inside the update
kernel, for comparison.

    torchrun --nproc-per-node 2 tools/burst_probe.py"""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import slam_b200  # noqa: E402
from slam_b200.synthetic import Q_DEFAULT, SIGMA_PHI, WHEELBASE, SyntheticDrive, synthetic_state  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    P, Nf, K, tile = 1 << 20, 48, 32, 4096
    Pg = P * world
    d = synthetic_state(tile, Nf, K, seed=1 + rank, map_seed=1)
    f = slam_b200.FastSLAM(slam_b200.FASTSLAM2, P, Nf, wheel_base=WHEELBASE, sigma_phi=SIGMA_PHI, n_effective=Pg + 1,
                           use_heading=True, scan_mode=slam_b200.SCAN_BLOCKED, device=local)
    idbuf = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        idbuf.copy_(torch.frombuffer(bytearray(slam_b200.FastSLAM.comm_unique_id()), dtype=torch.uint8))
    dist.broadcast(idbuf, 0)
    f.comm_init(rank, world, idbuf.cpu().numpy().tobytes())
    f.upload_tiled(d["xv"], d["Pv"], np.full(tile, np.float32(1.0 / Pg), np.float32), d["lm_mean"], d["lm_cov"], Nf)
    f.seed(7)
    drive = SyntheticDrive(d["lm"], K, d["R"], seed=12345)
    none = np.zeros((0, 2), np.float32)
    out = []
    for trial in range(6):
        for _ in range(8):
            V, G, dt, phi = drive.controls()
            f.predict(V, G, Q_DEFAULT, dt, phi)
        zf, idf = drive.observe()
        burst = trial % 2 == 1
        # before the update: weights that send (burst) every slot of the lower ranks to the last rank's particles
        w = np.full(P, np.float32(1e-30 if (burst and rank < world - 1) else 1.0), np.float32)
        f.upload(w=w)
        f.timing_enable(True)
        f.timing_reset()
        neff, did = f.update(zf, idf, none, d["R"])       # resamples (n_effective = N + 1); the NEXT update pays
        t1 = f.timing_read()
        f.timing_reset()
        for _ in range(8):
            V, G, dt, phi = drive.controls()
            f.predict(V, G, Q_DEFAULT, dt, phi)
        zf, idf = drive.observe()
        f.upload(w=np.full(P, np.float32(1.0), np.float32))
        neff, did = f.update(zf, idf, none, d["R"])
        t2 = f.timing_read()
        st = f.stats()
        out.append("trial %d %s: update after it %.3f ms, gather %.3f ms (rows fused %d, moved %d)"
                   % (trial, "BURST" if burst else "calm ", t2["update"]["ms"], t2["gather"]["ms"], st["rows_fused"], st["rows_moved"]))
    for r in range(world):
        dist.barrier()
        if r == rank:
            print("[rank %d] REMAP_MIN_ROWS=%s\n  " % (rank, os.environ.get("SLAMGPU_REMAP_MIN_ROWS")) + "\n  ".join(out), flush=True)
    f.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
