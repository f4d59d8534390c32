"""GPU, >= 2 devices: the sharded resample (in-kernel peer-memory exchanges + peer-memory migration) equals the
one-piece oracle at 2, 4 and 8 ranks.  Skipped where the box has fewer GPUs (the driver's 1-GPU box: bench.py's own
`parity` phase is what carries multi-rank evidence there, and profiles/ keeps the logs of the runs made on 2/4/8-GPU
boxes); tests/test_sharded_protocol.py covers the exchange protocol on CPU with gloo."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_sharded_equals_one_piece(world):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
                        "--master-addr", "127.0.0.1", "--master-port", str(29533 + world),
                        os.path.join(HERE, "sharded_worker.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "SHARDED_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
