"""GPU, >= 2 devices: the sharded resample (NCCL all-gathers + peer-memory migration) equals the one-piece oracle.
Skipped on a one-GPU box; tests/test_sharded_protocol.py covers the exchange protocol on CPU with gloo."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.mark.parametrize("world", [2])
def test_sharded_equals_one_piece(world):
    import torch
    if torch.cuda.device_count() < world:
        pytest.skip("needs %d GPUs" % world)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
                        "--master-addr", "127.0.0.1", "--master-port", "29533",
                        os.path.join(HERE, "sharded_worker.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "SHARDED_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
