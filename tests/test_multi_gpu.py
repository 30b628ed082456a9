"""Multi-GPU invariance through the library's own min-exchange of the placement keys (needs >= 2 GPUs; skipped
otherwise): the peer-memory kernel (k_exchange) and the NCCL all-reduce fallback.  Launches tests/mp_invariance.py
under torchrun, one rank per GPU."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("exchange", ["peer memory", "nccl"])
def test_min_exchange_matches_oracle(exchange):
    import torch
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs at least 2 GPUs")
    world = 2 if n < 4 else 4 if n < 8 else 8
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world),
           "--master-addr", "127.0.0.1", "--master-port", "29533" if exchange == "nccl" else "29534",
           os.path.join(ROOT, "tests", "mp_invariance.py")]
    env = dict(os.environ)
    if exchange == "nccl":
        env["GT_NO_P2P"] = "1"
    else:
        env.pop("GT_NO_P2P", None)
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=420, env=env)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "mp_invariance ok" in r.stdout and "exchange=%s" % exchange in r.stdout
