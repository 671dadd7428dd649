"""Multi-GPU check of the peer-memory gradient exchange (3dgrut_b200/dist.py PeerExchange / training_step_peer) against the NCCL
formulation.  Needs two GPUs on the box; skipped otherwise (the single-GPU parts are in test_gpu_parity.py)."""
import os
import re
import socket
import subprocess
import sys

import pytest
import torch

from tests.helpers import ROOT

pytestmark = pytest.mark.gpu


@pytest.mark.timeout(600)
def test_peer_exchange_equals_nccl_exchange_on_two_gpus():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(port), os.path.join(ROOT, "scripts", "multi_gpu_peer_check.py")]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=540, cwd=ROOT)
    # the two ranks share one pipe and their lines can run into each other: count the reports, not the lines
    reports = re.findall(r"rank \d+ step \d+: rel L2", res.stdout)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    assert len(reports) == 6, res.stdout[-2000:]   # 2 ranks x 3 steps, each within 1e-5 of the NCCL formulation (the script exits 1 otherwise)
