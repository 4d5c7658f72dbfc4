"""GPU, needs >= 2 devices (skips cleanly on one): the view-sharded window step of the real rasterizer on 2 ranks,
NCCL and the library's own peer-memory exchange, against one rank rendering every view (tests/multirank_window_check.py)."""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


@pytest.mark.parametrize("exchange", ["nccl", "auto", "peer-plain"])
def test_two_rank_window_gradients_equal_single_rank(exchange):
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
           "127.0.0.1", "--master-port", str(_free_port()), os.path.join(HERE, "multirank_window_check.py"), exchange]
    env = dict(os.environ)
    if exchange == "peer-plain":       # the peer load / store kernel instead of the NVSwitch multimem one
        cmd[-1] = "peer"
        env["SPLAT_NO_MULTIMEM"] = "1"
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, env=env)
    assert r.returncode == 0, (r.stdout[-2000:], r.stderr[-4000:])
    assert "MULTIRANK_OK world=2" in r.stdout, r.stdout[-2000:]
