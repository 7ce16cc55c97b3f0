"""Two ranks on two real GPUs (NCCL): the chunked work-stealing schedule of cgnn_b200/runtime.py - the replacement of
the reference's static `run % nb_gpu` placement (GNN.py:158-163) - on ragged pairs.  Skipped on a one-GPU box; the
gloo tests in test_distributed_host.py cover the same host logic with fake engines on CPU."""
import os
import subprocess
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.timeout(600)
def test_work_stealing_over_two_gpus():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
           "--master-port", str(_free_port()), os.path.join(ROOT, "tools", "check_stealing_nccl.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, cwd=ROOT, timeout=560)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "work stealing over 2 GPUs: OK" in r.stdout
