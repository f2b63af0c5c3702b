"""Multi-GPU parity (needs >= 2 visible GPUs, skipped otherwise): the fused output all-gather of the KV-head
tensor-parallel decode step, see tests/gpu_tp_peer.py."""
import os
import socket
import subprocess
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.gpu
def test_two_gpu_peer_gather_matches_nccl_and_oracle():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", str(port),
                        os.path.join(ROOT, "tests", "gpu_tp_peer.py")], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "TP_PEER_OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
