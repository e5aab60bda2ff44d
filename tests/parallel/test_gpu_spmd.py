"""Runs gpu_spmd_script.py with one rank per visible GPU (needs >= 2 GPUs; skipped otherwise)."""
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
        s.bind(('127.0.0.1', 0))
        return s.getsockname()[1]


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs >= 2 GPUs")
def test_multi_gpu_spmd():
    n = min(torch.cuda.device_count(), 8)
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node=%d' % n,
           '--master-addr', '127.0.0.1', '--master-port', str(_free_port()),
           os.path.join(HERE, 'gpu_spmd_script.py'), str(_free_port())]
    res = subprocess.run(cmd, env=dict(os.environ, SPARK_LOCAL_IP='127.0.0.1'), capture_output=True, text=True,
                         timeout=900)
    assert res.returncode == 0 and 'ALL OK' in res.stdout, res.stdout[-3000:] + res.stderr[-4000:]
