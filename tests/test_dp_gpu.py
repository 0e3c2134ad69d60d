"""
Data-parallel parity on real GPUs: N ranks (one process per GPU, torchrun, NCCL + the NVLink peer-memory sync-BN
exchange), each with its shard of one global batch, against the CPU oracle's single-device step on the whole batch; the
CUDA-graph trainer path keeps the replicas bit-identical; 200-replay soak of the peer-memory all-reduce with its time-out flag
asserted 0; the public API (`torchrun run.py`-style train()) shards the batch itself.  Skips below 2 visible GPUs (the driver's
`-m gpu` run is single-GPU); run it with `gpurun --gpus N`.
"""
import os
import socket
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _torchrun(world, script, timeout=600):
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', str(world), '--master-addr', '127.0.0.1',
           '--master-port', str(_free_port()), os.path.join(ROOT, script)]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=timeout, cwd=ROOT)
    print(r.stdout[-4000:])
    print(r.stderr[-2000:])
    return r


@pytest.mark.parametrize('world', [2, 4, 8])
def test_dp_step_parity_and_replica_identity(world):
    if not torch.cuda.is_available() or torch.cuda.device_count() < world:
        pytest.skip('needs %d visible GPUs' % world)
    r = _torchrun(world, 'tools/dp_gpu_check.py')
    assert r.returncode == 0
    assert 'ALL OK' in r.stdout


@pytest.mark.parametrize('world', [2])
def test_dp_train_through_the_public_api(world):
    if not torch.cuda.is_available() or torch.cuda.device_count() < world:
        pytest.skip('needs %d visible GPUs' % world)
    r = _torchrun(world, 'tools/dp_api_check.py')
    assert r.returncode == 0
    assert 'DP API OK' in r.stdout
