"""
Data-parallel exchange for the WGAN-GP step: one process per GPU, torch.distributed (NCCL over NVLink / NVSwitch on
the GPU box, gloo in the CPU tests of the host logic).  The reference is single-GPU (run.py:28-31); parity is defined
against the single-device step on the concatenated batch:

  * each rank runs B local utterances of the global batch, every mean is scaled by 1/global count,
  * BatchNorm statistics [sum x, sum x^2] and [sum dz, sum dz*xhat] are all-reduced (sync-BN), so that the generator
    sees the statistics of the whole batch, as the reference does,
  * the flat critic / generator gradient buffers are all-reduced (SUM) right before Adam.
"""
import os

import torch
import torch.distributed as dist


class DistSync(object):
    def __init__(self, group=None):
        if not dist.is_initialized():
            raise RuntimeError('torch.distributed is not initialised')
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.calls = 0

    def allreduce(self, t):
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
            self.calls += 1
        return t

    def warmup(self, device):
        """First collective outside any CUDA-graph capture (communicator setup is not capturable)."""
        t = torch.zeros(8, device=device)
        self.allreduce(t)
        if device.type == 'cuda':
            torch.cuda.synchronize(device)


def init_from_env(backend=None):
    """torchrun-style rendezvous (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_ADDR / MASTER_PORT) -> (sync, device)."""
    world = int(os.environ.get('WORLD_SIZE', '1'))
    local_rank = int(os.environ.get('LOCAL_RANK', '0'))
    if torch.cuda.is_available():
        torch.cuda.set_device(local_rank)
        device = torch.device('cuda', local_rank)
    else:
        device = torch.device('cpu')
    if world == 1:
        from .engine import LocalSync
        return LocalSync(), device
    if not dist.is_initialized():
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        os.environ.setdefault('MASTER_PORT', '29500')
        kw = {}
        if device.type == 'cuda':
            kw['device_id'] = device
        dist.init_process_group(backend or ('nccl' if device.type == 'cuda' else 'gloo'), **kw)
    sync = DistSync()
    sync.warmup(device)
    return sync, device


def shard_batch(X, rank, world):
    """Utterances [r*B/N, (r+1)*B/N) of a global batch; B must divide evenly."""
    B = X.shape[0]
    if B % world:
        raise ValueError('global batch of {} utterances does not split over {} ranks'.format(B, world))
    n = B // world
    return X[rank * n:(rank + 1) * n]
