"""
Backend shim: the module that replaces reference percivaltts/backend_tensorflow.py.  Where the reference entered one
global tf.Session at import (backend_tensorflow.py:38-39), this module hands out one process-wide CudaOps (libpctts.so
on the current CUDA device); there is no other backend and no CPU fallback.
"""
import numpy as np
import torch

_ops = None
_seed = 123            # backend_tensorflow.py:27-29 seeds numpy / TF with 123


def get_ops():
    """The process-wide kernel interface (raises without a B200 or without libpctts.so)."""
    global _ops
    if _ops is None:
        get_sync()                # under torchrun: rendezvous and torch.cuda.set_device(LOCAL_RANK) come first
        from .ops import CudaOps
        _ops = CudaOps()
    return _ops


def set_ops(ops):
    """Dependency injection for tests of the host logic (tests/cpu_ops_double.py)."""
    global _ops
    _ops = ops


_sync = None


def get_sync():
    """The process-wide data-parallel exchange: under torchrun (WORLD_SIZE > 1) torch.distributed is initialised on first use
    (NCCL, device = LOCAL_RANK) and a parallel.DistSync returned; otherwise None (single replica)."""
    global _sync
    import os
    if _sync is None and int(os.environ.get('WORLD_SIZE', '1')) > 1:
        from . import parallel
        _sync, _ = parallel.init_from_env()
    return _sync


def set_seed(seed):
    global _seed
    _seed = int(seed)
    np.random.seed(_seed)


def get_seed():
    return _seed


def cuda_available():
    """backend_tensorflow.py:55-57 (tf_cuda_available)"""
    return torch.cuda.is_available()


def gpu_memused():
    """MiB in use on the current device (backend_tensorflow.py:59-69, tf_gpu_memused)."""
    if not torch.cuda.is_available():
        return -1
    return int(torch.cuda.memory_allocated() // (1024 * 1024))


def nonlin_sigmoidparm(x, c=0.0, f=1.0):
    """backend_tensorflow.py:107-109: logistic with centre c and slope f (numpy, float64)."""
    x = np.asarray(x, dtype=np.float64)
    return 1.0 / (1.0 + np.exp(-(x - c) * f))
