"""
Data-parallel exchange for the WGAN-GP step: one process per GPU, torch.distributed (NCCL over NVLink / NVSwitch on
the GPU box, gloo in the CPU tests of the host logic).  The reference is single-GPU (run.py:28-31); parity is defined
against the single-device step on the concatenated batch:

  * each rank runs B local utterances of the global batch, every mean is scaled by 1/global count,
  * BatchNorm statistics [sum x, sum x^2] and [sum dz, sum dz*xhat] are all-reduced (sync-BN), so that the generator
    sees the statistics of the whole batch, as the reference does,
  * the flat critic / generator gradient buffers are all-reduced (SUM) right before Adam.
"""
import ctypes
import os

import torch
import torch.distributed as dist

from . import _lib

PEER_NMAX = 1024      # doubles per exchange through peer memory ([2, C] BatchNorm sums, C <= 512)


class DistSync(object):
    def __init__(self, group=None):
        if not dist.is_initialized():
            raise RuntimeError('torch.distributed is not initialised')
        self.group = group
        self.world = dist.get_world_size(group)
        self.rank = dist.get_rank(group)
        self.calls = 0
        self.peer_calls = 0
        self._peer = None
        self._bn = None            # (bases, err) of the sliced sync-BatchNorm protocol (second symmetric buffer)
        self._bn2 = None           # the same for the exchanges issued on the step's second side lane
        self._gs = {}              # id(ParamStore) -> (grad bases, npad): flat gradient buffers in symmetric memory
        self._gs_ctl = None        # control blocks of the fused all-reduce + Adam launch
        self.fused_adam_calls = 0
        self.peer_error = None

    def allreduce(self, t):
        if self.world > 1:
            if self._peer is not None and t.dtype == torch.float64 and t.is_contiguous() and 0 < t.numel() <= PEER_NMAX:
                # sync-BatchNorm sums: one-shot exchange through NVLink peer memory (csrc/peer.cu), not an NCCL launch
                lib, bases, err = self._peer
                _lib.check(lib.pctts_peer_allreduce_f64(t.data_ptr(), t.numel(), ctypes.cast(bases, ctypes.c_void_p), self.rank,
                                                        self.world, PEER_NMAX, err.data_ptr(),
                                                        torch.cuda.current_stream(t.device).cuda_stream),
                           'pctts_peer_allreduce_f64')
                self.peer_calls += 1
            else:
                dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
            self.calls += 1
        return t

    def enable_peer(self, device):
        """Map a symmetric buffer across the ranks (torch.distributed._symmetric_memory over NVLink) for the small fp64
        exchanges; returns False (and NCCL keeps doing them) when the platform cannot.  Collective: every rank calls it."""
        if self.world < 2 or device.type != 'cuda' or os.environ.get('PCTTS_PEER_AR', '1') == '0':
            return False
        ok = torch.zeros(1, device=device)
        try:
            import torch.distributed._symmetric_memory as symm_mem
            lib = _lib.load()
            nbytes = lib.pctts_peer_allreduce_bytes(self.world, PEER_NMAX)
            buf = symm_mem.empty((nbytes + 7) // 8, dtype=torch.float64, device=device)
            group = self.group if self.group is not None else dist.group.WORLD
            hdl = symm_mem.rendezvous(buf, group.group_name)
            buf.zero_()
            bases = (ctypes.c_ulonglong * self.world)(*[int(p) for p in hdl.buffer_ptrs])
            err = torch.zeros(1, dtype=torch.int32, device=device)
            # second symmetric buffer: BatchNorm second stage + exchange + finalize in one launch (one slot set per 32 channels)
            nb2 = lib.pctts_bn_sync_bytes(self.world)
            buf2 = symm_mem.empty((nb2 + 7) // 8, dtype=torch.float64, device=device)
            hdl2 = symm_mem.rendezvous(buf2, group.group_name)
            buf2.zero_()
            bases2 = (ctypes.c_ulonglong * self.world)(*[int(p) for p in hdl2.buffer_ptrs])
            # a second channel of the same protocol for the step's second side lane (engine.Generator: the noise-mask branch beside
            # the spectral branch): every rank meets the exchanges of ONE channel in the same order, two lanes need two channels
            buf3 = symm_mem.empty((nb2 + 7) // 8, dtype=torch.float64, device=device)
            hdl3 = symm_mem.rendezvous(buf3, group.group_name)
            buf3.zero_()
            bases3 = (ctypes.c_ulonglong * self.world)(*[int(p) for p in hdl3.buffer_ptrs])
            self._peer_keep = (buf, hdl, buf2, hdl2, buf3, hdl3)
            peer = (lib, bases, err)
            bn = (bases2, err)
            bn2 = (bases3, err)
            ok.fill_(1.0)
        except Exception as e:      # noqa: BLE001 -- any failure here means "use NCCL"
            peer = bn = bn2 = None
            self.peer_error = repr(e)
        torch.cuda.synchronize(device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)      # all ranks or none (also the barrier after zeroing)
        if float(ok) < 1.0:
            return False
        self._peer = peer
        # first exchange outside any capture: checks the mapping end to end
        t = torch.full((4,), float(self.rank + 1), dtype=torch.float64, device=device)
        self.allreduce(t)
        torch.cuda.synchronize(device)
        good = bool((t == self.world * (self.world + 1) / 2.0).all()) and int(peer[2]) == 0
        flag = torch.tensor([1.0 if good else 0.0], device=device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)
        if float(flag) < 1.0:
            self._peer = None
            return False
        if os.environ.get('PCTTS_BN_SYNC_FUSED', '1') != '0':
            self._bn = bn
            self._bn2 = bn2
        return True

    def enable_grad_sync(self, stores, device):
        """Move the flat gradient buffers of `stores` (engine.ParamStore) into symmetric memory so that the gradient all-reduce and
        the Adam update run as ONE launch over NVLink peer memory (csrc/peer.cu: pctts_grad_sync_adam) instead of an NCCL
        all-reduce followed by adam_kernel.  Collective; needs enable_peer() to have succeeded.  Returns False (NCCL + Adam stay)
        when the platform cannot or PCTTS_FUSED_ADAM=0."""
        if self.world < 2 or self._peer is None or getattr(device, 'type', 'cpu') != 'cuda' or os.environ.get('PCTTS_FUSED_ADAM', '1') == '0':
            return False
        ok = torch.zeros(1, device=device)
        made = None
        try:
            import torch.distributed._symmetric_memory as symm_mem
            lib = self._peer[0]
            group = self.group if self.group is not None else dist.group.WORLD
            keep, table = [], {}
            for st in stores:
                npad = max((st.ntrain + 3) // 4 * 4, 4)
                buf = symm_mem.empty(npad, dtype=torch.float32, device=device)
                hdl = symm_mem.rendezvous(buf, group.group_name)
                buf.zero_()
                keep.append((buf, hdl))
                table[id(st)] = ((ctypes.c_ulonglong * self.world)(*[int(p) for p in hdl.buffer_ptrs]), npad, buf)
            nctl = int(lib.pctts_grad_sync_ctl_bytes())
            ctl = symm_mem.empty((nctl + 7) // 8, dtype=torch.float64, device=device)
            hctl = symm_mem.rendezvous(ctl, group.group_name)
            ctl.zero_()
            keep.append((ctl, hctl))
            made = (table, (ctypes.c_ulonglong * self.world)(*[int(p) for p in hctl.buffer_ptrs]), keep)
            ok.fill_(1.0)
        except Exception as e:      # noqa: BLE001 -- any failure here means "NCCL + separate Adam"
            self.peer_error = repr(e)
        torch.cuda.synchronize(device)
        dist.all_reduce(ok, op=dist.ReduceOp.MIN, group=self.group)      # all ranks or none (also the barrier after zeroing)
        if float(ok) < 1.0:
            return False
        table, ctl_bases, keep = made
        for st in stores:
            st.use_grad_buffer(table[id(st)][2])
        self._gs, self._gs_ctl, self._gs_keep = table, ctl_bases, keep
        return True

    def grad_adam(self, store, lr, beta1, beta2, eps, losses=None):
        """All-reduce(sum) of store.grad over the replicas followed by store's Adam update: one fused launch over peer memory where
        enable_grad_sync() set it up, NCCL all-reduce + adam_kernel otherwise.  losses: a slice of the step's loss vector (local
        shares of the global means) summed over the replicas on the way."""
        ent = self._gs.get(id(store))
        if ent is None or store.ntrain == 0:
            if losses is not None:
                self.allreduce(losses)       # local shares -> global means (logging only)
            self.allreduce(store.grad)
            store.adam(lr, beta1, beta2, eps)
            return
        assert losses is None or (losses.dtype == torch.float32 and losses.is_contiguous() and losses.numel() <= 8)
        bases, npad, _ = ent
        lib, _, err = self._peer
        n = store.ntrain
        _lib.check(lib.pctts_grad_sync_adam(ctypes.cast(bases, ctypes.c_void_p), ctypes.cast(self._gs_ctl, ctypes.c_void_p), self.rank,
                                            self.world, store.flat.data_ptr(), store.m.data_ptr(), store.v.data_ptr(), n, npad,
                                            float(lr), float(beta1), float(beta2), float(eps), store.step.data_ptr(),
                                            losses.data_ptr() if losses is not None else None, losses.numel() if losses is not None else 0,
                                            err.data_ptr(),
                                            torch.cuda.current_stream(store.flat.device).cuda_stream), 'pctts_grad_sync_adam')
        store.ops.launches += 1
        store.ops.stamp('grad_sync_adam')
        self.fused_adam_calls += 1
        self.calls += 1
        store.ops.refresh_persistent(store.flat)

    def broadcast_scalar(self, x, src=0):
        """The python float of rank `src` on every rank (model-selection decisions must agree across the replicas)."""
        dev = torch.device('cuda', torch.cuda.current_device()) if dist.get_backend(self.group) == 'nccl' else torch.device('cpu')
        t = torch.tensor([float(x)], dtype=torch.float64, device=dev)
        dist.broadcast(t, src=src, group=self.group)
        return float(t.item())

    def _host_device(self):
        return torch.device('cuda', torch.cuda.current_device()) if dist.get_backend(self.group) == 'nccl' else torch.device('cpu')

    def broadcast_rng_state(self, state, src=0):
        """numpy's global MT19937 state (np.random.get_state()) of rank `src`, on every rank: the replicas draw the same shuffles
        and crop offsets, as the single process of the reference does (seed 123, data.py:32)."""
        import numpy as np
        name, keys, pos, has_gauss, cached = state
        assert name == 'MT19937' and len(keys) == 624
        t = torch.empty(627, dtype=torch.float64)                  # uint32 words are exact in a double
        t[:624] = torch.from_numpy(np.asarray(keys, dtype=np.float64))
        t[624], t[625], t[626] = float(pos), float(has_gauss), float(cached)
        t = t.to(self._host_device())
        dist.broadcast(t, src=src, group=self.group)
        t = t.cpu()
        return ('MT19937', t[:624].numpy().astype(np.uint32), int(t[624]), int(t[625]), float(t[626]))

    def assert_same_shape(self, shape, what=''):
        """Raise on EVERY rank when the replicas disagree on `shape` (sync-BN divides by rows * world and the peer exchanges run in
        lock-step, so unequal local shapes are silently wrong first and a time-out later)."""
        v = torch.tensor([float(d) for d in shape], dtype=torch.float64, device=self._host_device())
        lo, hi = v.clone(), v.clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=self.group)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=self.group)
        if not torch.equal(lo, hi):
            raise RuntimeError('replicas disagree on the shape of %s: rank %d has %s, min %s, max %s'
                               % (what, self.rank, tuple(shape), lo.tolist(), hi.tolist()))

    def barrier(self):
        dist.barrier(group=self.group)

    @property
    def bn_handle(self):
        """(peer bases, err flag) of the fused sync-BatchNorm launches, or None (exchange as separate launches)."""
        return self._bn

    @property
    def bn_handle2(self):
        """The channel of the fused sync-BatchNorm launches issued on the step's second side lane, or None."""
        return self._bn2

    @property
    def peer_err(self):
        """Device int32 [1] raised by a timed-out peer exchange (None without peer memory)."""
        return self._peer[2] if self._peer is not None else None

    def check_peer(self):
        """Raise if a peer-memory exchange timed out since the last check (one 4-byte device read: call it where the step's
        losses are read anyway)."""
        if self._peer is not None and int(self._peer[2]) != 0:
            raise _lib.PcttsError('a sync-BatchNorm exchange over NVLink peer memory timed out waiting for a peer rank: the '
                                  'statistics of this step are invalid (rank %d of %d)' % (self.rank, self.world))

    def peer_failed(self):
        """True if an exchange timed out waiting for a peer (the step's results are then invalid)."""
        return self._peer is not None and int(self._peer[2]) != 0

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
    sync.enable_peer(device)
    return sync, device


def shard_batch(X, rank, world):
    """Utterances [r*B/N, (r+1)*B/N) of a global batch; B must divide evenly."""
    B = X.shape[0]
    if B % world:
        raise ValueError('global batch of {} utterances does not split over {} ranks'.format(B, world))
    n = B // world
    return X[rank * n:(rank + 1) * n]
