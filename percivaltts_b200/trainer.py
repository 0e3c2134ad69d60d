"""
Host<->device plumbing around the WGAN-GP stepper: pinned staging buffers, CUDA-graph capture of the critic step and
of the generator step (one graph each per batch shape), and the per-batch policy of
OptimizerTTSWGAN.train_on_batch (reference percivaltts/optimizertts_wgan.py:216-241).

One process drives one GPU; under torch.distributed every rank runs the same code on its shard of the batch.
"""
import concurrent.futures
import os
from collections import OrderedDict

import numpy as np
import torch

from . import wgan as W

STAGE_THREADS = 4     # host threads that fill the pinned staging buffer of a synchronously staged batch (Trainer.stage)


class _ShapeState(object):
    """Static device / pinned buffers and the captured graphs for one (B, T)."""

    def __init__(self, ops, a, B, T):
        self.B, self.T = B, T
        self.X = ops.empty(B, T, a.ctxsize)
        self.Y = ops.empty(B, T, a.outsize)
        self.alpha = ops.zeros(B)
        cuda = ops.device.type == 'cuda'
        self.Xh = torch.empty(B, T, a.ctxsize, dtype=torch.float32, pin_memory=cuda)
        self.Yh = torch.empty(B, T, a.outsize, dtype=torch.float32, pin_memory=cuda)
        self.lossh = torch.zeros(W.NLOSS, dtype=torch.float32, pin_memory=cuda)
        self.errh = torch.zeros(1, dtype=torch.int32, pin_memory=cuda)       # peer-exchange time-out flag, read with the losses
        self.g_critic = self.g_critic_alpha = self.g_gen = self.g_both = self.g_both_alpha = None


class DeviceTrainer(object):
    """Per batch shape (B, T) the trainer owns ONE entry -- static input buffers, the stepper's and both networks' buffers, the
    kernels' scratch recorded under the shape tag, and (for shapes that recur) three captured CUDA graphs.  The entries live in an
    LRU bounded by count and by measured device bytes: a training run meets many T (T = shortest utterance of the batch capped
    at lengthmax, data.py:250-262; validation one T per utterance), and evicting an entry frees all of it together.
    A shape is run eagerly the first `graph_after` times it is seen and captured after that (capture = warm-up + 3 captures,
    seconds against a ~3 ms step)."""

    def __init__(self, stepper, use_graphs=True, max_shapes=None, max_bytes=None, graph_after=1):
        self.st = stepper
        self.ops = stepper.ops
        self.use_graphs = bool(use_graphs) and self.ops.device.type == 'cuda'
        self.max_shapes = int(max_shapes if max_shapes is not None else os.environ.get('PCTTS_SHAPE_CACHE', '128'))
        self.max_bytes = int(max_bytes if max_bytes is not None else float(os.environ.get('PCTTS_SHAPE_CACHE_GB', '32')) * 2 ** 30)
        self.graph_after = int(graph_after)
        # critic + generator update of one batch share the generator's forward pass (wgan.critic_generator_step); PCTTS_FUSE_STEPS=0: two graphs
        self.fuse_steps = os.environ.get('PCTTS_FUSE_STEPS', '1') != '0'
        self._shapes = OrderedDict()
        self._seen = {}
        self.evictions = 0
        self.generator_updates = 0
        self.h2d_bytes = 0
        self.d2h_bytes = 0
        # double-buffered prefetch of the NEXT batch (train_on_batch(..., next_batch=)): pinned + device staging, own copy stream
        self._pf = None
        self._pf_placed = None     # shape state whose input buffers already hold the prefetched batch
        self._pf_slot = 0
        self._pf_bufs = [None, None]
        self._copy_stream = None
        self._stage_pool = None

    # ---- shape state / graph capture ---------------------------------------------------------------------------
    def _mem(self):
        return torch.cuda.memory_allocated(self.ops.device) if self.ops.device.type == 'cuda' else 0

    def _evict(self, keep):
        """Drop least-recently-used shapes until the cache is within its limits (never the entry `keep`)."""
        while len(self._shapes) > 1 and (len(self._shapes) > self.max_shapes or
                                         sum(e.bytes for e in self._shapes.values()) > self.max_bytes):
            key = next(k for k in self._shapes if k != keep)
            old = self._shapes.pop(key)
            if self._pf_placed is old:
                self._pf_placed = None
            old.g_critic = old.g_critic_alpha = old.g_gen = old.g_both = old.g_both_alpha = None      # the graphs (and their private pools) go first
            del old
            self.st.release(*key)
            self.evictions += 1

    def shape_state(self, B, T, capture=True):
        """The entry of (B, T), created on first use; capture=True also makes sure its CUDA graphs exist."""
        key = (B, T)
        s = self._shapes.get(key)
        if s is None:
            m0 = self._mem()
            s = self._shapes[key] = _ShapeState(self.ops, self.st.a, B, T)
            s.m0, s.bytes = m0, 0
        else:
            self._shapes.move_to_end(key)
        if capture and self.use_graphs and s.g_critic is None:
            self._capture(s)
            self._account(s)
        return s

    def _account(self, s):
        s.bytes = max(s.bytes, self._mem() - s.m0)
        self._evict(keep=(s.B, s.T))

    def _capture(self, s):
        """Warm up eagerly on a side stream (allocates every plan buffer and workspace; parameters, BN moving
        statistics and the RNG counter are restored afterwards), then capture the two graphs."""
        st = self.st
        keep_state = st.G.params.state.clone()
        keep_counter = st.counter.clone()
        if not getattr(s, 'filled', False):      # a fresh entry: finite inputs for the warm-up; an entry that already ran eagerly (or
            s.X.zero_()                          # holds the prefetched next batch) keeps its contents
            s.Y.zero_()
        side = torch.cuda.Stream(device=self.ops.device)
        side.wait_stream(torch.cuda.current_stream(self.ops.device))
        with torch.cuda.stream(side):
            st.critic_step(s.X, s.Y, alpha=None, apply=False)
            st.generator_step(s.X, s.Y, apply=False)
        torch.cuda.current_stream(self.ops.device).wait_stream(side)
        torch.cuda.synchronize(self.ops.device)
        st.G.params.state.copy_(keep_state)
        st.counter.copy_(keep_counter)
        s.g_critic = torch.cuda.CUDAGraph()
        with torch.cuda.graph(s.g_critic):
            st.critic_step(s.X, s.Y, alpha=None, apply=True)
        s.g_critic_alpha = torch.cuda.CUDAGraph()
        with torch.cuda.graph(s.g_critic_alpha):
            st.critic_step(s.X, s.Y, alpha=s.alpha, apply=True)
        s.g_gen = torch.cuda.CUDAGraph()
        with torch.cuda.graph(s.g_gen):
            st.generator_step(s.X, s.Y, apply=True)
        s.g_both = s.g_both_alpha = None
        if self.fuse_steps:       # critic + generator update of one batch as ONE graph sharing the generator's forward pass
            s.g_both = torch.cuda.CUDAGraph()
            with torch.cuda.graph(s.g_both):
                st.critic_generator_step(s.X, s.Y)
            s.g_both_alpha = torch.cuda.CUDAGraph()
            with torch.cuda.graph(s.g_both_alpha):
                st.critic_generator_step(s.X, s.Y, alpha=s.alpha)

    # ---- steps on resident device data ---------------------------------------------------------------------------
    def critic_step_device(self, s, explicit_alpha=False):
        if s.g_critic is not None:
            (s.g_critic_alpha if explicit_alpha else s.g_critic).replay()
        else:
            self.st.critic_step(s.X, s.Y, alpha=s.alpha if explicit_alpha else None, apply=True)
            self._account(s)

    def step_device(self, s, with_generator, explicit_alpha=False):
        """One train_on_batch on resident data: the critic update and, when due, the generator update on the same batch."""
        if with_generator and self.fuse_steps:
            if getattr(s, 'g_both', None) is not None:
                (s.g_both_alpha if explicit_alpha else s.g_both).replay()
            else:
                self.st.critic_generator_step(s.X, s.Y, alpha=s.alpha if explicit_alpha else None)
                self._account(s)
            self.generator_updates += 1
            return
        self.critic_step_device(s, explicit_alpha=explicit_alpha)
        if with_generator:
            self.generator_step_device(s)

    def generator_step_device(self, s):
        if s.g_gen is not None:
            s.g_gen.replay()
        else:
            self.st.generator_step(s.X, s.Y, apply=True)
            self._account(s)
        self.generator_updates += 1

    # ---- validation on resident device data (optimizertts_wgan.py:244-268) ------------------------------------------------
    def evaluate_device(self, Xd, Yd, out_row):
        """Inference-mode losses of one utterance batch (device tensors [B,T,*]) -> out_row [NLOSS] (device), no host sync.
        Validation shapes are run eagerly (one T per utterance, each seen once per epoch) through the same shape cache."""
        B, T = int(Xd.shape[0]), int(Xd.shape[1])
        s = self.shape_state(B, T, capture=False)
        self.st.evaluate_losses(Xd, Yd)
        out_row.copy_(self.st.losses, non_blocking=True)
        self._account(s)

    # ---- the reference's per-batch entry point ---------------------------------------------------------------------
    def stage(self, s, X, Y, alpha=None):
        """numpy [B,T,*] float32 -> pinned -> device (async on the current stream).  A batch of a megabyte or more is cut into
        groups of utterances: a small pool of threads fills the pinned buffer (numpy's copy releases the GIL) while this thread
        sends each group to the device as soon as it is in place, so the host copy runs on several cores and the PCIe transfer
        hides behind it -- this is the synchronous path of the reference's 3-argument train_on_batch(batchid, X, Y)."""
        X = np.asarray(X, dtype=np.float32).reshape(s.Xh.shape)
        Y = np.asarray(Y, dtype=np.float32).reshape(s.Yh.shape)
        Xh, Yh = s.Xh.numpy(), s.Yh.numpy()
        B = int(X.shape[0])
        if self.ops.device.type != 'cuda' or B < 2 or X.nbytes + Y.nbytes < (1 << 20):
            np.copyto(Xh, X)
            np.copyto(Yh, Y)
            s.X.copy_(s.Xh, non_blocking=True)
            s.Y.copy_(s.Yh, non_blocking=True)
        else:
            if self._stage_pool is None:
                self._stage_pool = concurrent.futures.ThreadPoolExecutor(max_workers=STAGE_THREADS, thread_name_prefix='pctts-stage')
            jobs = []
            for host, src, dev_t, host_t, parts in ((Xh, X, s.X, s.Xh, min(B, STAGE_THREADS)), (Yh, Y, s.Y, s.Yh, min(B, 2))):
                edges = np.linspace(0, B, parts + 1).astype(int)
                for a, b in zip(edges[:-1], edges[1:]):
                    jobs.append((self._stage_pool.submit(np.copyto, host[a:b], src[a:b]), dev_t, host_t, int(a), int(b)))
            for fut, dev_t, host_t, a, b in jobs:
                fut.result()
                dev_t[a:b].copy_(host_t[a:b], non_blocking=True)
        s.filled = True
        self.h2d_bytes += s.Xh.numel() * 4 + s.Yh.numel() * 4
        if alpha is not None:
            s.alpha.copy_(torch.as_tensor(np.asarray(alpha, dtype=np.float32)), non_blocking=False)

    def _prefetch(self, X, Y):
        """Stage the next batch while the GPU runs the current one: numpy -> pinned (host memcpy) -> device staging buffer on
        the copy stream.  The batch is recognised by object identity at the next train_on_batch() call."""
        if self.ops.device.type != 'cuda':
            return
        X = np.ascontiguousarray(X, dtype=np.float32)
        Y = np.ascontiguousarray(Y, dtype=np.float32)
        j = self._pf_slot ^ 1
        need = (X.size, Y.size)
        b = self._pf_bufs[j]
        if b is None or b['cap'][0] < need[0] or b['cap'][1] < need[1]:
            b = self._pf_bufs[j] = dict(
                cap=need, Xh=torch.empty(need[0], dtype=torch.float32, pin_memory=True),
                Yh=torch.empty(need[1], dtype=torch.float32, pin_memory=True),
                Xd=torch.empty(need[0], dtype=torch.float32, device=self.ops.device),
                Yd=torch.empty(need[1], dtype=torch.float32, device=self.ops.device), ev=torch.cuda.Event())
        if self._copy_stream is None:
            self._copy_stream = torch.cuda.Stream(device=self.ops.device)
        # plain single-threaded memcpy (numpy, GIL released): it hides behind the step on the GPU anyway, whereas torch's copy_ fans
        # out over every core and leaves the OpenMP workers spinning beside the thread that has to launch the next step
        np.copyto(b['Xh'].numpy()[:X.size], X.reshape(-1))
        np.copyto(b['Yh'].numpy()[:Y.size], Y.reshape(-1))
        with torch.cuda.stream(self._copy_stream):
            b['Xd'][:X.size].copy_(b['Xh'][:X.size], non_blocking=True)
            b['Yd'][:Y.size].copy_(b['Yh'][:Y.size], non_blocking=True)
            b['ev'].record(self._copy_stream)
        self.h2d_bytes += (X.size + Y.size) * 4
        self._pf_slot = j
        self._pf = (X, Y, tuple(X.shape), tuple(Y.shape))

    def train_on_batch(self, batchid, X, Y, alpha=None, next_batch=None):
        """-> (critic_cost, generator_cost or None), python floats (one blocking read of the loss vector per batch,
        like the reference's float(critic_returns), optimizertts_wgan.py:220-221).
        next_batch = (X, Y) of the following call, if known: it is staged to the device while this batch computes."""
        B, T = int(X.shape[0]), int(X.shape[1])
        seen = self._seen.get((B, T), 0)
        self._seen[(B, T)] = seen + 1
        s = self.shape_state(B, T, capture=seen >= self.graph_after)
        pf = self._pf
        if pf is not None and alpha is None and (pf[0] is X or (pf[2] == tuple(X.shape) and np.shares_memory(pf[0], X))) \
                and (pf[1] is Y or (pf[3] == tuple(Y.shape) and np.shares_memory(pf[1], Y))):
            if self._pf_placed is not s:     # else: the previous call already moved it into this shape's input buffers
                b = self._pf_bufs[self._pf_slot]
                torch.cuda.current_stream(self.ops.device).wait_event(b['ev'])
                s.X.view(-1).copy_(b['Xd'][:s.X.numel()], non_blocking=True)
                s.Y.view(-1).copy_(b['Yd'][:s.Y.numel()], non_blocking=True)
        else:
            self.stage(s, X, Y, alpha)
        self._pf = None
        self._pf_placed = None
        trained_gen = batchid % W.critic_runs(self.generator_updates) == 0
        self.step_device(s, trained_gen, explicit_alpha=alpha is not None)
        s.lossh.copy_(self.st.losses, non_blocking=True)
        self.d2h_bytes += W.NLOSS * 4
        perr = getattr(self.st.sync, 'peer_err', None)
        if perr is not None:
            s.errh.copy_(perr, non_blocking=True)
            self.d2h_bytes += 4
        if next_batch is not None:
            self._prefetch(next_batch[0], next_batch[1])
            if self._pf is not None and tuple(self._pf[2][:2]) == (B, T):
                # same (B, T): the staged batch moves into the step's input buffers behind this step's kernels (stream order), so
                # the next call starts with its graph launch instead of two copies
                b = self._pf_bufs[self._pf_slot]
                torch.cuda.current_stream(self.ops.device).wait_event(b['ev'])
                s.X.view(-1).copy_(b['Xd'][:s.X.numel()], non_blocking=True)
                s.Y.view(-1).copy_(b['Yd'][:s.Y.numel()], non_blocking=True)
                s.filled = True
                self._pf_placed = s
        if self.ops.device.type == 'cuda':
            torch.cuda.current_stream(self.ops.device).synchronize()
        if perr is not None and int(s.errh[0]) != 0:
            # a sync-BatchNorm exchange gave up waiting for a peer: statistics, gradients and the update of this step are garbage
            self.st.sync.check_peer()
        critic_cost = float(s.lossh[W.L_CRITIC])
        gen_cost = float(s.lossh[W.L_GEN]) if trained_gen else None
        return critic_cost, gen_cost
