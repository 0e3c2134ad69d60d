"""
Host<->device plumbing around the WGAN-GP stepper: pinned staging buffers, CUDA-graph capture of the critic step and
of the generator step (one graph each per batch shape), and the per-batch policy of
OptimizerTTSWGAN.train_on_batch (reference percivaltts/optimizertts_wgan.py:216-241).

One process drives one GPU; under torch.distributed every rank runs the same code on its shard of the batch.
"""
import numpy as np
import torch

from . import wgan as W


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
        self.g_critic = self.g_critic_alpha = self.g_gen = None


class DeviceTrainer(object):
    def __init__(self, stepper, use_graphs=True, max_shapes=8):
        self.st = stepper
        self.ops = stepper.ops
        self.use_graphs = bool(use_graphs) and self.ops.device.type == 'cuda'
        self.max_shapes = max_shapes
        self._shapes = {}
        self.generator_updates = 0
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    # ---- shape state / graph capture ---------------------------------------------------------------------------
    def shape_state(self, B, T):
        key = (B, T)
        s = self._shapes.get(key)
        if s is None:
            if len(self._shapes) >= self.max_shapes:
                self._shapes.pop(next(iter(self._shapes)))
            s = self._shapes[key] = _ShapeState(self.ops, self.st.a, B, T)
            if self.use_graphs:
                self._capture(s)
        return s

    def _capture(self, s):
        """Warm up eagerly on a side stream (allocates every plan buffer and workspace; parameters, BN moving
        statistics and the RNG counter are restored afterwards), then capture the two graphs."""
        st = self.st
        keep_state = st.G.params.state.clone()
        keep_counter = st.counter.clone()
        s.X.zero_()
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

    # ---- steps on resident device data ---------------------------------------------------------------------------
    def critic_step_device(self, s, explicit_alpha=False):
        if self.use_graphs:
            (s.g_critic_alpha if explicit_alpha else s.g_critic).replay()
        else:
            self.st.critic_step(s.X, s.Y, alpha=s.alpha if explicit_alpha else None, apply=True)

    def generator_step_device(self, s):
        if self.use_graphs:
            s.g_gen.replay()
        else:
            self.st.generator_step(s.X, s.Y, apply=True)
        self.generator_updates += 1

    # ---- the reference's per-batch entry point ---------------------------------------------------------------------
    def stage(self, s, X, Y, alpha=None):
        """numpy [B,T,*] float32 -> pinned -> device (async on the current stream)."""
        s.Xh.copy_(torch.from_numpy(np.ascontiguousarray(X, dtype=np.float32)))
        s.Yh.copy_(torch.from_numpy(np.ascontiguousarray(Y, dtype=np.float32)))
        s.X.copy_(s.Xh, non_blocking=True)
        s.Y.copy_(s.Yh, non_blocking=True)
        self.h2d_bytes += s.Xh.numel() * 4 + s.Yh.numel() * 4
        if alpha is not None:
            s.alpha.copy_(torch.as_tensor(np.asarray(alpha, dtype=np.float32)), non_blocking=False)

    def train_on_batch(self, batchid, X, Y, alpha=None):
        """-> (critic_cost, generator_cost or None), python floats (one blocking read of the loss vector per batch,
        like the reference's float(critic_returns), optimizertts_wgan.py:220-221)."""
        B, T = int(X.shape[0]), int(X.shape[1])
        s = self.shape_state(B, T)
        self.stage(s, X, Y, alpha)
        self.critic_step_device(s, explicit_alpha=alpha is not None)
        trained_gen = False
        if batchid % W.critic_runs(self.generator_updates) == 0:
            self.generator_step_device(s)
            trained_gen = True
        s.lossh.copy_(self.st.losses, non_blocking=True)
        self.d2h_bytes += W.NLOSS * 4
        if self.ops.device.type == 'cuda':
            torch.cuda.current_stream(self.ops.device).synchronize()
        critic_cost = float(s.lossh[W.L_CRITIC])
        gen_cost = float(s.lossh[W.L_GEN]) if trained_gen else None
        return critic_cost, gen_cost
