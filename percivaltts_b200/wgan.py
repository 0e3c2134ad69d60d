"""
Device-side WGAN-GP step: the two compiled Keras graphs of OptimizerTTSWGAN.prepare()
(reference percivaltts/optimizertts_wgan.py:107-213) restated as explicit kernel schedules, plus the
train_on_batch() policy (:216-241).

  critic step      loss = mean(-D(real)) + mean(+D(fake)) + lambda * mean_b (1 - ||dD(xhat)/dxhat||_2)^2      (:131-154)
  generator step   loss = mean(w_wgan) * mean(-D(G(ctx))) + mean((y - G(ctx))^2 * w_ls)   ('WLSWGAN', :180-213)
                   loss = mean(-D(G(ctx)))                                                 ('WGAN',    :174-178)

Data parallel: every rank holds `B` utterances of the global batch; all means are over the global batch (local sums
are scaled by 1/global count and the flat gradient buffers are all-reduced with SUM before Adam).
"""
import os

import numpy as np
import torch

from . import arch as A
from .engine import Generator, Critic, LocalSync

# slots of the device loss vector
L_REAL, L_FAKE, L_GP, L_CRITIC, L_GW, L_GLS, L_GEN, L_SSE, L_GLS_F0 = 0, 1, 2, 3, 4, 5, 6, 7, 8
NLOSS = 9


def critic_runs(generator_updates):
    """optimizertts_wgan.py:225-228"""
    return 10 if (generator_updates < 25 or generator_updates % 500 == 0) else 5


class WGANStepper(object):
    def __init__(self, ops, arch, generator=None, critic=None, errtype='WLSWGAN', sync=None,
                 critic_lr=1e-4, critic_beta1=0.5, critic_beta2=0.9, gen_lr=1e-3, gen_beta1=0.5, gen_beta2=0.9,
                 gp_lambda=10.0, w_ls=None, w_wgan_mean=1.0, adam_eps=1e-7, seed=123,
                 critic_step_full_generator=False):
        if errtype not in ('WGAN', 'WLSWGAN'):
            raise ValueError('unknown errtype ' + str(errtype))
        self.ops, self.a, self.errtype = ops, arch, errtype
        self.sync = sync or LocalSync()
        self.G = generator or Generator(ops, arch)
        self.C = critic or Critic(ops, arch)
        if hasattr(self.sync, 'enable_grad_sync'):      # replicas: flat gradient buffers into NVLink symmetric memory (collective)
            self.sync.enable_grad_sync([self.G.params, self.C.params], ops.device)
        self.copt = (critic_lr, critic_beta1, critic_beta2)
        self.gopt = (gen_lr, gen_beta1, gen_beta2)
        self.adam_eps = adam_eps
        self.lam = float(gp_lambda)
        # False: the frozen generator of the critic step evaluates only what the critic reads (the spectral branch)
        self.critic_step_full_generator = bool(critic_step_full_generator)
        # PCTTS_REAL_EARLY=1: D(real) on the side lane beside the frozen generator instead of batched with the two other
        # applications -- measured on B200: 16 more launches, same step time (2.729 vs 2.722 ms), so it stays off
        self.real_rows_early = os.environ.get('PCTTS_REAL_EARLY', '0') != '0'
        # PCTTS_STAGGER_CONV1D=1: the critic's ctx branch (side lane) forked behind the generator's first Conv1D instead of beside it
        # -- measured on B200: critic step 1.962 -> 1.995 ms (the side lane then ends later than the frozen generator), so it stays off
        self.stagger_conv1d = os.environ.get('PCTTS_STAGGER_CONV1D', '0') != '0'
        # PCTTS_F0_FREE=1: generator step with the F0 lane (LSTM forward, its LSE column, BPTT) never joined between forward and backward
        # -- measured on B200: generator step 2.45-2.48 vs 2.43-2.50 ms, inside the noise (what the LSTM pair costs the step is the 80 SMs
        # it holds for 1 ms, not the position of its join), so it stays off
        self.f0_free = os.environ.get('PCTTS_F0_FREE', '0') != '0'
        self.wbar = float(w_wgan_mean)
        self.w_ls = None
        if errtype == 'WLSWGAN':
            if w_ls is None:
                raise ValueError('WLSWGAN needs the LSE weight vector')
            self.w_ls = ops.zeros(arch.outsize)
            self.w_ls.copy_(torch.as_tensor(np.asarray(w_ls, dtype=np.float64)).to(self.w_ls.dtype))
        self.losses = ops.zeros(NLOSS)
        self.seed = int(seed) + 7919 * self.sync.rank     # independent alpha streams per replica
        self.counter = ops.zeros(1, dtype=torch.int64)
        self._bufs = {}

    def buffers(self, B, T):
        key = (B, T)
        b = self._bufs.get(key)
        if b is None:
            o, a = self.ops, self.a
            b = self._bufs[key] = dict(
                fake=o.empty(B, T, a.outsize), s0=o.empty(3 * B, T, a.spec_size), alpha=o.zeros(B),
                g=o.empty(B, T, a.spec_size, 1), ss=o.zeros(B), dpred=o.empty(B, T, a.outsize))
        return b

    def release(self, B, T):
        """Free every per-shape buffer of (B, T): step buffers, both networks' training plans, the kernels' scratch recorded under
        the shape tag.  The caller drops the CUDA graphs captured for the shape first."""
        self._bufs.pop((B, T), None)
        self.G.release(B, T, with_backward=True)
        self.C.release(B, T)
        self.ops.release_shape((B, T))

    def _scoped(self, fn, X, *args, **kw):
        with self.ops.shape_scope((int(X.shape[0]), int(X.shape[1]))), self.ops.step_scope():
            return fn(X, *args, **kw)

    # ------------------------------------------------------------------------------------------------------------
    def critic_step(self, X, *args, **kw):
        return self._scoped(self._critic_step, X, *args, **kw)

    def critic_generator_step(self, X, *args, **kw):
        return self._scoped(self._critic_generator_step, X, *args, **kw)

    def _critic_generator_step(self, X, Y, alpha=None):
        """One critic update followed by one generator update on the SAME batch (what train_on_batch does whenever the generator is
        due, optimizertts_wgan.py:216-241).  The generator's weights do not move in the critic step, so G(ctx) and every activation
        of its forward pass are the same in both graphs of the reference: the critic step evaluates the whole generator once
        (materialised, moving statistics updated) and the generator step starts from it instead of repeating the forward pass --
        same values, one forward pass and one round of sync-BatchNorm exchanges less."""
        self._critic_step(X, Y, alpha=alpha, apply=True, for_generator=True)
        self._generator_step(X, Y, apply=True, reuse_forward=True)

    def _critic_step(self, X, Y, alpha=None, apply=True, for_generator=False):
        """One critic update on device tensors X [B,T,ctx], Y [B,T,O].  alpha [B] overrides the random
        interpolation weights (RandomWeightedAverage, optimizertts_wgan.py:44-51)."""
        o, a, sync = self.ops, self.a, self.sync
        B, T, _ = X.shape
        M, S = B * T, a.spec_size
        bf = self.buffers(B, T)
        gpl = self.G.plan(B, T)
        cpl = self.C.plan(3, B, T)
        # the critic's ctx branch reads the labels and the critic weights only: side lane, beside the frozen generator
        s0 = bf['s0'].view(3 * B, T, S, 1)
        early = self.real_rows_early

        def critic_ctx_branch():
            with o.side_lane():
                self.C.forward_ctx(cpl, X)
                if early:
                    # D(real) needs nothing from the generator either: real block of the packed input, then its forward pass
                    o.copy2d(bf['s0'][:B].view(M, S), Y.view(M, a.outsize)[:, 1:1 + S])
                    self.C.forward_spec(cpl, s0, samples=slice(0, B))
        stagger = self.stagger_conv1d
        if not stagger:
            critic_ctx_branch()
        # frozen generator, learning phase 1: batch statistics, no moving-average update (:121-128)
        if for_generator:       # the generator step that follows on this batch starts from this forward pass (critic_generator_step)
            self.G.forward(gpl, X, bf['fake'], training=True, update_stats=True, sync=sync, spec_only=False, materialize=True,
                           after_first_conv=critic_ctx_branch if stagger else None)
        else:
            self.G.forward(gpl, X, bf['fake'], training=True, update_stats=False, sync=sync,
                           spec_only=not self.critic_step_full_generator, materialize=False,
                           after_first_conv=critic_ctx_branch if stagger else None)
        if alpha is None:
            o.uniform(bf['alpha'], self.seed, self.counter)
            alpha = bf['alpha']
        o.critic_pack_inputs(Y, bf['fake'], alpha, bf['s0'], S, 1, skip_real=early)
        o.join_lanes()
        d = self.C.forward_spec(cpl, s0, samples=slice(B, 3 * B) if early else None)
        ng = float(M * sync.world)
        o.sum_scaled(d[0:M], -1.0 / ng, self.losses, L_REAL)          # wasserstein_loss with y_true = -1 (:152)
        o.sum_scaled(d[M:2 * M], 1.0 / ng, self.losses, L_FAKE)       # y_true = +1 (:153)
        # d(objective)/d(D): -1/N on real rows, +1/N on fake rows; the interpolated rows carry K.gradients' implicit
        # all-ones seed of sum(D(xhat)) (:57)
        o.fill_segments(cpl['dd'].view(-1), M, -1.0 / ng, M, 1.0 / ng, M, 1.0)
        rows = slice(2 * B, 3 * B)
        self.C.backward_to_input(cpl, bf['g'], g_rows=rows, ctx_grads_apps=2)
        o.gp_norms(bf['g'].view(B, T * S), bf['ss'])                   # ||g_b||^2 (:59-64)
        # tangent seed = lambda * d(mean_b (1-||g_b||)^2)/dg, written over the (no longer needed) xhat rows of s0
        o.gp_scale(bf['g'].view(B, T * S), bf['ss'], bf['s0'][rows].view(B, T * S), self.lam, B * sync.world, self.losses, L_GP)
        self.C.tangent_forward(cpl, rows)
        self.C.param_grads(cpl, first_order_apps=2)
        o.loss_combine(self.losses, L_CRITIC, L_REAL, 1.0, L_FAKE, 1.0, L_GP, self.lam)   # loss_weights [1,1,lambda] (:150)
        lv = self.losses[L_REAL:L_CRITIC + 1]      # local shares -> global means (logging only)
        if apply:       # gradient all-reduce over the replicas + Adam (one fused peer-memory launch where available)
            sync.grad_adam(self.C.params, *self.copt, eps=self.adam_eps, losses=lv if sync.world > 1 else None)
        elif sync.world > 1:
            sync.allreduce(lv)

    def generator_step(self, X, *args, **kw):
        return self._scoped(self._generator_step, X, *args, **kw)

    def _generator_step(self, X, Y, apply=True, reuse_forward=False):
        """One generator update (BN moving statistics move here, like in the reference's generator_model).  reuse_forward: the
        generator's forward pass on this batch is the one the preceding critic step left in the plan (critic_generator_step)."""
        o, a, sync = self.ops, self.a, self.sync
        B, T, _ = X.shape
        M, S, O = B * T, a.spec_size, a.outsize
        bf = self.buffers(B, T)
        gpl = self.G.plan(B, T)
        pred = bf['fake']
        cpl = self.C.plan(1, B, T)
        pl_ctx_ev = [None]

        def critic_ctx_branch():
            with o.side_lane():
                self.C.forward_ctx(cpl, X)
                pl_ctx_ev[0] = o.lane_event()
        if not self.stagger_conv1d:
            critic_ctx_branch()
        # The F0 lane (LSTM) is not joined here: the critic reads the spectral columns only, so the recurrence and, behind it, the
        # F0 column's loss term and the BPTT run beside the critic's forward / backward and the other branches' backward.
        f0_free = self.f0_free and self.errtype == 'WLSWGAN' and not reuse_forward
        f0_issued = None
        if reuse_forward:
            if self.stagger_conv1d:
                critic_ctx_branch()
        else:
            self.G.forward(gpl, X, pred, training=True, update_stats=True, sync=sync,
                           after_first_conv=critic_ctx_branch if self.stagger_conv1d else None, join_f0=not f0_free)
        if f0_free:
            o.join_lanes(2)         # (the critic's ctx branch ran on lane 1 in FRONT of the F0 branch; forward_spec needs it)
            o.wait_lane_event(pl_ctx_ev[0])
        else:
            o.join_lanes()
        d = self.C.forward_spec(cpl, pred[:, :, 1:1 + S].unsqueeze(-1))
        ng = float(M * sync.world)
        o.sum_scaled(d, -1.0 / ng, self.losses, L_GW)
        dpred = bf['dpred']
        if self.errtype == 'WLSWGAN' and f0_free:
            # specweighted_lse_loss (:73-79) in two parts: every column but F0 here, the F0 column on the F0 lane behind the LSTM
            o.lse_loss_grad_cols(pred, Y, self.w_ls, dpred, 1, O - 1, 1.0 / (ng * O), 1.0 / (ng * O), self.losses, L_GLS)
            with o.side_lane(nowait=True):      # behind the LSTM forward on its own lane; nothing here needs the main lane
                o.lse_loss_grad_cols(pred, Y, self.w_ls, dpred, 0, 1, 1.0 / (ng * O), 1.0 / (ng * O), self.losses, L_GLS_F0)
                f0_issued = (self.G.backward_f0(gpl, pred, dpred),)
            seed = -self.wbar / ng
        elif self.errtype == 'WLSWGAN':
            o.lse_loss_grad(pred, Y, self.w_ls, dpred, 1.0 / (ng * O), 1.0 / (ng * O), self.losses, L_GLS)   # (:73-79)
            o.loss_combine(self.losses, L_GEN, L_GW, self.wbar, L_GLS, 1.0)                                  # (:213)
            seed = -self.wbar / ng
        else:
            o.fill_segments(dpred.view(-1), dpred.numel(), 0.0)
            o.fill_segments(self.losses[L_GLS:L_GLS + 1], 1, 0.0)
            o.loss_combine(self.losses, L_GEN, L_GW, 1.0)
            seed = -1.0 / ng
        o.fill_segments(cpl['dd'].view(-1), M, seed)
        self.C.backward_to_input(cpl, dpred[:, :, 1:1 + S].unsqueeze(-1), accumulate=True)
        self.G.backward(gpl, pred, dpred, sync=sync, f0_issued=f0_issued)
        if self.errtype == 'WLSWGAN' and f0_free:      # the two parts of the LSE term (backward joined the lanes), then (:213)
            o.loss_combine(self.losses, L_GLS, L_GLS, 1.0, L_GLS_F0, 1.0)
            o.loss_combine(self.losses, L_GEN, L_GW, self.wbar, L_GLS, 1.0)
        lv = self.losses[L_GW:L_GEN + 1]
        if apply:
            sync.grad_adam(self.G.params, *self.gopt, eps=self.adam_eps, losses=lv if sync.world > 1 else None)
        elif sync.world > 1:
            sync.allreduce(lv)

    def evaluate_losses(self, X, *args, **kw):
        return self._scoped(self._evaluate_losses, X, *args, **kw)

    def _evaluate_losses(self, X, Y, alpha=None):
        """Critic and generator objectives on one batch in inference mode (BN moving statistics), no update:
        what critic_model.evaluate / generator_model.evaluate return in update_validation_cost (:244-262); losses[L_SSE] is the
        sum of squared prediction errors of the batch (data.cost_model_prediction_rmse, data.py:360-375) from the same forward."""
        o, a = self.ops, self.a
        B, T, _ = X.shape
        M, S, O = B * T, a.spec_size, a.outsize
        bf = self.buffers(B, T)
        gpl = self.G.plan(B, T)
        nosync = LocalSync()
        self.G.forward(gpl, X, bf['fake'], training=False, sync=nosync)
        if alpha is None:
            o.uniform(bf['alpha'], self.seed, self.counter)
            alpha = bf['alpha']
        o.critic_pack_inputs(Y, bf['fake'], alpha, bf['s0'], S, 1)
        cpl = self.C.plan(3, B, T)
        d = self.C.forward(cpl, bf['s0'].view(3 * B, T, S, 1), X)
        ng = float(M)
        o.sum_scaled(d[0:M], -1.0 / ng, self.losses, L_REAL)
        o.sum_scaled(d[M:2 * M], 1.0 / ng, self.losses, L_FAKE)
        o.sum_scaled(d[M:2 * M], -1.0 / ng, self.losses, L_GW)
        o.fill_segments(cpl['dd'].view(-1), M, 0.0, M, 0.0, M, 1.0)
        rows = slice(2 * B, 3 * B)
        self.C.backward_to_input(cpl, bf['g'], g_rows=rows)       # no parameter gradients in a validation pass
        o.gp_norms(bf['g'].view(B, T * S), bf['ss'])
        o.gp_scale(bf['g'].view(B, T * S), bf['ss'], bf['s0'][rows].view(B, T * S), self.lam, B, self.losses, L_GP)
        o.loss_combine(self.losses, L_CRITIC, L_REAL, 1.0, L_FAKE, 1.0, L_GP, self.lam)
        if self.errtype == 'WLSWGAN':
            o.lse_loss_grad(bf['fake'], Y, self.w_ls, None, 1.0 / (ng * O), 0.0, self.losses, L_GLS)
            o.loss_combine(self.losses, L_GEN, L_GW, self.wbar, L_GLS, 1.0)
        else:
            o.loss_combine(self.losses, L_GEN, L_GW, 1.0)
        o.lse_loss_grad(bf['fake'], Y, None, None, 1.0, 0.0, self.losses, L_SSE)

    def predict(self, X, *args, **kw):
        with self.ops.shape_scope(('predict', int(X.shape[0]), int(X.shape[1]))), self.ops.step_scope():
            return self._predict(X, *args, **kw)

    def _predict(self, X, out=None):
        """Generator inference forward (BN moving statistics): ModelTTS.predict (modeltts.py:68-69)."""
        B, T, _ = X.shape
        gpl = self.G.plan(B, T, with_backward=False)
        if out is None:
            out = self.ops.empty(B, T, self.a.outsize)
        return self.G.forward(gpl, X, out, training=False, sync=LocalSync())
