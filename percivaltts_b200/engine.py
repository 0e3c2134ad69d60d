"""
Host-side execution engine of the WGAN-GP step: explicit forward / backward / double-backward schedules of the
generator and the critic, written against the kernel interface of percivaltts_b200/ops.py (libpctts.so).

torch is used for device memory only: no autograd, no torch operators on the data path.  Every gradient below is
derived by hand and checked against the CPU oracle in tests/.

Reference graph:  generator percivaltts/modeltts_common.py:61-126, critic percivaltts/networks_critic.py:44-96,
layer builders percivaltts/networktts.py:59-126, losses / step graphs percivaltts/optimizertts_wgan.py:44-79,107-213.
"""
from collections import OrderedDict

import numpy as np
import torch

from . import arch as A
from .ops import EPI_LRELU, EPI_SIGMOID, EPI_ACCUM


class LocalSync(object):
    """Single-replica stand-in for the data-parallel exchange (see parallel.DistSync)."""
    world = 1
    rank = 0

    def allreduce(self, t):
        return t

    def grad_adam(self, store, lr, beta1, beta2, eps, losses=None):
        store.adam(lr, beta1, beta2, eps)


class ParamStore(object):
    """Flat fp32 buffers (parameters, gradients, Adam slots) with named views; non-trainable BN moving
    statistics live in a second flat buffer.  One flat gradient buffer = one NCCL all-reduce per step."""

    def __init__(self, ops, shapes):
        self.ops = ops
        self.shapes = list(shapes)
        ntrain = sum(int(np.prod(s)) for _, s, t in self.shapes if t)
        nstate = sum(int(np.prod(s)) for _, s, t in self.shapes if not t)
        self.flat = ops.zeros(max(ntrain, 1))
        self.grad = ops.zeros(max(ntrain, 1))
        self.m = ops.zeros(max(ntrain, 1))
        self.v = ops.zeros(max(ntrain, 1))
        self.state = ops.zeros(max(nstate, 1))
        self.step = ops.zeros(1, dtype=torch.int32)
        self.ntrain, self.nstate = ntrain, nstate
        ops.register_persistent(self.flat)      # GEMM operand shadows of the weights follow load() / adam()
        self.p, self.g = OrderedDict(), OrderedDict()
        ot = os_ = 0
        for name, shape, trainable in self.shapes:
            n = int(np.prod(shape))
            if trainable:
                self.p[name] = self.flat[ot:ot + n].view(*shape)
                self.g[name] = self.grad[ot:ot + n].view(*shape)
                ot += n
            else:
                self.p[name] = self.state[os_:os_ + n].view(*shape)
                os_ += n

    def use_grad_buffer(self, buf):
        """Re-home the flat gradient buffer (data parallel: into NVLink symmetric memory, parallel.DistSync.enable_grad_sync) and
        rebuild the named views.  Before any plan or CUDA graph has captured the old addresses."""
        assert buf.dtype == self.grad.dtype and buf.numel() >= max(self.ntrain, 1) and buf.is_contiguous()
        self.grad = buf[:max(self.ntrain, 1)]
        ot = 0
        for name, shape, trainable in self.shapes:
            n = int(np.prod(shape))
            if trainable:
                self.g[name] = self.grad[ot:ot + n].view(*shape)
                ot += n

    def count_params(self):
        return self.ntrain + self.nstate

    def load(self, weights):
        """weights: name -> array-like (numpy or torch); every inventory entry must be present."""
        for name, shape, _ in self.shapes:
            w = torch.as_tensor(np.asarray(weights[name])).to(self.flat.dtype).reshape(*shape)
            self.p[name].copy_(w)
        self.ops.refresh_persistent(self.flat)

    def export(self):
        return OrderedDict((name, self.p[name].detach().cpu().numpy().copy()) for name, _, _ in self.shapes)

    def export_grads(self):
        return OrderedDict((name, self.g[name].detach().cpu().numpy().copy()) for name in self.g)

    def adam(self, lr, beta1, beta2, eps=1e-7):
        if self.ntrain:
            self.ops.adam(self.flat[:self.ntrain], self.grad[:self.ntrain], self.m[:self.ntrain], self.v[:self.ntrain],
                          lr, beta1, beta2, eps, self.step)
            self.ops.refresh_persistent(self.flat)


def _im2col_view(xp, T, W):
    """Virtual im2col of a time-padded [B, T+W-1, C] buffer: row (b,t) = the W*C floats starting at xp[b,t,0]."""
    B, Tp, C = xp.shape
    return xp.as_strided((B, T, W * C), (Tp * C, C, 1))


class _BN(object):
    """Buffers of one BatchNormalization(axis=-1)+LeakyReLU application on a [rows, C] pre-activation."""

    def __init__(self, ops, prefix, rows, C):
        self.prefix, self.rows, self.C = prefix, rows, C
        self.mean = ops.zeros(C)
        self.invstd = ops.zeros(C)


class _BNNet(object):
    """BatchNormalization(axis=-1) + LeakyReLU forward / backward shared by the generator-type networks
    (pFC / pCNN1D / pCNN2D with bn=True, networktts.py:59-63,116-126)."""

    def plan(self, B, T, with_backward=True):
        key = (B, T, bool(with_backward))
        pl = self._plans.get(key)
        if pl is None:
            pl = self._plans[key] = self._make_plan(B, T, with_backward)
        return pl

    def release(self, B, T, with_backward=None):
        """Drop the buffers of one batch shape (the owner's LRU; with_backward selects the training or the inference plan)."""
        for bwd in ((False, True) if with_backward is None else (bool(with_backward),)):
            self._plans.pop((B, T, bwd), None)

    def _fused_sync(self, sync):
        """Second reduction stage, replica exchange and finalize ride in ONE launch: on a single replica, or when the replicas
        exchange through NVLink peer memory (parallel.DistSync.enable_peer)."""
        o = self.ops
        if not getattr(o, 'bn_sync', False) or not getattr(o, 'bn_single_launch', True):
            return False
        return sync.world == 1 or getattr(sync, 'bn_handle', None) is not None

    def _bn_fwd(self, z2, y2, bn, training, update_stats, sync, ws=None, apply=True, planes=False):
        """ws: the batch sums when the producer of z2 already reduced them (Conv2D epilogue).  apply=False: only the
        statistics (bn.mean / bn.invstd) -- the consumer applies BatchNorm + LeakyReLU on load."""
        o, p = self.ops, self.params.p
        g, b = p[bn.prefix + '.gamma'], p[bn.prefix + '.beta']
        mm, mv = p[bn.prefix + '.moving_mean'], p[bn.prefix + '.moving_var']
        pk = dict(planes=True) if planes else {}      # y2 feeds a GEMM: its bf16 operand planes come out of the same kernel
        if training:
            fused = self._fused_sync(sync)
            if ws is None:
                ws = o.bn_stats(z2, final=not fused and (sync.world > 1 or not getattr(o, 'bn_single_launch', True)))
            if fused and sync.world > 1 and getattr(ws, 'pending', 0) > 0:
                o.bn_sync_finalize(ws, bn.rows * sync.world, A.BN_EPS, A.BN_MOMENTUM, bn.mean, bn.invstd,
                                   mm if update_stats else None, mv if update_stats else None, sync)
            else:
                sync.allreduce(ws.sums)
                o.bn_finalize(ws, bn.rows * sync.world, A.BN_EPS, A.BN_MOMENTUM, bn.mean, bn.invstd,
                              mm if update_stats else None, mv if update_stats else None)
            if apply:
                o.bn_apply(z2, y2, bn.mean, bn.invstd, g, b, A.BN_EPS, 0, 1, **pk)
        else:
            o.bn_apply(z2, y2, mm, mv, g, b, A.BN_EPS, 1, 1, **pk)

    def _bn_bwd(self, z2, dy2, dz2, bn, sync, planes=False):
        o, p, g = self.ops, self.params.p, self.params.g
        pk = dict(planes=True) if planes else {}
        gam, bet = p[bn.prefix + '.gamma'], p[bn.prefix + '.beta']
        # dgamma / dbeta come from the all-reduced sums, i.e. every rank holds the global value: pre-divide by world so
        # that the all-reduce(sum) of the flat gradient buffer does not count them `world` times
        if self._fused_sync(sync):
            ws = o.bn_bwd_reduce(z2, dy2, bn.mean, bn.invstd, gam, bet, 1, final=False)
            o.bn_sync_bwd(ws, g[bn.prefix + '.gamma'], g[bn.prefix + '.beta'], 1.0 / sync.world, False, sync)
            o.bn_bwd_apply(ws, z2, dy2, dz2, bn.mean, bn.invstd, gam, bet, bn.rows * sync.world, 1, None, None, **pk)
            return
        ws = o.bn_bwd_reduce(z2, dy2, bn.mean, bn.invstd, gam, bet, 1)
        sync.allreduce(ws.sums)
        o.bn_bwd_apply(ws, z2, dy2, dz2, bn.mean, bn.invstd, gam, bet, bn.rows * sync.world, 1,
                       g[bn.prefix + '.gamma'], g[bn.prefix + '.beta'], grad_scale=1.0 / sync.world)


class GenericFC(_BNNet):
    """Generic(['FC', ...]) + network_final for the PML vocoder without MLPG (modeltts_common.py:36-55):
    pFC = Dense(no bias) + BN + LeakyReLU per layer (networktts.py:59-63,138-141), then
    concat[Dense(1+spec), Dense(nm, sigmoid)] (networktts.py:196-197,225).  ctx [B,T,ctxsize] -> [B,T,1+spec+nm]."""

    def __init__(self, ops, arch, params=None):
        self.ops, self.a = ops, arch
        self.params = params if params is not None else ParamStore(ops, A.generic_param_shapes(
            arch.ctxsize, arch.spec_size, arch.nm_size, arch.hiddenwidth, arch.layertypes))
        self._plans = {}

    def _make_plan(self, B, T, bwd):
        o, a = self.ops, self.a
        M, H, L = B * T, a.hiddenwidth, len(a.layertypes)
        pl = dict(B=B, T=T, M=M)
        pl['z'] = [o.empty(M, H) for _ in range(L)]
        pl['y'] = [o.empty(M, H) for _ in range(L)]
        pl['bn'] = [_BN(o, 'g.fc%d.bn' % i, M, H) for i in range(L)]
        if bwd:
            pl['dA'], pl['dB'] = o.empty(M, H), o.empty(M, H)
            pl['dfs'] = o.empty(M, 1 + a.spec_size)
            pl['dz_nm'] = o.empty(M, a.nm_size)
        return pl

    def forward(self, pl, ctx, out, training=True, update_stats=False, sync=None):
        o, a, p = self.ops, self.a, self.params.p
        sync = sync or LocalSync()
        M, S = pl['M'], a.spec_size
        x2 = ctx.reshape(M, a.ctxsize)
        for i in range(len(a.layertypes)):
            o.gemm(x2, p['g.fc%d.kernel' % i], pl['z'][i])
            self._bn_fwd(pl['z'][i], pl['y'][i], pl['bn'][i], training, update_stats, sync, planes=True)
            x2 = pl['y'][i]
        out2 = out.view(M, a.outsize)
        o.gemm(x2, p['g.f0spec.kernel'], out2[:, 0:1 + S], bias=p['g.f0spec.bias'])
        o.gemm(x2, p['g.nm.kernel'], out2[:, 1 + S:], bias=p['g.nm.bias'], flags=EPI_SIGMOID)
        pl['ctx_in'], pl['last'] = ctx, x2
        return out

    def backward(self, pl, out, dout, sync=None):
        """dout = d(loss)/d(out) [B,T,O]; fills params.g for every trainable parameter."""
        o, a, p, g = self.ops, self.a, self.params.p, self.params.g
        sync = sync or LocalSync()
        M, S, O, L = pl['M'], a.spec_size, a.outsize, len(a.layertypes)
        out2, dout2 = out.view(M, O), dout.view(M, O)
        xlast = pl['last']
        o.sigmoid_bwd(dout2[:, 1 + S:], out2[:, 1 + S:], pl['dz_nm'])
        o.gemm(xlast, pl['dz_nm'], g['g.nm.kernel'], transA=True)
        o.colsum(pl['dz_nm'], g['g.nm.bias'])
        o.copy2d(pl['dfs'], dout2[:, 0:1 + S])
        o.gemm(xlast, pl['dfs'], g['g.f0spec.kernel'], transA=True)
        o.colsum(pl['dfs'], g['g.f0spec.bias'])
        if L == 0:
            return
        dcur = pl['dA']
        o.gemm(pl['dz_nm'], p['g.nm.kernel'], dcur, transB=True)
        o.gemm(pl['dfs'], p['g.f0spec.kernel'], dcur, transB=True, flags=EPI_ACCUM)
        for i in range(L - 1, -1, -1):
            xin = pl['y'][i - 1] if i > 0 else pl['ctx_in'].reshape(M, a.ctxsize)
            dz = pl['dB'] if dcur is pl['dA'] else pl['dA']
            self._bn_bwd(pl['z'][i], dcur, dz, pl['bn'][i], sync, planes=True)
            o.gemm(xin, dz, g['g.fc%d.kernel' % i], transA=True)
            if i > 0:
                dnext = pl['dA'] if dz is pl['dB'] else pl['dB']
                o.gemm(dz, p['g.fc%d.kernel' % i], dnext, transB=True)
                dcur = dnext


class Generator(_BNNet):
    """DCNNF0SpecNoiseFeatures (modeltts_common.py:61-126): ctx [B,T,ctxsize] -> [B,T,1+spec+nm]."""

    def __init__(self, ops, arch, params=None):
        if arch.gen_nbcnnlayers < 1:
            raise ValueError('arch_gen_nbcnnlayers must be >= 1 on the B200 path')
        self.ops, self.a = ops, arch
        self.params = params if params is not None else ParamStore(ops, A.generator_param_shapes(arch))
        self._plans = {}

    # ---- buffers ----------------------------------------------------------------------------------------------
    def _make_plan(self, B, T, bwd):
        o, a = self.ops, self.a
        H, S, NM, O, Fl, W = a.hiddenwidth, a.spec_size, a.nm_size, a.outsize, a.gen_nbfilters, a.ctx_winlen
        M = B * T
        pl = dict(B=B, T=T, M=M)
        cin = a.ctxsize
        pl['ctx_xp'], pl['ctx_z'], pl['ctx_y'], pl['ctx_bn'] = [], [], [], []
        for i in range(a.ctx_nbcnnlayers):
            pl['ctx_xp'].append(o.zeros(B, T + W - 1, cin))
            pl['ctx_z'].append(o.empty(M, H))
            pl['ctx_y'].append(o.empty(M, H))
            pl['ctx_bn'].append(_BN(o, 'g.ctx.cnn%d.bn' % i, M, H))
            cin = H
        pl['fc_z'] = [o.empty(M, H) for _ in range(2)]
        pl['fc_y'] = [o.empty(M, H) for _ in range(2)]
        pl['fc_bn'] = [_BN(o, 'g.ctx.fc%d.bn' % i, M, H) for i in range(2)]
        pl['zx'] = [o.empty(B, T, 4 * H) for _ in range(2)]
        pl['hout'] = o.empty(B, T, 2 * H)
        pl['gates'] = [o.empty(B, T, 4 * H) for _ in range(2)] if bwd else None
        pl['cstate'] = [o.empty(B, T, H) for _ in range(2)] if bwd else None
        pl['s'] = o.empty(M, S)
        L = a.gen_nbcnnlayers
        pl['cv_z'] = [o.empty(B, T, S, Fl) for _ in range(L)]
        pl['cv_a'] = [o.empty(B, T, S, Fl) for _ in range(L)]
        pl['cv_bn'] = [_BN(o, 'g.spec.cnn%d.bn' % i, M * S, Fl) for i in range(L)]
        Ln = L // 2
        pl['nm_z'] = [o.empty(M, H) for _ in range(Ln)]
        pl['nm_y'] = [o.empty(M, H) for _ in range(Ln)]
        pl['nm_bn'] = [_BN(o, 'g.nm.fc%d.bn' % i, M, H) for i in range(Ln)]
        if bwd:
            pl['d_lctx'] = o.empty(M, H)
            pl['d_lctx_f0'] = o.empty(M, H)
            pl['dA'] = o.empty(M, H)           # ping-pong [M,H] gradient buffers
            pl['dB'] = o.empty(M, H)
            pl['dz_nm'] = o.empty(M, NM)
            pl['dspec'] = o.empty(M, S)
            pl['df0'] = o.empty(M, 1)
            pl['dmapA'] = o.empty(B, T, S, Fl)
            pl['dmapB'] = o.empty(B, T, S, Fl)
            pl['cv_dz'] = [o.empty(B, T, S, Fl) for _ in range(L)]     # kept per layer: the weight gradients of the stack are one launch
            pl['d_lctx_nm'] = o.empty(M, H)
            pl['ds'] = o.empty(M, S)
            pl['dhout'] = o.empty(B, T, 2 * H)
            pl['dzl'] = [o.empty(B, T, 4 * H) for _ in range(2)]
            pl['wflip'] = [o.empty(a.gen_winlen, a.spec_freqlen, Fl, 1 if i == 0 else Fl) for i in range(L)]
            pl['wflip_out'] = o.empty(a.gen_winlen, a.spec_freqlen, 1, Fl)
            if a.ctx_nbcnnlayers > 1:
                pl['ctx_dzp'] = o.zeros(B, T + W - 1, H)
                pl['ctx_kflip'] = o.empty(W, 1, H, H)
        return pl

    def _nm_lane(self, sync, training=True):
        """The noise-mask branch may run on the second side lane: always on a single replica; with replicas its BatchNorm
        exchanges need their own peer-memory channel (DistSync.bn_handle2)."""
        if not training or sync.world == 1:
            return True
        return self._fused_sync(sync) and getattr(sync, 'bn_handle2', None) is not None

    # ---- forward ----------------------------------------------------------------------------------------------
    def forward(self, pl, ctx, out, training=True, update_stats=False, sync=None, spec_only=False, materialize=True,
                after_first_conv=None, join_f0=True):
        """ctx [B,T,ctxsize] -> out [B,T,O] (both contiguous device tensors).  training: BN uses batch statistics
        (all-reduced through `sync`); update_stats: also move the BN moving averages (generator step only).
        spec_only: evaluate the ctx trunk and the spectral branch only and leave the F0 / noise-mask columns of `out`
        untouched -- the critic reads nothing but out[:,:,1:1+S] (networks_critic.py:58), so in the critic step the two
        other branches cannot influence any returned loss or updated weight."""
        o, a, p = self.ops, self.a, self.params.p
        sync = sync or LocalSync()
        B, T, M = pl['B'], pl['T'], pl['M']
        H, S, O, Fl, W = a.hiddenwidth, a.spec_size, a.outsize, a.gen_nbfilters, a.ctx_winlen
        left = (W - 1) // 2
        x = ctx
        for i in range(a.ctx_nbcnnlayers):                                   # pCNN1D (networktts.py:116-120)
            xp = pl['ctx_xp'][i]
            o.pad_time(xp, x, left, W - 1 - left)
            o.conv1d(xp, p['g.ctx.cnn%d.kernel' % i], pl['ctx_z'][i], T)
            if i == 0 and after_first_conv is not None:
                # the caller's independent branch (the critic's ctx branch: another machine-filling Conv1D) is forked HERE, behind
                # this Conv1D: two 132-CTA GEMMs launched together only slow each other down on the lane that has no slack
                after_first_conv()
                after_first_conv = None
            self._bn_fwd(pl['ctx_z'][i], pl['ctx_y'][i], pl['ctx_bn'][i], training, update_stats, sync, planes=True)
            x = pl['ctx_y'][i].view(B, T, H)
        if after_first_conv is not None:
            after_first_conv()
        x2 = x.reshape(M, -1)
        for i in range(2):                                                   # pFC (networktts.py:59-63)
            o.gemm(x2, p['g.ctx.fc%d.kernel' % i], pl['fc_z'][i])
            self._bn_fwd(pl['fc_z'][i], pl['fc_y'][i], pl['fc_bn'][i], training, update_stats, sync, planes=True)
            x2 = pl['fc_y'][i]
        lctx = x2
        out2 = out.view(M, O)
        # F0: Bidirectional LSTM + Dense(1)   (modeltts_common.py:82-84, networktts.py:85-96)
        # The recurrence is latency-bound and occupies 80 of the 148 SMs: it runs on the side lane while the spectral and
        # noise-mask branches run on the main one (the three branches only share their input lctx).
        if not spec_only:
            o.preshare(lctx)
            with o.side_lane():
                for d, dn in enumerate(('fw', 'bw')):
                    o.gemm(lctx, p['g.f0.blstm.%s.kernel' % dn], pl['zx'][d].view(M, 4 * H), bias=p['g.f0.blstm.%s.bias' % dn])
                o.lstm_fwd(pl['zx'], [p['g.f0.blstm.fw.recurrent'], p['g.f0.blstm.bw.recurrent']], pl['hout'], pl['gates'],
                           pl['cstate'])
                o.gemm(pl['hout'].view(M, 2 * H), p['g.f0.out.kernel'], out2[:, 0:1], bias=p['g.f0.out.bias'])
        # NM: pFC x (L//2), Dense(sigmoid)   (modeltts_common.py:117-121): a chain of small GEMM + BatchNorm launches that shares
        # nothing but lctx with the spectral branch -- second side lane (with replicas: only when the sync-BN exchange has a
        # channel of its own for that lane, the ranks must meet the exchanges of a channel in one order)
        def nm_branch():
            n2 = lctx
            for i in range(a.gen_nbcnnlayers // 2):
                o.gemm(n2, p['g.nm.fc%d.kernel' % i], pl['nm_z'][i])
                self._bn_fwd(pl['nm_z'][i], pl['nm_y'][i], pl['nm_bn'][i], training, update_stats, sync, planes=True)
                n2 = pl['nm_y'][i]
            o.gemm(n2, p['g.nm.out.kernel'], out2[:, 1 + S:], bias=p['g.nm.out.bias'], flags=EPI_SIGMOID)
        nm_lane = not spec_only and self._nm_lane(sync, training)
        if nm_lane:
            with o.side_lane(lane=2):
                nm_branch()
        # Spec: Dense projection, [B,T,S,1] map, pCNN2D stack, Conv2D(1)   (modeltts_common.py:95-101)
        o.gemm(lctx, p['g.spec.proj.kernel'], pl['s'], bias=p['g.spec.proj.bias'])
        amap = pl['s'].view(B, T, S, 1)
        L = a.gen_nbcnnlayers
        pending = None      # BatchNorm of the previous layer still to be applied (on load) by the next convolution
        for i in range(L):
            ws = o.conv2d(amap, p['g.spec.cnn%d.kernel' % i], pl['cv_z'][i], stats=training, in_bn=pending,
                          stats_final=not self._fused_sync(sync))
            # forward-only use (frozen generator of the critic step): nothing reads the activation map but the next
            # convolution, which forms it on load -- it is written for the last layer only (input of the output Conv2D)
            keep = materialize or not training or i == L - 1
            bn = pl['cv_bn'][i]
            self._bn_fwd(pl['cv_z'][i].view(M * S, Fl), pl['cv_a'][i].view(M * S, Fl), bn, training, update_stats, sync,
                         ws=ws, apply=keep)
            if keep:
                pending, amap = None, pl['cv_a'][i]
            else:
                pending = (bn.mean, bn.invstd, p[bn.prefix + '.gamma'], p[bn.prefix + '.beta'], pl['cv_a'][i])
                amap = pl['cv_z'][i]
        o.conv2d(amap, p['g.spec.out.kernel'], out[:, :, 1:1 + S].unsqueeze(-1), bias=p['g.spec.out.bias'])
        if not spec_only and not nm_lane:
            nm_branch()
        if not spec_only:
            # join_f0=False (generator step): the F0 lane is left running -- only out[:, :, 0] comes from it, and the caller keeps
            # everything that reads that column (its LSE term, the F0 backward) on that lane
            o.join_lanes(None if join_f0 else 2)
        pl['lctx'], pl['ctx_in'] = lctx, ctx
        return out

    # ---- backward ---------------------------------------------------------------------------------------------
    def backward_f0(self, pl, out, dout):
        """Backward of the F0 branch (Dense(1), BLSTM) on the CURRENT lane: needs dout[:, :, 0] only; leaves its share of d(lctx) in
        pl['d_lctx_f0'] and returns the lane event behind it (None in line); the branch's weight gradients follow the event."""
        o, a, p, g = self.ops, self.a, self.params.p, self.params.g
        B, T, M = pl['B'], pl['T'], pl['M']
        H, O = a.hiddenwidth, a.outsize
        lctx = pl['lctx']
        out2, dout2 = out.view(M, O), dout.view(M, O)
        f0_ev = None
        o.copy2d(pl['df0'], dout2[:, 0:1])
        hout2 = pl['hout'].view(M, 2 * H)
        o.gemm(hout2, pl['df0'], g['g.f0.out.kernel'], transA=True)
        o.colsum(pl['df0'], g['g.f0.out.bias'])
        o.gemm(pl['df0'], p['g.f0.out.kernel'], pl['dhout'].view(M, 2 * H), transB=True)
        U = [p['g.f0.blstm.fw.recurrent'], p['g.f0.blstm.bw.recurrent']]
        o.lstm_bwd(pl['dhout'], U, pl['gates'], pl['cstate'], pl['dzl'])
        # the branch's share of d(lctx) first: it is all the ctx trunk waits for; the weight gradients follow behind the event
        for d, dn in enumerate(('fw', 'bw')):
            o.gemm(pl['dzl'][d].view(M, 4 * H), p['g.f0.blstm.%s.kernel' % dn], pl['d_lctx_f0'], transB=True,
                   flags=EPI_ACCUM if d > 0 else 0)
        f0_ev = o.lane_event()
        for d, dn in enumerate(('fw', 'bw')):
            dz2 = pl['dzl'][d].view(M, 4 * H)
            o.gemm(lctx, dz2, g['g.f0.blstm.%s.kernel' % dn], transA=True)
            o.colsum(dz2, g['g.f0.blstm.%s.bias' % dn])
            gU = g['g.f0.blstm.%s.recurrent' % dn]
            if T > 1:
                hd = pl['hout'][:, :, d * H:(d + 1) * H]
                if d == 0:      # forward direction: the state feeding step t is h[t-1]
                    o.gemm(hd[:, 0:T - 1], pl['dzl'][d][:, 1:T], gU, transA=True)
                else:           # backward direction: the state feeding step t is h[t+1]
                    o.gemm(hd[:, 1:T], pl['dzl'][d][:, 0:T - 1], gU, transA=True)
            else:
                o.fill_segments(gU.view(-1), gU.numel(), 0.0)
        return f0_ev

    def backward(self, pl, out, dout, sync=None, f0_issued=None):
        """dout = d(loss)/d(out) [B,T,O]; fills params.g for every trainable generator parameter."""
        o, a, p, g = self.ops, self.a, self.params.p, self.params.g
        sync = sync or LocalSync()
        B, T, M = pl['B'], pl['T'], pl['M']
        H, S, NM, O, Fl, W = a.hiddenwidth, a.spec_size, a.nm_size, a.outsize, a.gen_nbfilters, a.ctx_winlen
        L, Ln = a.gen_nbcnnlayers, a.gen_nbcnnlayers // 2
        lctx, dl = pl['lctx'], pl['d_lctx']
        out2, dout2 = out.view(M, O), dout.view(M, O)
        # ---- F0 branch on the side lane (BPTT is latency-bound); its share of d(lctx) goes to its own buffer.  f0_issued: the caller
        # has already queued backward_f0() on that lane (generator step: behind the F0 column's loss term, without a fork)
        o.preshare(lctx)
        if f0_issued is not None:
            f0_ev = f0_issued[0]
        else:
            with o.side_lane():
                f0_ev = self.backward_f0(pl, out, dout)
        # ---- NM branch: shares nothing but lctx with the spectral branch -> second side lane, its share of d(lctx) in its own buffer
        dl_nm = pl['d_lctx_nm']

        def nm_branch():
            nlast = pl['nm_y'][Ln - 1] if Ln > 0 else lctx
            o.sigmoid_bwd(dout2[:, 1 + S:], out2[:, 1 + S:], pl['dz_nm'])
            o.gemm(nlast, pl['dz_nm'], g['g.nm.out.kernel'], transA=True)
            o.colsum(pl['dz_nm'], g['g.nm.out.bias'])
            dcur = dl_nm if Ln == 0 else pl['dA']
            o.gemm(pl['dz_nm'], p['g.nm.out.kernel'], dcur, transB=True)
            for i in range(Ln - 1, -1, -1):
                xin = pl['nm_y'][i - 1] if i > 0 else lctx
                dz = pl['dB'] if dcur is pl['dA'] else pl['dA']
                self._bn_bwd(pl['nm_z'][i], dcur, dz, pl['nm_bn'][i], sync, planes=True)
                o.gemm(xin, dz, g['g.nm.fc%d.kernel' % i], transA=True)
                dnext = dl_nm if i == 0 else dcur
                o.gemm(dz, p['g.nm.fc%d.kernel' % i], dnext, transB=True)
                dcur = dnext
        nm_lane, nm_ev = self._nm_lane(sync), None
        if nm_lane:
            with o.side_lane(lane=2):
                nm_branch()
                nm_ev = o.lane_event()
        # ---- Spec branch
        o.copy2d(pl['dspec'], dout2[:, 1:1 + S])
        dmap = pl['dspec'].view(B, T, S, 1)
        o.conv2d_wgrad(pl['cv_a'][L - 1], dmap, g['g.spec.out.kernel'])
        o.colsum(pl['dspec'].view(M * S, 1), g['g.spec.out.bias'])
        o.conv2d_flip(p['g.spec.out.kernel'], pl['wflip_out'])
        da = pl['dmapA']
        o.conv2d(dmap, pl['wflip_out'], da)
        wlayers = []        # weight gradients of the stack: one launch after the data-gradient chain (every dz is kept)
        for i in range(L - 1, -1, -1):
            dz = pl['cv_dz'][i]
            self._bn_bwd(pl['cv_z'][i].view(M * S, Fl), da.view(M * S, Fl), dz.view(M * S, Fl), pl['cv_bn'][i], sync)
            ain = pl['cv_a'][i - 1] if i > 0 else pl['s'].view(B, T, S, 1)
            wlayers.append(dict(x=ain, dy=dz, dw=g['g.spec.cnn%d.kernel' % i]))
            dnext = (pl['dmapB'] if da is pl['dmapA'] else pl['dmapA']) if i > 0 else pl['ds'].view(B, T, S, 1)
            o.conv2d_dgrad(dz, p['g.spec.cnn%d.kernel' % i], dnext, pl['wflip'][i])
            da = dnext
        multi = hasattr(o, 'conv2d_wgrad_multi_ok') and o.conv2d_wgrad_multi_ok(wlayers)
        if multi:       # beside the rest of the backward pass (second side lane, behind the noise-mask branch)
            with o.side_lane(lane=2):
                o.conv2d_wgrad_multi(wlayers)
        else:
            for Lw in wlayers:
                o.conv2d_wgrad(Lw['x'], Lw['dy'], Lw['dw'])
        o.gemm(lctx, pl['ds'], g['g.spec.proj.kernel'], transA=True)
        o.colsum(pl['ds'], g['g.spec.proj.bias'])
        o.gemm(pl['ds'], p['g.spec.proj.kernel'], dl, transB=True)
        if not nm_lane:
            nm_branch()
        # the ctx trunk needs the branches' shares of d(lctx): the F0 lane as a whole, of lane 2 only the noise-mask branch (the
        # weight gradients queued behind it are joined at the end)
        if f0_ev is not None:
            o.wait_lane_event(f0_ev)
        else:
            o.join_lanes(1)
        o.wait_lane_event(nm_ev)
        o.add3(dl, dl, pl['d_lctx_f0'], dl_nm)
        # ---- ctx trunk: pFC x2 then pCNN1D stack
        dcur = dl
        for i in (1, 0):
            if i > 0:
                xin = pl['fc_y'][i - 1]
            else:
                xin = pl['ctx_y'][-1] if a.ctx_nbcnnlayers > 0 else pl['ctx_in'].view(M, -1)
            dz = pl['dB'] if dcur is not pl['dB'] else pl['dA']
            self._bn_bwd(pl['fc_z'][i], dcur, dz, pl['fc_bn'][i], sync, planes=True)
            o.gemm(xin, dz, g['g.ctx.fc%d.kernel' % i], transA=True)
            if i > 0 or a.ctx_nbcnnlayers > 0:
                dnext = pl['dA'] if dz is pl['dB'] else pl['dB']
                o.gemm(dz, p['g.ctx.fc%d.kernel' % i], dnext, transB=True)
                dcur = dnext
        left = (W - 1) // 2
        for i in range(a.ctx_nbcnnlayers - 1, -1, -1):
            dz = pl['dB'] if dcur is not pl['dB'] else pl['dA']
            self._bn_bwd(pl['ctx_z'][i], dcur, dz, pl['ctx_bn'][i], sync, planes=True)
            k = p['g.ctx.cnn%d.kernel' % i]
            o.conv1d_wgrad(pl['ctx_xp'][i], dz, g['g.ctx.cnn%d.kernel' % i], T)
            if i > 0:
                # data gradient = 'same' correlation of dz with the time-flipped, in/out-transposed kernel
                o.pad_time(pl['ctx_dzp'], dz.view(B, T, H), W - 1 - left, left)
                o.conv2d_flip(k.view(W, 1, H, H), pl['ctx_kflip'])
                dnext = pl['dA'] if dz is pl['dB'] else pl['dB']
                o.gemm(_im2col_view(pl['ctx_dzp'], T, W), pl['ctx_kflip'].view(W * H, H), dnext)
                dcur = dnext
        o.join_lanes()


class Critic(object):
    """Critic (networks_critic.py:44-96): (features [B,T,O], ctx [B,T,ctxsize]) -> [B,T,1].

    `napp` applications that share the ctx input are batched along the sample axis of the spectral branch
    ([real | fake | interpolated] in the critic step); the ctx branch is evaluated once and its contribution to the
    first post layer is added per application (the reference recomputes it for each application,
    optimizertts_wgan.py:131-137 -- same values, one third of the work)."""

    def __init__(self, ops, arch, params=None):
        if arch.gen_nbcnnlayers < 1:
            raise ValueError('arch_gen_nbcnnlayers must be >= 1 on the B200 path')
        self.ops, self.a = ops, arch
        self.params = params if params is not None else ParamStore(ops, A.critic_param_shapes(arch))
        self._plans = {}

    def plan(self, napp, B, T):
        key = (napp, B, T)
        pl = self._plans.get(key)
        if pl is None:
            pl = self._plans[key] = self._make_plan(napp, B, T)
        return pl

    def release(self, B, T):
        """Drop the buffers of one batch shape (trainer's LRU)."""
        for key in [k for k in self._plans if k[1:] == (B, T)]:
            del self._plans[key]

    def _chain(self, T):
        """The spectral Conv2D stack runs as one multi-layer launch per chain (forward, data gradient, tangent sweep)."""
        a = self.a
        return a.gen_nbcnnlayers <= 8 and self.ops.conv2d_chain_ok(T, a.spec_size, a.gen_winlen, a.spec_freqlen, a.gen_nbfilters)

    def _make_plan(self, napp, B, T):
        o, a = self.ops, self.a
        H, S, Fl, W, L = a.hiddenwidth, a.spec_size, a.gen_nbfilters, a.ctx_winlen, a.gen_nbcnnlayers
        M, nB = B * T, napp * B
        nM = nB * T
        pl = dict(B=B, T=T, M=M, napp=napp, nB=nB, nM=nM)
        cin = a.ctxsize
        pl['ctx_xp'], pl['ctx_y'] = [], []
        for i in range(a.ctx_nbcnnlayers):
            pl['ctx_xp'].append(o.zeros(B, T + W - 1, cin))
            pl['ctx_y'].append(o.empty(M, H))
            cin = H
        pl['fc_y'] = [o.empty(M, H) for _ in range(2)]
        # the ctx branch's term of the first post layer: rows [0, M) are computed, the copies behind them let a row range that
        # starts inside an application (split forward) address its addend without a modulo
        pl['ctxterm_rep'] = o.empty(napp * M, H)
        pl['ctxterm'] = pl['ctxterm_rep'][:M]
        pl['a'] = [o.empty(nB, T, S, Fl) for _ in range(L)]
        pl['p'] = [o.empty(nM, H) for _ in range(3)]
        pl['d'] = o.empty(nM, 1)
        pl['dd'] = o.empty(nM, 1)
        pl['dp'] = [o.empty(nM, H) for _ in range(3)]
        pl['da'] = [o.empty(nB, T, S, Fl) for _ in range(L)]
        pl['wflip'] = [o.empty(a.gen_winlen, a.spec_freqlen, Fl, 1 if i == 0 else Fl) for i in range(L)]
        if napp > 1:       # row ranges (applications) of these matrices are GEMM operands of their own
            for t in pl['p'] + pl['dp'] + [pl['a'][L - 1].view(nM, S * Fl)]:
                o.share_rows(t)
        pl['dct'] = o.empty(M, H)
        pl['dcA'] = o.empty(M, H)
        pl['dcB'] = o.empty(M, H)
        if a.ctx_nbcnnlayers > 1:
            pl['ctx_dzp'] = o.zeros(B, T + W - 1, H)
            pl['ctx_kflip'] = o.empty(W, 1, H, H)
        return pl

    def forward(self, pl, s0, ctx):
        """s0: spectral input map view [napp*B, T, S, 1]; ctx [B,T,ctxsize].  -> pl['d'] [napp*B*T, 1]."""
        self.forward_ctx(pl, ctx)
        return self.forward_spec(pl, s0)

    def forward_ctx(self, pl, ctx):
        """The ctx branch (networks_critic.py:80-85), once per batch: -> pl['ctxterm'], its contribution to the first post
        layer.  It depends on the labels and the critic weights only, so the critic step runs it on the side lane while the
        frozen generator produces the fake features."""
        o, a, p = self.ops, self.a, self.params.p
        B, T, M = pl['B'], pl['T'], pl['M']
        H, S, Fl, W = a.hiddenwidth, a.spec_size, a.gen_nbfilters, a.ctx_winlen
        SF = S * Fl
        left = (W - 1) // 2
        x = ctx
        for i in range(a.ctx_nbcnnlayers):
            xp = pl['ctx_xp'][i]
            o.pad_time(xp, x, left, W - 1 - left)
            o.conv1d(xp, p['c.ctx.cnn%d.kernel' % i], pl['ctx_y'][i], T, bias=p['c.ctx.cnn%d.bias' % i], flags=EPI_LRELU)
            x = pl['ctx_y'][i].view(B, T, H)
        x2 = x.reshape(M, -1)
        for i in range(2):
            o.gemm(x2, p['c.ctx.fc%d.kernel' % i], pl['fc_y'][i], bias=p['c.ctx.fc%d.bias' % i], flags=EPI_LRELU)
            x2 = pl['fc_y'][i]
        w0 = p['c.post.fc0.kernel']
        o.gemm(x2, w0[SF:], pl['ctxterm'], bias=p['c.post.fc0.bias'])
        if self._split_forward(pl) is not None:
            for k in range(1, pl['napp']):
                o.copy2d(pl['ctxterm_rep'][k * M:(k + 1) * M], pl['ctxterm'])
        pl['ctx_in'], pl['ctx_last'] = ctx, x2

    def _split_forward(self, pl):
        """Samples of the first chain launch when the batched forward needs two rounds of clusters (None: one launch)."""
        o = self.ops
        if not (getattr(o, 'split_fwd', False) and self._chain(pl['T'])):      # (lanes off: the same launches, in line)
            return None
        nc = o.lib.pctts_conv2d_chain_supported(int(pl['T']), int(self.a.spec_size), o.nterms)
        per_round = 148 // nc if nc > 0 else 0
        if per_round < 1 or not (per_round < pl['nB'] <= 2 * per_round):
            return None
        return per_round

    def _spec_chain(self, pl, s0, sm):
        """The spectral Conv2D stack (networks_critic.py:58-70) on the samples `sm` (slice over the napp*B axis)."""
        o, a, p = self.ops, self.a, self.params.p
        L = a.gen_nbcnnlayers
        amap = s0[sm]
        if self._chain(pl['T']):
            # the BatchNorm-free stack (networks_critic.py:66-68) as ONE launch: tiles stay in shared memory from layer to layer
            o.conv2d_chain(amap, [dict(w=p['c.spec.cnn%d.kernel' % i], bias=p['c.spec.cnn%d.bias' % i], y=pl['a'][i][sm],
                                       flags=EPI_LRELU) for i in range(L)])
        else:
            for i in range(L):
                o.conv2d(amap, p['c.spec.cnn%d.kernel' % i], pl['a'][i][sm], bias=p['c.spec.cnn%d.bias' % i], flags=EPI_LRELU)
                amap = pl['a'][i][sm]

    def _post_layers(self, pl, r):
        """Post layers (networks_critic.py:89-96) on the rows `r` of the batched matrices; concat([spec, ctx]) . W0 =
        spec . W0[:SF] + ctx . W0[SF:], the second term being pl['ctxterm'] of the row's frame."""
        o, a, p = self.ops, self.a, self.params.p
        M, nM = pl['M'], pl['nM']
        S, Fl, L = a.spec_size, a.gen_nbfilters, a.gen_nbcnnlayers
        SF = S * Fl
        w0 = p['c.post.fc0.kernel']
        if r == slice(None) or (r.start % M == 0):
            addmod = pl['ctxterm']                              # row m of the range uses ctxterm[m % M]
        else:
            off, n = r.start % M, r.stop - r.start
            addmod = pl['ctxterm_rep'][off:off + n]             # replicated rows: no modulo needed
        o.gemm(pl['a'][L - 1].view(nM, SF)[r], w0[:SF], pl['p'][0][r], addmod=addmod, flags=EPI_LRELU)
        for i in (1, 2):
            o.gemm(pl['p'][i - 1][r], p['c.post.fc%d.kernel' % i], pl['p'][i][r], bias=p['c.post.fc%d.bias' % i], flags=EPI_LRELU)
        o.gemm(pl['p'][2][r], p['c.out.kernel'], pl['d'][r], bias=p['c.out.bias'])

    def forward_spec(self, pl, s0, samples=None):
        """Spectral branch (networks_critic.py:58-70) and the post layers on the applications `samples` (a slice over the
        napp*B sample axis, whole applications; default all); needs forward_ctx()."""
        o, a, p = self.ops, self.a, self.params.p
        T, nM = pl['T'], pl['nM']
        S, Fl, L = a.spec_size, a.gen_nbfilters, a.gen_nbcnnlayers
        SF = S * Fl
        w0 = p['c.post.fc0.kernel']
        if samples is None or (samples.start == 0 and samples.stop == pl['nB']):
            sm, r = slice(None), slice(None)
        else:
            assert samples.start % pl['B'] == 0 and samples.stop % pl['B'] == 0
            sm, r = samples, slice(samples.start * T, samples.stop * T)
        split = self._split_forward(pl) if (sm == slice(None)) else None
        if split is not None:
            # More clusters than the machine holds at once (30 samples x 8 CTAs on 148 SMs = two rounds): the first round's samples go
            # first, and their post layers run on the side lane UNDER the second round's chain launch instead of behind it.
            n1 = split
            self._spec_chain(pl, s0, slice(0, n1))
            with o.side_lane():
                self._post_layers(pl, slice(0, n1 * T))
            self._spec_chain(pl, s0, slice(n1, pl['nB']))
            self._post_layers(pl, slice(n1 * T, nM))
            o.join_lanes(1)
        else:
            self._spec_chain(pl, s0, sm)
            self._post_layers(pl, r)
        pl['s0'] = s0
        return pl['d']

    def backward_to_input(self, *args, **kw):
        with self.ops.backward_terms():      # gradient passes: linear in their seeds (ops.bwd_nterms)
            return self._backward_to_input(*args, **kw)

    def _backward_to_input(self, pl, g_out, g_rows=None, accumulate=False, ctx_grads_apps=0):
        """Back-propagate pl['dd'] (d objective / d critic output, [nM,1]) to the spectral input.  The per-layer
        pre-activation gradients stay in pl['dp'] / pl['da'] for param_grads().  g_out: [n,T,S,1] view receiving the
        input gradient of the applications `g_rows` (slice over the napp*B sample axis; default all)."""
        o, a, p = self.ops, self.a, self.params.p
        nM = pl['nM']
        S, Fl, L = a.spec_size, a.gen_nbfilters, a.gen_nbcnnlayers
        SF = S * Fl
        chain = self._chain(pl['T'])
        if not chain:      # the chain kernel forms the data-gradient kernels on the fly
            for i in range(L - 1, -1, -1):
                o.conv2d_flip(p['c.spec.cnn%d.kernel' % i], pl['wflip'][i])
        if g_rows is not None and ctx_grads_apps and g_rows.start > 0 and g_rows.stop == pl['nB']:
            # critic step.  Only the interpolated applications' input gradient is on the critical path (norm -> penalty ->
            # tangent sweep); the other applications' pre-activation gradients are needed by the weight gradients only:
            # side lane, followed there by everything else that hangs off them (ctx branch, bias sums).
            with o.side_lane():
                first = slice(0, g_rows.start)
                self._dgrad_chain(pl, first, part='fc')
                # the ctx-branch gradients hang off dp[0] alone: a second side lane beside this one's convolutions
                self.param_grads_ctx(pl, ctx_grads_apps)
                self._dgrad_chain(pl, first, part='conv')
                # the spectral kernels' gradient over these applications' rows; the lane that carries the tangent sweep adds
                # the gradient-penalty rows' share afterwards (param_grads), so its long chain ends with a third of the work
                self._conv_wgrads(pl, slice(0, g_rows.start), first=True)
                pl['conv_wgrads_apps'] = g_rows.start // pl['B']
                pl['side_event'] = o.lane_event()
            # the interpolated rows' input gradient feeds 1 - ||g|| (a difference of nearly equal numbers once the critic is
            # trained): forward precision.  Everything else on the gradient passes is summed, never differenced.
            with o.full_terms():
                self._dgrad_chain(pl, g_rows, final=(g_out, accumulate) if chain else None)
        elif chain and (g_rows is None or (g_rows.start == 0 and g_rows.stop == pl['nB'])):
            self._dgrad_chain(pl, slice(0, pl['nB']), final=(g_out, accumulate))
            if ctx_grads_apps:
                self.param_grads_ctx(pl, ctx_grads_apps)
        else:
            with o.full_terms():      # in-line schedule of the critic step / validation pass: the penalty rows are among these
                self._dgrad_chain(pl, slice(0, pl['nB']))
            if ctx_grads_apps:
                self.param_grads_ctx(pl, ctx_grads_apps)
            if chain:
                with o.full_terms():
                    o.conv2d_chain(pl['da'][0][g_rows], [dict(w=p['c.spec.cnn0.kernel'], flip=True, y=g_out,
                                                              flags=EPI_ACCUM if accumulate else 0)])
                return
        if chain:
            return
        src = pl['da'][0] if g_rows is None else pl['da'][0][g_rows]
        o.conv2d(src, pl['wflip'][0], g_out, flags=EPI_ACCUM if accumulate else 0)

    def _dgrad_chain(self, pl, samples, final=None, part='all'):
        """Pre-activation gradients dp[2..0], da[L-1..0] of the applications `samples` (slice over the napp*B axis).
        final = (g_out, accumulate): also the gradient w.r.t. the spectral input of these applications (chain kernel only).
        part = 'fc' / 'conv': only the post layers (down to da[L-1]) / only the Conv2D stack."""
        o, a, p = self.ops, self.a, self.params.p
        T, nM = pl['T'], pl['nM']
        S, Fl, L = a.spec_size, a.gen_nbfilters, a.gen_nbcnnlayers
        SF = S * Fl
        whole = samples.start == 0 and samples.stop == pl['nB']
        r = slice(None) if whole else slice(samples.start * T, samples.stop * T)
        if part != 'conv':
            o.gemm(pl['dd'][r], p['c.out.kernel'], pl['dp'][2][r], transB=True, aux=pl['p'][2][r])
            o.gemm(pl['dp'][2][r], p['c.post.fc2.kernel'], pl['dp'][1][r], transB=True, aux=pl['p'][1][r])
            o.gemm(pl['dp'][1][r], p['c.post.fc1.kernel'], pl['dp'][0][r], transB=True, aux=pl['p'][0][r])
            o.gemm(pl['dp'][0][r], p['c.post.fc0.kernel'][:SF], pl['da'][L - 1].view(nM, SF)[r], transB=True,
                   aux=pl['a'][L - 1].view(nM, SF)[r])
        if part == 'fc':
            return
        sm = slice(None) if whole else samples
        if self._chain(T):
            layers = [dict(w=p['c.spec.cnn%d.kernel' % i], flip=True, aux=pl['a'][i - 1][sm], y=pl['da'][i - 1][sm])
                      for i in range(L - 1, 0, -1)]
            if final is not None:
                layers.append(dict(w=p['c.spec.cnn0.kernel'], flip=True, y=final[0], flags=EPI_ACCUM if final[1] else 0))
            if layers:
                o.conv2d_chain(pl['da'][L - 1][sm], layers)
            return
        assert final is None
        for i in range(L - 1, 0, -1):
            o.conv2d(pl['da'][i][sm], pl['wflip'][i], pl['da'][i - 1][sm], aux=pl['a'][i - 1][sm])

    def tangent_forward(self, *args, **kw):
        with self.ops.backward_terms():      # gradient passes: linear in their seeds (ops.bwd_nterms)
            return self._tangent_forward(*args, **kw)

    def _tangent_forward(self, pl, rows):
        """Forward-mode sweep of the linearised critic (LeakyReLU masks frozen) on the applications `rows`, whose
        input tangent has been written over pl['s0'][rows].  The tangents overwrite the stored activations of those
        rows in place, so that param_grads() sees [a | a | tangent] as one batched layer input."""
        o, a, p = self.ops, self.a, self.params.p
        T = pl['T']
        S, Fl, L = a.spec_size, a.gen_nbfilters, a.gen_nbcnnlayers
        SF = S * Fl
        tmap = pl['s0'][rows]
        n = tmap.shape[0] * T
        r2 = slice(rows.start * T, rows.stop * T)
        if self._chain(T):
            o.conv2d_chain(tmap, [dict(w=p['c.spec.cnn%d.kernel' % i], aux=pl['a'][i][rows], y=pl['a'][i][rows]) for i in range(L)])
            tmap = pl['a'][L - 1][rows]
        else:
            for i in range(L):
                y = pl['a'][i][rows]
                o.conv2d(tmap, p['c.spec.cnn%d.kernel' % i], y, aux=y)
                tmap = y
        o.gemm(tmap.view(n, SF), p['c.post.fc0.kernel'][:SF], pl['p'][0][r2], aux=pl['p'][0][r2])
        for i in (1, 2):
            o.gemm(pl['p'][i - 1][r2], p['c.post.fc%d.kernel' % i], pl['p'][i][r2], aux=pl['p'][i][r2])

    def param_grads_ctx(self, *args, **kw):
        with self.ops.backward_terms():      # gradient passes: linear in their seeds (ops.bwd_nterms)
            return self._param_grads_ctx(*args, **kw)

    def _param_grads_ctx(self, pl, first_order_apps=2):
        """Gradients of the ctx branch and of its rows of the first post layer, on the side lane.  They depend on pl['dp'][0]
        only, so the critic step forks them as soon as backward_to_input() has produced it (ctx_grads_apps), long before the
        gradient-penalty sweep that the spectral weight gradients wait for; param_grads() joins."""
        o, a, p, g = self.ops, self.a, self.params.p, self.params.g
        B, T, M = pl['B'], pl['T'], pl['M']
        H, S, Fl, W = a.hiddenwidth, a.spec_size, a.gen_nbfilters, a.ctx_winlen
        SF = S * Fl
        # ctx branch: its term enters the first post layer of every first-order application
        dct = pl['dct']
        if first_order_apps == 1:
            o.copy2d(dct, pl['dp'][0][:M])
        else:
            o.add(dct, pl['dp'][0][:M], pl['dp'][0][M:2 * M])
            for k in range(2, first_order_apps):
                o.copy2d(dct, pl['dp'][0][k * M:(k + 1) * M], accumulate=True)
        with o.side_lane(second=True):      # ctx-branch gradients (Conv1D wgrad, FCs) beside the spectral-branch weight gradients
            o.colsum(dct, g['c.post.fc0.bias'])
            o.gemm(pl['ctx_last'], dct, g['c.post.fc0.kernel'][SF:], transA=True)
            dcur = pl['dcA']
            o.gemm(dct, p['c.post.fc0.kernel'][SF:], dcur, transB=True, aux=pl['fc_y'][1])
            for i in (1, 0):
                if i > 0:
                    xin = pl['fc_y'][0]
                else:
                    xin = pl['ctx_y'][-1] if a.ctx_nbcnnlayers > 0 else pl['ctx_in'].view(M, -1)
                o.gemm(xin, dcur, g['c.ctx.fc%d.kernel' % i], transA=True)
                o.colsum(dcur, g['c.ctx.fc%d.bias' % i])
                if i > 0 or a.ctx_nbcnnlayers > 0:
                    dnext = pl['dcB'] if dcur is pl['dcA'] else pl['dcA']
                    o.gemm(dcur, p['c.ctx.fc%d.kernel' % i], dnext, transB=True, aux=xin)
                    dcur = dnext
            left = (W - 1) // 2
            for i in range(a.ctx_nbcnnlayers - 1, -1, -1):
                o.conv1d_wgrad(pl['ctx_xp'][i], dcur, g['c.ctx.cnn%d.kernel' % i], T)
                o.colsum(dcur, g['c.ctx.cnn%d.bias' % i])
                if i > 0:
                    o.pad_time(pl['ctx_dzp'], dcur.view(B, T, H), W - 1 - left, left)
                    o.conv2d_flip(p['c.ctx.cnn%d.kernel' % i].view(W, 1, H, H), pl['ctx_kflip'])
                    dnext = pl['dcB'] if dcur is pl['dcA'] else pl['dcA']
                    o.gemm(_im2col_view(pl['ctx_dzp'], T, W), pl['ctx_kflip'].view(W * H, H), dnext, aux=pl['ctx_y'][i - 1])
                    dcur = dnext
            # bias gradients of the post layers: first-order applications only, straight from dp / dd
            fM = first_order_apps * M
            for i in (1, 2):
                o.colsum(pl['dp'][i][:fM], g['c.post.fc%d.bias' % i])
            o.colsum(pl['dd'][:fM], g['c.out.bias'])
        pl['ctx_grads_done'] = True

    def _conv_wgrads(self, pl, samples, first):
        """Weight gradients of the spectral Conv2D stack over the applications `samples` (slice over the napp*B axis): written
        together with the bias gradients (`first`: first-order applications), or accumulated without them (tangent rows)."""
        o, g = self.ops, self.params.g
        layers = [dict(x=(pl['a'][i - 1] if i > 0 else pl['s0'])[samples], dy=pl['da'][i][samples], dw=g['c.spec.cnn%d.kernel' % i],
                       dbias=g['c.spec.cnn%d.bias' % i] if first else None) for i in range(self.a.gen_nbcnnlayers)]
        if hasattr(o, 'conv2d_wgrad_multi_ok') and o.conv2d_wgrad_multi_ok(layers):
            o.conv2d_wgrad_multi(layers, accumulate=not first)       # the whole stack in one launch
            return
        for L in layers:
            o.conv2d_wgrad(L['x'], L['dy'], L['dw'], dbias=L['dbias'], accumulate=not first)

    def param_grads(self, *args, **kw):
        with self.ops.backward_terms():      # gradient passes: linear in their seeds (ops.bwd_nterms)
            return self._param_grads(*args, **kw)

    def _param_grads(self, pl, first_order_apps=2):
        """Gradients of every critic parameter from the stored layer inputs (activations, or tangents on the GP rows)
        and pre-activation gradients.  Kernels use all napp applications; biases and the ctx branch only the first
        `first_order_apps` (the gradient penalty reaches neither: optimizertts_wgan.py:57-68 differentiates through
        piecewise-constant masks)."""
        o, a, p, g = self.ops, self.a, self.params.p, self.params.g
        B, T, M, nM = pl['B'], pl['T'], pl['M'], pl['nM']
        H, S, Fl, W, L = a.hiddenwidth, a.spec_size, a.gen_nbfilters, a.ctx_winlen, a.gen_nbcnnlayers
        SF = S * Fl
        fM = first_order_apps * M
        if not pl.get('ctx_grads_done'):
            self.param_grads_ctx(pl, first_order_apps)
        pl['ctx_grads_done'] = False
        o.wait_lane_event(pl.pop('side_event', None))      # the other applications' dp / da (side lane) are complete
        # spectral kernels: first-order applications (with the biases), then the gradient-penalty rows accumulated on top --
        # always as these two launches per layer, whichever lane ran the first, so that both schedules give the same bits
        done = pl.pop('conv_wgrads_apps', 0)
        assert done in (0, first_order_apps)
        if not done:
            self._conv_wgrads(pl, slice(0, first_order_apps * B), first=True)
        if pl['nB'] > first_order_apps * B:
            self._conv_wgrads(pl, slice(first_order_apps * B, pl['nB']), first=False)
        o.gemm(pl['a'][L - 1].view(nM, SF), pl['dp'][0], g['c.post.fc0.kernel'][:SF], transA=True)
        for i in (1, 2):
            o.gemm(pl['p'][i - 1], pl['dp'][i], g['c.post.fc%d.kernel' % i], transA=True)
        o.gemm(pl['p'][2], pl['dd'], g['c.out.kernel'], transA=True)
        o.join_lanes()
