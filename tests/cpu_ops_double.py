"""
TEST DOUBLE of percivaltts_b200.ops.CudaOps -- plain torch on the CPU, same method signatures and semantics.

It exists so that the host-side schedules of percivaltts_b200/engine.py and wgan.py (hand-derived backward and
double-backward passes, buffer views, data-parallel scaling) can be checked against the oracle in this GPU-less
container.  It lives under tests/ and is never importable from the product package: the product path has no CPU
fallback.  On the GPU box, tests/test_kernels_gpu.py checks every CUDA entry point against these same definitions.
"""
import torch
import torch.nn.functional as F

EPI_BIAS, EPI_LRELU, EPI_SIGMOID, EPI_DLRELU_AUX, EPI_ACCUM, EPI_ATOMIC, EPI_ADDMOD = 1, 2, 4, 8, 16, 32, 64
ALPHA = 0.3


def _lrelu(x):
    return torch.where(x > 0, x, ALPHA * x)


def _dlrelu_from_out(y):
    return torch.where(y > 0, torch.ones_like(y), torch.full_like(y, ALPHA))


def _rows(t):
    return t.reshape(-1, t.shape[-1])


class _Ws(object):
    def __init__(self, C, dtype):
        self.sums = torch.zeros(2, C, dtype=torch.float64)


class CpuOpsDouble(object):
    name = 'cpu-double'

    def __init__(self, dtype=torch.float32):
        self.dtype = dtype
        self.device = torch.device('cpu')
        self.launches = 0
        self._red = {}

    def empty(self, *shape, dtype=None):
        return torch.full(shape, float('nan') if (dtype or self.dtype).is_floating_point else 0, dtype=dtype or self.dtype)

    def zeros(self, *shape, dtype=None):
        return torch.zeros(*shape, dtype=dtype or self.dtype)

    def reduce_ws(self, rows, C):
        key = (int(rows), int(C))
        if key not in self._red:
            self._red[key] = _Ws(C, self.dtype)
        return self._red[key]

    # ---- operand-shadow bookkeeping of the CUDA path: nothing to do here -----------------------------------------
    def step_scope(self):
        import contextlib
        return contextlib.nullcontext(self)

    def shape_scope(self, tag):
        import contextlib
        return contextlib.nullcontext(self)

    def release_shape(self, tag):
        pass

    def side_lane(self, second=False, lane=None, nowait=False):
        import contextlib
        return contextlib.nullcontext(self)

    def full_terms(self):
        import contextlib
        return contextlib.nullcontext(self)

    def backward_terms(self):
        import contextlib
        return contextlib.nullcontext(self)

    def join_lanes(self, which=None):
        pass

    def share_rows(self, t):
        pass

    def lane_event(self):
        return None

    def wait_lane_event(self, ev):
        pass

    def preshare(self, t):
        pass

    def register_persistent(self, flat):
        pass

    def refresh_persistent(self, flat):
        pass

    # ---- conv1d (virtual im2col) ---------------------------------------------------------------------------------
    def conv1d(self, xp, kernel, y, T, bias=None, aux=None, flags=0):
        B, Tp, CIN = xp.shape
        W, _, COUT = kernel.shape
        self.gemm(xp.as_strided((B, T, W * CIN), (Tp * CIN, CIN, 1)), kernel.reshape(W * CIN, COUT), y, bias=bias, aux=aux,
                  flags=flags)

    def conv1d_wgrad(self, xp, dz, dw, T):
        B, Tp, CIN = xp.shape
        W, _, COUT = dw.shape
        self.gemm(xp.as_strided((B, T, W * CIN), (Tp * CIN, CIN, 1)), dz, dw.view(W * CIN, COUT), transA=True)

    # ---- gemm ------------------------------------------------------------------------------------------------
    def gemm(self, A, B, C, transA=False, transB=False, alpha=1.0, bias=None, aux=None, addmod=None, flags=0):
        self.launches += 1
        Am, Bm = _rows(A), _rows(B)
        if transA:
            Am = Am.t()
        if transB:
            Bm = Bm.t()
        R = alpha * (Am @ Bm)
        M, N = R.shape
        assert C.numel() == M * N and C.shape[-1] == N, (C.shape, R.shape)
        if bias is not None:
            R = R + bias.reshape(1, N)
        if addmod is not None:
            idx = torch.arange(M) % addmod.shape[0]
            R = R + addmod[idx]
        if flags & EPI_LRELU:
            R = _lrelu(R)
        if flags & EPI_SIGMOID:
            R = torch.sigmoid(R)
        if aux is not None:
            R = R * _dlrelu_from_out(_rows(aux))
        R = R.reshape(C.shape)
        if flags & EPI_ACCUM:
            C.add_(R)
        else:
            C.copy_(R)

    # ---- conv2d ----------------------------------------------------------------------------------------------
    @staticmethod
    def _conv(x, w, pad_t, pad_f):
        KH, KW = w.shape[0], w.shape[1]
        xp = F.pad(x.permute(0, 3, 1, 2), (pad_f, KW - 1 - pad_f, pad_t, KH - 1 - pad_t))
        return F.conv2d(xp, w.permute(3, 2, 0, 1)).permute(0, 2, 3, 1)

    def conv2d(self, x, w, y, bias=None, aux=None, pad=None, flags=0, stats=False, in_bn=None, stats_final=True):
        if in_bn is not None:
            mean, invstd, gamma, beta, scratch = in_bn
            C = x.shape[-1]
            self.bn_apply(x.reshape(-1, C), scratch.view(-1, C), mean, invstd, gamma, beta, 0.0, 0, 1)
            x = scratch
        self.launches += 1
        KH, KW = w.shape[0], w.shape[1]
        pad_t, pad_f = ((KH - 1) // 2, (KW - 1) // 2) if pad is None else pad
        R = self._conv(x, w, pad_t, pad_f)
        if bias is not None:
            R = R + bias
        if flags & EPI_LRELU:
            R = _lrelu(R)
        if aux is not None:
            R = R * _dlrelu_from_out(aux)
        if flags & EPI_ACCUM:
            y.add_(R)
        else:
            y.copy_(R)
        if stats:
            return self.bn_stats(y.reshape(-1, y.shape[-1]))
        return None

    def conv2d_chain_ok(self, T, F, KH, KW, channels):
        return getattr(self, 'conv_chain', False) and KH == 5 and KW == 5 and channels == 4

    def conv2d_chain(self, x, layers):
        """Same semantics as CudaOps.conv2d_chain, layer by layer."""
        cur = x
        for L in layers:
            w = L['w']
            if L.get('flip', False):
                wt = torch.empty(w.shape[0], w.shape[1], w.shape[3], w.shape[2], dtype=w.dtype)
                self.conv2d_flip(w, wt)
                w = wt
            self.conv2d(cur, w, L['y'], bias=L.get('bias'), aux=L.get('aux'), flags=int(L.get('flags', 0)))
            cur = L['y']

    def conv2d_flip(self, w, wt):
        self.launches += 1
        wt.copy_(torch.flip(w, dims=(0, 1)).permute(0, 1, 3, 2))

    def conv2d_dgrad(self, dy, w, dx, scratch, aux=None, flags=0):
        self.conv2d_flip(w, scratch)
        self.conv2d(dy, scratch, dx, aux=aux, flags=flags)

    def conv2d_wgrad(self, x, dy, dw, scale=1.0, accumulate=False, pad=None, dbias=None, bias_samples=None):
        self.launches += 1
        if dbias is not None:
            nb = x.shape[0] if bias_samples is None else bias_samples
            dbias.copy_(dy[:nb].sum(dim=(0, 1, 2)))
        KH, KW, CIN, COUT = dw.shape
        pad_t, pad_f = ((KH - 1) // 2, (KW - 1) // 2) if pad is None else pad
        T, Fq = x.shape[1], x.shape[2]
        xp = F.pad(x, (0, 0, pad_f, KW - 1 - pad_f, pad_t, KH - 1 - pad_t))
        R = torch.zeros(KH, KW, CIN, COUT, dtype=x.dtype)
        for kt in range(KH):
            for kf in range(KW):
                R[kt, kf] = torch.einsum('ntfi,ntfo->io', xp[:, kt:kt + T, kf:kf + Fq], dy)
        R = R * scale
        if accumulate:
            dw.add_(R)
        else:
            dw.copy_(R)

    # ---- batch norm ------------------------------------------------------------------------------------------
    def bn_stats(self, x, final=True):
        self.launches += 1
        ws = self.reduce_ws(*x.shape)
        xd = x.double()
        ws.sums[0] = xd.sum(0)
        ws.sums[1] = (xd * xd).sum(0)
        return ws

    def bn_finalize(self, ws, count, eps, momentum, mean, invstd, moving_mean=None, moving_var=None):
        self.launches += 1
        m = ws.sums[0] / count
        v = torch.clamp(ws.sums[1] / count - m * m, min=0.0)
        mean.copy_(m.to(mean.dtype))
        invstd.copy_((1.0 / torch.sqrt(v + eps)).to(invstd.dtype))
        if moving_mean is not None:
            moving_mean.copy_(momentum * moving_mean + (1.0 - momentum) * m.to(mean.dtype))
            moving_var.copy_(momentum * moving_var + (1.0 - momentum) * v.to(mean.dtype))

    def bn_apply(self, x, y, mean, invstd_or_var, gamma, beta, eps, use_var, lrelu, planes=False):
        self.launches += 1
        istd = torch.rsqrt(invstd_or_var + eps) if use_var else invstd_or_var
        v = (x - mean) * istd * gamma + beta
        y.copy_(_lrelu(v) if lrelu else v)

    def _bn_dz(self, x, dy, mean, invstd, gamma, beta, lrelu):
        xh = (x - mean) * invstd
        dz = dy
        if lrelu:
            z = gamma * xh + beta
            dz = dy * torch.where(z > 0, torch.ones_like(z), torch.full_like(z, ALPHA))
        return xh, dz

    def bn_bwd_reduce(self, x, dy, mean, invstd, gamma, beta, lrelu):
        self.launches += 1
        ws = self.reduce_ws(*x.shape)
        xh, dz = self._bn_dz(x, dy, mean, invstd, gamma, beta, lrelu)
        ws.sums[0] = dz.double().sum(0)
        ws.sums[1] = (dz.double() * xh.double()).sum(0)
        return ws

    def bn_bwd_apply(self, ws, x, dy, dx, mean, invstd, gamma, beta, count, lrelu, dgamma, dbeta, grad_scale=1.0,
                     accumulate=False, planes=False):
        self.launches += 1
        xh, dz = self._bn_dz(x, dy, mean, invstd, gamma, beta, lrelu)
        m1 = (ws.sums[0] / count).to(x.dtype)
        m2 = (ws.sums[1] / count).to(x.dtype)
        dx.copy_(gamma * invstd * (dz - m1 - xh * m2))
        for gtensor, s in ((dbeta, ws.sums[0]), (dgamma, ws.sums[1])):
            if gtensor is not None:
                v = (s * grad_scale).to(gtensor.dtype)
                if accumulate:
                    gtensor.add_(v)
                else:
                    gtensor.copy_(v)

    def colsum(self, x, out, accumulate=False):
        self.launches += 1
        v = x.double().sum(0).to(out.dtype).reshape(out.shape)
        if accumulate:
            out.add_(v)
        else:
            out.copy_(v)

    # ---- LSTM ------------------------------------------------------------------------------------------------
    def lstm_fwd(self, zx, U, hout, gates=None, cstate=None, cluster_size=8, batch_chunk=2):
        self.launches += 1
        ndir = len(zx)
        B, T, H4 = zx[0].shape
        H = H4 // 4
        for d in range(ndir):
            h = torch.zeros(B, H, dtype=zx[0].dtype)
            c = torch.zeros(B, H, dtype=zx[0].dtype)
            order = range(T - 1, -1, -1) if d == 1 else range(T)
            for t in order:
                z = zx[d][:, t] + h @ U[d]
                i, f = torch.sigmoid(z[:, :H]), torch.sigmoid(z[:, H:2 * H])
                g, o = torch.tanh(z[:, 2 * H:3 * H]), torch.sigmoid(z[:, 3 * H:])
                c = f * c + i * g
                h = o * torch.tanh(c)
                hout[:, t, d * H:(d + 1) * H] = h
                if gates is not None:
                    gates[d][:, t] = torch.cat([i, f, g, o], dim=1)
                    cstate[d][:, t] = c

    def lstm_bwd(self, dhout, U, gates, cstate, dz, cluster_size=8, batch_chunk=2):
        self.launches += 1
        ndir = len(U)
        B, T, H4 = gates[0].shape
        H = H4 // 4
        for d in range(ndir):
            dhrec = torch.zeros(B, H, dtype=dhout.dtype)
            dccar = torch.zeros(B, H, dtype=dhout.dtype)
            order = range(T) if d == 1 else range(T - 1, -1, -1)
            for t in order:
                tprev = t + 1 if d == 1 else t - 1
                gt = gates[d][:, t]
                i, f, g, o = gt[:, :H], gt[:, H:2 * H], gt[:, 2 * H:3 * H], gt[:, 3 * H:]
                c = cstate[d][:, t]
                cprev = cstate[d][:, tprev] if 0 <= tprev < T else torch.zeros_like(c)
                dh = dhout[:, t, d * H:(d + 1) * H] + dhrec
                tc = torch.tanh(c)
                dc = dccar + dh * o * (1 - tc * tc)
                dzt = torch.cat([dc * g * i * (1 - i), dc * cprev * f * (1 - f), dc * i * (1 - g * g), dh * tc * o * (1 - o)], 1)
                dz[d][:, t] = dzt
                dccar = dc * f
                dhrec = dzt @ U[d].t()

    # ---- WGAN-GP pieces --------------------------------------------------------------------------------------
    def critic_pack_inputs(self, real, fake, alpha, s0, S, off, skip_real=False):
        self.launches += 1
        B = real.shape[0]
        r, k = real[:, :, off:off + S], fake[:, :, off:off + S]
        a = alpha.reshape(B, 1, 1)
        if not skip_real:
            s0[0:B] = r
        s0[B:2 * B] = k
        s0[2 * B:] = a * r + (1 - a) * k

    def fill_segments(self, p, n0, v0, n1=0, v1=0.0, n2=0, v2=0.0):
        self.launches += 1
        assert p.numel() == n0 + n1 + n2
        p[:n0] = v0
        p[n0:n0 + n1] = v1
        p[n0 + n1:] = v2

    def sum_scaled(self, x, scale, losses, slot):
        self.launches += 1
        losses[slot] = float(x.double().sum() * scale)

    def gp_norms(self, g, ss):
        self.launches += 1
        B = g.shape[0]
        ss[:B] = (g.reshape(B, -1).double() ** 2).sum(1).to(ss.dtype)

    def gp_scale(self, g, ss, r0, lam, Bglobal, losses, slot):
        self.launches += 1
        B = g.shape[0]
        nb = torch.sqrt(ss[:B])
        r0.copy_((lam * 2.0 * (nb - 1.0) / nb / Bglobal).reshape(B, 1) * g.reshape(B, -1))
        losses[slot] = float(((1.0 - nb) ** 2).sum() / Bglobal)

    def loss_combine(self, losses, out, a, ca, b=-1, cb=0.0, c=-1, cc=0.0):
        self.launches += 1
        v = ca * losses[a]
        if b >= 0:
            v = v + cb * losses[b]
        if c >= 0:
            v = v + cc * losses[c]
        losses[out] = v

    def lse_loss_grad(self, pred, real, w, dpred, scale_loss, scale_grad, losses, slot):
        self.launches += 1
        d = pred - real
        wv = w if w is not None else torch.ones(pred.shape[-1], dtype=pred.dtype)
        losses[slot] = float((d * d * wv).double().sum() * scale_loss)
        if dpred is not None:
            dpred.copy_(scale_grad * 2.0 * d * wv)

    def lse_loss_grad_cols(self, pred, real, w, dpred, c0, nc, scale_loss, scale_grad, losses, slot):
        self.launches += 1
        d = (pred - real)[..., c0:c0 + nc]
        wv = (w if w is not None else torch.ones(pred.shape[-1], dtype=pred.dtype))[c0:c0 + nc]
        losses[slot] = float((d * d * wv).double().sum() * scale_loss)
        if dpred is not None:
            dpred[..., c0:c0 + nc] = scale_grad * 2.0 * d * wv

    def adam(self, p, g, m, v, lr, beta1, beta2, eps, step):
        self.launches += 1
        t = int(step[0]) + 1
        lr_t = lr * (1.0 - beta2 ** t) ** 0.5 / (1.0 - beta1 ** t)
        m.copy_(beta1 * m + (1 - beta1) * g)
        v.copy_(beta2 * v + (1 - beta2) * g * g)
        p.sub_(lr_t * m / (torch.sqrt(v) + eps))
        step += 1

    # ---- small movers ----------------------------------------------------------------------------------------
    def add3(self, out, a, b, c):
        self.launches += 1
        out.copy_((a + b) + c)

    def add(self, out, a, b):
        self.launches += 1
        out.copy_(a + b)

    def sigmoid_bwd(self, dy, y, dz):
        self.launches += 1
        dz.copy_(dy * y * (1 - y))

    def copy2d(self, dst, src, accumulate=False):
        self.launches += 1
        if accumulate:
            dst.add_(src)
        else:
            dst.copy_(src)

    def pad_time(self, dst, src, left, right):
        self.launches += 1
        T = src.shape[1]
        dst.zero_()
        dst[:, left:left + T] = src

    def uniform(self, out, seed, counter):
        self.launches += 1
        g = torch.Generator().manual_seed(int(seed) * 1000003 + int(counter[0]))
        out.copy_(torch.rand(out.shape, generator=g).to(out.dtype))
        counter += 1

    # ---- PML -------------------------------------------------------------------------------------------------
    def denorm_split(self, cmp, mean, std, S, NM, out=None, f0=None, spec=None, nm=None):
        v = cmp if mean is None else cmp * std + mean
        if out is not None:
            out.copy_(v)
        if f0 is not None:
            f0.copy_(torch.exp(v[:, 0]))
        if spec is not None:
            spec.copy_(v[:, 1:1 + S])
        if nm is not None:
            nm.copy_(v[:, 1 + S:1 + S + NM])
        self.launches += 1

    def pml_linbnd2fwbnd(self, x, idx, wgt, out, log_abs):
        self.launches += 1
        j = idx.long()
        j1 = torch.clamp(j + 1, max=x.shape[1] - 1)
        a, b = x[:, j], x[:, j1]
        if log_abs:
            a, b = torch.log(torch.abs(a)), torch.log(torch.abs(b))
        out.copy_(a + wgt * (b - a))

    def pml_fwbnd2linbnd(self, x, idx, wgt, out, smooth, do_exp):
        self.launches += 1
        K = x.shape[1]
        j = idx.long()
        p1, p2 = x[:, j], x[:, torch.clamp(j + 1, max=K - 1)]
        w = wgt
        if smooth:
            p0, p3 = x[:, torch.clamp(j - 1, min=0)], x[:, torch.clamp(j + 2, max=K - 1)]
            v = 0.5 * (2 * p1 + (p2 - p0) * w + (2 * p0 - 5 * p1 + 4 * p2 - p3) * w * w + (3 * p1 - p0 - 3 * p2 + p3) * w ** 3)
        else:
            v = p1 + w * (p2 - p1)
        out.copy_(torch.exp(v) if do_exp else v)
