"""
CPU ORACLE -- TEST INFRASTRUCTURE ONLY (never imported by the product package).

A PyTorch-CPU restatement of the PercivalTTS WGAN-GP training step, written from the
reference's graph definition.  Only tests/, __graft_entry__.smoke() and bench.py's
cpu_baseline / --impl reference legs may import this file.

PARITY UNPINNED (vs TensorFlow): the reference needs Python 2.7 + TensorFlow 1.9 + Keras
2.2.2 + h5py + the (missing) `sigproc` submodule, none of which exist in this image, and the
reference's own tests pin no loss / output / gradient value for this path.
What *is* pinned by the reference's tests, and checked in tests/test_oracle.py:
  * count_params()==2195            tests/test_smoke_tensorflowkeras.py:51-53
  * feature sizes 83/84/249/252     tests/test_smoke_tensorflowkeras.py:23,133,149,158
  * label frames 667 / 5694         tests/test_smoke.py:80-85
The oracle is additionally validated by (a) an independent numpy forward (oracle/numpy_forward.py),
(b) fp64-vs-fp32 agreement and (c) finite differences of the gradient penalty.

Keras / TF-1.9 semantics hard-coded here:
  Dense acts on the last axis; Conv = zero-padded 'same' cross-correlation, HWIO / WIO kernels;
  BatchNormalization(axis=-1, momentum=0.99, eps=1e-3): training mode uses the biased batch
  variance, moving stats <- 0.99*moving + 0.01*batch (biased variance);  LeakyReLU(alpha=0.3);
  LSTM gates in order i,f,c,o with tanh / sigmoid, unit_forget_bias; Bidirectional concat
  [fwd, bwd]; Adam  p -= lr*sqrt(1-b2^t)/(1-b1^t) * m/(sqrt(v)+eps), eps=1e-7.

Reference lines followed:
  generator   percivaltts/modeltts_common.py:61-126   layer builders percivaltts/networktts.py:59-126
  critic      percivaltts/networks_critic.py:44-96
  step        percivaltts/optimizertts_wgan.py:44-79 (losses), :107-213 (graphs), :216-241 (schedule)
  LS weights  percivaltts/optimizertts_wgan.py:186-210, percivaltts/backend_tensorflow.py:107-109
"""
import math
from collections import OrderedDict

import numpy as np
import torch
import torch.nn.functional as F

LRELU_ALPHA = 0.3      # networktts.py:62,119,125 ; networks_critic.py:68
BN_EPS = 1e-3          # Keras BatchNormalization default
BN_MOMENTUM = 0.99     # Keras BatchNormalization default


class Arch(object):
    """The cfgarch fields the graph builders read (modeltts_common.py:76-121, networks_critic.py:62-94)."""

    def __init__(self, ctxsize=425, spec_size=129, nm_size=33, hiddenwidth=256, ctx_nbcnnlayers=1,
                 ctx_winlen=21, gen_nbcnnlayers=8, gen_nbfilters=4, gen_winlen=5, spec_freqlen=5):
        self.ctxsize = ctxsize
        self.spec_size = spec_size
        self.nm_size = nm_size
        self.hiddenwidth = hiddenwidth
        self.ctx_nbcnnlayers = ctx_nbcnnlayers
        self.ctx_winlen = ctx_winlen
        self.gen_nbcnnlayers = gen_nbcnnlayers
        self.gen_nbfilters = gen_nbfilters
        self.gen_winlen = gen_winlen
        self.spec_freqlen = spec_freqlen

    @property
    def outsize(self):
        return 1 + self.spec_size + self.nm_size     # vocoders.py:176-177


def demo_arch():
    """percivaltts/run.py:76,89,114-121"""
    return Arch()


def tiny_arch():
    """tests/test_smoke_tensorflowkeras.py:183-190"""
    return Arch(ctxsize=425, spec_size=65, nm_size=17, hiddenwidth=2, ctx_nbcnnlayers=2, ctx_winlen=3,
                gen_nbcnnlayers=2, gen_nbfilters=2, gen_winlen=3, spec_freqlen=3)


# ----------------------------------------------------------------------------------------------
# parameter inventories (name -> shape), Keras layouts
# ----------------------------------------------------------------------------------------------

def _bn_entries(prefix, ch):
    return [(prefix + '.gamma', (ch,)), (prefix + '.beta', (ch,)),
            (prefix + '.moving_mean', (ch,)), (prefix + '.moving_var', (ch,))]


def generator_param_shapes(a):
    """Ordered (name, shape, trainable) for DCNNF0SpecNoiseFeatures (modeltts_common.py:65-126)."""
    H, S, NMS, Fl = a.hiddenwidth, a.spec_size, a.nm_size, a.gen_nbfilters
    out = []
    cin = a.ctxsize
    for i in range(a.ctx_nbcnnlayers):                                  # :76-77  pCNN1D (no bias, BN)
        out.append(('g.ctx.cnn%d.kernel' % i, (a.ctx_winlen, cin, H)))
        out += _bn_entries('g.ctx.cnn%d.bn' % i, H)
        cin = H
    for i in range(2):                                                  # :78-79  pFC x2
        out.append(('g.ctx.fc%d.kernel' % i, (cin, H)))
        out += _bn_entries('g.ctx.fc%d.bn' % i, H)
        cin = H
    for d in ('fw', 'bw'):                                              # :82-83  pBLSTM
        out.append(('g.f0.blstm.%s.kernel' % d, (cin, 4 * H)))
        out.append(('g.f0.blstm.%s.recurrent' % d, (H, 4 * H)))
        out.append(('g.f0.blstm.%s.bias' % d, (4 * H,)))
    out.append(('g.f0.out.kernel', (2 * H, 1)))                         # :84
    out.append(('g.f0.out.bias', (1,)))
    out.append(('g.spec.proj.kernel', (cin, S)))                        # :95
    out.append(('g.spec.proj.bias', (S,)))
    c2 = 1
    for i in range(a.gen_nbcnnlayers):                                  # :97-98 pCNN2D (no bias, BN)
        out.append(('g.spec.cnn%d.kernel' % i, (a.gen_winlen, a.spec_freqlen, c2, Fl)))
        out += _bn_entries('g.spec.cnn%d.bn' % i, Fl)
        c2 = Fl
    out.append(('g.spec.out.kernel', (a.gen_winlen, a.spec_freqlen, c2, 1)))   # :100
    out.append(('g.spec.out.bias', (1,)))
    cn = cin
    for i in range(a.gen_nbcnnlayers // 2):                             # :119 (python-2 integer division)
        out.append(('g.nm.fc%d.kernel' % i, (cn, H)))
        out += _bn_entries('g.nm.fc%d.bn' % i, H)
        cn = H
    out.append(('g.nm.out.kernel', (cn, NMS)))                          # :121
    out.append(('g.nm.out.bias', (NMS,)))
    return [(n, s, not (n.endswith('moving_mean') or n.endswith('moving_var'))) for n, s in out]


def critic_param_shapes(a):
    """Ordered (name, shape, trainable) for Critic (networks_critic.py:44-96); bn=False everywhere (:50)."""
    H, S, Fl = a.hiddenwidth, a.spec_size, a.gen_nbfilters
    out = []
    if a.gen_nbcnnlayers > 0:
        c2 = 1
        for i in range(a.gen_nbcnnlayers):                              # :66-68
            out.append(('c.spec.cnn%d.kernel' % i, (a.gen_winlen, a.spec_freqlen, c2, Fl)))
            out.append(('c.spec.cnn%d.bias' % i, (Fl,)))
            c2 = Fl
        specfeat = S * c2                                               # :70
    else:
        cin = S
        for i in range(3):                                              # :73-75
            out.append(('c.spec.fc%d.kernel' % i, (cin, H)))
            out.append(('c.spec.fc%d.bias' % i, (H,)))
            cin = H
        specfeat = H
    cin = a.ctxsize
    for i in range(a.ctx_nbcnnlayers):                                  # :82-83
        out.append(('c.ctx.cnn%d.kernel' % i, (a.ctx_winlen, cin, H)))
        out.append(('c.ctx.cnn%d.bias' % i, (H,)))
        cin = H
    for i in range(2):                                                  # :84-85
        out.append(('c.ctx.fc%d.kernel' % i, (cin, H)))
        out.append(('c.ctx.fc%d.bias' % i, (H,)))
        cin = H
    cin = specfeat + cin                                                # :89 concat [spec, ctx]
    for i in range(3):                                                  # :91-93
        out.append(('c.post.fc%d.kernel' % i, (cin, H)))
        out.append(('c.post.fc%d.bias' % i, (H,)))
        cin = H
    out.append(('c.out.kernel', (cin, 1)))                              # :96
    out.append(('c.out.bias', (1,)))
    return [(n, s, True) for n, s in out]


def generic_fc_count_params(ctxsize, spec_size, nm_size, hiddenwidth, nlayers=3):
    """Keras count_params() of Generic(['FC']*nlayers) + network_final for PML without MLPG
    (modeltts_common.py:36-55, networktts.py:59-63,192-197).  Pinned: 2195 for (425,65,17,4,3)."""
    n, cin = 0, ctxsize
    for _ in range(nlayers):
        n += cin * hiddenwidth + 4 * hiddenwidth        # Dense(no bias) + BN(gamma,beta,mean,var)
        cin = hiddenwidth
    n += cin * (1 + spec_size) + (1 + spec_size)        # lo_f0spec
    n += cin * nm_size + nm_size                        # lo_nm
    return n


def count_params(shapes):
    return int(sum(int(np.prod(s)) for _, s, _ in shapes))


# ----------------------------------------------------------------------------------------------
# initialisation (Keras defaults: glorot_uniform kernels, zeros biases, orthogonal recurrent,
# unit_forget_bias, BN gamma=1 beta=0 mean=0 var=1)
# ----------------------------------------------------------------------------------------------

def _glorot(rng, shape):
    if len(shape) == 2:
        fan_in, fan_out = shape
    else:
        rf = int(np.prod(shape[:-2]))
        fan_in, fan_out = rf * shape[-2], rf * shape[-1]
    lim = math.sqrt(6.0 / (fan_in + fan_out))
    return rng.uniform(-lim, lim, size=shape)


def _orthogonal(rng, shape):
    a = rng.normal(0.0, 1.0, size=(shape[1], shape[0])) if shape[0] < shape[1] else rng.normal(0.0, 1.0, size=shape)
    q, r = np.linalg.qr(a)
    q = q * np.sign(np.diag(r))
    return q.T if shape[0] < shape[1] else q


def init_params(shapes, seed=123, dtype=np.float32):
    rng = np.random.RandomState(seed)
    w = OrderedDict()
    for name, shape, _ in shapes:
        if name.endswith('.recurrent'):
            v = _orthogonal(rng, shape)
        elif name.endswith('.kernel'):
            v = _glorot(rng, shape)
        elif name.endswith('.gamma') or name.endswith('.moving_var'):
            v = np.ones(shape)
        elif name.endswith('.bias') and '.blstm.' in name:
            v = np.zeros(shape)
            H = shape[0] // 4
            v[H:2 * H] = 1.0                      # unit_forget_bias
        else:
            v = np.zeros(shape)
        w[name] = np.ascontiguousarray(v, dtype=dtype)
    return w


def perturb_params(w, seed=7, scale=0.1):
    """Move biases / BN affine / moving stats off their trivial initial values so parity tests exercise them."""
    rng = np.random.RandomState(seed)
    out = OrderedDict()
    for k, v in w.items():
        if k.endswith('.kernel') or k.endswith('.recurrent'):
            out[k] = v.copy()
        elif k.endswith('.moving_var'):
            out[k] = (v + scale * rng.uniform(0, 1, size=v.shape)).astype(v.dtype)
        else:
            out[k] = (v + scale * rng.normal(size=v.shape)).astype(v.dtype)
    return out


def to_torch(w, dtype=torch.float32, requires_grad=False):
    out = OrderedDict()
    for k, v in w.items():
        t = torch.tensor(np.asarray(v), dtype=dtype)
        t.requires_grad_(requires_grad and not (k.endswith('moving_mean') or k.endswith('moving_var')))
        out[k] = t
    return out


# ----------------------------------------------------------------------------------------------
# layers
# ----------------------------------------------------------------------------------------------

def lrelu(x):
    return F.leaky_relu(x, LRELU_ALPHA)


def conv1d_same(x, k, bias=None):
    """x [B,T,Cin], k [W,Cin,Cout] (Keras) -> [B,T,Cout]; TF 'same': left=(W-1)//2, right=W-1-left."""
    W = k.shape[0]
    left = (W - 1) // 2
    xp = F.pad(x.permute(0, 2, 1), (left, W - 1 - left))
    y = F.conv1d(xp.contiguous(), k.permute(2, 1, 0).contiguous(), bias)
    return y.permute(0, 2, 1)


def conv2d_same(x, k, bias=None):
    """x [B,T,Fq,Cin] (NHWC), k [KH(time),KW(freq),Cin,Cout] (HWIO) -> [B,T,Fq,Cout]."""
    KH, KW = k.shape[0], k.shape[1]
    lt, lf = (KH - 1) // 2, (KW - 1) // 2
    xp = F.pad(x.permute(0, 3, 1, 2), (lf, KW - 1 - lf, lt, KH - 1 - lt))
    y = F.conv2d(xp.contiguous(), k.permute(3, 2, 0, 1).contiguous(), bias)
    return y.permute(0, 2, 3, 1)


def batchnorm(x, w, prefix, training, new_stats=None):
    """BatchNormalization(axis=-1).  Training: batch stats over all non-channel axes (biased variance)."""
    g, b = w[prefix + '.gamma'], w[prefix + '.beta']
    if training:
        red = tuple(range(x.dim() - 1))
        mean = x.mean(dim=red)
        var = ((x - mean) ** 2).mean(dim=red)
        if new_stats is not None:
            mm, mv = w[prefix + '.moving_mean'], w[prefix + '.moving_var']
            new_stats[prefix + '.moving_mean'] = (BN_MOMENTUM * mm + (1 - BN_MOMENTUM) * mean).detach()
            new_stats[prefix + '.moving_var'] = (BN_MOMENTUM * mv + (1 - BN_MOMENTUM) * var).detach()
    else:
        mean, var = w[prefix + '.moving_mean'], w[prefix + '.moving_var']
    return (x - mean) / torch.sqrt(var + BN_EPS) * g + b


def lstm_dir(x, W, U, b, reverse=False):
    """Keras LSTM(tanh, sigmoid, return_sequences) -- networktts.py:90-94.  x [B,T,D] -> [B,T,H]."""
    B, T, _ = x.shape
    H = U.shape[0]
    zx = x @ W + b
    h = x.new_zeros(B, H)
    c = x.new_zeros(B, H)
    outs = [None] * T
    order = range(T - 1, -1, -1) if reverse else range(T)
    for t in order:
        z = zx[:, t] + h @ U
        i = torch.sigmoid(z[:, :H])
        f = torch.sigmoid(z[:, H:2 * H])
        g = torch.tanh(z[:, 2 * H:3 * H])
        o = torch.sigmoid(z[:, 3 * H:])
        c = f * c + i * g
        h = o * torch.tanh(c)
        outs[t] = h
    return torch.stack(outs, dim=1)


def generator_forward(a, w, ctx, training, new_stats=None):
    """DCNNF0SpecNoiseFeatures graph (modeltts_common.py:65-126).  ctx [B,T,ctxsize] -> [B,T,1+S+NM]."""
    x = ctx
    for i in range(a.ctx_nbcnnlayers):
        x = lrelu(batchnorm(conv1d_same(x, w['g.ctx.cnn%d.kernel' % i]), w, 'g.ctx.cnn%d.bn' % i, training, new_stats))
    for i in range(2):
        x = lrelu(batchnorm(x @ w['g.ctx.fc%d.kernel' % i], w, 'g.ctx.fc%d.bn' % i, training, new_stats))
    l_ctx = x
    hf = lstm_dir(l_ctx, w['g.f0.blstm.fw.kernel'], w['g.f0.blstm.fw.recurrent'], w['g.f0.blstm.fw.bias'])
    hb = lstm_dir(l_ctx, w['g.f0.blstm.bw.kernel'], w['g.f0.blstm.bw.recurrent'], w['g.f0.blstm.bw.bias'], reverse=True)
    f0 = torch.cat([hf, hb], dim=-1) @ w['g.f0.out.kernel'] + w['g.f0.out.bias']
    s = l_ctx @ w['g.spec.proj.kernel'] + w['g.spec.proj.bias']
    s = s.unsqueeze(-1)
    for i in range(a.gen_nbcnnlayers):
        s = lrelu(batchnorm(conv2d_same(s, w['g.spec.cnn%d.kernel' % i]), w, 'g.spec.cnn%d.bn' % i, training, new_stats))
    s = conv2d_same(s, w['g.spec.out.kernel'], w['g.spec.out.bias']).squeeze(-1)
    n = l_ctx
    for i in range(a.gen_nbcnnlayers // 2):
        n = lrelu(batchnorm(n @ w['g.nm.fc%d.kernel' % i], w, 'g.nm.fc%d.bn' % i, training, new_stats))
    n = torch.sigmoid(n @ w['g.nm.out.kernel'] + w['g.nm.out.bias'])
    return torch.cat([f0, s, n], dim=-1)


def critic_forward(a, w, feats, ctx):
    """Critic graph (networks_critic.py:44-96).  (feats [B,T,1+S+NM], ctx [B,T,ctxsize]) -> [B,T,1]."""
    S = a.spec_size
    s = feats[:, :, 1:1 + S]                                            # :58
    if a.gen_nbcnnlayers > 0:
        s = s.unsqueeze(-1)
        for i in range(a.gen_nbcnnlayers):
            s = lrelu(conv2d_same(s, w['c.spec.cnn%d.kernel' % i], w['c.spec.cnn%d.bias' % i]))
        s = s.reshape(s.shape[0], s.shape[1], -1)                       # :70  freq-major, channel-minor
    else:
        for i in range(3):
            s = lrelu(s @ w['c.spec.fc%d.kernel' % i] + w['c.spec.fc%d.bias' % i])
    c = ctx
    for i in range(a.ctx_nbcnnlayers):
        c = lrelu(conv1d_same(c, w['c.ctx.cnn%d.kernel' % i], w['c.ctx.cnn%d.bias' % i]))
    for i in range(2):
        c = lrelu(c @ w['c.ctx.fc%d.kernel' % i] + w['c.ctx.fc%d.bias' % i])
    p = torch.cat([s, c], dim=-1)
    for i in range(3):
        p = lrelu(p @ w['c.post.fc%d.kernel' % i] + w['c.post.fc%d.bias' % i])
    return p @ w['c.out.kernel'] + w['c.out.bias']


# ----------------------------------------------------------------------------------------------
# losses and steps
# ----------------------------------------------------------------------------------------------

def mel(f):
    return 1127.0 * np.log1p(np.asarray(f, dtype=np.float64) / 700.0)


def freq2fwspecidx(freq, fs, nbbnds):
    """Position (fractional band index) of `freq` on the mel-warped band axis -- see oracle/pml_oracle.py."""
    return (nbbnds - 1) * mel(freq) / mel(0.5 * fs)


def wlswgan_weights(a, fs=16000, LScoef=0.25, transfreqcutoff=4000, transcoef=1.0 / 8.0, incnoise=False):
    """optimizertts_wgan.py:186-210 -> (w_wgan[out], w_ls[out], mean(w_wgan)) built in float64 like numpy does."""
    els = [np.zeros(1)]
    specvs = np.arange(a.spec_size, dtype=np.float32).astype(np.float64)
    if LScoef == 0.0:
        els.append(np.ones(a.spec_size))
    else:
        c = freq2fwspecidx(transfreqcutoff, fs, a.spec_size)
        els.append(1.0 / (1.0 + np.exp(-(specvs - c) * transcoef)))    # backend_tensorflow.py:107-109
    if a.nm_size > 0 and incnoise:
        if LScoef == 0.0:
            els.append(np.ones(a.nm_size))
        else:
            nv = np.arange(a.nm_size, dtype=np.float32).astype(np.float64)
            c = freq2fwspecidx(transfreqcutoff, fs, a.nm_size)
            els.append(1.0 / (1.0 + np.exp(-(nv - c) * transcoef)))
    else:
        els.append(np.zeros(a.nm_size))
    wg = np.hstack(els) * (1.0 - LScoef)
    return wg, 1.0 - wg, float(np.mean(wg))


def gradient_penalty(dout_hat, xhat, create_graph):
    """optimizertts_wgan.py:53-68.  K.gradients(y_pred, x) differentiates sum(y_pred); no eps in the sqrt."""
    g, = torch.autograd.grad(dout_hat.sum(), xhat, create_graph=create_graph)
    n = torch.sqrt((g ** 2).sum(dim=tuple(range(1, g.dim()))))
    return ((1.0 - n) ** 2).mean(), g, n


def critic_step(a, wg, wc, ctx, real, alpha, gp_lambda=10.0, gen_training=True):
    """Critic graph of prepare() (optimizertts_wgan.py:115-154): returns dict with the total loss, its three
    parts, the critic outputs and d(loss)/d(critic params).  alpha [B,1,1] replaces K.random_uniform (:50).
    The frozen generator runs with learning_phase=1, i.e. BN batch statistics."""
    with torch.no_grad():
        fake = generator_forward(a, wg, ctx, training=gen_training)
    xhat = (alpha * real + (1.0 - alpha) * fake).detach().requires_grad_(True)      # :51
    d_real = critic_forward(a, wc, real, ctx)
    d_fake = critic_forward(a, wc, fake, ctx)
    d_hat = critic_forward(a, wc, xhat, ctx)
    l_real = (-1.0 * d_real).mean()                                     # :70-71 with y_true=-1 (:152)
    l_fake = (+1.0 * d_fake).mean()
    gp, g, n = gradient_penalty(d_hat, xhat, create_graph=True)
    loss = l_real + l_fake + gp_lambda * gp                             # loss_weights=[1,1,lambda] (:150)
    names = [k for k, v in wc.items() if v.requires_grad]
    grads = torch.autograd.grad(loss, [wc[k] for k in names], allow_unused=True)
    gd = OrderedDict()
    for k, gr in zip(names, grads):
        gd[k] = torch.zeros_like(wc[k]) if gr is None else gr.detach()
    return dict(loss=loss.detach(), l_real=l_real.detach(), l_fake=l_fake.detach(), gp=gp.detach(),
                fake=fake, xhat=xhat.detach(), d_real=d_real.detach(), d_fake=d_fake.detach(),
                d_hat=d_hat.detach(), gp_grad=g.detach(), gp_norm=n.detach(), grads=gd)


def generator_step(a, wg, wc, ctx, real, errtype='WLSWGAN', fs=16000, LScoef=0.25, transfreqcutoff=4000,
                   transcoef=1.0 / 8.0, incnoise=False):
    """Generator graph of prepare() (optimizertts_wgan.py:160-213)."""
    new_stats = OrderedDict()
    pred = generator_forward(a, wg, ctx, training=True, new_stats=new_stats)
    valid = critic_forward(a, wc, pred, ctx)
    l_w = (-1.0 * valid).mean()
    if errtype == 'WGAN':
        loss, l_ls = l_w, torch.zeros(())
    else:
        w_wgan, w_ls, wbar = wlswgan_weights(a, fs, LScoef, transfreqcutoff, transcoef, incnoise)
        w_ls_t = torch.tensor(w_ls, dtype=pred.dtype)
        l_ls = (((real - pred) ** 2) * w_ls_t).mean()                   # :73-79
        loss = wbar * l_w + 1.0 * l_ls                                  # :213
    names = [k for k, v in wg.items() if v.requires_grad]
    grads = torch.autograd.grad(loss, [wg[k] for k in names], allow_unused=True)
    gd = OrderedDict()
    for k, gr in zip(names, grads):
        gd[k] = torch.zeros_like(wg[k]) if gr is None else gr.detach()
    return dict(loss=loss.detach(), l_w=l_w.detach(), l_ls=l_ls.detach(), pred=pred.detach(),
                valid=valid.detach(), grads=gd, new_stats=new_stats)


def generic_param_shapes(ctxsize, spec_size, nm_size, hiddenwidth, nlayers=3):
    """Ordered [(name, shape, trainable)] of Generic(['FC']*nlayers) + network_final for PML without MLPG
    (modeltts_common.py:36-55; pFC networktts.py:59-63; lo_f0spec / lo_nm networktts.py:196-197)."""
    out, cin = [], ctxsize
    for i in range(nlayers):
        out.append(('g.fc%d.kernel' % i, (cin, hiddenwidth)))
        out += _bn_entries('g.fc%d.bn' % i, hiddenwidth)
        cin = hiddenwidth
    out += [('g.f0spec.kernel', (cin, 1 + spec_size)), ('g.f0spec.bias', (1 + spec_size,)),
            ('g.nm.kernel', (cin, nm_size)), ('g.nm.bias', (nm_size,))]
    return [(n, s, not (n.endswith('moving_mean') or n.endswith('moving_var'))) for n, s in out]


def generic_forward(w, ctx, training, new_stats=None):
    """Generic FC graph: pFC = Dense(no bias) + BN + LeakyReLU (networktts.py:59-63) per layer, then
    concat[Dense(1+S), Dense(NM, sigmoid)] (networktts.py:196-197,225).  The layer count is read off the weights."""
    x, i = ctx, 0
    while 'g.fc%d.kernel' % i in w:
        x = lrelu(batchnorm(x @ w['g.fc%d.kernel' % i], w, 'g.fc%d.bn' % i, training, new_stats))
        i += 1
    f0spec = x @ w['g.f0spec.kernel'] + w['g.f0spec.bias']
    nm = torch.sigmoid(x @ w['g.nm.kernel'] + w['g.nm.bias'])
    return torch.cat([f0spec, nm], dim=-1)


def lse_step(a, wg, ctx, real):
    """OptimizerTTS LSE graph (optimizertts.py:56-57,386-402): loss = mean((y - model(x))^2), training-mode BatchNorm.
    The model is the DCNN generator, or the Generic FC stack when `a` is None."""
    new_stats = OrderedDict()
    if a is None:
        pred = generic_forward(wg, ctx, training=True, new_stats=new_stats)
    else:
        pred = generator_forward(a, wg, ctx, training=True, new_stats=new_stats)
    loss = ((real - pred) ** 2).mean()
    names = [k for k, v in wg.items() if v.requires_grad]
    grads = torch.autograd.grad(loss, [wg[k] for k in names], allow_unused=True)
    gd = OrderedDict((k, torch.zeros_like(wg[k]) if g is None else g.detach()) for k, g in zip(names, grads))
    return dict(loss=loss.detach(), pred=pred.detach(), grads=gd, new_stats=new_stats)


class LSETrainer(object):
    """OptimizerTTS.train_on_batch of the LSE scheme (optimizertts.py:378-402): Adam(lr=10^-3.39794, 0.9, 0.999, eps=1e-8),
    returned cost = sqrt(loss)."""

    def __init__(self, a, wg, dtype=torch.float32, lr=10 ** -3.39794, beta1=0.9, beta2=0.999, eps=1e-8):
        self.a = a
        self.wg = to_torch(wg, dtype, requires_grad=True)
        self.opt = Adam(lr, beta1, beta2, eps)
        self.dtype = dtype

    def train_on_batch(self, X, Y):
        r = lse_step(self.a, self.wg, torch.as_tensor(X, dtype=self.dtype), torch.as_tensor(Y, dtype=self.dtype))
        self.opt.update(self.wg, r['grads'])
        with torch.no_grad():
            for k, v in r['new_stats'].items():
                self.wg[k].copy_(v)
        return math.sqrt(float(r['loss'])), r


class Adam(object):
    """Keras 2.2 Adam (call sites optimizertts_wgan.py:145,172): decay=0, amsgrad=False."""

    def __init__(self, lr, beta1, beta2, eps=1e-7):
        self.lr, self.b1, self.b2, self.eps = lr, beta1, beta2, eps
        self.t = 0
        self.m, self.v = {}, {}

    def update(self, w, grads):
        self.t += 1
        lr_t = self.lr * math.sqrt(1.0 - self.b2 ** self.t) / (1.0 - self.b1 ** self.t)
        with torch.no_grad():
            for k, g in grads.items():
                if k not in self.m:
                    self.m[k] = torch.zeros_like(g)
                    self.v[k] = torch.zeros_like(g)
                self.m[k] = self.b1 * self.m[k] + (1.0 - self.b1) * g
                self.v[k] = self.b2 * self.v[k] + (1.0 - self.b2) * g * g
                w[k] -= lr_t * self.m[k] / (torch.sqrt(self.v[k]) + self.eps)


def critic_runs(generator_updates):
    """optimizertts_wgan.py:225-228"""
    return 10 if (generator_updates < 25 or generator_updates % 500 == 0) else 5


class OracleTrainer(object):
    """train_on_batch restated (optimizertts_wgan.py:216-241): one critic update per batch, one generator
    update when batchid % critic_runs == 0."""

    def __init__(self, a, wg, wc, errtype='WLSWGAN', dtype=torch.float32, gp_lambda=10.0, LScoef=0.25, fs=16000):
        self.a, self.errtype, self.gp_lambda, self.LScoef, self.fs = a, errtype, gp_lambda, LScoef, fs
        self.wg = to_torch(wg, dtype, requires_grad=True)
        self.wc = to_torch(wc, dtype, requires_grad=True)
        self.opt_c = Adam(1e-4, 0.5, 0.9)        # :92-94,145
        self.opt_g = Adam(1e-3, 0.5, 0.9)        # :95-97,172
        self.generator_updates = 0
        self.dtype = dtype

    def train_on_batch(self, batchid, X, Y, alpha):
        X = torch.as_tensor(X, dtype=self.dtype)
        Y = torch.as_tensor(Y, dtype=self.dtype)
        alpha = torch.as_tensor(alpha, dtype=self.dtype).reshape(-1, 1, 1)
        rc = critic_step(self.a, self.wg, self.wc, X, Y, alpha, self.gp_lambda)
        self.opt_c.update(self.wc, rc['grads'])
        cost = None
        rg = None
        if batchid % critic_runs(self.generator_updates) == 0:
            rg = generator_step(self.a, self.wg, self.wc, X, Y, self.errtype, fs=self.fs, LScoef=self.LScoef)
            self.opt_g.update(self.wg, rg['grads'])
            with torch.no_grad():
                for k, v in rg['new_stats'].items():
                    self.wg[k].copy_(v)
            cost = float(rg['loss'])
            self.generator_updates += 1
        return float(rc['loss']), cost, rc, rg


# ----------------------------------------------------------------------------------------------
# synthetic demo-shaped batches
# ----------------------------------------------------------------------------------------------

def synthetic_batch(a, B, T, seed=123):
    rng = np.random.RandomState(seed)
    X = np.empty((B, T, a.ctxsize), dtype=np.float32)
    nbin = max(a.ctxsize - 9, 0)
    X[:, :, :nbin] = np.where(rng.uniform(size=(B, T, nbin)) < 0.1, 1.0, -1.0)
    X[:, :, nbin:] = rng.uniform(-1, 1, size=(B, T, a.ctxsize - nbin))
    Y = np.empty((B, T, a.outsize), dtype=np.float32)
    Y[:, :, :1 + a.spec_size] = rng.normal(size=(B, T, 1 + a.spec_size))
    Y[:, :, 1 + a.spec_size:] = rng.uniform(0, 1, size=(B, T, a.nm_size))
    alpha = rng.uniform(0, 1, size=(B,)).astype(np.float32)
    return X, Y, alpha
