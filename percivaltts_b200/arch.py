"""
Architecture description and parameter inventories of the two networks on the WGAN-GP path.

  generator  DCNNF0SpecNoiseFeatures   reference percivaltts/modeltts_common.py:61-126
  critic     Critic                    reference percivaltts/networks_critic.py:44-96
  layer builders pFC / pCNN1D / pCNN2D / pBLSTM   reference percivaltts/networktts.py:59-63,72-96,116-126

Parameters keep the Keras layouts (Dense [in,out]; Conv1D [W,in,out]; Conv2D [KH(time),KW(freq),in,out]; LSTM kernel
[in,4H], recurrent [H,4H], bias [4H] with gate order i,f,c,o) so that weights exported from the reference map 1:1.
"""
import math
from collections import OrderedDict

import numpy as np

LRELU_ALPHA = 0.3      # networktts.py:62,119,125 ; networks_critic.py:68
BN_EPS = 1e-3          # keras BatchNormalization default
BN_MOMENTUM = 0.99     # keras BatchNormalization default


class Arch(object):
    """The cfgarch fields read by the graph builders (modeltts_common.py:76-121, networks_critic.py:62-94)."""

    FIELDS = ('ctxsize', 'spec_size', 'nm_size', 'hiddenwidth', 'ctx_nbcnnlayers', 'ctx_winlen', 'gen_nbcnnlayers',
              'gen_nbfilters', 'gen_winlen', 'spec_freqlen')

    def __init__(self, ctxsize=425, spec_size=129, nm_size=33, hiddenwidth=256, ctx_nbcnnlayers=1, ctx_winlen=21,
                 gen_nbcnnlayers=8, gen_nbfilters=4, gen_winlen=5, spec_freqlen=5):
        self.ctxsize = int(ctxsize)
        self.spec_size = int(spec_size)
        self.nm_size = int(nm_size)
        self.hiddenwidth = int(hiddenwidth)
        self.ctx_nbcnnlayers = int(ctx_nbcnnlayers)
        self.ctx_winlen = int(ctx_winlen)
        self.gen_nbcnnlayers = int(gen_nbcnnlayers)
        self.gen_nbfilters = int(gen_nbfilters)
        self.gen_winlen = int(gen_winlen)
        self.spec_freqlen = int(spec_freqlen)

    @property
    def outsize(self):
        return 1 + self.spec_size + self.nm_size     # vocoders.py:97-99 (f0 + spec + noise, no vuv for PML)

    @classmethod
    def from_cfg(cls, ctxsize, vocoder, cfgarch):
        """Read the `arch_*` attributes the reference's constructors read from their cfgarch argument."""
        return cls(ctxsize=ctxsize, spec_size=vocoder.specsize(), nm_size=vocoder.noisesize(),
                   hiddenwidth=cfgarch.arch_hiddenwidth, ctx_nbcnnlayers=cfgarch.arch_ctx_nbcnnlayers,
                   ctx_winlen=cfgarch.arch_ctx_winlen, gen_nbcnnlayers=cfgarch.arch_gen_nbcnnlayers,
                   gen_nbfilters=cfgarch.arch_gen_nbfilters, gen_winlen=cfgarch.arch_gen_winlen,
                   spec_freqlen=cfgarch.arch_spec_freqlen)

    def as_dict(self):
        return dict((k, getattr(self, k)) for k in self.FIELDS)


class GenericArch(object):
    """What Generic(ctxsize, vocoder, layertypes=[...], cfgarch) reads (modeltts_common.py:36-55): only 'FC' layers are on the
    B200 path (the recurrent / gated / dropout builders of networktts.py:130-175 are not)."""

    FIELDS = ('ctxsize', 'spec_size', 'nm_size', 'hiddenwidth', 'layertypes')

    def __init__(self, ctxsize, spec_size, nm_size, hiddenwidth, layertypes=('FC', 'FC', 'FC')):
        self.ctxsize, self.spec_size, self.nm_size = int(ctxsize), int(spec_size), int(nm_size)
        self.hiddenwidth = int(hiddenwidth)
        self.layertypes = list(layertypes)
        for lt in self.layertypes:
            if lt != 'FC':
                raise ValueError('Unknown layer type ' + str(lt))     # networktts.py:171,177

    @property
    def outsize(self):
        return 1 + self.spec_size + self.nm_size

    def as_dict(self):
        return dict((k, getattr(self, k)) for k in self.FIELDS)


def _bn(prefix, ch):
    return [(prefix + '.gamma', (ch,), True), (prefix + '.beta', (ch,), True),
            (prefix + '.moving_mean', (ch,), False), (prefix + '.moving_var', (ch,), False)]


def generator_param_shapes(a):
    """Ordered [(name, shape, trainable)] of DCNNF0SpecNoiseFeatures (modeltts_common.py:65-126)."""
    H, S, NMS, Fl = a.hiddenwidth, a.spec_size, a.nm_size, a.gen_nbfilters
    out = []
    cin = a.ctxsize
    for i in range(a.ctx_nbcnnlayers):                                  # :76-77 pCNN1D: Conv1D(no bias)+BN+LeakyReLU
        out.append(('g.ctx.cnn%d.kernel' % i, (a.ctx_winlen, cin, H), True))
        out += _bn('g.ctx.cnn%d.bn' % i, H)
        cin = H
    for i in range(2):                                                  # :78-79 pFC x2
        out.append(('g.ctx.fc%d.kernel' % i, (cin, H), True))
        out += _bn('g.ctx.fc%d.bn' % i, H)
        cin = H
    for d in ('fw', 'bw'):                                              # :82-83 pBLSTM
        out.append(('g.f0.blstm.%s.kernel' % d, (cin, 4 * H), True))
        out.append(('g.f0.blstm.%s.recurrent' % d, (H, 4 * H), True))
        out.append(('g.f0.blstm.%s.bias' % d, (4 * H,), True))
    out.append(('g.f0.out.kernel', (2 * H, 1), True))                   # :84
    out.append(('g.f0.out.bias', (1,), True))
    out.append(('g.spec.proj.kernel', (cin, S), True))                  # :95
    out.append(('g.spec.proj.bias', (S,), True))
    c2 = 1
    for i in range(a.gen_nbcnnlayers):                                  # :97-98 pCNN2D: Conv2D(no bias)+BN+LeakyReLU
        out.append(('g.spec.cnn%d.kernel' % i, (a.gen_winlen, a.spec_freqlen, c2, Fl), True))
        out += _bn('g.spec.cnn%d.bn' % i, Fl)
        c2 = Fl
    out.append(('g.spec.out.kernel', (a.gen_winlen, a.spec_freqlen, c2, 1), True))   # :100
    out.append(('g.spec.out.bias', (1,), True))
    cn = cin
    for i in range(a.gen_nbcnnlayers // 2):                             # :119 (python-2 integer division)
        out.append(('g.nm.fc%d.kernel' % i, (cn, H), True))
        out += _bn('g.nm.fc%d.bn' % i, H)
        cn = H
    out.append(('g.nm.out.kernel', (cn, NMS), True))                    # :121
    out.append(('g.nm.out.bias', (NMS,), True))
    return out


def critic_param_shapes(a):
    """Ordered [(name, shape, trainable)] of Critic (networks_critic.py:44-96); bn=False everywhere (:50)."""
    H, S, Fl = a.hiddenwidth, a.spec_size, a.gen_nbfilters
    out = []
    if a.gen_nbcnnlayers > 0:
        c2 = 1
        for i in range(a.gen_nbcnnlayers):                              # :66-68
            out.append(('c.spec.cnn%d.kernel' % i, (a.gen_winlen, a.spec_freqlen, c2, Fl), True))
            out.append(('c.spec.cnn%d.bias' % i, (Fl,), True))
            c2 = Fl
        specfeat = S * c2                                               # :70
    else:
        cin = S
        for i in range(3):                                              # :73-75
            out.append(('c.spec.fc%d.kernel' % i, (cin, H), True))
            out.append(('c.spec.fc%d.bias' % i, (H,), True))
            cin = H
        specfeat = H
    cin = a.ctxsize
    for i in range(a.ctx_nbcnnlayers):                                  # :82-83
        out.append(('c.ctx.cnn%d.kernel' % i, (a.ctx_winlen, cin, H), True))
        out.append(('c.ctx.cnn%d.bias' % i, (H,), True))
        cin = H
    for i in range(2):                                                  # :84-85
        out.append(('c.ctx.fc%d.kernel' % i, (cin, H), True))
        out.append(('c.ctx.fc%d.bias' % i, (H,), True))
        cin = H
    cin = specfeat + cin                                                # :89 concat [spec, ctx]
    for i in range(3):                                                  # :91-93
        out.append(('c.post.fc%d.kernel' % i, (cin, H), True))
        out.append(('c.post.fc%d.bias' % i, (H,), True))
        cin = H
    out.append(('c.out.kernel', (cin, 1), True))                        # :96
    out.append(('c.out.bias', (1,), True))
    return out


def generic_param_shapes(ctxsize, spec_size, nm_size, hiddenwidth, layertypes):
    """Generic(['FC',...]) + network_final for a vocoder without MLPG (modeltts_common.py:36-55,
    networktts.py:59-63,192-197).  Only 'FC' layers are on the supported path."""
    out = []
    cin = ctxsize
    for i, lt in enumerate(layertypes):
        if lt != 'FC':
            raise ValueError('Unknown layer type ' + str(lt))     # networktts.py:171,177
        out.append(('g.fc%d.kernel' % i, (cin, hiddenwidth), True))
        out += _bn('g.fc%d.bn' % i, hiddenwidth)
        cin = hiddenwidth
    out.append(('g.f0spec.kernel', (cin, 1 + spec_size), True))
    out.append(('g.f0spec.bias', (1 + spec_size,), True))
    out.append(('g.nm.kernel', (cin, nm_size), True))
    out.append(('g.nm.bias', (nm_size,), True))
    return out


def count_params(shapes):
    """keras count_params(): trainable + non-trainable (pinned 2195 by tests/test_smoke_tensorflowkeras.py:51-53)."""
    return int(sum(int(np.prod(s)) for _, s, _ in shapes))


# ---- keras default initialisers (glorot_uniform kernels, orthogonal recurrent, zero biases, unit_forget_bias) --------

def _glorot(rng, shape):
    if len(shape) == 2:
        fan_in, fan_out = shape
    else:
        rf = int(np.prod(shape[:-2]))
        fan_in, fan_out = rf * shape[-2], rf * shape[-1]
    lim = math.sqrt(6.0 / (fan_in + fan_out))
    return rng.uniform(-lim, lim, size=shape)


def _orthogonal(rng, shape):
    rows, cols = shape
    a = rng.normal(0.0, 1.0, size=(max(rows, cols), min(rows, cols)))
    q, r = np.linalg.qr(a)
    q = q * np.sign(np.diag(r))
    return q if rows >= cols else q.T


def init_params(shapes, seed=123):
    """-> OrderedDict name -> float32 array with the keras default initial values."""
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
            v[H:2 * H] = 1.0
        else:
            v = np.zeros(shape)
        w[name] = np.ascontiguousarray(v, dtype=np.float32)
    return w


def mel(f):
    return 1127.0 * np.log1p(np.asarray(f, dtype=np.float64) / 700.0)


def wlswgan_weights(a, fs, fwspecidx, LScoef=0.25, transfreqcutoff=4000, transcoef=1.0 / 8.0, incnoise=False):
    """Spectral weighting of the mixed LSE / WGAN loss (optimizertts_wgan.py:186-210, backend_tensorflow.py:107-109).
    `fwspecidx(freq, fs, nbbnds)` is sigproc's freq2fwspecidx.  -> (w_wgan, w_ls, mean(w_wgan)), float64."""
    els = [np.zeros(1)]
    specvs = np.arange(a.spec_size, dtype=np.float32).astype(np.float64)
    if LScoef == 0.0:
        els.append(np.ones(a.spec_size))
    else:
        c = fwspecidx(transfreqcutoff, fs, a.spec_size)
        els.append(1.0 / (1.0 + np.exp(-(specvs - c) * transcoef)))
    if a.nm_size > 0 and incnoise:
        if LScoef == 0.0:
            els.append(np.ones(a.nm_size))
        else:
            nv = np.arange(a.nm_size, dtype=np.float32).astype(np.float64)
            c = fwspecidx(transfreqcutoff, fs, a.nm_size)
            els.append(1.0 / (1.0 + np.exp(-(nv - c) * transcoef)))
    else:
        els.append(np.zeros(a.nm_size))
    wg = np.hstack(els) * (1.0 - LScoef)
    return wg, 1.0 - wg, float(np.mean(wg))
