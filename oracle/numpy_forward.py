"""
CPU ORACLE -- TEST INFRASTRUCTURE ONLY.  PARITY UNPINNED (see percival_oracle.py header).

Independent plain-numpy forward of the generator and critic (no torch, no library convolutions), used
only to cross-check oracle/percival_oracle.py on small shapes.  Convolutions are written as explicit
tap loops over zero-padded arrays; the LSTM as an explicit time loop.

Follows percivaltts/modeltts_common.py:65-126, percivaltts/networks_critic.py:44-96,
percivaltts/networktts.py:59-63,85-96,116-126.
"""
import numpy as np

ALPHA, EPS = 0.3, 1e-3


def _lrelu(x):
    return np.where(x > 0, x, ALPHA * x)


def _sigmoid(x):
    return 1.0 / (1.0 + np.exp(-x))


def _conv1d(x, k, b=None):
    W = k.shape[0]
    left = (W - 1) // 2
    T = x.shape[1]
    xp = np.pad(x, ((0, 0), (left, W - 1 - left), (0, 0)))
    y = np.zeros(x.shape[:2] + (k.shape[2],), dtype=x.dtype)
    for j in range(W):
        y += np.einsum('btc,cn->btn', xp[:, j:j + T], k[j])
    return y if b is None else y + b


def _conv2d(x, k, b=None):
    KH, KW = k.shape[:2]
    lt, lf = (KH - 1) // 2, (KW - 1) // 2
    T, Fq = x.shape[1], x.shape[2]
    xp = np.pad(x, ((0, 0), (lt, KH - 1 - lt), (lf, KW - 1 - lf), (0, 0)))
    y = np.zeros(x.shape[:3] + (k.shape[3],), dtype=x.dtype)
    for i in range(KH):
        for j in range(KW):
            y += np.einsum('btfc,cn->btfn', xp[:, i:i + T, j:j + Fq], k[i, j])
    return y if b is None else y + b


def _bn(x, w, p, training):
    if training:
        ax = tuple(range(x.ndim - 1))
        m, v = x.mean(axis=ax), x.var(axis=ax)
    else:
        m, v = w[p + '.moving_mean'], w[p + '.moving_var']
    return (x - m) / np.sqrt(v + EPS) * w[p + '.gamma'] + w[p + '.beta']


def _lstm(x, W, U, b, reverse):
    B, T, _ = x.shape
    H = U.shape[0]
    h = np.zeros((B, H), dtype=x.dtype)
    c = np.zeros((B, H), dtype=x.dtype)
    out = np.zeros((B, T, H), dtype=x.dtype)
    for t in (range(T - 1, -1, -1) if reverse else range(T)):
        z = x[:, t] @ W + h @ U + b
        i, f = _sigmoid(z[:, :H]), _sigmoid(z[:, H:2 * H])
        g, o = np.tanh(z[:, 2 * H:3 * H]), _sigmoid(z[:, 3 * H:])
        c = f * c + i * g
        h = o * np.tanh(c)
        out[:, t] = h
    return out


def generator_forward(a, w, ctx, training):
    x = ctx
    for i in range(a.ctx_nbcnnlayers):
        x = _lrelu(_bn(_conv1d(x, w['g.ctx.cnn%d.kernel' % i]), w, 'g.ctx.cnn%d.bn' % i, training))
    for i in range(2):
        x = _lrelu(_bn(x @ w['g.ctx.fc%d.kernel' % i], w, 'g.ctx.fc%d.bn' % i, training))
    hf = _lstm(x, w['g.f0.blstm.fw.kernel'], w['g.f0.blstm.fw.recurrent'], w['g.f0.blstm.fw.bias'], False)
    hb = _lstm(x, w['g.f0.blstm.bw.kernel'], w['g.f0.blstm.bw.recurrent'], w['g.f0.blstm.bw.bias'], True)
    f0 = np.concatenate([hf, hb], -1) @ w['g.f0.out.kernel'] + w['g.f0.out.bias']
    s = (x @ w['g.spec.proj.kernel'] + w['g.spec.proj.bias'])[..., None]
    for i in range(a.gen_nbcnnlayers):
        s = _lrelu(_bn(_conv2d(s, w['g.spec.cnn%d.kernel' % i]), w, 'g.spec.cnn%d.bn' % i, training))
    s = _conv2d(s, w['g.spec.out.kernel'], w['g.spec.out.bias'])[..., 0]
    n = x
    for i in range(a.gen_nbcnnlayers // 2):
        n = _lrelu(_bn(n @ w['g.nm.fc%d.kernel' % i], w, 'g.nm.fc%d.bn' % i, training))
    n = _sigmoid(n @ w['g.nm.out.kernel'] + w['g.nm.out.bias'])
    return np.concatenate([f0, s, n], -1)


def critic_forward(a, w, feats, ctx):
    s = feats[:, :, 1:1 + a.spec_size]
    if a.gen_nbcnnlayers > 0:
        s = s[..., None]
        for i in range(a.gen_nbcnnlayers):
            s = _lrelu(_conv2d(s, w['c.spec.cnn%d.kernel' % i], w['c.spec.cnn%d.bias' % i]))
        s = s.reshape(s.shape[0], s.shape[1], -1)
    else:
        for i in range(3):
            s = _lrelu(s @ w['c.spec.fc%d.kernel' % i] + w['c.spec.fc%d.bias' % i])
    c = ctx
    for i in range(a.ctx_nbcnnlayers):
        c = _lrelu(_conv1d(c, w['c.ctx.cnn%d.kernel' % i], w['c.ctx.cnn%d.bias' % i]))
    for i in range(2):
        c = _lrelu(c @ w['c.ctx.fc%d.kernel' % i] + w['c.ctx.fc%d.bias' % i])
    p = np.concatenate([s, c], -1)
    for i in range(3):
        p = _lrelu(p @ w['c.post.fc%d.kernel' % i] + w['c.post.fc%d.bias' % i])
    return p @ w['c.out.kernel'] + w['c.out.bias']
