"""
Device-side least-squares training step: the compiled Keras model of OptimizerTTS.prepare()
(reference percivaltts/optimizertts.py:386-402: `kerasmodel.compile(loss=lse_loss, optimizer=Adam(...))`,
`lse_loss = mean((y_true - y_pred)^2)`, :56-57) restated as an explicit kernel schedule on libpctts.so.

The model is any generator-type engine module (engine.GenericFC for modeltts_common.Generic, engine.Generator for
DCNNF0SpecNoiseFeatures); BatchNorm runs in training mode and moves its moving statistics, like Keras' train_on_batch.
Data parallel: local sums scaled by 1/global count, sync-BN, flat gradient all-reduced (SUM) before Adam.
"""
from collections import OrderedDict

from .engine import LocalSync

L_LSE = 0
NLOSS = 4


class LSEStepper(object):
    def __init__(self, ops, module, lr=10 ** -3.39794, beta1=0.9, beta2=0.999, eps=1e-8, sync=None):
        self.ops, self.M = ops, module
        self.a = module.a
        self.sync = sync or LocalSync()
        self.opt = (float(lr), float(beta1), float(beta2))
        self.eps = float(eps)
        self.losses = ops.zeros(NLOSS)
        self._bufs = {}
        self._lru = OrderedDict()      # batch shapes in order of last use (T varies from batch to batch)
        self.max_shapes = 16

    def buffers(self, B, T):
        b = self._bufs.get((B, T))
        if b is None:
            o = self.ops
            b = self._bufs[(B, T)] = dict(pred=o.empty(B, T, self.a.outsize), dpred=o.empty(B, T, self.a.outsize))
        return b

    def release(self, B, T):
        self._bufs.pop((B, T), None)
        self.M.release(B, T, with_backward=True)
        self.ops.release_shape((B, T))

    def step(self, X, Y, apply=True):
        """One update on device tensors X [B,T,ctx], Y [B,T,O]; losses[L_LSE] = mean((Y - model(X))^2) over the global batch."""
        with self.ops.shape_scope((int(X.shape[0]), int(X.shape[1]))), self.ops.step_scope():
            o, sync = self.ops, self.sync
            B, T, _ = X.shape
            O = self.a.outsize
            self._lru[(B, T)] = True
            self._lru.move_to_end((B, T))
            while len(self._lru) > self.max_shapes:
                old, _ = self._lru.popitem(last=False)
                self.release(*old)
            bf = self.buffers(B, T)
            pl = self.M.plan(B, T)
            self.M.forward(pl, X, bf['pred'], training=True, update_stats=True, sync=sync)
            n = float(B * T * O * sync.world)
            o.lse_loss_grad(bf['pred'], Y, None, bf['dpred'], 1.0 / n, 1.0 / n, self.losses, L_LSE)
            self.M.backward(pl, bf['pred'], bf['dpred'], sync=sync)
            if sync.world > 1:
                sync.allreduce(self.losses[L_LSE:L_LSE + 1])
            if apply:
                sync.allreduce(self.M.params.grad)
                self.M.params.adam(*self.opt, eps=self.eps)
