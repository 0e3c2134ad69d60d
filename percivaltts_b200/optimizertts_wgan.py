"""
WGAN-GP optimizer plugin: the B200 counterpart of reference percivaltts/optimizertts_wgan.py:82-328
(OptimizerTTSWGAN).  Same constructor, same six hooks, same defaults; prepare() builds the two step schedules
(critic step / generator step) as CUDA graphs on libpctts.so instead of two compiled Keras models.

    opti = OptimizerTTSWGAN(cfg, model, errtype='WLSWGAN', critic=Critic(vocoder, ctxsize, cfg))
    opti.train(indir, outdir, wdir, fid_lst_tra, fid_lst_val, 'model.h5')          # full loop (base class)
    opti.prepare(); cost = opti.train_on_batch(batchid, X_trab, Y_trab)            # the hot path alone

Under torchrun (WORLD_SIZE > 1) each rank passes its shard of the global batch to train_on_batch().
"""
import numpy as np

from . import arch as A
from . import backend_b200
from . import data
from . import optimizertts
from . import sigproc as sp
from .backend_b200 import nonlin_sigmoidparm
from .trainer import DeviceTrainer
from .wgan import WGANStepper, NLOSS, L_CRITIC, L_GEN


class OptimizerTTSWGAN(optimizertts.OptimizerTTS):
    supports_prefetch = True

    def __init__(self, cfgtomerge, model, errtype='WGAN', critic=None, **kwargs):
        optimizertts.OptimizerTTS.__init__(self, cfgtomerge, model, errtype, **kwargs)
        self.critic = critic
        self.costs_tra_critic_batches = []
        self.generator_updates = 0
        self.sync = None
        self.trainer = None
        self.use_graphs = True

    def default_options(self, cfg):
        """optimizertts_wgan.py:91-104"""
        cfg.train_wgan_critic_learningrate_log10 = -4
        cfg.train_wgan_critic_adam_beta1 = 0.5
        cfg.train_wgan_critic_adam_beta2 = 0.9
        cfg.train_wgan_gen_learningrate_log10 = -3
        cfg.train_wgan_gen_adam_beta1 = 0.5
        cfg.train_wgan_gen_adam_beta2 = 0.9
        cfg.train_wgan_pg_lambda = 10
        cfg.train_wgan_LScoef = 0.25
        cfg.train_wgan_validation_ltm_winlen = 20
        cfg.train_wgan_critic_LSWGANtransfreqcutoff = 4000
        cfg.train_wgan_critic_LSWGANtranscoef = 1.0 / 8.0
        cfg.train_wgan_critic_use_WGAN_incnoisefeature = False
        return cfg

    # ---- graph construction -----------------------------------------------------------------------------------------
    def wganls_weights(self):
        """The spectral weighting vectors of the WLSWGAN mix (optimizertts_wgan.py:186-210), built in float64 numpy:
        -> (w_wgan, w_ls, mean(w_wgan))."""
        cfg, voc = self.cfg, self._model.vocoder
        els = [[0.0]]                                                      # f0
        specvs = np.arange(voc.specsize(), dtype=np.float32)
        if cfg.train_wgan_LScoef == 0.0:
            els.append(np.ones(voc.specsize()))
        else:
            els.append(nonlin_sigmoidparm(specvs, sp.freq2fwspecidx(cfg.train_wgan_critic_LSWGANtransfreqcutoff, voc.fs, voc.specsize()),
                                          cfg.train_wgan_critic_LSWGANtranscoef))
        if voc.noisesize() > 0 and cfg.train_wgan_critic_use_WGAN_incnoisefeature:
            if cfg.train_wgan_LScoef == 0.0:
                els.append(np.ones(voc.noisesize()))
            else:
                noisevs = np.arange(voc.noisesize(), dtype=np.float32)
                els.append(nonlin_sigmoidparm(noisevs, sp.freq2fwspecidx(cfg.train_wgan_critic_LSWGANtransfreqcutoff, voc.fs, voc.noisesize()),
                                              cfg.train_wgan_critic_LSWGANtranscoef))
        else:
            els.append(np.zeros(max(voc.noisesize(), 0)))
        if voc.vuvsize() > 0:
            els.append([0.0])
        w = np.hstack(els) * (1.0 - cfg.train_wgan_LScoef)
        return w, 1.0 - w, float(np.mean(w))

    def prepare(self, sync=None, ops=None):
        """Build the critic-step and generator-step schedules (optimizertts_wgan.py:107-213)."""
        print('    Prepare {} training...'.format(self._errtype))
        if self.critic is None:
            raise ValueError('OptimizerTTSWGAN needs a critic')
        cfg = self.cfg
        ops = ops or backend_b200.get_ops()
        if sync is None:
            sync = backend_b200.get_sync()          # torchrun: rendezvous from the environment (WORLD_SIZE > 1), else None
        if sync is None:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
                from .parallel import DistSync
                sync = DistSync()
                sync.warmup(ops.device)
                sync.enable_peer(ops.device)
        self.sync = sync
        generator = self._model.kerasmodel.device_module(ops)
        critic = self.critic.kerasmodel.device_module(ops)
        print('    critic architecture:')
        self.critic.kerasmodel.summary()
        w_ls, wbar = None, 1.0
        if self._errtype == 'WLSWGAN':
            print('        use WLSWGAN optimization')
            _, w_ls, wbar = self.wganls_weights()
        elif self._errtype == 'WGAN':
            print('        use WGAN optimization')
        else:
            raise ValueError('OptimizerTTSWGAN handles errtype WGAN or WLSWGAN, not ' + str(self._errtype))
        stepper = WGANStepper(
            ops, generator.a, generator=generator, critic=critic, errtype=self._errtype, sync=sync,
            critic_lr=float(10 ** cfg.train_wgan_critic_learningrate_log10), critic_beta1=float(cfg.train_wgan_critic_adam_beta1),
            critic_beta2=float(cfg.train_wgan_critic_adam_beta2), gen_lr=float(10 ** cfg.train_wgan_gen_learningrate_log10),
            gen_beta1=float(cfg.train_wgan_gen_adam_beta1), gen_beta2=float(cfg.train_wgan_gen_adam_beta2),
            gp_lambda=float(cfg.train_wgan_pg_lambda), w_ls=w_ls, w_wgan_mean=wbar, seed=backend_b200.get_seed())
        self.stepper = stepper
        self.trainer = DeviceTrainer(stepper, use_graphs=self.use_graphs)
        self.trainer.generator_updates = self.generator_updates

    # ---- the hot path -----------------------------------------------------------------------------------------------
    def train_on_batch(self, batchid, X_trab, Y_trab, alpha=None, next_batch=None):
        """One critic update, plus one generator update when batchid % critic_runs == 0 (optimizertts_wgan.py:216-241).
        X_trab [B,T,ctxsize], Y_trab [B,T,out] host float32 arrays (not retained).  Returns the generator cost or None.
        `alpha` [B] fixes the interpolation weights of the gradient penalty (tests); default: device RNG.
        `next_batch` = (X, Y) of the following call when the caller already holds it (the base loop does): its host->device
        copy then overlaps this batch's kernels."""
        critic_cost, cost_tra = self.trainer.train_on_batch(batchid, X_trab, Y_trab, alpha=alpha, next_batch=next_batch)
        self.costs_tra_critic_batches.append(critic_cost)
        self.generator_updates = self.trainer.generator_updates
        return cost_tra

    def _validation_set_on_device(self, X_vals, Y_vals):
        """The validation utterances are loaded once per training run (optimizertts.py:153-156): they are uploaded once too and
        stay resident ([1,T,ctx] / [1,T,out] per utterance, cropped to the common length like data.cost_model_mfn's inputs)."""
        import torch
        key = (id(X_vals), id(Y_vals), len(X_vals))
        cache = getattr(self, '_val_cache', None)
        if cache is None or cache[0] != key:
            dev = self.stepper.ops.device
            pairs = []
            for X, Y in zip(X_vals, Y_vals):
                n = min(X.shape[0], Y.shape[0])
                pairs.append((torch.from_numpy(np.ascontiguousarray(X[np.newaxis, :n], dtype=np.float32)).to(dev),
                              torch.from_numpy(np.ascontiguousarray(Y[np.newaxis, :n], dtype=np.float32)).to(dev)))
            cache = self._val_cache = (key, pairs, self.stepper.ops.zeros(max(len(pairs), 1), NLOSS), (X_vals, Y_vals))
        return cache[1], cache[2]

    def update_validation_cost(self, costs, X_vals, Y_vals):
        """Per-utterance validation (batch size 1, inference-mode BatchNorm): prediction RMSE, generator objective, critic
        objective and its long-term mean, which is the model-selection criterion (optimizertts_wgan.py:244-268).
        Device-side pass: the utterances are resident, each one is ONE inference-mode evaluation (the prediction feeds the RMSE,
        the generator objective and the critic objective -- the reference runs predict + two evaluate calls per utterance), the
        per-utterance loss vectors are collected on the device and read back with a single copy per epoch."""
        from .wgan import L_SSE
        pairs, rows = self._validation_set_on_device(X_vals, Y_vals)
        for k, (Xd, Yd) in enumerate(pairs):
            if Xd.shape[1] == 0:
                rows[k].zero_()
                continue
            self.trainer.evaluate_device(Xd, Yd, rows[k])
        losses = rows[:len(pairs)].cpu().numpy().astype(np.float64)          # the one device->host copy (and sync) of the pass
        nelem = float(sum(Yd.numel() for _, Yd in pairs))
        costs['model_rmse_validation'].append(float(np.sqrt(losses[:, L_SSE].sum() / max(nelem, 1.0))))
        gen_costs, critic_costs = losses[:, L_GEN], losses[:, L_CRITIC]
        costs['model_validation'].append(float(np.mean(gen_costs)))
        mean_critic_tra = float(np.mean(self.costs_tra_critic_batches)) if self.costs_tra_critic_batches else float('nan')
        costs['critic_training'].append(mean_critic_tra)
        costs['critic_validation'].append(float(np.mean(critic_costs)))
        costs['critic_validation_ltm'].append(float(np.mean(costs['critic_validation'][-self.cfg.train_wgan_validation_ltm_winlen:])))
        if mean_critic_tra <= 0.0:
            print('Average critic loss is negative: Training is likely to take ages to converge or not converge at all. ')
        self.costs_tra_critic_batches = []
        return costs['critic_validation_ltm'][-1]

    # ---- training state ------------------------------------------------------------------------------------------------
    def saveOptimizer(self, store, fname):
        """Adam slots and step counter of one network (reference :270-290 wrote them with h5py)."""
        np.savez(fname, m=store.m.cpu().numpy(), v=store.v.cpu().numpy(), step=store.step.cpu().numpy())

    def loadOptimizer(self, store, fname):
        import torch
        with np.load(fname if fname.endswith('.npz') else fname + '.npz') as z:
            if z['m'].shape[0] != store.m.shape[0]:
                print('Restoring optimizer failed from ' + fname + '. Fresh optimizer used instead (i.e. momentums, etc., might be wrong)')
                return
            store.m.copy_(torch.from_numpy(z['m']))
            store.v.copy_(torch.from_numpy(z['v']))
            store.step.copy_(torch.from_numpy(z['step']))

    def saveTrainingStateLossSpecific(self, fstate):
        """Same file stems as the reference (:308-316) plus the critic weights, which the reference omits
        ("TODO That's not enough", :310)."""
        if self.sync is not None and self.sync.rank != 0:
            return
        self.saveOptimizer(self.stepper.G.params, fstate + '.generator.optimizer.npz')
        self.saveOptimizer(self.stepper.C.params, fstate + '.critic.optimizer.npz')
        self._model.kerasmodel.save_weights(fstate + '.model.weights.npz')
        self.critic.kerasmodel.save_weights(fstate + '.critic.weights.npz')
        np.savez(fstate + '.wgan.npz', generator_updates=self.generator_updates,
                 rng_counter=self.stepper.counter.cpu().numpy())

    def loadTrainingStateLossSpecific(self, fstate):
        import torch
        if self.trainer is None:
            self.prepare()
        self.loadOptimizer(self.stepper.G.params, fstate + '.generator.optimizer.npz')
        self.loadOptimizer(self.stepper.C.params, fstate + '.critic.optimizer.npz')
        self._model.kerasmodel.load_weights(fstate + '.model.weights.npz')
        self.critic.kerasmodel.load_weights(fstate + '.critic.weights.npz')
        with np.load(fstate + '.wgan.npz') as z:
            self.generator_updates = int(z['generator_updates'])
            self.stepper.counter.copy_(torch.from_numpy(z['rng_counter']))
        self.trainer.generator_updates = self.generator_updates
