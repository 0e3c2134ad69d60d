"""
Base optimizer: configuration defaults and the epoch / batch loop that calls the hot path.
Counterpart of reference percivaltts/optimizertts.py:60-372 (OptimizerTTS).  Template-method hooks, same names and
contracts as the reference (:375-436): default_options(cfg), prepare(), train_on_batch(batchid, X, Y),
update_validation_cost(costs, X_vals, Y_vals), saveTrainingStateLossSpecific(fstate),
loadTrainingStateLossSpecific(fstate).

The loop stays host python (it is the caller of the hot path, not part of it); plots are out of scope.
The base class itself is the reference's plain least-squares ('LSE') optimizer (:378-436): its step runs on the GPU through
lse.LSEStepper (forward, loss + gradient, backward, Adam: the same kernels as the WGAN-GP path).
"""
import collections
import concurrent.futures
import copy
import glob
import os
import pickle
import sys
import time
from collections import defaultdict

import numpy as np

from . import data
from .backend_b200 import gpu_memused
from .percivaltts import configuration, numpy_force_random_seed, print_log, print_tty, time2str

numpy_force_random_seed()      # reference optimizertts.py:37

_DP_CHECK = os.environ.get('PCTTS_DP_CHECK', '0') not in ('', '0')   # compare the local (B,T) across the replicas before EVERY batch


def proc_memresident():
    """Resident memory of this process in MiB (reference percivaltts.py, used in the epoch log line)."""
    try:
        with open('/proc/self/statm') as f:
            return int(f.read().split()[1]) * os.sysconf('SC_PAGE_SIZE') // (1024 * 1024)
    except (OSError, ValueError, IndexError):
        return -1


class OptimizerTTS(object):
    _model = None
    _errtype = 'LSE'

    def __init__(self, cfgtomerge, model, errtype='LSE', **kwargs):
        self._model = model
        self._errtype = errtype
        cfg = configuration()
        # defaults common to every optimisation scheme (optimizertts.py:71-92)
        cfg.train_min_nbepochs = 200
        cfg.train_max_nbepochs = 300
        cfg.train_nbepochs_scalewdata = True
        cfg.train_cancel_nodecepochs = 50
        cfg.train_cancel_validthresh = 10.0
        cfg.train_batch_size = 5
        cfg.train_batch_padtype = 'randshift'
        cfg.train_batch_cropmode = 'begendbigger'
        cfg.train_batch_length = None
        cfg.train_batch_lengthmax = None
        cfg.train_nbtrials = 1
        cfg.train_hypers = []
        cfg = self.default_options(cfg)
        cfg.train_log_plot = False          # plots (matplotlib) are out of scope on this path
        if cfgtomerge is not None:
            cfg.merge(cfgtomerge)
        for k, v in kwargs.items():
            setattr(cfg, k, v)
        self.cfg = cfg
        print('Training configuration')
        self.cfg.print_content()

    # ---- hooks; the defaults are the LSE (MSE) scheme of the reference (optimizertts.py:375-436) ---------------------------
    supports_prefetch = False      # train_on_batch() takes next_batch=(X, Y) (B200 addition; the reference loads synchronously)

    def default_options(self, cfg):
        """optimizertts.py:378-384"""
        cfg.train_lse_learningrate_log10 = -3.39794      # 10**-3.39794 = 0.0004
        cfg.train_lse_adam_beta1 = 0.9
        cfg.train_lse_adam_beta2 = 0.999
        cfg.train_lse_adam_epsilon_log10 = -8
        return cfg

    def prepare(self, sync=None, ops=None):
        """optimizertts.py:386-397: Adam on lse_loss = mean((y - model(x))^2)."""
        print('    Prepare LSE training')
        from . import backend_b200
        from .lse import LSEStepper
        ops = ops or backend_b200.get_ops()
        module = self._model.kerasmodel.device_module(ops)
        cfg = self.cfg
        self.lse = LSEStepper(ops, module, lr=float(10 ** cfg.train_lse_learningrate_log10), beta1=float(cfg.train_lse_adam_beta1),
                              beta2=float(cfg.train_lse_adam_beta2), eps=float(10 ** cfg.train_lse_adam_epsilon_log10),
                              sync=sync if sync is not None else backend_b200.get_sync())
        print('    optimizer: Adam')

    def train_on_batch(self, batchid, X_trab, Y_trab):
        """optimizertts.py:399-404: one update, returns sqrt(loss) (host numpy in, python float out)."""
        import torch
        dev = self.lse.ops.device
        Xd = torch.from_numpy(np.ascontiguousarray(X_trab, dtype=np.float32)).to(dev)
        Yd = torch.from_numpy(np.ascontiguousarray(Y_trab, dtype=np.float32)).to(dev)
        self.lse.step(Xd, Yd)
        return float(np.sqrt(float(self.lse.losses[0])))

    def update_validation_cost(self, costs, X_vals, Y_vals):
        """optimizertts.py:406-412"""
        costs['model_rmse_validation'].append(data.cost_model_prediction_rmse(self._model, [X_vals], Y_vals))
        return costs['model_rmse_validation'][-1]

    def saveTrainingStateLossSpecific(self, fstate):
        """optimizertts.py:414-424 saved the compiled Keras model with its optimizer; here weights + Adam slots (rank 0)."""
        if getattr(self.lse.sync, 'rank', 0) != 0:
            return
        store = self.lse.M.params
        self._model.kerasmodel.save_weights(fstate + '.model.weights.npz')
        np.savez(fstate + '.model.optimizer.npz', m=store.m.cpu().numpy(), v=store.v.cpu().numpy(), step=store.step.cpu().numpy())

    def loadTrainingStateLossSpecific(self, fstate):
        import torch
        if getattr(self, 'lse', None) is None:
            self.prepare()
        self._model.kerasmodel.load_weights(fstate + '.model.weights.npz')
        store = self.lse.M.params
        with np.load(fstate + '.model.optimizer.npz') as z:
            if z['m'].shape[0] == store.m.shape[0]:
                store.m.copy_(torch.from_numpy(z['m']))
                store.v.copy_(torch.from_numpy(z['v']))
                store.step.copy_(torch.from_numpy(z['step']))

    # ---- training state (optimizertts.py:103-145) -------------------------------------------------------------------
    def saveTrainingState(self, fstate, extras=None, printfn=print):
        if extras is None:
            extras = dict()
        printfn('    saving training state in {} ...'.format(fstate), end='')
        self.saveTrainingStateLossSpecific(fstate)
        with open(fstate + '.cfgextras.pkl', 'wb') as f:
            pickle.dump([self.cfg, extras, np.random.get_state()], f)
        print(' done')
        sys.stdout.flush()

    def loadTrainingState(self, fstate):
        print('    reloading training state from {} ...'.format(fstate), end='')
        self.loadTrainingStateLossSpecific(fstate)
        with open(fstate + '.cfgextras.pkl', 'rb') as f:
            DATA = pickle.load(f)
        print(' done')
        return DATA

    # ---- training loop ------------------------------------------------------------------------------------------------
    def train_oneparamset(self, indir, outdir, wdir, fid_lst_tra, fid_lst_val, params_savefile, trialstr='', cont=None):
        """One full training run (optimizertts.py:149-303): load validation set, loop epochs x batches calling
        train_on_batch(), validate, checkpoint, early-stop."""
        cfg = self.cfg
        print('Loading all validation data at once ...')
        X_vals = data.load(indir, fid_lst_val, verbose=1, label='Context labels: ')
        Y_vals = data.load(outdir, fid_lst_val, verbose=1, label='Output features: ')
        X_vals, Y_vals = data.croplen([X_vals, Y_vals])
        print('    {} validation files'.format(len(fid_lst_val)))
        worst_val = data.cost_0pred_rmse(Y_vals)
        print('    0-pred validation RMSE = {} (100%)'.format(worst_val))
        init_val = data.cost_model_prediction_rmse(self._model, [X_vals], Y_vals)
        print('    initial validation RMSE = {} ({:.4f}%)'.format(init_val, 100.0 * init_val / worst_val))
        best_val = None

        nbbatches = int(len(fid_lst_tra) / cfg.train_batch_size)
        print('    using {} batches of {} sentences each'.format(nbbatches, cfg.train_batch_size))
        print('    model #parameters={}'.format(self._model.count_params()))
        nbtrainframes = sum(data.loadfile(outdir, fid).shape[0] for fid in fid_lst_tra)
        print('    Training set: {} sentences, #frames={}'.format(len(fid_lst_tra), nbtrainframes))
        if cfg.train_nbepochs_scalewdata and cfg.train_batch_lengthmax is not None:
            # an "epoch" only sees lengthmax frames per sentence: rescale the epoch counts accordingly (:178-187)
            epochcoef = nbtrainframes / float(cfg.train_batch_lengthmax * len(fid_lst_tra))
            cfg.train_min_nbepochs = int(cfg.train_min_nbepochs * epochcoef)
            cfg.train_max_nbepochs = int(cfg.train_max_nbepochs * epochcoef)
            print('    scaled number of epochs: min={} max={}'.format(cfg.train_min_nbepochs, cfg.train_max_nbepochs))

        self.prepare()
        # data parallel (torchrun, one process per GPU): every rank draws the SAME global batch (same numpy RNG stream) and keeps
        # its utterances [r*B/N, (r+1)*B/N); files are written by rank 0 only; validation decisions are rank 0's
        sync = getattr(self, 'sync', None) or getattr(getattr(self, 'lse', None), 'sync', None)
        world, rank = getattr(sync, 'world', 1), getattr(sync, 'rank', 0)
        if cfg.train_batch_size % world:
            raise ValueError('train_batch_size={} does not split over {} ranks'.format(cfg.train_batch_size, world))

        def shard(A):
            n = A.shape[0] // world
            return A[rank * n:(rank + 1) * n]

        costs = defaultdict(list)
        epochs_modelssaved, epochs_durs = [], []
        nbnodecepochs = 0
        epochstart = 1
        fbase = os.path.splitext(params_savefile)[0]
        fstate = fbase + '-trainingstate-last.h5'
        if cont and len(glob.glob(fstate + '*')) > 0:
            print('    reloading previous training state ...')
            savedcfg, extras, rngstate = self.loadTrainingState(fstate)
            np.random.set_state(rngstate)
            costs = extras['costs']
            epochs_modelssaved = extras['epochs_modelssaved']
            epochs_durs = extras['epochs_durs']
            epochstart = extras['epoch'] + 1
            if (savedcfg.train_min_nbepochs == cfg.train_min_nbepochs and savedcfg.train_max_nbepochs == cfg.train_max_nbepochs
                    and savedcfg.train_cancel_nodecepochs == cfg.train_cancel_nodecepochs):
                best_val = extras['best_val']
                nbnodecepochs = extras['nbnodecepochs']

        print_log('    start training ...')
        epoch = -1
        self.epoch_timings = []
        loader = concurrent.futures.ThreadPoolExecutor(max_workers=1, thread_name_prefix='pctts-loader')
        for epoch in range(epochstart, 1 + cfg.train_max_nbepochs):
            timeepochstart = time.time()
            if world > 1:
                # the replicas must draw the same global batches (shuffle, crop offsets) or their (B,T) and their sync-BN counts
                # diverge: rank 0's numpy RNG state is authoritative at every epoch start (which also covers a resume)
                np.random.set_state(sync.broadcast_rng_state(np.random.get_state()))
            rndidx = np.arange(int(nbbatches * cfg.train_batch_size))
            np.random.shuffle(rndidx)
            rndidxb = np.split(rndidx, nbbatches)
            costs_tra_batches, load_times, train_times = [], [], []
            frames_epoch = 0
            def load_batch(bid):
                fid_lst_trab = [fid_lst_tra[bidx] for bidx in rndidxb[bid]]
                kw = dict(length=cfg.train_batch_length, lengthmax=cfg.train_batch_lengthmax, maskpadtype=cfg.train_batch_padtype,
                          cropmode=cfg.train_batch_cropmode)
                # same arrays and the same RNG draws as data.load_inoutset(...) (tests/test_loader_cpu.py), but only the window's
                # frames of this rank's utterances are copied out of the files: one pass instead of four, 1/world of the reads
                return data.load_inoutset_shard(indir, outdir, wdir, fid_lst_trab, rank, world, **kw)
            # The reference loads batch by batch on the main thread (optimizertts.py:231-232) and the GPU waits for the disk.  Here a
            # single loader thread runs two batches ahead: batch i+2 is read and cropped while batch i trains and the copy of batch
            # i+1 to the device overlaps its kernels.  One thread, batches in order, and nobody else draws from numpy's global RNG
            # during the epoch: the sequence of window draws is the reference's.
            pending = collections.deque(loader.submit(load_batch, b) for b in range(min(2, nbbatches)))
            t0 = time.time()
            nxt = pending.popleft().result() if nbbatches > 0 else None
            for batchid in range(nbbatches):
                print_tty('\r    Training batch {}/{}'.format(1 + batchid, nbbatches))
                X_trab, Y_trab, W_trab = nxt
                frames_epoch += int(X_trab.shape[0]) * int(X_trab.shape[1])
                if world > 1 and (batchid == 0 or _DP_CHECK):
                    sync.assert_same_shape(X_trab.shape[:2], 'batch %d of epoch %d' % (batchid, epoch))
                if batchid + 2 < nbbatches:
                    pending.append(loader.submit(load_batch, batchid + 2))
                nxt = pending.popleft().result() if batchid + 1 < nbbatches else None
                load_times.append(time.time() - t0)          # time the step loop WAITED for data, not the loader's own time
                t1 = time.time()
                if self.supports_prefetch and nxt is not None:
                    cost_tra = self.train_on_batch(batchid, X_trab, Y_trab, next_batch=(nxt[0], nxt[1]))
                else:
                    cost_tra = self.train_on_batch(batchid, X_trab, Y_trab)
                t0 = time.time()
                train_times.append(t0 - t1)
                if cost_tra is not None:
                    if np.isnan(cost_tra):
                        print_log('    E{} Batch {}/{} train cost = {}'.format(epoch, 1 + batchid, nbbatches, cost_tra))
                        raise ValueError('ERROR: Training cost is nan!')
                    costs_tra_batches.append(cost_tra)
            print_tty('\r' + ' ' * 60 + '\r')
            costs['model_training'].append(np.mean(costs_tra_batches) if costs_tra_batches else float('nan'))
            # what the log line below prints, kept for callers that measure the loop (bench.py's e2e_files leg)
            self.epoch_timings.append({'epoch': epoch, 'batches': nbbatches, 'wait_for_data_s': float(np.sum(load_times)),
                                       'train_s': float(np.sum(train_times)), 'frames': int(frames_epoch)})

            cost_val = self.update_validation_cost(costs, X_vals, Y_vals)
            if world > 1:
                cost_val = sync.broadcast_scalar(cost_val)      # the random interpolation weights differ per replica

            print_log('    E{}/{} {}  cost_tra={:.6f} (load:{}s train:{}s)  cost_val={:.6f} ({:.4f}% RMSE)  {} MiB GPU {} MiB RAM'.format(
                epoch, cfg.train_max_nbepochs, trialstr, costs['model_training'][-1], time2str(np.sum(load_times)),
                time2str(np.sum(train_times)), cost_val, 100 * costs['model_rmse_validation'][-1] / worst_val, gpu_memused(),
                proc_memresident()))
            if np.isnan(cost_val):
                raise ValueError('ERROR: Validation cost is nan!')

            if rank == 0:
                self._model.save(fbase + '-last.h5', printfn=print_log, extras={'cost_val': cost_val})
            if epoch >= cfg.train_min_nbepochs:
                if best_val is None or cost_val < best_val:
                    best_val = cost_val
                    if rank == 0:
                        self._model.save(params_savefile, printfn=print_log, extras={'cost_val': cost_val},
                                         infostr='(E{} C{:.4f})'.format(epoch, best_val))
                    epochs_modelssaved.append(epoch)
                    nbnodecepochs = 0
                else:
                    nbnodecepochs += 1

            epochs_durs.append(time.time() - timeepochstart)
            print_log('    ET: {}   train ~time left: {}'.format(
                time2str(epochs_durs[-1]), time2str(np.median(epochs_durs[-10:]) * (cfg.train_max_nbepochs - epoch))))
            if rank == 0:
                self.saveTrainingState(fstate, printfn=print_log, extras={
                    'cost_val': cost_val, 'best_val': best_val, 'costs': costs, 'epochs_modelssaved': epochs_modelssaved,
                    'epochs_durs': epochs_durs, 'nbnodecepochs': nbnodecepochs, 'epoch': epoch})
            if world > 1:
                sync.barrier()          # the files of this epoch exist before any rank goes on (or returns)
            if nbnodecepochs >= cfg.train_cancel_nodecepochs:
                print_log('WARNING: validation error did not decrease for {} epochs. Early stop!'.format(cfg.train_cancel_nodecepochs))
                break

        loader.shutdown(wait=False)
        if best_val is None:
            raise ValueError('No model has been saved during training!')
        return {'epoch_stopped': epoch, 'worst_val': worst_val,
                'best_epoch': epochs_modelssaved[-1] if epochs_modelssaved else -1, 'best_val': best_val}

    @classmethod
    def randomize_hyper(cls, cfg):
        """Draw every (name, min, max) of cfg.train_hypers uniformly; integer bounds give integers (optimizertts.py:306-321)."""
        cfg = copy.copy(cfg)
        hyperstr = ''
        for name, lo, hi in cfg.train_hypers:
            if isinstance(lo, int) and isinstance(hi, int):
                val = int(np.random.randint(lo, hi))
                hyperstr += name + '=' + str(val) + ','
            else:
                val = float(np.random.uniform(lo, hi))
                hyperstr += name + '=' + str(val) + ','
            setattr(cfg, name, val)
        return cfg, hyperstr.rstrip(',')

    def train(self, indir, outdir, wdir, fid_lst_tra, fid_lst_val, params_savefile, cont=None):
        """Single run, or random hyper-parameter search over cfg.train_nbtrials trials (optimizertts.py:324-372)."""
        cfg = self.cfg
        if cfg.train_nbtrials > 1:
            cfg.train_hypers = [h for h in cfg.train_hypers]
        cfg_base = copy.copy(cfg)
        best = None
        for trial in range(1, 1 + cfg.train_nbtrials):
            try:
                trialstr = 'trial{}'.format(trial) if cfg.train_nbtrials > 1 else ''
                if cfg.train_nbtrials > 1 and len(cfg_base.train_hypers) > 0:
                    self.cfg, hyperstr = self.randomize_hyper(cfg_base)
                    print('    randomized hyper-parameters: ' + hyperstr)
                res = self.train_oneparamset(indir, outdir, wdir, fid_lst_tra, fid_lst_val, params_savefile,
                                             trialstr=trialstr, cont=cont)
                if best is None or res['best_val'] < best['best_val']:
                    best = res
            except ValueError as e:
                if cfg.train_nbtrials > 1:
                    print_log('WARNING: Training crashed: ' + str(e))
                else:
                    raise
        return best
