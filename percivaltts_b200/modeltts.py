"""
Model wrapper: the python-3 / B200 counterpart of reference percivaltts/modeltts.py:46-205 (ModelTTS).

`kerasmodel` keeps its name (the optimizer touches it, optimizertts_wgan.py:111,121,316,328) but holds a
NativeNetwork: parameter inventory + host weights, materialised on the GPU (engine.Generator / engine.Critic on
libpctts.so) the first time it is used.  Construction and count_params() need no GPU.
"""
import json
import os
import pickle
import sys
from collections import OrderedDict

import numpy as np
import torch

from . import arch as A
from . import backend_b200


class NativeNetwork(object):
    """Stands where the reference has a keras.Model."""

    def __init__(self, kind, arch, shapes, seed=None):
        self.kind = kind                  # 'generator' | 'critic' | 'generic'
        self.arch = arch
        self.shapes = list(shapes)
        self._host = A.init_params(self.shapes, backend_b200.get_seed() if seed is None else seed)
        self._module = None

    # -- keras.Model-like surface -------------------------------------------------------------------------------
    def count_params(self):
        return A.count_params(self.shapes)

    def summary(self):
        print('    {} network: {} parameter tensors, {} parameters'.format(self.kind, len(self.shapes), self.count_params()))
        for name, shape, trainable in self.shapes:
            print('        {:<34}{:<22}{}'.format(name, str(tuple(shape)), '' if trainable else '(non-trainable)'))

    def get_weights(self):
        """OrderedDict name -> float32 numpy array (read back from the device when materialised)."""
        if self._module is not None:
            return self._module.params.export()
        return OrderedDict((k, v.copy()) for k, v in self._host.items())

    def set_weights(self, weights):
        for name, shape, _ in self.shapes:
            w = np.asarray(weights[name], dtype=np.float32)
            if tuple(w.shape) != tuple(shape):
                raise ValueError('shape mismatch for {}: {} vs {}'.format(name, w.shape, shape))
            self._host[name] = w.copy()
        if self._module is not None:
            self._module.params.load(self._host)

    def device_module(self, ops=None):
        """engine.Generator / engine.Critic holding the parameters on the GPU (created once)."""
        if self._module is None:
            from . import engine
            ops = ops or backend_b200.get_ops()
            if self.kind == 'generator':
                self._module = engine.Generator(ops, self.arch)
            elif self.kind == 'critic':
                self._module = engine.Critic(ops, self.arch)
            else:
                raise NotImplementedError('the Generic (FC/LSE) model is not on the B200 WGAN-GP path')
            self._module.params.load(self._host)
        return self._module

    def predict(self, x):
        """numpy [B,T,ctxsize] -> numpy [B,T,out], BatchNorm in inference mode (moving statistics)."""
        if self.kind != 'generator':
            raise NotImplementedError('predict() is defined for the generator')
        G = self.device_module()
        x = np.ascontiguousarray(x, dtype=np.float32)
        if x.ndim != 3 or x.shape[2] != self.arch.ctxsize:
            raise ValueError('expected input of shape [batch, frames, {}]'.format(self.arch.ctxsize))
        B, T, _ = x.shape
        if T == 0 or B == 0:
            return np.zeros((B, T, self.arch.outsize), dtype=np.float32)
        xd = torch.from_numpy(x).to(G.ops.device)
        out = G.ops.empty(B, T, self.arch.outsize)
        G.forward(G.plan(B, T, with_backward=False), xd, out, training=False)
        return out.cpu().numpy()

    def to_json(self):
        return json.dumps(dict(kind=self.kind, arch=self.arch.as_dict() if self.arch is not None else None,
                               shapes=[[n, list(s), bool(t)] for n, s, t in self.shapes]), indent=1)

    def save_weights(self, fname):
        np.savez(fname, **self.get_weights())

    def load_weights(self, fname):
        with np.load(fname if fname.endswith('.npz') else fname + '.npz') as z:
            self.set_weights(dict((k, z[k]) for k in z.files))


def network_from_json(txt):
    d = json.loads(txt)
    arch = A.Arch(**d['arch']) if d.get('arch') else None
    return NativeNetwork(d['kind'], arch, [(n, tuple(s), bool(t)) for n, s, t in d['shapes']])


class ModelTTS(object):
    ctxsize = -1
    vocoder = None
    kerasmodel = None

    def __init__(self, ctxsize, vocoder, kerasmodel=None):
        print('Building the TTS-dedicated model')
        self.ctxsize = ctxsize
        self.vocoder = vocoder
        if kerasmodel is not None:
            self.kerasmodel = kerasmodel
            self.kerasmodel.summary()

    def predict(self, x):
        return self.kerasmodel.predict(x)                                 # modeltts.py:68-69

    def count_params(self):
        return self.kerasmodel.count_params()                             # modeltts.py:72-73

    def save(self, fmodel, cfg=None, extras=None, printfn=print, infostr=''):
        """Same three files as modeltts.py:75-105, with a numpy container instead of HDF5:
        <fmodel>.arch.json, <fmodel>.weights.npz, <fmodel>.cfgextras.pkl"""
        if extras is None:
            extras = dict()
        printfn('    saving parameters in {} ...'.format(fmodel), end='')
        with open(fmodel + '.arch.json', 'w') as f:
            f.write(self.kerasmodel.to_json())
        self.kerasmodel.save_weights(fmodel + '.weights.npz')
        with open(fmodel + '.cfgextras.pkl', 'wb') as f:
            pickle.dump([cfg, extras], f)
        print(' done ' + infostr)
        sys.stdout.flush()

    def load(self, fmodel, printfn=print, compile=True):
        printfn('    reloading parameters from {} ...'.format(fmodel), end='')
        if self.kerasmodel is None:
            with open(fmodel + '.arch.json') as f:
                self.kerasmodel = network_from_json(f.read())
        self.kerasmodel.load_weights(fmodel + '.weights.npz')
        with open(fmodel + '.cfgextras.pkl', 'rb') as f:
            DATA = pickle.load(f)
        print(' done')
        sys.stdout.flush()
        return DATA

    def generate_cmp(self, inpath, outpath, fid_lst):
        """One utterance at a time, raw float32 [T, out] files (modeltts.py:133-141)."""
        from . import data
        outdir = os.path.dirname(outpath)
        if outdir and not os.path.isdir(outdir):
            os.makedirs(outdir)
        X = data.load(inpath, fid_lst, verbose=1)
        for vi, fid in enumerate(fid_lst):
            CMP = self.predict(X[vi][np.newaxis])
            CMP[0].astype('float32').tofile(outpath.replace('*', fid))

    def generate_wav(self, *args, **kwargs):
        raise NotImplementedError('waveform synthesis (pulsemodel) is outside the B200 hot path; use generate_cmp() '
                                  'and VocoderPML.decompress_features()')
