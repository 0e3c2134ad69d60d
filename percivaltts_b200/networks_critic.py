"""
Critic of the WGAN (reference percivaltts/networks_critic.py:44-96); its graph is executed by engine.Critic.
`input_features`, `input_ctx` and `output` were Keras tensors in the reference (:36-38); here they are descriptive
handles only -- OptimizerTTSWGAN.prepare() wires the critic through `kerasmodel`.
"""
from . import arch as A
from . import modeltts


class Critic(object):
    def __init__(self, vocoder, ctxsize, cfgarch):
        self.vocoder = vocoder
        self.ctxsize = ctxsize
        self.cfgarch = cfgarch
        arch = A.Arch.from_cfg(ctxsize, vocoder, cfgarch)
        self.kerasmodel = modeltts.NativeNetwork('critic', arch, A.critic_param_shapes(arch), seed=None)
        self.input_features = ('input_features', (None, None, vocoder.featuressize()))
        self.input_ctx = ('input_ctx', (None, None, ctxsize))
        self.output = ('critic_output', (None, None, 1))

    def count_params(self):
        return self.kerasmodel.count_params()
