"""
`torchrun --nproc-per-node N tools/dp_api_check.py`: data-parallel training through the PUBLIC API only -- what `torchrun run.py`
does: nobody calls parallel.init_from_env by hand; backend_b200 rendezvous from the environment, every rank lands
on its own GPU (LOCAL_RANK), optimizertts.train_oneparamset shards every global batch, rank 0 alone writes the files, the
replicas end bit-identical.  Works on CPU too (gloo + the test double of the kernel interface) for tests/test_dp_cpu.py.
"""
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, 'tests'))

import numpy as np
import torch
import torch.distributed as dist


def main():
    import faulthandler
    faulthandler.dump_traceback_later(300, exit=True)
    from percivaltts_b200 import backend_b200, modeltts_common, networks_critic, optimizertts_wgan, vocoders
    from test_api_cpu import make_files, tiny_cfg
    rank, world = int(os.environ.get('RANK', '0')), int(os.environ.get('WORLD_SIZE', '1'))
    cpu = not torch.cuda.is_available()
    if cpu:
        torch.set_num_threads(1)
        from cpu_ops_double import CpuOpsDouble
        backend_b200.get_sync()
        backend_b200.set_ops(CpuOpsDouble(torch.float32))
    workdir = os.environ.get('PCTTS_DP_WORKDIR') or os.path.join(tempfile.gettempdir(), 'pctts_dp_api_%s' % os.environ.get('MASTER_PORT', '0'))
    if rank == 0:
        os.makedirs(workdir, exist_ok=True)

    class P(object):                       # make_files wants a pathlib-like object
        def __init__(self, p):
            self.p = p

        def __truediv__(self, o):
            return P(os.path.join(self.p, o))

        def __str__(self):
            return self.p

    lens = [60, 64, 57, 53, 49, 56, 61, 50, 55, 58]
    sync = backend_b200.get_sync()
    assert sync is not None and sync.world == world
    if rank == 0:
        make_files(P(workdir), lens)
    sync.barrier()
    fids = ['utt%02d' % i for i in range(len(lens))]
    indir, outdir, wdir = (os.path.join(workdir, 'lab', '*.lab') + ':(-1,425)', os.path.join(workdir, 'cmp', '*.cmp') + ':(-1,83)',
                           os.path.join(workdir, 'w', '*.w') + ':(-1,1)')
    cfg = tiny_cfg()
    cfg.train_batch_size = 2 * world
    cfg.train_min_nbepochs = 1
    cfg.train_max_nbepochs = 2
    cfg.train_cancel_nodecepochs = 2
    cfg.train_batch_lengthmax = 32
    if os.environ.get('PCTTS_DP_MIXED_LENGTHS'):
        # T = the shortest utterance of each global batch (lengthmax None), and the ranks' numpy RNGs deliberately apart: only
        # the epoch-start broadcast of rank 0's state keeps the replicas on the same batches
        cfg.train_batch_lengthmax = None
        np.random.seed(1000 + rank)
    cfg.train_nbepochs_scalewdata = False
    voc = vocoders.VocoderPML(16000, 0.005, 65, 17)
    quiet = open(os.devnull, 'w') if rank else None
    if quiet:
        sys.stdout = quiet
    model = modeltts_common.DCNNF0SpecNoiseFeatures(425, voc, cfg)
    critic = networks_critic.Critic(voc, 425, cfg)
    opti = optimizertts_wgan.OptimizerTTSWGAN(cfg, model, errtype='WLSWGAN', critic=critic)
    if cpu:
        opti.use_graphs = False
    fmodel = os.path.join(workdir, 'model.h5')
    res = opti.train(indir, outdir, wdir, fids[:8], fids[8:], fmodel)
    sys.stdout = sys.__stdout__
    ops = backend_b200.get_ops()
    dev = ops.device
    if not cpu:
        assert dev.index == int(os.environ.get('LOCAL_RANK', '0')), ('rank on the wrong GPU', dev)
    st = opti.stepper
    flat = torch.cat([st.C.params.flat, st.G.params.flat, st.G.params.state]).clone()
    ref = flat.clone()
    dist.broadcast(ref, 0)
    same = torch.tensor([int(torch.equal(ref, flat) and np.isfinite(res['best_val']) and res['epoch_stopped'] == 2)], device=dev)
    dist.all_reduce(same, op=dist.ReduceOp.MIN)
    files_ok = os.path.exists(fmodel + '.weights.npz') and os.path.exists(os.path.join(workdir, 'model-trainingstate-last.h5.wgan.npz'))
    if rank == 0:
        print('world %d: replicas identical and finite: %s; rank-0 files written: %s  %s'
              % (world, bool(int(same)), files_ok, 'DP API OK' if int(same) and files_ok else 'DP API FAILED'), flush=True)
    sys.stdout.flush()
    os._exit(0 if int(same) and files_ok else 1)


if __name__ == '__main__':
    main()
