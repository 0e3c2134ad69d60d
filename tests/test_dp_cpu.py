"""
Data-parallel host logic on the CPU (gloo, world_size 2): sharding the batch by utterance with sync-BN and SUM
all-reduce of the flat gradient buffers reproduces the single-device step on the concatenated batch.
"""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

KW = dict(ctxsize=14, spec_size=11, nm_size=3, hiddenwidth=6, ctx_nbcnnlayers=1, ctx_winlen=5, gen_nbcnnlayers=2,
          gen_nbfilters=2, gen_winlen=3, spec_freqlen=3)


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, outdir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, 'tests'))
    os.environ.update(MASTER_ADDR='127.0.0.1', MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world), LOCAL_RANK=str(rank))
    torch.set_num_threads(1)
    dist.init_process_group('gloo', rank=rank, world_size=world)
    from cpu_ops_double import CpuOpsDouble
    from oracle import percival_oracle as O
    from percivaltts_b200 import arch as A
    from percivaltts_b200.parallel import DistSync, shard_batch
    from percivaltts_b200.wgan import WGANStepper
    ao, a = O.Arch(**KW), A.Arch(**KW)
    wg = O.perturb_params(O.init_params(O.generator_param_shapes(ao), seed=1))
    wc = O.perturb_params(O.init_params(O.critic_param_shapes(ao), seed=2))
    _, wls, wbar = O.wlswgan_weights(ao)
    dt = torch.float64
    st = WGANStepper(CpuOpsDouble(dt), a, w_ls=wls, w_wgan_mean=wbar, sync=DistSync())
    st.G.params.load(wg)
    st.C.params.load(wc)
    for it in range(2):
        X, Y, alpha = O.synthetic_batch(ao, 4, 17, seed=40 + it)
        Xl, Yl, al = (torch.tensor(shard_batch(v, rank, world), dtype=dt) for v in (X, Y, alpha))
        st.critic_step(Xl, Yl, alpha=al)
        st.generator_step(Xl, Yl)
    np.savez(os.path.join(outdir, 'rank%d.npz' % rank), losses=st.losses.numpy(),
             **dict(('C.' + k, v) for k, v in st.C.params.export().items()),
             **dict(('G.' + k, v) for k, v in st.G.params.export().items()))
    dist.destroy_process_group()


def test_two_rank_step_equals_single_device_step(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    from oracle import percival_oracle as O
    ao = O.Arch(**KW)
    wg = O.perturb_params(O.init_params(O.generator_param_shapes(ao), seed=1))
    wc = O.perturb_params(O.init_params(O.critic_param_shapes(ao), seed=2))
    tr = O.OracleTrainer(ao, wg, wc, dtype=torch.float64)
    for it in range(2):
        X, Y, alpha = O.synthetic_batch(ao, 4, 17, seed=40 + it)
        oc, og, _, _ = tr.train_on_batch(0, X, Y, alpha)
    r0 = np.load(str(tmp_path / 'rank0.npz'))
    r1 = np.load(str(tmp_path / 'rank1.npz'))
    assert abs(r0['losses'][3] - oc) < 1e-9 and abs(r0['losses'][6] - og) < 1e-9      # all-reduced global means
    for k, v in tr.wc.items():
        v = v.detach().numpy()
        assert np.array_equal(r0['C.' + k], r1['C.' + k]), k                             # replicas stay identical
        assert np.abs(r0['C.' + k] - v).max() < 1e-8 * (np.abs(v).max() + 1e-9), k
    for k, v in tr.wg.items():
        v = v.detach().numpy()
        assert np.array_equal(r0['G.' + k], r1['G.' + k]), k
        assert np.abs(r0['G.' + k] - v).max() < 1e-8 * (np.abs(v).max() + 1e-9), k


def test_shard_batch_rejects_uneven_split():
    from percivaltts_b200.parallel import shard_batch
    with pytest.raises(ValueError):
        shard_batch(np.zeros((5, 3, 2)), 0, 2)
    assert shard_batch(np.arange(8).reshape(4, 2), 1, 2).tolist() == [[4, 5], [6, 7]]


def test_two_rank_train_through_the_public_api(tmp_path):
    """`torchrun run.py`-style data parallelism through the public API alone (gloo, 2 ranks, CPU test double): the backend
    rendezvous from the environment, the epoch loop shards every global batch, rank 0 writes the files, replicas end identical."""
    import subprocess
    env = dict(os.environ, PCTTS_DP_WORKDIR=str(tmp_path / 'work'), CUDA_VISIBLE_DEVICES='')
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr', '127.0.0.1',
           '--master-port', str(_free_port()), os.path.join(ROOT, 'tools', 'dp_api_check.py')]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)
    assert r.returncode == 0 and 'DP API OK' in r.stdout, (r.stdout[-3000:], r.stderr[-3000:])


def test_two_rank_train_with_mixed_lengths_and_unsynchronised_rngs(tmp_path):
    """lengthmax=None: T is the shortest utterance of each global batch, so replicas whose numpy RNGs differ would shuffle and
    crop differently and disagree on (B,T).  The ranks start from different seeds; rank 0's state, broadcast at every epoch start,
    must keep them on the same batches (the reference is one process seeded with 123, data.py:32)."""
    import subprocess
    env = dict(os.environ, PCTTS_DP_WORKDIR=str(tmp_path / 'work'), CUDA_VISIBLE_DEVICES='', PCTTS_DP_MIXED_LENGTHS='1',
               PCTTS_DP_CHECK='1')
    cmd = [sys.executable, '-m', 'torch.distributed.run', '--nnodes=1', '--nproc-per-node', '2', '--master-addr', '127.0.0.1',
           '--master-port', str(_free_port()), os.path.join(ROOT, 'tools', 'dp_api_check.py')]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)
    assert r.returncode == 0 and 'DP API OK' in r.stdout, (r.stdout[-3000:], r.stderr[-3000:])


def test_import_seeds_numpy_like_the_reference():
    """reference data.py:32 / optimizertts.py:37: importing the package leaves numpy's global RNG at seed 123."""
    import subprocess
    code = ("import numpy as np; np.random.seed(5); import percivaltts_b200.data, percivaltts_b200.optimizertts; a = np.random.rand(3);"
            "np.random.seed(123); assert (a == np.random.rand(3)).all(); print('SEEDED')")
    r = subprocess.run([sys.executable, '-c', code], capture_output=True, text=True, timeout=300, cwd=ROOT,
                       env=dict(os.environ, CUDA_VISIBLE_DEVICES=''))
    assert 'SEEDED' in r.stdout, r.stderr[-2000:]
