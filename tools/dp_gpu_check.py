"""
Data-parallel parity on real GPUs (launch: torchrun --nproc-per-node N tools/dp_gpu_check.py): N ranks, each with B/N utterances
of one global batch, sync-BN + NCCL all-reduce of the flat gradients, against the CPU oracle's single-device step on the whole
batch.  Also runs the CUDA-graph trainer path for a few steps to check that replicas stay bit-identical.
"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import torch.distributed as dist
from oracle import percival_oracle as O
from percivaltts_b200 import arch as A, parallel, wgan as W
from percivaltts_b200.ops import CudaOps
from percivaltts_b200.trainer import DeviceTrainer
from percivaltts_b200.wgan import WGANStepper

KW = dict(ctxsize=60, spec_size=33, nm_size=9, hiddenwidth=32, ctx_nbcnnlayers=1, ctx_winlen=21, gen_nbcnnlayers=4,
          gen_nbfilters=4, gen_winlen=5, spec_freqlen=5)


def main():
    import faulthandler
    faulthandler.dump_traceback_later(300, exit=True)
    sync, device = parallel.init_from_env()
    rank, world = sync.rank, sync.world
    print('rank %d: peer-memory exchange %s' % (rank, 'on' if sync._peer is not None else 'off (%s)' % sync.peer_error), flush=True)
    ops = CudaOps()
    ao, a = O.Arch(**KW), A.Arch(**KW)
    wg = O.perturb_params(O.init_params(O.generator_param_shapes(ao), seed=1))
    wc = O.perturb_params(O.init_params(O.critic_param_shapes(ao), seed=2))
    _, wls, wbar = O.wlswgan_weights(ao)
    st = WGANStepper(ops, a, w_ls=wls, w_wgan_mean=wbar, sync=sync)
    st.G.params.load(wg)
    st.C.params.load(wc)
    Bg, T = 2 * world, 48
    X, Y, alpha = O.synthetic_batch(ao, Bg, T, seed=11)
    Xl, Yl, al = (torch.tensor(parallel.shard_batch(v, rank, world)).to(device) for v in (X, Y, alpha))
    st.critic_step(Xl, Yl, alpha=al, apply=False)
    sync.allreduce(st.C.params.grad)
    lc = float(st.losses[W.L_CRITIC])
    gc = st.C.params.export_grads()
    st.generator_step(Xl, Yl, apply=False)
    sync.allreduce(st.G.params.grad)
    lg = float(st.losses[W.L_GEN])
    gg = st.G.params.export_grads()
    ok = True
    if rank == 0:
        dt = torch.float64
        wgt, wct = O.to_torch(wg, dt, True), O.to_torch(wc, dt, True)
        Xt, Yt = torch.tensor(X, dtype=dt), torch.tensor(Y, dtype=dt)
        rc = O.critic_step(ao, wgt, wct, Xt, Yt, torch.tensor(alpha, dtype=dt).reshape(-1, 1, 1))
        rg = O.generator_step(ao, wgt, wct, Xt, Yt)
        def rel(got, want):
            want = want.detach().double().numpy()
            return float(np.abs(got - want).max() / (np.abs(want).max() + 1e-30))
        worst = max([rel(gc[k], v) for k, v in rc['grads'].items() if float(v.abs().max()) > 1e-10] +
                    [rel(gg[k], v) for k, v in rg['grads'].items() if float(v.abs().max()) > 1e-10])
        print('dp%d: critic loss %.6f (oracle %.6f)  generator loss %.6f (oracle %.6f)  worst grad rel err %.2e'
              % (world, lc, float(rc['loss']), lg, float(rg['loss']), worst))
        ok = abs(lc - float(rc['loss'])) < 1e-3 * max(abs(float(rc['loss'])), 1e-2) and abs(lg - float(rg['loss'])) < 1e-3 * abs(float(rg['loss'])) and worst < 1e-3
    # graph path: replicas must stay bit-identical after updates
    tr = DeviceTrainer(st)
    for i in range(6):
        Xb, Yb, _ = O.synthetic_batch(ao, Bg, T, seed=20 + i)
        tr.train_on_batch(i * 5, parallel.shard_batch(Xb, rank, world), parallel.shard_batch(Yb, rank, world))
    # a rank that arrives late (a host pause of 3 s before one step) is waited for, not timed out on: the exchange's spin limit is ~60 s
    # and its time-out flag is read back with the losses of every step (DistSync.check_peer)
    if world > 1 and sync._peer is not None:
        import time
        if rank == world - 1:
            time.sleep(3.0)
        Xb, Yb, _ = O.synthetic_batch(ao, Bg, T, seed=40)
        tr.train_on_batch(30, parallel.shard_batch(Xb, rank, world), parallel.shard_batch(Yb, rank, world))
        torch.cuda.synchronize()
        late_ok = not sync.peer_failed()
        print('rank %d: step with rank %d arriving 3 s late: %s' % (rank, world - 1, 'ok' if late_ok else 'FAILED (time-out flag set)'), flush=True)
        ok = ok and late_ok
    # soak of the peer-memory exchange: 200 replays of a captured graph holding 8 exchanges of different sizes (the parity
    # double-buffering of slots / flags under replay at this world size), results checked, time-out flag must stay 0
    soak_ok = True
    if world > 1 and sync._peer is not None:
        sizes = [8, 512, 64, 1024, 8, 8, 256, 16]
        bufs = [torch.zeros(n, dtype=torch.float64, device=device) for n in sizes]
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for b in bufs:
                sync.allreduce(b)
        torch.cuda.current_stream().wait_stream(side)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for j, b in enumerate(bufs):
                b.copy_(torch.arange(b.numel(), dtype=torch.float64, device=device) * (rank + 1) + j)
                sync.allreduce(b)
        for it in range(200):
            g.replay()
            if it % 50 == 49:
                torch.cuda.synchronize()
                for j, b in enumerate(bufs):
                    want = torch.arange(b.numel(), dtype=torch.float64, device=device) * (world * (world + 1) / 2.0) + j * world
                    soak_ok = soak_ok and bool(torch.equal(b, want))
        torch.cuda.synchronize()
        soak_ok = soak_ok and not sync.peer_failed()
        print('rank %d: peer all-reduce soak (200 graph replays x %d exchanges) %s, err flag %d'
              % (rank, len(sizes), 'ok' if soak_ok else 'FAILED', int(sync.peer_err)), flush=True)
    ok = ok and soak_ok
    # fused gradient all-reduce + Adam over peer memory (csrc/peer.cu) against NCCL all-reduce + adam_kernel on copies
    fused_ok = True
    if world > 1 and sync._gs:
        for store in (st.C.params, st.G.params):
            n = store.ntrain
            gen = torch.Generator(device=device).manual_seed(100 + rank)
            store.grad.copy_(torch.randn(n, device=device, generator=gen) * 0.01)
            want_g = store.grad.clone()
            dist.all_reduce(want_g)
            p0, m0, v0, s0 = store.flat.clone(), store.m.clone(), store.v.clone(), store.step.clone()
            ops.adam(p0[:n], want_g, m0[:n], v0[:n], 1e-3, 0.5, 0.9, 1e-7, s0)
            before = sync.fused_adam_calls
            for rep in range(3):          # replays: sequence numbers / flags advance
                if rep:
                    store.grad.copy_(torch.randn(n, device=device, generator=gen) * 0.01)
                    want_g = store.grad.clone()
                    dist.all_reduce(want_g)
                    ops.adam(p0[:n], want_g, m0[:n], v0[:n], 1e-3, 0.5, 0.9, 1e-7, s0)
                lv = torch.arange(4, dtype=torch.float32, device=device) + rank
                sync.grad_adam(store, 1e-3, 0.5, 0.9, 1e-7, losses=lv)
                good_l = bool(torch.equal(lv, torch.arange(4, dtype=torch.float32, device=device) * world + world * (world - 1) / 2.0))
                fused_ok = fused_ok and good_l
            torch.cuda.synchronize()
            err_p = float((store.flat[:n] - p0[:n]).abs().max() / (p0[:n].abs().max() + 1e-30))
            err_g = float((store.grad - want_g).abs().max() / (want_g.abs().max() + 1e-30))
            good = (sync.fused_adam_calls == before + 3 and err_p < 1e-6 and err_g < 1e-6 and int(store.step) == int(s0)
                    and not sync.peer_failed())
            fused_ok = fused_ok and good
            print('rank %d: fused all-reduce + Adam (%d params, 3 calls): rel err params %.1e, summed gradient %.1e, step %d/%d %s'
                  % (rank, n, err_p, err_g, int(store.step), int(s0), 'ok' if good else 'FAILED'), flush=True)
    elif world > 1:
        print('rank %d: fused all-reduce + Adam not set up (%s)' % (rank, sync.peer_error), flush=True)
    ok = ok and fused_ok
    flat = torch.cat([st.C.params.flat, st.G.params.flat, st.G.params.state])
    ref = flat.clone()
    if world > 1:
        dist.broadcast(ref, 0)
    same = bool(torch.equal(ref, flat))
    t = torch.tensor([int(ok and same)], device=device)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        print('replicas identical after 6 graph steps:', same, ' ALL OK' if int(t) else ' FAILED')
    # destroy_process_group() blocks while captured CUDA graphs still hold NCCL work: leave through os._exit
    sys.stdout.flush()
    os._exit(0 if int(t) else 1)


if __name__ == '__main__':
    main()
