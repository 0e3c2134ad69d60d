#!/usr/bin/env python
"""
bench.py -- WGAN-GP training throughput of the B200 path (BASELINE.json metric: train frames/s; critic step ms).

    python bench.py [--gpus N] [--steps K] [--warmup W]            # N>1: launched by torchrun, one rank per GPU
    python bench.py --impl reference [...]                          # the CPU arm: oracle restatement on the host cores

Workload (config #2 of BASELINE.json): SLT-Arctic demo-shaped Percival CNN-WGAN, synthetic labels / PML
frames, per-GPU batch B=10 utterances x T=400 frames, ctx=425, out=163 (spec 129 + nm 33 + f0), fp32.
A *step* is one OptimizerTTSWGAN.train_on_batch() call: one critic update, plus one generator update on every 5th
step (the reference's steady-state schedule, optimizertts_wgan.py:225-231 with generator_updates >= 25).
`value`  : frames/s with the batches already resident in HBM (graph replays only).
`e2e`    : frames/s through the public API with host numpy batches (pinned staging + H2D + loss D2H inside the timing; like
           the package's own epoch loop the call hands over the next host batch too, so its copy overlaps this step).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

B_PER_GPU, T_FRAMES = 10, 400
METRIC, UNIT = 'wgan_gp_train_frames_per_s', 'frames/s'
WORKLOAD = 'slt_arctic_demo_cnn_wgan_gp_step B=10/GPU T=400 ctx=425 out=163 H=256 fp32 synthetic (BASELINE config #2)'


def demo_cfg():
    from percivaltts_b200.percivaltts import configuration
    cfg = configuration()               # reference run.py:76,89,114-131
    cfg.fs = 16000
    cfg.arch_hiddenwidth = 256
    cfg.arch_ctx_nbcnnlayers = 1
    cfg.arch_ctx_winlen = 21
    cfg.arch_gen_nbcnnlayers = 8
    cfg.arch_gen_nbfilters = 4
    cfg.arch_gen_winlen = 5
    cfg.arch_spec_freqlen = 5
    cfg.train_batch_size = B_PER_GPU
    cfg.train_batch_lengthmax = T_FRAMES
    cfg.train_wgan_LScoef = 0.25
    return cfg


def synthetic_batch(B, T, seed):
    """sparse +-1 label columns + 9 continuous ones; N(0,1) lf0/spec, U(0,1) noise mask."""
    rng = np.random.RandomState(seed)
    X = np.empty((B, T, 425), dtype=np.float32)
    X[:, :, :416] = np.where(rng.uniform(size=(B, T, 416)) < 0.1, 1.0, -1.0)
    X[:, :, 416:] = rng.uniform(-1, 1, size=(B, T, 9))
    Y = np.empty((B, T, 163), dtype=np.float32)
    Y[:, :, :130] = rng.normal(size=(B, T, 130))
    Y[:, :, 130:] = rng.uniform(0, 1, size=(B, T, 33))
    return X, Y


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs (profiling recipe)."""
    Q = ('clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q,
                                          '--format=csv,noheader,nounits', '-lms', os.environ.get('PCTTS_BENCH_CLOCK_MS', '200')], stdout=subprocess.PIPE, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if self.proc is None:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['nvidia-smi unavailable']}
        time.sleep(0.25)
        self.proc.terminate()
        self.thread.join(timeout=2)
        sm, smax, reasons = [], None, set()
        for ln in self.lines:
            f = [x.strip() for x in ln.split(',')]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                smax = float(f[1])
            except ValueError:
                continue
            for name, val in zip(('hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap'), f[3:7]):
                if val.lower().startswith('active'):
                    reasons.add(name)
        return {'sm_mhz': float(np.median(sm)) if sm else None, 'sm_max_mhz': smax, 'reasons': sorted(reasons),
                'samples': len(sm)}


def max_over_ranks(x, device):
    import torch
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        t = torch.tensor([x], dtype=torch.float64, device=device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())
    return x


def barrier():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()


def time_kernel(fn, device, iters=20, flush_mb=256):
    """Average device time of one launch, CUDA events on the launching stream, L2 flushed before every launch."""
    import torch
    flush = torch.empty(flush_mb * 1024 * 1024 // 4, dtype=torch.float32, device=device)
    for _ in range(3):
        fn()
    tot = 0.0
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        tot += e0.elapsed_time(e1)
    return tot / iters


def cpu_baseline_cycle(threads):
    """The oracle (torch-CPU fp32 restatement of the reference graph) on the host cores: one steady-state cycle of
    5 batches (5 critic updates + 1 generator update) at the full B=10, T=400 after one warm-up batch."""
    import torch
    from oracle import percival_oracle as O
    torch.set_num_threads(threads)
    a = O.demo_arch()
    tr = O.OracleTrainer(a, O.init_params(O.generator_param_shapes(a), seed=1), O.init_params(O.critic_param_shapes(a), seed=2),
                         dtype=torch.float32)
    tr.generator_updates = 25
    X, Y, alpha = O.synthetic_batch(a, B_PER_GPU, T_FRAMES, seed=123)
    tr.train_on_batch(1, X, Y, alpha)
    t0 = time.perf_counter()
    for i in range(5):
        tr.train_on_batch(i, X, Y, alpha)
    dt = time.perf_counter() - t0
    return 5 * B_PER_GPU * T_FRAMES / dt, dt


def bench_config(world, B, precision='fp32', global_batch=None):
    """The `config` object of the JSON line -- the same function, hence the same dict, for both arms."""
    wl = WORKLOAD
    if precision != 'fp32' or global_batch:
        wl = WORKLOAD.replace('fp32', precision).replace('B=10/GPU', 'B=%d/GPU' % B).replace(
            'config #2', 'config #3' if precision == 'bf16' else 'config #2 variant')
    return {'workload': wl, 'global_batch_utterances': B * world, 'frames_per_step': B * T_FRAMES * world,
            'schedule': '1 critic update per step + 1 generator update every 5th step (generator_updates>=25)',
            'l2': 'per-step working set (activations of 3x batched critic + generator, ~0.6 GB) exceeds the 126 MB L2; '
                  'inputs cycle over 4 resident batches',
            'parallelism': 'dp%d sync-BN + flat-gradient all-reduce' % world if world > 1 else 'single GPU'}


def run_reference(args):
    """--impl reference: the reference's TensorFlow-1.9 / python-2 step cannot run here, so the timed CPU arm
    is the oracle port with all host threads.  A step is one train_on_batch on the full B=10 batch (~0.7 s on 16 cores), the
    same unit as the B200 arm; only a run asked for more than 150 steps drops to 4 utterances per step to stay within minutes."""
    rank = int(os.environ.get('RANK', '0'))
    if rank != 0:
        return
    import torch
    from oracle import percival_oracle as O
    threads = os.cpu_count() or 1
    torch.set_num_threads(threads)
    a = O.demo_arch()
    tr = O.OracleTrainer(a, O.init_params(O.generator_param_shapes(a), seed=1), O.init_params(O.critic_param_shapes(a), seed=2),
                         dtype=torch.float32)
    tr.generator_updates = 25
    Bs = B_PER_GPU if args.steps <= 150 else 4
    X, Y, alpha = O.synthetic_batch(a, Bs, T_FRAMES, seed=123)
    for i in range(max(args.warmup, 1)):
        tr.train_on_batch(i + 1, X, Y, alpha)
    t0 = time.perf_counter()
    for i in range(args.steps):
        tr.train_on_batch(i, X, Y, alpha)
    dt = time.perf_counter() - t0
    value = args.steps * Bs * T_FRAMES / dt
    sample = '%d steps of train_on_batch on %d of the 10 utterances per batch (T=400), generator every 5th step' % (args.steps, Bs)
    print(json.dumps({
        'impl': 'reference', 'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': args.gpus, 'steps': args.steps,
        'warmup': args.warmup, 'ms_per_step': 1e3 * dt / args.steps, 'higher_is_better': True, 'scaling': 'weak',
        'vs_baseline': None, 'dtype': 'f32', 'data': 'synthetic', 'config': bench_config(args.gpus, B_PER_GPU),
        'arm_note': 'CPU port of the reference graph on rank 0: one replica\'s batch per step whatever --gpus says',
        'cpu_baseline': {'value': value, 'unit': UNIT, 'cores': threads, 'kind': 'port', 'sample': sample},
        'e2e': {'value': value, 'unit': UNIT, 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0}}))


def step_min_flops(a, B, T):
    """Algorithmic flops (2 x multiply-adds of every contraction) of one critic step and one generator step in the minimum-work
    variant this build executes: the critic's ctx branch once per batch, the frozen generator's spectral branch
    only in the critic step.  Backward = one data-gradient + one weight-gradient contraction per forward contraction where the
    gradient is needed; the gradient penalty adds the input-gradient chain, the tangent sweep and the weight gradients of the
    interpolated application."""
    M, H, S, Fl, W, L = B * T, a.hiddenwidth, a.spec_size, a.gen_nbfilters, a.ctx_winlen, a.gen_nbcnnlayers
    KK = a.gen_winlen * a.spec_freqlen
    conv1d = 2.0 * M * W * a.ctxsize * H
    fc = 2.0 * M * H * H
    c2_first, c2_mid, c2_last = 2.0 * M * S * KK * 1 * Fl, 2.0 * M * S * KK * Fl * Fl, 2.0 * M * S * KK * Fl * 1
    g_ctx = a.ctx_nbcnnlayers * conv1d + 2 * fc
    g_spec = 2.0 * M * H * S + c2_first + (L - 1) * c2_mid + c2_last
    g_f0 = 2 * (2.0 * M * H * 4 * H) + 2 * (2.0 * M * H * 4 * H) + 2.0 * M * 2 * H
    g_nm = (L // 2) * fc + 2.0 * M * H * a.nm_size
    g_fwd = g_ctx + g_spec + g_f0 + g_nm
    c_ctx = a.ctx_nbcnnlayers * conv1d + 2 * fc + fc                       # + the ctx rows of the first post layer
    c_app = c2_first + (L - 1) * c2_mid + 2.0 * M * S * Fl * H + 2 * fc + 2.0 * M * H          # spectral stack + post layers, one application
    critic = (g_ctx + g_spec) + c_ctx + 3 * c_app            # forward: frozen generator (spectral) + ctx once + 3 applications
    critic += 2 * c_ctx - conv1d                               # ctx-branch backward: weight gradients everywhere, data gradients but into the labels
    critic += 3 * (2 * c_app) + c_app                          # 3 applications: data + weight gradients; + the tangent sweep of the penalty
    gen = g_fwd + c_ctx + c_app + c_app + 2 * g_fwd - conv1d   # G fwd, D fwd, D input gradient, G backward (data + weight)
    return {'critic': critic, 'generator': gen}


def kernel_rooflines(st, ops, B, T, device, peak_tf, peak_hbm, critic_ms, gen_ms):
    """Each hot kernel timed ALONE in this run (CUDA events on the launching stream, L2 flushed by a 1 GiB fill before every
    launch), against the roofline that bounds it; algorithmic bytes / flops per launch as stated in DESIGN.md section 4.
    `share_of_critic_step` = kernel_ms x launches per critic step / measured critic step (two lanes overlap, so the shares are
    upper bounds of the exposed time); the ncu launch list under profiles/ holds the serialised per-launch times."""
    import torch
    a = st.a
    S, Fl, H, W = a.spec_size, a.gen_nbfilters, a.hiddenwidth, a.ctx_winlen
    L = a.gen_nbcnnlayers
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.float32, device=device)

    def timed(fn, iters=12):
        for _ in range(3):
            fn()
        tot = 0.0
        for _ in range(iters):
            flush.zero_()
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a0.record()
            fn()
            a1.record()
            a1.synchronize()
            tot += a0.elapsed_time(a1)
        return tot / iters

    out = []
    products = max(ops.nterms, 1)
    with ops.shape_scope((B, T)), ops.step_scope():
        gpl, cpl = st.G.plan(B, T), st.C.plan(3, B, T)
        bf = st.buffers(B, T)
        cp, gp = st.C.params.p, st.G.params.p
        # -- Conv2D 5x5 4->4 on the batched critic map [3B,T,S,4] (forward: bias + LeakyReLU) -- HBM bound
        x3, y3 = cpl['a'][1], cpl['a'][2]
        t = timed(lambda: ops.conv2d(x3, cp['c.spec.cnn2.kernel'], y3, bias=cp['c.spec.cnn2.bias'], flags=2))
        nbytes = 2.0 * x3.numel() * 4
        out.append({'name': 'conv2d_4to4_critic_3B', 'kernel': 'conv2d_tc_kernel<%d> on [3B,T,%d,4] (critic forward layer)' % ({6: 3, 3: 2, 1: 1}.get(ops.nterms, 0), S),
                    'bound': 'hbm', 'achieved': nbytes / (t * 1e-3) / 1e9, 'peak': peak_hbm, 'unit': 'GB/s',
                    'frac': nbytes / (t * 1e-3) / 1e9 / peak_hbm, 'kernel_ms': t, 'bytes_per_launch': nbytes, 'traffic': None,
                    'launches_per_critic_step': 0 if st.C._chain(T) else 2 * (L - 1), 'share_of_critic_step': 0.0 if st.C._chain(T) else t * 2 * (L - 1) / critic_ms,
                    'algorithmic': 'one fp32 map read + one written: 2 x 3B x T x %d x 4 ch x 4 B' % S})
        # -- the critic's 8-layer stack as ONE launch per chain (conv2d_chain.cu): forward on [3B,..], data gradient (+ input
        #    gradient) and tangent sweep on [B,..]; bytes = every map the chain must read or write once (fp32)
        if st.C._chain(T):
            px = B * T * S * 4.0                                  # bytes of one single-channel map of B samples
            s0 = bf['s0'].view(3 * B, T, S, 1)
            fwd_layers = [dict(w=cp['c.spec.cnn%d.kernel' % i], bias=cp['c.spec.cnn%d.bias' % i], y=cpl['a'][i], flags=2) for i in range(L)]
            t = timed(lambda: ops.conv2d_chain(s0, fwd_layers))
            nb = 3 * px * (1 + 4 * L)
            out.append({'name': 'conv2d_chain_fwd_3B', 'kernel': 'conv2d_chain_kernel<3>: %d Conv2D layers (1->4, 4->4 x%d) + bias + LeakyReLU on [3B,T,%d,.], 8-CTA cluster per sample' % (L, L - 1, S),
                        'bound': 'hbm', 'achieved': nb / (t * 1e-3) / 1e9, 'peak': peak_hbm, 'unit': 'GB/s', 'frac': nb / (t * 1e-3) / 1e9 / peak_hbm,
                        'kernel_ms': t, 'bytes_per_launch': nb, 'traffic': None, 'launches_per_critic_step': 1,
                        'share_of_critic_step': t / critic_ms, 'layers_per_launch': L, 'us_per_layer': 1e3 * t / L,
                        'algorithmic': 'input map (1 ch) read + %d output maps (4 ch) written once: (1 + 4 x %d) x 3B x T x %d x 4 B; tiles stay in shared memory between layers' % (L, L, S)})
            if (B, T, S, L) == (10, 400, 129, 8):
                # one `ncu --set full` capture of this launch (grid 240 = 3B x 8 CTAs) inside the step at the same shape:
                # dram__bytes_read.sum 6 287 872 + dram__bytes_write.sum 140 961 280 (profiles/r03/r03_d_ncu_full_chain_raw.csv);
                # below the algorithmic 204.3 MB because the rest of the maps is still in the 126 MB L2 when the launch ends
                out[-1]['traffic'] = 6287872 + 140961280
                out[-1]['traffic_source'] = 'profiles/r03/r03_d_ncu_full_chain_raw.csv (ncu --set full, round 3, same shape, launch inside the critic step)'
                out[-1]['bound_note'] = ('per byte the kernel is HBM-class (25 flop/B) and is reported against HBM as the contract asks, but DRAM '
                                         'traffic is 147 MB in 0.24 ms = 0.6 TB/s: what limits it is the tcgen05.mma issue floor named in `limiter`')
            rows = slice(2 * B, 3 * B)
            limiter = ('tcgen05 operand reads from shared memory in the block-Toeplitz form: ~52 cycles per 128x(48|32|16)x16 MMA, 390 MMAs '
                       'per layer and CTA + ~5 k cycles of halo hand-over per layer (clock trace: profiles/r02_b_chain_trace.txt)')
            out[-1]['limiter'] = limiter
            dg_layers = [dict(w=cp['c.spec.cnn%d.kernel' % i], flip=True, aux=cpl['a'][i - 1][rows], y=cpl['da'][i - 1][rows]) for i in range(L - 1, 0, -1)]
            dg_layers.append(dict(w=cp['c.spec.cnn0.kernel'], flip=True, y=bf['g']))
            t = timed(lambda: ops.conv2d_chain(cpl['da'][L - 1][rows], dg_layers))
            nb = px * (4 + 8 * (L - 1) + 1)
            out.append({'name': 'conv2d_chain_dgrad_B', 'kernel': 'conv2d_chain_kernel<3>: data-gradient chain of the interpolated rows (%d layers 4->4 with LeakyReLU\' masks + 4->1 input gradient) on [B,T,%d,.], six products' % (L - 1, S),
                        'bound': 'hbm', 'achieved': nb / (t * 1e-3) / 1e9, 'peak': peak_hbm, 'unit': 'GB/s', 'frac': nb / (t * 1e-3) / 1e9 / peak_hbm,
                        'kernel_ms': t, 'bytes_per_launch': nb, 'traffic': None, 'launches_per_critic_step': 1,
                        'share_of_critic_step': t / critic_ms, 'layers_per_launch': L, 'us_per_layer': 1e3 * t / L, 'limiter': limiter,
                        'algorithmic': 'top gradient map + %d mask maps read, %d gradient maps + the 1-ch input gradient written' % (L - 1, L - 1)})
            # the first-order applications' data-gradient chain [2B] and the tangent sweep [B]: gradient-pass precision (three products)
            r2 = slice(0, 2 * B)
            dg2 = [dict(w=cp['c.spec.cnn%d.kernel' % i], flip=True, aux=cpl['a'][i - 1][r2], y=cpl['da'][i - 1][r2]) for i in range(L - 1, 0, -1)]
            tg = [dict(w=cp['c.spec.cnn%d.kernel' % i], aux=cpl['a'][i][rows], y=cpl['a'][i][rows]) for i in range(L)]
            with ops.backward_terms():
                t2 = timed(lambda: ops.conv2d_chain(cpl['da'][L - 1][r2], dg2))
                t3 = timed(lambda: ops.conv2d_chain(s0[rows], tg))
            nb2, nb3 = 2 * px * (4 + 8 * (L - 1)), px * (1 + 8 * L)
            parts_b = {6: 3, 3: 2, 1: 1}.get(ops.bwd_nterms, 0)
            out.append({'name': 'conv2d_chain_dgrad_2B', 'kernel': 'conv2d_chain_kernel<%d>: data-gradient chain of the real + fake rows on [2B,T,%d,4], side lane' % (parts_b, S),
                        'bound': 'hbm', 'achieved': nb2 / (t2 * 1e-3) / 1e9, 'peak': peak_hbm, 'unit': 'GB/s', 'frac': nb2 / (t2 * 1e-3) / 1e9 / peak_hbm,
                        'kernel_ms': t2, 'bytes_per_launch': nb2, 'traffic': None, 'launches_per_critic_step': 1, 'share_of_critic_step': t2 / critic_ms,
                        'algorithmic': 'top gradient map + %d mask maps read, %d gradient maps written, 2B samples' % (L - 1, L - 1)})
            out.append({'name': 'conv2d_chain_tangent_B', 'kernel': 'conv2d_chain_kernel<%d>: forward-mode sweep of the gradient penalty, in place over the stored activations of the interpolated rows' % parts_b,
                        'bound': 'hbm', 'achieved': nb3 / (t3 * 1e-3) / 1e9, 'peak': peak_hbm, 'unit': 'GB/s', 'frac': nb3 / (t3 * 1e-3) / 1e9 / peak_hbm,
                        'kernel_ms': t3, 'bytes_per_launch': nb3, 'traffic': None, 'launches_per_critic_step': 1, 'share_of_critic_step': t3 / critic_ms,
                        'algorithmic': 'seed map (1 ch) + %d activation maps read (masks), %d tangent maps written over them' % (L, L)})
            fam = [r for r in out if r['name'].startswith('conv2d_chain')]
            for r in fam:
                r['family_share_of_critic_step'] = sum(q['share_of_critic_step'] for q in fam)
        # -- one 4->4 layer on a B-sized map (the generator's stack: BatchNorm between the layers, one launch per layer)
        x1, y1 = gpl['cv_a'][1], gpl['cv_z'][2]
        t = timed(lambda: ops.conv2d(x1, gp['g.spec.cnn2.kernel'], y1))
        nb1 = 2.0 * x1.numel() * 4
        out.append({'name': 'conv2d_4to4_B', 'kernel': 'conv2d_tc_kernel on [B,T,%d,4]' % S, 'bound': 'hbm',
                    'achieved': nb1 / (t * 1e-3) / 1e9, 'peak': peak_hbm, 'unit': 'GB/s', 'frac': nb1 / (t * 1e-3) / 1e9 / peak_hbm,
                    'kernel_ms': t, 'bytes_per_launch': nb1, 'traffic': None, 'launches_per_critic_step': (L - 1) if st.C._chain(T) else 3 * (L - 1),
                    'share_of_critic_step': t * ((L - 1) if st.C._chain(T) else 3 * (L - 1)) / critic_ms,
                    'algorithmic': 'one fp32 map read + one written: 2 x B x T x %d x 4 ch x 4 B' % S})
        # -- weight + bias gradients of the critic's whole Conv2D stack in ONE launch (conv2d_wgrad44.cu): [2B] first-order rows on the
        #    side lane, [B] tangent rows accumulated on the main lane.  50 flop per byte read: bound by the fp32 FMA pipe, not by HBM
        #    (both fractions are given; `bound` keeps the contract's vocabulary)
        fma_peak = 148 * 128 * 1.965e9
        for tag, rws, nsm, first in (('2B', slice(0, 2 * B), 2 * B, True), ('B', slice(2 * B, 3 * B), B, False)):
            lay = [dict(x=(cpl['a'][i - 1] if i > 0 else bf['s0'].view(3 * B, T, S, 1))[rws], dy=cpl['da'][i][rws], dw=st.C.params.g['c.spec.cnn%d.kernel' % i],
                        dbias=st.C.params.g['c.spec.cnn%d.bias' % i] if first else None) for i in range(L)]
            if not ops.conv2d_wgrad_multi_ok(lay):
                continue
            t = timed(lambda: ops.conv2d_wgrad_multi(lay, accumulate=not first))
            nbw = nsm * T * S * 4.0 * (1 + 4 + (L - 1) * 8)
            fma = nsm * T * S * 400.0 * L
            out.append({'name': 'conv2d_wgrad_stack_%s' % tag, 'kernel': 'conv2d_wgrad44_kernel: weight (+ bias) gradients of all %d Conv2D layers on [%s,T,%d,.] in one launch' % (L, tag, S),
                        'bound': 'hbm', 'achieved': nbw / (t * 1e-3) / 1e9, 'peak': peak_hbm, 'unit': 'GB/s', 'frac': nbw / (t * 1e-3) / 1e9 / peak_hbm,
                        'kernel_ms': t, 'bytes_per_launch': nbw, 'traffic': None, 'launches_per_critic_step': 1, 'share_of_critic_step': t / critic_ms,
                        'fma_per_launch': fma, 'fma_frac_of_fp32_peak': fma / (t * 1e-3) / fma_peak,
                        'limiter': 'fp32 FMA pipe (CUDA cores): 400 FMA per pixel and layer against 32 bytes read; peak = 148 SMs x 128 FMA/clk x 1.965 GHz',
                        'algorithmic': 'every layer input map + every gradient map read once (1-ch input of the first layer)'})
        # -- Conv1D k=21 425->256 forward: virtual-im2col tcgen05 GEMM -- tensor bound
        kern, xp, z = gp['g.ctx.cnn0.kernel'], gpl['ctx_xp'][0], gpl['ctx_z'][0]
        t = timed(lambda: ops.conv1d(xp, kern, z, T), iters=20)
        fl = 2.0 * B * T * W * a.ctxsize * H
        ach = fl / (t * 1e-3) / 1e12
        out.append({'name': 'conv1d_fwd', 'kernel': 'gemm_tc_kernel<128,PARTS> + split-K reduce (Conv1D k=%d %d->%d forward, M=%d N=%d K=%d)' % (W, a.ctxsize, H, B * T, H, W * a.ctxsize),
                    'bound': 'tensor', 'achieved': ach, 'peak': peak_tf, 'unit': 'TFLOP/s', 'frac': ach / peak_tf, 'kernel_ms': t,
                    'flop_per_launch': fl, 'traffic': None, 'precision': ops.precision, 'tensor_products_per_flop': products,
                    'tensor_pipe_frac': products * ach / peak_tf, 'launches_per_critic_step': 2,
                    'share_of_critic_step': 2 * t / critic_ms,
                    'note': 'achieved counts ALGORITHMIC fp32 flops (2MNK); precision %s issues %d bf16 tcgen05 products per algorithmic '
                            'product, so the tensor pipe itself runs at tensor_pipe_frac of the measured bf16 peak' % (ops.precision, products)})
        # -- Conv1D weight gradient
        dz = gpl['dB']
        gkern = st.G.params.g['g.ctx.cnn0.kernel']
        t = timed(lambda: ops.conv1d_wgrad(xp, dz, gkern, T))
        with ops.backward_terms():      # the critic's ctx-branch weight gradient runs at the gradient-pass precision
            t_b = timed(lambda: ops.conv1d_wgrad(xp, dz, gkern, T))
        ach = fl / (t * 1e-3) / 1e12
        out.append({'name': 'conv1d_wgrad', 'kernel': 'gemm_tc_kernel<128,PARTS> tail-split (Conv1D weight gradient, M=%d N=%d K=%d)' % (W * a.ctxsize, H, B * T),
                    'bound': 'tensor', 'achieved': ach, 'peak': peak_tf, 'unit': 'TFLOP/s', 'frac': ach / peak_tf, 'kernel_ms': t,
                    'flop_per_launch': fl, 'traffic': None, 'tensor_pipe_frac': products * ach / peak_tf, 'launches_per_critic_step': 1,
                    'kernel_ms_gradient_pass_precision': t_b, 'gradient_pass_products': max(ops.bwd_nterms, 1),
                    'share_of_critic_step': t_b / critic_ms})
    del flush
    return out


def run_b200(args):
    import torch
    from percivaltts_b200 import modeltts_common, networks_critic, optimizertts_wgan, parallel, vocoders, _lib

    sync, device = parallel.init_from_env()
    rank, world = sync.rank, sync.world
    if world != args.gpus:
        raise SystemExit('--gpus %d but WORLD_SIZE=%d (launch with torchrun for N>1)' % (args.gpus, world))
    quiet = open(os.devnull, 'w')
    stdout, sys.stdout = sys.stdout, quiet             # the API prints its configuration like the reference does
    try:
        cfg = demo_cfg()
        voc = vocoders.VocoderPML(16000, 0.005, 129, 33)
        model = modeltts_common.DCNNF0SpecNoiseFeatures(425, voc, cfg)
        critic = networks_critic.Critic(voc, 425, cfg)
        opti = optimizertts_wgan.OptimizerTTSWGAN(cfg, model, errtype='WLSWGAN', critic=critic)
        opti.prepare(sync=sync if world > 1 else None)
    finally:
        sys.stdout = stdout
    tr, st, ops, lib = opti.trainer, opti.stepper, opti.stepper.ops, _lib.load()
    tr.generator_updates = opti.generator_updates = 25                 # steady-state schedule: critic_runs == 5
    B, T = B_PER_GPU, T_FRAMES
    if args.global_batch:
        if args.global_batch % world:
            raise SystemExit('--global-batch %d does not split evenly over %d ranks' % (args.global_batch, world))
        B = args.global_batch // world
    R = 4
    host = [synthetic_batch(B, T, 123 + 1000 * rank + i) for i in range(R)]
    s = tr.shape_state(B, T)                                             # eager warm-up + CUDA-graph capture
    dev = [(torch.from_numpy(x).to(device), torch.from_numpy(y).to(device)) for x, y in host]
    # kernels per step = launches the library records while the same schedule runs eagerly once
    n1 = lib.pctts_launch_count()
    st.critic_step(s.X, s.Y, apply=True)
    n_critic = lib.pctts_launch_count() - n1
    n2 = lib.pctts_launch_count()
    st.generator_step(s.X, s.Y, apply=True)
    n_gen = lib.pctts_launch_count() - n2
    n3 = lib.pctts_launch_count()
    if tr.fuse_steps:
        st.critic_generator_step(s.X, s.Y)
    n_both = (lib.pctts_launch_count() - n3) if tr.fuse_steps else n_critic + n_gen
    torch.cuda.synchronize(device)

    def resident_step(i):
        x, y = dev[i % R]
        s.X.copy_(x, non_blocking=True)
        s.Y.copy_(y, non_blocking=True)
        tr.step_device(s, i % 5 == 0)       # what train_on_batch runs: critic update, + generator update on the same batch every 5th step

    # ---- value: device-resident loop ---------------------------------------------------------------------------------
    # the sampler starts in front of the warm-up: nvidia-smi's own start-up (NVML initialisation takes driver locks and was seen
    # to stall the launching thread for 100s of ms) then falls into the untimed steps, its polling into the timed ones
    clocks = ClockSampler(device.index or 0)
    if rank == 0:
        clocks.start()
        time.sleep(0.5)
    for i in range(args.warmup):
        resident_step(i)
    torch.cuda.synchronize(device)
    barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(device)
    e0.record()
    for i in range(args.steps):
        resident_step(i)
    e1.record()
    torch.cuda.synchronize(device)
    barrier()
    ms = max_over_ranks(e0.elapsed_time(e1), device)
    frames = args.steps * B * T * world
    value = frames / (ms * 1e-3)

    # ---- critic-only / generator-only step times (BASELINE metric "critic step ms") -------------------------------------
    def loop_ms(fn, n):
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize(device)
        a0.record()
        for _ in range(n):
            fn()
        a1.record()
        torch.cuda.synchronize(device)
        return a0.elapsed_time(a1) / n
    critic_ms = max_over_ranks(loop_ms(lambda: tr.critic_step_device(s), 10), device)
    gen_ms = max_over_ranks(loop_ms(lambda: tr.generator_step_device(s), 5), device)
    both_ms = max_over_ranks(loop_ms(lambda: tr.step_device(s, True), 5), device)      # critic + generator update of one batch, one graph
    tr.generator_updates = 25

    # ---- e2e: the public API with host buffers --------------------------------------------------------------------------
    for i in range(args.warmup):      # same call as the timed loop: the prefetch path allocates its pinned / device staging buffers
        opti.train_on_batch(i + 1, host[i % R][0], host[i % R][1], next_batch=host[(i + 1) % R])     # and copy stream on first use
    tr.generator_updates = opti.generator_updates = 25
    tr.h2d_bytes = tr.d2h_bytes = 0
    torch.cuda.synchronize(device)
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        # the base loop's call (optimizertts.py): this batch plus the already loaded next one, whose H2D overlaps this step
        opti.train_on_batch(i, host[i % R][0], host[i % R][1], next_batch=host[(i + 1) % R])
    torch.cuda.synchronize(device)
    e2e_s = time.perf_counter() - t0
    barrier()
    e2e_s = max_over_ranks(e2e_s, device)
    e2e_h2d, e2e_d2h = tr.h2d_bytes // args.steps, tr.d2h_bytes // args.steps
    # the same loop through the reference's 3-argument hook (no next_batch hint): synchronous staging of every batch
    tr.generator_updates = opti.generator_updates = 25
    torch.cuda.synchronize(device)
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        opti.train_on_batch(i, host[i % R][0], host[i % R][1])
    torch.cuda.synchronize(device)
    e2e_sync_s = max_over_ranks(time.perf_counter() - t0, device)
    barrier()
    clk = clocks.stop() if rank == 0 else None
    e2e_value = frames / e2e_s

    peaks = {}
    try:
        with open(os.path.join(ROOT, 'MEASURED_PEAKS.json')) as f:
            peaks = json.load(f)
    except (OSError, ValueError):
        pass
    peak_tf = float(peaks.get('bf16_tflops', 1590.0))
    peak_hbm = float(peaks.get('hbm_gbs', 6650.0))
    peak_src = 'MEASURED_PEAKS.json (burst: each kernel is timed alone)' if peaks else 'B200_PROFILING.md fallback (1590 TFLOP/s, 6650 GB/s)'
    rooflines = kernel_rooflines(st, ops, B, T, device, peak_tf, peak_hbm, critic_ms, gen_ms) if not args.no_extras else []
    for r in rooflines:
        r['peak_source'] = peak_src
    # `roofline` = the kernel with the largest measured share of the critic step; the others follow in `rooflines`
    roofline = max(rooflines, key=lambda r: r.get('share_of_critic_step', 0.0)) if rooflines else None
    flops = step_min_flops(st.a, B, T)
    whole = {'variant': 'minimum work: critic ctx branch once per batch, frozen generator spectral branch only in the critic step '
                        '(SURVEY.md 8d); algorithmic 2MNK flops of every contraction, forward + data/weight gradients + GP tangent sweep',
             'critic_step_gflop': flops['critic'] / 1e9, 'generator_step_gflop': flops['generator'] / 1e9,
             'critic_step_tflops': flops['critic'] / (critic_ms * 1e-3) / 1e12, 'generator_step_tflops': flops['generator'] / (gen_ms * 1e-3) / 1e12,
             'critic_step_frac_of_bf16_peak': flops['critic'] / (critic_ms * 1e-3) / 1e12 / peak_tf,
             'generator_step_frac_of_bf16_peak': flops['generator'] / (gen_ms * 1e-3) / 1e12 / peak_tf}
    conv1d = next((r for r in rooflines if r['name'] == 'conv1d_fwd'), None)

    if rank != 0:
        return
    ngen_steps = (args.steps + 4) // 5
    launches = (args.steps - ngen_steps) * n_critic + ngen_steps * n_both
    out = {
        'metric': METRIC, 'value': value, 'unit': UNIT, 'n_gpus': world, 'steps': args.steps, 'warmup': args.warmup,
        'ms_per_step': ms / args.steps, 'higher_is_better': True, 'scaling': 'strong' if args.global_batch else 'weak', 'vs_baseline': None,
        'dtype': 'bf16' if ops.precision == 'bf16' else 'f32',
        'data': 'synthetic',
        'config': bench_config(world, B, ops.precision, args.global_batch),
        'schedule_note': 'on the steps with a generator update both updates run as one graph sharing the generator forward pass of the batch',
        # disclosure: what "f32" means here -- forward contractions and the penalty's input gradient use three bf16 planes / six
        # products (fp32-grade); the critic's other gradient passes two planes / three products
        'gradient_pass_products': int(max(ops.bwd_nterms, 1)), 'forward_products': int(max(ops.nterms, 1)),
        'precision_note': 'PCTTS_BWD_TERMS=6 restores six products in every pass; worst gradient error against the fp64 oracle is '
                          'printed by __graft_entry__.smoke() (2.4e-5 at three products, bar 1e-3)',
        'critic_step_ms': critic_ms, 'generator_step_ms': gen_ms, 'critic_plus_generator_step_ms': both_ms,
        'conv_tc_util_pct': 100.0 * conv1d['tensor_pipe_frac'] if conv1d else None, 'gemm_precision': ops.precision,
        'e2e': {'value': e2e_value, 'unit': UNIT, 'h2d_bytes_per_step': e2e_h2d,
                'd2h_bytes_per_step': e2e_d2h, 'ms_per_step': 1e3 * e2e_s / args.steps,
                'call': 'OptimizerTTSWGAN.train_on_batch(batchid, X, Y, next_batch=(X1, Y1)) -- the package\'s own epoch loop'},
        'e2e_sync': {'value': frames / e2e_sync_s, 'unit': UNIT, 'ms_per_step': 1e3 * e2e_sync_s / args.steps,
                     'call': 'OptimizerTTSWGAN.train_on_batch(batchid, X, Y) -- the reference\'s 3-argument hook, synchronous staging'},
        'gpu_launches': int(launches), 'kernels_per_critic_step': int(n_critic), 'kernels_per_generator_step': int(n_gen),
        'kernels_per_critic_plus_generator_step': int(n_both),
        'clocks': clk, 'roofline': roofline, 'rooflines': rooflines, 'whole_step': whole,
    }
    if world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        v, dt = cpu_baseline_cycle(threads)
        out['cpu_baseline'] = {'value': v, 'unit': UNIT, 'cores': threads, 'kind': 'port',
                               'sample': 'one steady-state cycle (5 batches of B=10 x T=400: 5 critic + 1 generator update) of the '
                                         'torch-CPU fp32 oracle, %.1f s' % dt}
    if world == 1 and not args.no_extras:
        # BASELINE configs #4 and #5 (their own workloads; `value` above stays config #2)
        sys.path.insert(0, os.path.join(ROOT, 'tools'))
        import bench_gp
        import bench_pml
        import bench_files
        out['e2e_files'] = bench_files.run(demo_cfg())     # the same step driven by train() on files (loader thread + prefetch)
        out['config4_gp_microbench'] = bench_gp.run(ops, iters=6)
        out['config5_pml_1M_frames'] = bench_pml.run(ops, cpu_frames=4000)
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=50)
    ap.add_argument('--warmup', type=int, default=10)
    ap.add_argument('--impl', default='b200', choices=['b200', 'reference'])
    ap.add_argument('--no-cpu-baseline', action='store_true')
    ap.add_argument('--precision', default=None, choices=['fp32', 'fp32x3', 'bf16', 'simt'],
                    help='GEMM arithmetic (default fp32 = config #2; bf16 = BASELINE config #3)')
    ap.add_argument('--global-batch', type=int, default=None,
                    help='strong scaling: total utterances per step, split evenly over the ranks (default: 10 per GPU, weak scaling)')
    ap.add_argument('--no-extras', action='store_true', help='skip the config #4 / #5 microbenches and the per-kernel rooflines')
    args = ap.parse_args()
    if args.precision:
        os.environ['PCTTS_PRECISION'] = args.precision
    args.warmup = max(args.warmup, 3)
    if args.impl == 'reference':
        run_reference(args)
    else:
        run_b200(args)


if __name__ == '__main__':
    main()
