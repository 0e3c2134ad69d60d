"""Developer probe: time line of one eager critic / generator step at the demo size.  A %globaltimer stamp is launched behind every
C-ABI call on the call's own stream (ops.timeline_*), so the listing shows when each kernel FINISHED in the concurrent
three-lane execution (the stamps add ~2 us of launch each; the shape of the schedule is what matters).
python tools/timeline.py [critic|generator]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from percivaltts_b200 import arch as A
from percivaltts_b200.ops import CudaOps
from percivaltts_b200.wgan import WGANStepper

which = sys.argv[1] if len(sys.argv) > 1 else 'critic'
B, T = int(os.environ.get('B', 10)), int(os.environ.get('T', 400))
sync = None
if int(os.environ.get('WORLD_SIZE', '1')) > 1:       # torchrun: the data-parallel step (rank 0 prints)
    from percivaltts_b200 import parallel
    sync, _ = parallel.init_from_env()
ops = CudaOps()
a = A.Arch()
st = WGANStepper(ops, a, w_ls=np.ones(a.outsize), w_wgan_mean=0.5, sync=sync)
st.G.params.load(A.init_params(A.generator_param_shapes(a), 1))
st.C.params.load(A.init_params(A.critic_param_shapes(a), 2))
rng = np.random.RandomState(0)
X = torch.tensor(rng.normal(size=(B, T, a.ctxsize)).astype(np.float32)).cuda()
Y = torch.tensor(rng.normal(size=(B, T, a.outsize)).astype(np.float32)).cuda()
step = st.critic_step if which == 'critic' else st.generator_step
for _ in range(3):
    step(X, Y)
torch.cuda.synchronize()
# eager launches would be CPU-bound (the stamps show launch gaps, not the schedule): hold the GPU back until everything is queued
ops.timeline_start()
hold = torch.cuda.Event()
torch.cuda._sleep(int(4e7))          # ~20 ms of spin on the main stream while the host queues the step
step(X, Y)
ops.stamp('end')
tl = ops.timeline_stop()
if sync is not None and sync.rank != 0:
    sys.stdout.flush()
    os._exit(0)
t0 = min(t for w, l, t in tl[1:])
last = {}
print('%-4s %-26s %4s %9s %8s' % ('#', 'call', 'lane', 'end us', 'dur us'))
for i, (w, l, t) in enumerate(tl[1:]):
    e = (t - t0) / 1e3
    d = e - last.get(l, 0.0)
    last[l] = e
    print('%-4d %-26s %4d %9.1f %8.1f' % (i, w.replace('pctts_', ''), l, e, d))
if sync is not None:
    sys.stdout.flush()
    os._exit(0)
