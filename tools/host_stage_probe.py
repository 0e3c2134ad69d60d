"""Developer probe: cost of the host side of the e2e path (numpy batch -> pinned staging buffer), per call, in this process."""
import time, os, sys
import numpy as np
import torch
torch.cuda.init()
B, T = 10, 400
host = [(np.random.rand(B, T, 425).astype(np.float32), np.random.rand(B, T, 163).astype(np.float32)) for _ in range(4)]
Xh = torch.empty(B * T * 425, dtype=torch.float32, pin_memory=True)
Yh = torch.empty(B * T * 163, dtype=torch.float32, pin_memory=True)
ts = []
for i in range(300):
    X, Y = host[i % 4]
    t0 = time.perf_counter()
    Xh.copy_(torch.from_numpy(X.reshape(-1)))
    Yh.copy_(torch.from_numpy(Y.reshape(-1)))
    ts.append(time.perf_counter() - t0)
    time.sleep(0.002)
ts = np.array(ts[20:]) * 1e3
print('threads %d  affinity %d cpus  staging ms: median %.3f  mean %.3f  p95 %.3f  max %.3f' % (
    torch.get_num_threads(), len(os.sched_getaffinity(0)), np.median(ts), ts.mean(), np.percentile(ts, 95), ts.max()))
