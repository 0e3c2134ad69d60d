"""Developer probe: cycles per tcgen05.mma (M=128, K=16, bf16) for the shared-memory layouts of the A operand and several N."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from percivaltts_b200.ops import CudaOps
ops = CudaOps()
out = torch.zeros(2, dtype=torch.int64, device='cuda')
names = {0: 'chain form (2 parity arrays, no swizzle)', 1: 'SWIZZLE_32B', 2: 'SWIZZLE_64B', 3: 'SWIZZLE_128B', 4: 'no swizzle, K chunks 2 KB apart'}
nm = 400
for layout in (0, 4, 1, 2, 3):
    for shift in (0, 33):
        row = []
        for N in (16, 32, 48, 96, 128, 256):
            for _ in range(2):
                rc = ops.lib.pctts_umma_probe(layout, N, nm, shift, out.data_ptr(), None)
                torch.cuda.synchronize()
            row.append('N=%d: %.1f' % (N, float(out[1]) / nm))
        print('%-42s shift %2d rows | cycles per MMA  %s' % (names[layout], shift, '  '.join(row)), flush=True)
