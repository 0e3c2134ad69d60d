"""Developer probe: can this box map a symmetric (peer-addressable) buffer across the ranks of one node?"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem

rank, world, lr = int(os.environ['RANK']), int(os.environ['WORLD_SIZE']), int(os.environ['LOCAL_RANK'])
torch.cuda.set_device(lr)
dev = torch.device('cuda', lr)
dist.init_process_group('nccl', device_id=dev)
t = symm_mem.empty(4096, dtype=torch.float64, device=dev)
h = symm_mem.rendezvous(t, dist.group.WORLD.group_name)
t.fill_(float(rank + 1))
torch.cuda.synchronize()
dist.barrier()
print('rank', rank, 'world', h.world_size, 'buffer_ptrs', [hex(p) for p in h.buffer_ptrs], 'signal_pads', [hex(p) for p in h.signal_pad_ptrs],
      'pad size', getattr(h, 'signal_pad_size', None), 'buffer size', getattr(h, 'buffer_size', None), flush=True)
peer = h.get_buffer((rank + 1) % world, (8,), torch.float64)
print('rank', rank, 'reads peer', peer.tolist()[:2], flush=True)
dist.barrier()
dist.destroy_process_group()
