"""Does the symmetric-memory allocator hand out an NVSwitch multicast mapping on this box?  (torchrun, >= 2 ranks)"""
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ['LLMT_COMM'] = 'nvlink'
rank = int(os.environ['RANK'])
torch.cuda.set_device(int(os.environ['LOCAL_RANK']))
dist.init_process_group('nccl', device_id=torch.device('cuda', int(os.environ['LOCAL_RANK'])))
from llm_training_b200.parallel import init_mesh  # noqa: E402
from llm_training_b200.parallel.comm import get_comm  # noqa: E402
comm = get_comm(init_mesh(dist.get_world_size(), 1).dp)
buf = comm.alloc(1 << 20, torch.bfloat16)
print(f'rank {rank}: multicast_ptr = {buf.multicast_ptr:#x}', flush=True)
dist.barrier()
dist.destroy_process_group()
