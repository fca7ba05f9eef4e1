"""All-reduce timing of the step's gradient buckets (run under torchrun): CUDA events, max over ranks.

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/bench_allreduce.py
"""
import os

import torch
import torch.distributed as dist

local = int(os.environ.get('LOCAL_RANK', 0))
torch.cuda.set_device(local)
dist.init_process_group('nccl', device_id=torch.device('cuda', local))
world = dist.get_world_size()
for n in (47_744, 5_382_144, 8_252_992, 12_258_208, 31_232_224):
    x = torch.ones(n, device='cuda')
    for _ in range(5):
        dist.all_reduce(x)
    torch.cuda.synchronize()
    dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        dist.all_reduce(x)
    e1.record()
    e1.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / 20], device='cuda')
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if dist.get_rank() == 0:
        ms = float(t)
        print('world %d  all_reduce fp32 x %9d (%6.1f MB): %7.3f ms  algbw %6.1f GB/s  busbw %6.1f GB/s'
              % (world, n, n * 4 / 1e6, ms, n * 4 / ms / 1e6, n * 4 / ms / 1e6 * 2 * (world - 1) / world), flush=True)
dist.barrier()
torch.cuda.synchronize()
os._exit(0)
