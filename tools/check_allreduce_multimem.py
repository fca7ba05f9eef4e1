"""Multi-GPU check of csrc/allreduce.cu against NCCL (run under torchrun, >= 2 GPUs on one NVSwitch node).

    python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/check_allreduce_multimem.py

Fills a symmetric fp32 buffer with rank-dependent data, reduces several (offset, count) slices with
vnb_allreduce_multimem (mean over ranks) and compares with NCCL all_reduce / world on a copy: bit-exact at
world 2 (one addition), <= 1e-6 relative otherwise (the switch's summation order is its own).  Also times the
31.2 M-element buffer.  Prints 'allreduce_multimem ok' on rank 0 and exits 0; any mismatch exits 1.
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import torch.distributed._symmetric_memory as symm  # noqa: E402

from vae_npvc_b200 import _C  # noqa: E402

local = int(os.environ.get('LOCAL_RANK', 0))
torch.cuda.set_device(local)
dev = torch.device('cuda', local)
dist.init_process_group('nccl', device_id=dev)
rank, world = dist.get_rank(), dist.get_world_size()
n = 31_232_224
buf = symm.empty(n, dtype=torch.float32, device=dev)
hb = symm.rendezvous(buf, dist.group.WORLD)
flags = symm.empty(128 * world, dtype=torch.int32, device=dev)
flags.zero_()
hf = symm.rendezvous(flags, dist.group.WORLD)
mc = int(hb.multicast_ptr)
if rank == 0:
    print('world %d multicast_ptr %#x' % (world, mc), flush=True)
if mc == 0:
    print('no multicast mapping on this machine')
    os._exit(2)
flag_ptrs = torch.tensor([int(p) for p in hf.buffer_ptrs], dtype=torch.int64, device=dev)
g = torch.Generator(device=dev).manual_seed(1234 + rank)
src = torch.randn(n, device=dev, generator=g)
dist.barrier()
torch.cuda.synchronize()
bad = 0
for nblocks in (1, 32):
    for (off, cnt) in ((0, n), (18926272, 12258208), (4, 8), (1024, 4 * 777), (5382144, 5291136), (n - 4, 4)):
        buf.copy_(src)
        ref = src.clone()
        dist.all_reduce(ref[off:off + cnt])
        ref[off:off + cnt] /= world
        _C.call('allreduce_multimem', mc, off, cnt, flag_ptrs, rank, world, 1.0 / world, nblocks)
        torch.cuda.synchronize()
        if world == 2:
            same = torch.equal(buf, ref)
        else:
            same = torch.allclose(buf, ref, rtol=1e-6, atol=1e-6)
        err = float((buf - ref).abs().max())
        # every rank must hold the same bits
        chk = buf[off:off + cnt].double().sum().reshape(1)
        lo_, hi_ = chk.clone(), chk.clone()
        dist.all_reduce(lo_, op=dist.ReduceOp.MIN)
        dist.all_reduce(hi_, op=dist.ReduceOp.MAX)
        ident = bool(lo_ == hi_)
        if rank == 0:
            print('nblocks %2d offset %9d count %9d: max |diff| %.3e  %s  replicas identical: %s'
                  % (nblocks, off, cnt, err, 'ok' if same else 'MISMATCH', ident), flush=True)
        if not same or not ident:
            bad += 1
for nblocks in (8, 16, 32, 64):
    dist.barrier()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        _C.call('allreduce_multimem', mc, 0, n, flag_ptrs, rank, world, 1.0, nblocks)
    e1.record()
    e1.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / 20], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        ms = float(t)
        print('nblocks %2d: %d fp32 in %.3f ms  algbw %.1f GB/s' % (nblocks, n, ms, n * 4 / ms / 1e6), flush=True)
tot = torch.tensor([bad], device=dev)
dist.all_reduce(tot)
if rank == 0:
    print('allreduce_multimem ok' if int(tot) == 0 else 'allreduce_multimem FAILED', flush=True)
dist.barrier()
torch.cuda.synchronize()
sys.stdout.flush()
os._exit(0 if int(tot) == 0 else 1)
