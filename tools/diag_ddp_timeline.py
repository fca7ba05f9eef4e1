"""Where the data-parallel step's time goes: GPU-timer stamps (vnb_debug_stamp) at the milestones of a REPLAYED step.

    torchrun --nproc-per-node N tools/diag_ddp_timeline.py [precision]        # or plain python for one GPU
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

import bench  # noqa: E402
from vae_npvc_b200 import synth  # noqa: E402
from vae_npvc_b200.trainer.basic import Trainer  # noqa: E402

rank, world = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1))
prec = sys.argv[1] if len(sys.argv) > 1 else 'bf16x2'
wl = bench._wl('vcc2020')
cfg = dict(wl['arch'], model_type='vae_npvc_b200.model.vqvae', precision=prec, cuda_graph=True, graph_warmup=3,
           learning_rate=1e-3, max_grad_norm=10, debug_stamps=True,
           allreduce_buckets=tuple(int(v) for v in os.environ.get('VNB_BUCKETS', '3,3').split(',')))
tr = Trainer(cfg)
tr.model.load_state_dict(synth.synthetic_state_dict(wl['arch'], 777))
x, y = synth.synthetic_batch(wl['arch'], wl['B'], wl['T'], 777 + rank)
x, y = x.cuda(), y.cuda()
for _ in range(12):
    tr.train_step_device(x, y)
torch.cuda.synchronize()
st = tr._stamps.cpu().tolist()
t0 = st[0]
names = {0: 'step start', 1: 'forward issued/done', 2: 'backward done (main stream)', 3: 'collectives joined',
         4: 'weight-norm backward done', 5: 'optimiser done'}
if rank == 0:
    print('precision %s, world %d, buckets %s' % (prec, world, sorted(tr._buckets.items(), key=lambda kv: -kv[1][0])))
    for k in (1, 2, 3, 4, 5):
        print('%-30s +%8.1f us' % (names[k], (st[k] - t0) / 1e3))
    k = 10
    while k < 62 and st[k]:
        print('all-reduce #%d                  start +%8.1f us   end +%8.1f us   (%.1f us)'
              % ((k - 10) // 2, (st[k] - t0) / 1e3, (st[k + 1] - t0) / 1e3, (st[k + 1] - st[k]) / 1e3))
        k += 2
if world > 1:
    import torch.distributed as dist
    dist.barrier()
    torch.cuda.synchronize()
    sys.stdout.flush()
    os._exit(0)
