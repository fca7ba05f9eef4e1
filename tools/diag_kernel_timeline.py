"""In-graph duration of every primitive call of one training step (GPU-timer stamps between the calls, captured into
the step's CUDA graph): what each kernel costs where it runs - warm L2, no host gaps - as opposed to the cold, serialised
ncu launch list.  Side streams are switched off so that consecutive stamps bracket exactly one call.

    python tools/diag_kernel_timeline.py [precision]      -> table by primitive, gpurun_out/kernel_timeline_<precision>.json
"""
import collections
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from vae_npvc_b200 import _C, synth  # noqa: E402
from vae_npvc_b200 import prims as Pm  # noqa: E402
from vae_npvc_b200.trainer.basic import Trainer  # noqa: E402

prec = sys.argv[1] if len(sys.argv) > 1 else 'bf16x2'
arch = synth.arch_config(X=80, C=512, D=64, K=512, Y=128, S=128, y_num=117, stacks=10, use_ema=False, embed_norm=True)
cfg = dict(arch, model_type='vae_npvc_b200.model.vqvae', precision=prec, cuda_graph=True, graph_warmup=2, wgrad_stream=False,
           bank_overlap=False, learning_rate=1e-3, max_grad_norm=10)
tr = Trainer(cfg)
tr.model.load_state_dict(synth.synthetic_state_dict(arch, 777))
x, y = synth.synthetic_batch(arch, 16, 128, 777)
x, y = x.cuda(), y.cuda()
NSLOT = 8192
buf = torch.zeros(NSLOT, dtype=torch.int64, device='cuda')
calls = []
state = {'on': False, 'slot': 0}
NAMES = [n for n in dir(Pm) if callable(getattr(Pm, n)) and not n.startswith('_') and n not in (
    'as_cl', 'cl_empty', 'cl_zeros', 'is_cl', 'to_nct', 'pitch', 'Planes', 'ConvWeights')]


def wrap(name, fn):
    def inner(*a, **k):
        out = fn(*a, **k)
        if state['on'] and state['slot'] < NSLOT:
            tag = name
            if name in ('conv_fwd', 'conv_wgrad'):
                tag = '%s taps%s' % (name, a[2])
            _C.call('debug_stamp', buf, state['slot'])
            calls.append((state['slot'], tag))
            state['slot'] += 1
        return out
    return inner


for n in NAMES:
    f = getattr(Pm, n)
    if isinstance(f, type):
        continue
    setattr(Pm, n, wrap(n, f))
for _ in range(2):
    tr.train_step_device(x, y)                 # eager warm-up, stamps off
state['on'] = True
_C.call('debug_stamp', buf, 0)
state['slot'] = 1
calls.append((0, 'start'))
tr.train_step_device(x, y)                     # this call captures the graph (python runs once, stamps are captured)
state['on'] = False
for _ in range(5):
    tr.train_step_device(x, y)                 # replays
torch.cuda.synchronize()
assert tr.last_step_on_graph
t = buf.cpu().tolist()
agg = collections.OrderedDict()
prev = t[0]
for slot, tag in calls[1:]:
    d = (t[slot] - prev) * 1e-3
    prev = t[slot]
    a = agg.setdefault(tag, [0, 0.0])
    a[0] += 1
    a[1] += d
# calibration: what one stamp itself costs
c0 = torch.zeros(64, dtype=torch.int64, device='cuda')
g = torch.cuda.CUDAGraph()
with torch.cuda.graph(g):
    for i in range(32):
        _C.call('debug_stamp', c0, i)
g.replay()
torch.cuda.synchronize()
cc = c0.cpu().tolist()
stamp_us = (cc[31] - cc[0]) * 1e-3 / 31
total = sum(v[1] for v in agg.values())
print('precision %s: %d stamped calls, %.1f us between first and last stamp; one stamp alone costs %.2f us (included in every row)'
      % (prec, len(calls) - 1, total, stamp_us))
rows = sorted(agg.items(), key=lambda kv: -kv[1][1])
for tag, (n, us) in rows:
    print('%-28s n=%4d  total %8.1f us  avg %6.2f us  (net of stamp %6.2f)  %5.1f %%' % (tag, n, us, us / n, us / n - stamp_us, 100 * us / total))
os.makedirs(os.path.join(ROOT, 'gpurun_out'), exist_ok=True)
json.dump({'precision': prec, 'stamp_us': stamp_us, 'rows': [[k, v[0], v[1]] for k, v in rows]},
          open(os.path.join(ROOT, 'gpurun_out', 'kernel_timeline_%s.json' % prec), 'w'))
