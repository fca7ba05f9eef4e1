"""Conversion / token-extraction throughput at the VCC2020 shape (BASELINE config 5's
"batched conversion inference"): frames/s end to end (host numpy features in, host numpy features out), for

  loop    the reference's calling pattern on the drop-in Model: model.infer at batch 1 per utterance
          (vae_npvc/decoder/basic.py:58-75), weight-norm re-derived per conv per call
  single  ConvertEngine, one utterance per batch (frozen operand planes only)
  packed  ConvertEngine, length-sorted zero-padded batches (masked GroupNorm statistics)
  cpu     the oracle (reference restatement, torch CPU fp32) on a bounded sample of the same utterances

    python tools/bench_convert.py [--utts 96] [--precision bf16] [--out gpurun_out/convert.json]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--utts', type=int, default=96)
    ap.add_argument('--precision', default='bf16')
    ap.add_argument('--min-len', type=int, default=120)
    ap.add_argument('--max-len', type=int, default=640)
    ap.add_argument('--cpu-utts', type=int, default=3)
    ap.add_argument('--out', default='')
    a = ap.parse_args()
    from vae_npvc_b200 import synth
    from oracle import vqvae_oracle as O
    import vae_npvc_b200.functional as Fn
    from vae_npvc_b200 import _C
    from vae_npvc_b200.convert import ConvertEngine
    from vae_npvc_b200.model.vqvae import Model
    _C.require_cuda()
    Fn.set_precision(a.precision)
    arch = synth.arch_config()
    sd = synth.synthetic_state_dict(arch, 777)
    m = Model(arch)
    m.load_state_dict(sd)
    m = m.cuda().eval()
    rng = np.random.RandomState(0)
    lengths = [int(v) for v in rng.randint(a.min_len, a.max_len + 1, size=a.utts)]
    gen = torch.Generator().manual_seed(1)
    feats = [torch.randn(n, 80, generator=gen).numpy() for n in lengths]
    targets = [i % arch['y_num'] for i in range(a.utts)]
    frames = sum(lengths)

    def timed(fn, reps=3):
        fn()                                                   # warm-up (allocator, first-use module loads)
        best = float('inf')
        for _ in range(reps):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            fn()
            torch.cuda.synchronize()
            best = min(best, time.perf_counter() - t0)
        return best

    def loop():
        out = []
        with torch.no_grad():
            for f, t in zip(feats, targets):
                x = torch.from_numpy(f).float().cuda().t().unsqueeze(0)
                y = torch.tensor([t]).long().cuda().view(1, -1)
                out.append(m.infer((x, y))[0].t().detach().cpu().numpy())
        return out

    res = {'workload': 'VCC2020-shape VQ-VAE conversion, %d utterances of %d-%d frames (%d frames), %s operands'
                       % (a.utts, a.min_len, a.max_len, frames, a.precision), 'frames': frames}
    t = timed(loop)
    res['loop_model_infer_batch1'] = {'s': t, 'frames_per_s': frames / t}
    single = ConvertEngine(m, precision=a.precision, max_batch=1)
    packed = ConvertEngine(m, precision=a.precision, max_batch=64, max_batch_frames=1 << 15, max_pad_frac=0.25)
    for name, eng in (('single', single), ('packed', packed)):
        eng.launches = 0
        t = timed(lambda: eng.convert(feats, targets))
        res['engine_' + name] = {'s': t, 'frames_per_s': frames / t, 'launches_per_pass': eng.launches // 4}
        t = timed(lambda: eng.encode(feats, 'csid'))
        res['engine_' + name + '_csid'] = {'s': t, 'frames_per_s': frames / t}
    from vae_npvc_b200.convert import plan_batches
    plan = plan_batches(lengths, 64, 1 << 15, 0.25)
    padded = sum(len(b) * lengths[b[0]] for b in plan)
    res['packing'] = {'batches': len(plan), 'padded_frames': padded, 'pad_fraction': 1.0 - frames / padded}
    a_, b_ = packed.convert(feats, targets), single.convert(feats, targets)
    res['packed_vs_single_max_rel'] = max(float(np.linalg.norm(x - y) / np.linalg.norm(y)) for x, y in zip(a_, b_))
    # CPU baseline: the oracle per utterance (the reference's loop) on a bounded sample
    params = O.params_from_state_dict(sd, requires_grad=False)
    torch.set_num_threads(os.cpu_count() or 1)
    k = min(a.cpu_utts, a.utts)
    t0 = time.perf_counter()
    with torch.no_grad():
        for f, tg in list(zip(feats, targets))[:k]:
            O.model_infer(params, arch, torch.from_numpy(f).t().unsqueeze(0), torch.tensor([[tg]]))
    dt = time.perf_counter() - t0
    res['cpu_oracle_loop'] = {'s': dt, 'frames_per_s': sum(lengths[:k]) / dt, 'cores': os.cpu_count(),
                              'sample': '%d utterances (%d frames)' % (k, sum(lengths[:k]))}
    line = json.dumps(res)
    print(line)
    if a.out:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
        with open(a.out, 'w') as f:
            f.write(line + '\n')


if __name__ == '__main__':
    main()
