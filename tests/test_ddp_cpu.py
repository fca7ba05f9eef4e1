"""world_size-2 `gloo` test of the batch-sharded data-parallel step on CPU.

Each rank runs the REAL Trainer (flat buffers, weight bank, gradient all-reduce, EMA-statistics
all-reduce) with the CUDA primitives replaced by their torch-CPU contract restatement, on its half of
a global batch.  Because every rank normalises its loss by its own B*T and shards are equal, the
AVERAGED gradient equals the single-process gradient of the global batch, so after N steps the
parameters must match a 1-process run on the full batch (plain VQ: to fp32 round-off; EMA: the
all-reduced code sums / counts give every rank the same codebook as the 1-process run).
"""
import os
import socket
import sys

import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_port():
    s = socket.socket()
    s.bind(('127.0.0.1', 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _make_trainer(arch, model_type='vae_npvc_b200.model.vqvae'):
    from oracle import prims_ref
    import vae_npvc_b200.functional as Fn
    import vae_npvc_b200.model.layers_vq as LVQ
    import vae_npvc_b200.trainer.basic as TB
    Fn.Pm = LVQ.Pm = TB.Pm = prims_ref
    Fn.set_precision('fp32')
    cfg = dict(arch)
    cfg.update(model_type=model_type, device='cpu', cuda_graph=False, precision='fp32',
               learning_rate=1e-3, max_grad_norm=10)
    return TB.Trainer(cfg)


def _run(rank, world, port, use_ema, steps, out_dir, kind='vqvae'):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from oracle import synth
    torch.set_num_threads(1)
    if world > 1:
        os.environ['MASTER_ADDR'], os.environ['MASTER_PORT'] = '127.0.0.1', str(port)
        dist.init_process_group('gloo', rank=rank, world_size=world)
    v2 = kind == 'vqvae2'
    arch = synth.vqvae2_small_config() if v2 else synth.small_config(use_ema=use_ema)
    torch.manual_seed(777)
    tr = _make_trainer(arch, 'vae_npvc_b200.model.vqvae2' if v2 else 'vae_npvc_b200.model.vqvae')
    tr.model.load_state_dict(synth.vqvae2_state_dict(arch, 777) if v2 else synth.synthetic_state_dict(arch, 777))
    if use_ema:
        # start from an initialised EMA state with every code well alive, so that neither the random
        # init nor the dead-code restart (rank-local random draws, layers_vq.py:212-221) is involved
        q = tr.model.quantizer
        q.emb_init.fill_(True)
        q.emb_elem.fill_(10.0)
        q.emb_sum.copy_(q.embeddings * 10.0)
    B, T = (4, 64) if v2 else (4, 20)
    details = []
    for s in range(steps):
        x, y = synth.vqvae2_batch(arch, B, T, 100 + s) if v2 else synth.synthetic_batch(arch, B, T, 100 + s)
        if world > 1:
            h = B // world
            x, y = x[rank * h:(rank + 1) * h], y[rank * h:(rank + 1) * h]
        _, d = tr.train_step([x, y], s)
        details.append(d)
    state = {n: p.detach().clone() for n, p in tr.model.named_parameters()}
    state.update({'buf.' + n: b.detach().clone() for n, b in tr.model.named_buffers()})
    torch.save({'state': state, 'details': details}, os.path.join(out_dir, 'w%d_r%d.pt' % (world, rank)))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


@pytest.mark.parametrize('use_ema', [False, True])
def test_two_rank_gloo_step_equals_single_process(tmp_path, use_ema):
    steps = 2
    out = str(tmp_path)
    _run(0, 1, 0, use_ema, steps, out)
    mp.spawn(_run, args=(2, _free_port(), use_ema, steps, out), nprocs=2, join=True)
    one = torch.load(os.path.join(out, 'w1_r0.pt'))
    r0 = torch.load(os.path.join(out, 'w2_r0.pt'))
    r1 = torch.load(os.path.join(out, 'w2_r1.pt'))
    for k, v in one['state'].items():
        a, b = r0['state'][k], r1['state'][k]
        assert torch.equal(a, b), 'ranks diverged on ' + k               # replicas stay bit-identical
        if v.dtype.is_floating_point:
            err = float((a.double() - v.double()).norm() / (v.double().norm() + 1e-30))
            assert err < 2e-5, (k, err)
        else:
            assert torch.equal(a, v), k
    # the mean of the per-rank losses is the global-batch loss
    for s in range(steps):
        for key in ('Total', 'X like', 'VQ loss'):
            m = 0.5 * (r0['details'][s][key] + r1['details'][s][key])
            assert abs(m - one['details'][s][key]) <= 2e-5 * max(1.0, abs(one['details'][s][key])), (s, key)


def test_two_rank_gloo_step_hierarchical_model(tmp_path):
    """The same for model/vqvae2.py: strided convs outside the weight bank (their gradients come
    through autograd and are averaged with the rest of the flat buffer), style-token level, per-level quantisers."""
    steps = 2
    out = str(tmp_path)
    _run(0, 1, 0, False, steps, out, 'vqvae2')
    mp.spawn(_run, args=(2, _free_port(), False, steps, out, 'vqvae2'), nprocs=2, join=True)
    one = torch.load(os.path.join(out, 'w1_r0.pt'))
    r0 = torch.load(os.path.join(out, 'w2_r0.pt'))
    r1 = torch.load(os.path.join(out, 'w2_r1.pt'))
    for k, v in one['state'].items():
        a, b = r0['state'][k], r1['state'][k]
        assert torch.equal(a, b), 'ranks diverged on ' + k
        err = float((a.double() - v.double()).norm() / (v.double().norm() + 1e-30))
        assert err < 5e-5, (k, err)
    for s_ in range(steps):
        for key in ('Total', 'X like', 'VQ loss'):
            m = 0.5 * (r0['details'][s_][key] + r1['details'][s_][key])
            assert abs(m - one['details'][s_][key]) <= 2e-5 * max(1.0, abs(one['details'][s_][key])), (s_, key)
