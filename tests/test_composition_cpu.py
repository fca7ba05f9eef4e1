"""Host-side composition check on CPU: the REAL vae_npvc_b200 modules and autograd Functions
(forward sequencing + every hand-written backward), with the CUDA primitives replaced by their
torch-CPU contract restatement (oracle/prims_ref.py), against the reference goldens.

This validates everything above the C ABI without a GPU; the `-m gpu` tests then validate the
kernels below it (each primitive against the same prims_ref function, and the whole model).
"""
import numpy as np
import pytest
import torch

from conftest import load_npz, rel_err
from oracle import prims_ref, synth
from oracle import vqvae_oracle as O

TOL = 2e-5


@pytest.fixture()
def cpu_prims(monkeypatch):
    import vae_npvc_b200.functional as Fn
    import vae_npvc_b200.model.layers_vq as LVQ
    monkeypatch.setattr(Fn, 'Pm', prims_ref)
    monkeypatch.setattr(LVQ, 'Pm', prims_ref)
    old = Fn.get_precision()
    Fn.set_precision('fp32')
    yield Fn
    Fn.set_precision(old)


def _model(arch, sd):
    from vae_npvc_b200.model.vqvae import Model
    m = Model(arch)
    m.load_state_dict({k: v.clone() for k, v in sd.items()})
    return m


@pytest.mark.parametrize('name', ['small_vqnorm', 'small_vqplain', 'small_2layer_dil', 'small_jitter'])
def test_model_forward_backward_matches_reference(cpu_prims, golden_meta, name):
    case = golden_meta['cases'][name]
    arch = case['arch']
    sd = synth.synthetic_state_dict(arch, 777)
    x, y = synth.synthetic_batch(arch, case['B'], case['T'], 777)
    m = _model(arch, sd).train()
    if 'jitter_src' in case:
        src = case['jitter_src']
        m.jitter.sources = lambda length: np.asarray(src)
    xhat, loss, losses = m([x, y])
    loss.backward()
    gold = load_npz(name)
    assert rel_err(xhat, gold['xhat']) < TOL
    for k, v in case['losses'].items():
        assert abs(losses[k] - v) <= TOL * max(1.0, abs(v)), k
    params = dict(m.named_parameters())
    n = 0
    for key, g in gold.items():
        if key.startswith('grad:'):
            n += 1
            assert params[key[5:]].grad is not None, key
            assert rel_err(params[key[5:]].grad, g) < 5 * TOL, key
    assert n == len(case['grad_norms'])
    # the in-place embed_norm of the codebook Parameter (layers_vq.py:96)
    if arch.get('embed_norm', True):
        assert abs(float(m.quantizer.embeddings.detach().norm(dim=1).mean()) - 1.0) < 1e-6


def test_two_plane_mode_on_the_contract(cpu_prims, golden_meta):
    """'bf16x2' (hi + mid operand planes): the composition with 16-bit operands stays within 1e-4 of the reference."""
    cpu_prims.set_precision('bf16x2')
    case = golden_meta['cases']['small_vqnorm']
    arch = case['arch']
    x, y = synth.synthetic_batch(arch, case['B'], case['T'], 777)
    m = _model(arch, synth.synthetic_state_dict(arch, 777)).train()
    xhat, loss, losses = m([x, y])
    loss.backward()
    gold = load_npz('small_vqnorm')
    e = rel_err(xhat, gold['xhat'])
    assert 1e-8 < e < 1e-4, e                    # not bit-identical to the 3-plane mode, but close
    for key, g in gold.items():
        if key.startswith('grad:'):
            assert rel_err(dict(m.named_parameters())[key[5:]].grad, g) < 3e-4, key


def test_convert_paths(cpu_prims, golden_meta):
    arch = golden_meta['cases']['small_convert']['arch']
    sd = synth.synthetic_state_dict(arch, 777)
    x, y = synth.synthetic_batch(arch, 3, 20, 777)
    gold = load_npz('small_convert')
    m = _model(arch, sd).eval()
    with torch.no_grad():
        ids = m.encode(x)
        assert ids.dtype == torch.int64 and tuple(ids.shape) == (3, 20)
        assert np.array_equal(ids.numpy(), gold['ids'])
        assert rel_err(m.decode((ids, y)), gold['decode']) < TOL
        assert rel_err(m.infer((x, y)), gold['infer']) < TOL
        assert np.array_equal(m.encode([x, y]).numpy(), gold['ids'])          # list input (vqvae.py:46-49)


def _ema_preset(arch, sd, x):
    """Initialised EMA buffers with every code alive, so no random restart is drawn."""
    p = O.params_from_state_dict(sd, requires_grad=False)
    z = O.encoder_forward(p, arch['encoder'], x)
    zf = z.transpose(1, 2).reshape(-1, z.shape[1])
    K = arch['z_num']
    rows = zf[torch.arange(K) % zf.shape[0]].clone() + 0.01 * torch.arange(K).view(K, 1) / K
    sd = dict(sd)
    sd['quantizer.emb_init'] = torch.tensor(True)
    sd['quantizer.embeddings'] = rows
    sd['quantizer.emb_sum'] = rows * 3.0
    sd['quantizer.emb_elem'] = torch.full((K,), 3.0)
    return sd


def test_ema_model_matches_oracle(cpu_prims, golden_meta):
    arch = golden_meta['cases']['small_ema']['arch']
    sd0 = synth.synthetic_state_dict(arch, 777)
    x, y = synth.synthetic_batch(arch, 3, 20, 777)
    sd = _ema_preset(arch, sd0, x)
    m = _model(arch, sd).train()
    xhat, loss, losses = m([x, y])
    loss.backward()
    po = O.params_from_state_dict(sd)
    oxhat, oloss, olosses, oidx = O.model_forward(po, arch, x, y, True)
    oloss.backward()
    assert rel_err(xhat, oxhat) < TOL
    for k, v in olosses.items():
        assert abs(losses[k] - v) <= TOL * max(1.0, abs(v)), k
    for n, p in m.named_parameters():
        og = po[n].grad
        if og is None:
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, n
        else:
            assert rel_err(p.grad, og) < 5 * TOL, n
    msd = m.state_dict()
    for b in ('emb_sum', 'emb_elem', 'embeddings'):
        assert rel_err(msd['quantizer.' + b], po['quantizer.' + b]) < TOL, b
    # eval: no update, no init
    m.eval()
    before = m.quantizer.embeddings.clone()
    with torch.no_grad():
        m([x, y])
    assert torch.equal(before, m.quantizer.embeddings)


def test_ema_first_call_initialises_from_batch(cpu_prims, golden_meta):
    for name in ('small_ema', 'small_ema_bigK'):                      # bigK: N < K -> tiled + noise
        arch = golden_meta['cases'][name]['arch']
        sd = synth.synthetic_state_dict(arch, 777)
        x, y = synth.synthetic_batch(arch, 3, 20, 777)
        m = _model(arch, sd).train()
        assert not bool(m.quantizer.emb_init)
        xhat, loss, losses = m([x, y])
        assert bool(m.quantizer.emb_init) and set(losses) >= {'entropy', 'used_curr', 'usage', 'diff_emb'}
        z = O.encoder_forward(O.params_from_state_dict(sd, requires_grad=False), arch['encoder'], x)
        zf = z.transpose(1, 2).reshape(-1, z.shape[1])
        # after init + one update every codebook row is (close to) a batch frame or a mean of frames
        assert torch.isfinite(m.quantizer.embeddings).all()
        assert float(m.quantizer.emb_elem.min()) >= 0.9 - 1e-6
        assert m.quantizer.embeddings.abs().max() <= zf.abs().max() * 1.1 + 0.1


@pytest.mark.parametrize('normalize', [False, True])
@pytest.mark.parametrize('red', ['sum', 'mean', 'batch_mean', 'frame_mean', 'none'])
def test_vq_layer_all_reductions(cpu_prims, normalize, red):
    from vae_npvc_b200.model.layers_vq import VectorQuantizer
    gold = load_npz('vq_layer')
    key = f'vq_n{int(normalize)}_{red}'
    q = VectorQuantizer(16, 8, normalize=normalize, reduction=red)
    with torch.no_grad():
        q.embeddings.copy_(torch.tensor(gold['E']))
    z = torch.tensor(gold['z']).requires_grad_(True)
    zv, lq, le, det = q(z)
    (zv.sum() * 0.5 + lq.sum() + 0.3 * le.sum()).backward()
    assert rel_err(zv, gold[key + ':z_vq']) < 1e-6
    assert rel_err(lq, gold[key + ':lq']) < 1e-5 and rel_err(le, gold[key + ':le']) < 1e-5
    assert rel_err(z.grad, gold[key + ':dz']) < 2e-5 and rel_err(q.embeddings.grad, gold[key + ':dE']) < 2e-5
    assert rel_err(q.embeddings.detach(), gold[key + ':E_after']) < 1e-6
    assert abs(det['entropy'] - float(gold[key + ':entropy'])) < 1e-4
    # bypass switch (layers_vq.py:80-82)
    q.quantize = False
    out = q(z)
    assert out[0] is z and float(out[1]) == 0.0


def test_layers_accept_time_varying_cond(cpu_prims):
    """Layer-level drop-in: DeConv1d_Layernorm_GLU_ResSkip.forward(x, c) with a full (B,Y,T)
    conditioning tensor equals the oracle's block, and equals the broadcast fast path when c is
    constant over time."""
    from vae_npvc_b200.model.layers import DeConv1d_Layernorm_GLU_ResSkip
    torch.manual_seed(0)
    blk = DeConv1d_Layernorm_GLU_ResSkip(kernel_size=3, in_channels=8, cond_channels=6, skip_channels=5, dilation=1)
    for m in (blk.conv_in, blk.conv_cond, blk.res_skip_layers):
        torch.nn.utils.weight_norm(m)
    x = torch.randn(2, 8, 12, requires_grad=True)
    c = torch.randn(2, 6, 12, requires_grad=True)
    xo, sk = blk(x, c)
    (xo.sum() + 2 * sk.sum()).backward()
    p = {'b.' + k: v.detach().clone().requires_grad_(True) for k, v in blk.state_dict().items()}
    xr, cr = x.detach().clone().requires_grad_(True), c.detach().clone().requires_grad_(True)
    oxo, osk = O.dec_block(p, 'b', xr, cr, 1)
    (oxo.sum() + 2 * osk.sum()).backward()
    assert rel_err(xo, oxo) < TOL and rel_err(sk, osk) < TOL
    assert rel_err(x.grad, xr.grad) < 5 * TOL and rel_err(c.grad, cr.grad) < 5 * TOL
    for k, v in blk.named_parameters():
        assert rel_err(v.grad, p['b.' + k].grad) < 5 * TOL, k
    c1 = torch.randn(2, 6, 1)
    a = blk(x.detach(), c1.expand(-1, -1, 12).contiguous())
    b = blk(x.detach(), c1)
    assert rel_err(a[0], b[0]) < TOL and rel_err(a[1], b[1]) < TOL


def test_remove_weight_norm_keeps_outputs(cpu_prims, golden_meta):
    arch = golden_meta['cases']['small_vqnorm']['arch']
    sd = synth.synthetic_state_dict(arch, 777)
    x, y = synth.synthetic_batch(arch, 3, 20, 777)
    m = _model(arch, sd).eval()
    with torch.no_grad():
        a = m.infer((x, y))
        m.remove_weight_norm()
        assert 'encoder.encode.0.weight' in m.state_dict() and 'encoder.encode.0.weight_g' not in m.state_dict()
        b = m.infer((x, y))
    assert rel_err(b, a) < TOL


def _trainer(cpu_prims, monkeypatch, arch, extra):
    import vae_npvc_b200.trainer.basic as TB
    monkeypatch.setattr(TB, 'Pm', prims_ref)
    cfg = dict(arch)
    cfg.update(model_type='vae_npvc_b200.model.vqvae', device='cpu', cuda_graph=False, precision='fp32')
    cfg.update(extra)
    return TB.Trainer(cfg)


def test_trainer_three_clipped_adam_steps(cpu_prims, monkeypatch, golden_meta, tmp_path):
    """trainer/basic.py:55-79 semantics: clip_grad_norm_ -> Adam(0.5, 0.999) -> StepLR, on the flat
    buffers, against parameters recorded from the reference after 3 steps."""
    case = golden_meta['cases']['small_train3']
    arch = case['arch']
    tr = _trainer(cpu_prims, monkeypatch, arch, dict(
        learning_rate=case['lr'], max_grad_norm=case['max_grad_norm'], lr_scheduler='StepLR',
        lr_param={'step_size': case['step_size'], 'gamma': case['gamma'], 'last_epoch': -1}))
    tr.model.load_state_dict(synth.synthetic_state_dict(arch, 777))
    it = 0
    for i, seed in enumerate(case['batch_seeds']):
        x, y = synth.synthetic_batch(arch, case['B'], case['T'], seed)
        it, detail = tr.train_step([x, y], it)
        assert it == i + 1 and all(isinstance(v, float) for v in detail.values())
        for k, v in case['losses'][i].items():
            assert abs(detail[k] - v) <= TOL * max(1.0, abs(v)), (i, k)
    gold = load_npz('small_train3')
    params = dict(tr.model.named_parameters())
    for key, p in gold.items():
        assert rel_err(params[key[6:]], p) < TOL, key
    # parameters are views of one flat buffer; checkpoint round trip in the reference's format
    assert all(p.data_ptr() >= tr._flat_p.data_ptr() for p in tr._params)
    ck = tmp_path / 'iter.3'
    tr.save_checkpoint(str(ck))
    data = torch.load(str(ck), map_location='cpu')
    assert set(data) == {'model', 'optimizer', 'iteration'} and data['iteration'] == 3
    assert set(data['optimizer']) == {'state', 'param_groups'} and len(data['optimizer']['state']) == len(tr._params)
    assert float(data['optimizer']['state'][0]['step']) == 3.0
    tr2 = _trainer(cpu_prims, monkeypatch, arch, dict(learning_rate=case['lr'], max_grad_norm=case['max_grad_norm']))
    assert tr2.load_checkpoint(str(ck)) == 3
    assert torch.equal(tr2._flat_p, tr._flat_p) and torch.equal(tr2.optimizer.m, tr.optimizer.m)
    assert float(tr2.optimizer.state[0]) == 3.0 and float(tr2.optimizer.state[1]) == 0.0
    # validation leaves the model in train mode; the only write is the reference's own in-place
    # codebook renormalisation, which also runs in eval mode (layers_vq.py:96)
    before = {n: p.detach().clone() for n, p in tr.model.named_parameters()}
    x, y = synth.synthetic_batch(arch, case['B'], case['T'], 5)
    out = tr.valid([[x, y], [x, y]])
    assert tr.model.training and len(out['Total']) == 2
    for n, p in tr.model.named_parameters():
        if n != 'quantizer.embeddings':
            assert torch.equal(p.detach(), before[n]), n
    assert abs(float(tr.model.quantizer.embeddings.detach().norm(dim=1).mean()) - 1.0) < 1e-6


def test_resume_keeps_decayed_lr_like_torch(cpu_prims, monkeypatch, tmp_path):
    """the reference builds Trainer (and StepLR) BEFORE load_checkpoint; optimizer.load_state_dict then
    installs the checkpoint's DECAYED lr and the fresh StepLR keeps chaining gamma off it
    (trainer/basic.py:43-46,117-121).  Checked against torch.optim.Adam + StepLR driven the same way."""
    arch = synth.small_config(embed_norm=False)      # (the in-place codebook renormalisation is not an optimiser op)
    extra = dict(learning_rate=1e-3, max_grad_norm=10, lr_scheduler='StepLR',
                 lr_param={'step_size': 3, 'gamma': 0.5, 'last_epoch': -1})
    tr = _trainer(cpu_prims, monkeypatch, arch, extra)
    tr.model.load_state_dict(synth.synthetic_state_dict(arch, 777))
    # torch's own optimiser + scheduler over a clone of the same parameters, fed OUR gradients
    tp = [p.detach().clone().requires_grad_(True) for p in tr._params]
    opt = torch.optim.Adam(tp, lr=1e-3, betas=(0.5, 0.999), weight_decay=0.0)
    sch = torch.optim.lr_scheduler.StepLR(opt, step_size=3, gamma=0.5)

    def torch_step(trn, params, o, s_):
        g = trn._flat_g.clone()
        total = float(g.pow(2).sum().sqrt())
        coef = min(1.0, 10 / (total + 1e-6))
        for q, off in zip(params, trn._offsets):
            n = q.numel()
            q.grad = (g[off:off + n] * coef).view_as(q).clone()
        o.step()
        s_.step()

    it = 0
    for i in range(7):
        x, y = synth.synthetic_batch(arch, 3, 20, 300 + i)
        it, _ = tr.train_step([x, y], it)
        torch_step(tr, tp, opt, sch)
    assert abs(tr.optimizer.current_lr() - 2.5e-4) < 1e-12 and abs(opt.param_groups[0]['lr'] - 2.5e-4) < 1e-12
    for a, b in zip(tr._params, tp):
        assert rel_err(a, b) < 1e-5
    ck = tmp_path / 'iter.7'
    tr.save_checkpoint(str(ck))
    data = torch.load(str(ck), map_location='cpu')
    assert abs(data['optimizer']['param_groups'][0]['lr'] - 2.5e-4) < 1e-12          # what torch would have saved
    assert abs(data['optimizer']['param_groups'][0]['initial_lr'] - 1e-3) < 1e-12
    # resume: fresh trainer + fresh torch optimiser/scheduler, both load the checkpoint
    tr2 = _trainer(cpu_prims, monkeypatch, arch, extra)
    assert tr2.load_checkpoint(str(ck)) == 7
    tp2 = [p.detach().clone().requires_grad_(True) for p in tr2._params]
    opt2 = torch.optim.Adam(tp2, lr=1e-3, betas=(0.5, 0.999), weight_decay=0.0)
    sch2 = torch.optim.lr_scheduler.StepLR(opt2, step_size=3, gamma=0.5)
    opt2.load_state_dict(opt.state_dict())
    lrs, want = [], []
    for i in range(4):
        x, y = synth.synthetic_batch(arch, 3, 20, 400 + i)
        lrs.append(tr2.optimizer.current_lr())
        want.append(opt2.param_groups[0]['lr'])
        it, _ = tr2.train_step([x, y], it)
        torch_step(tr2, tp2, opt2, sch2)
    assert want == pytest.approx([2.5e-4, 2.5e-4, 2.5e-4, 1.25e-4])
    assert lrs == pytest.approx(want)
    for a, b in zip(tr2._params, tp2):
        assert rel_err(a, b) < 1e-5


def _cdvae_small():
    arch = synth.cdvae_config(Xa=10, Xb=6, C=16, D=8, K=32, Y=8, S=12, y_num=5, stacks=2)
    return arch, synth.cdvae_state_dict(arch, 777), synth.cdvae_batch(arch, 3, 20, 777)


def test_cdvae_composition_matches_oracle_parts(cpu_prims, monkeypatch):
    """BASELINE configs[2] (cross-domain, two encoders / decoders over one shared codebook): the product composition
    against the same wiring of the oracle's parts - outputs, losses, every gradient; then the Trainer takes steps on
    its three-tensor batch."""
    from oracle import cdvae_oracle as CO
    from vae_npvc_b200.model.cdvae import Model
    arch, sd, (xs, y) = _cdvae_small()
    m = Model(arch)
    assert list(m.state_dict().keys()) == list(sd.keys())
    m.load_state_dict({k: v.clone() for k, v in sd.items()})
    m.train()
    outs, loss, losses = m(xs + [y])
    loss.backward()
    po = O.params_from_state_dict(sd)
    oouts, oloss, olosses, _ = CO.cdvae_forward(po, arch, xs, y)
    oloss.backward()
    for a, b in zip(outs, oouts):
        assert a.shape == b.shape and rel_err(a, b) < TOL
    assert set(losses) == set(olosses)
    for k, v in olosses.items():
        assert abs(losses[k] - v) <= TOL * max(1.0, abs(v)), k
    n = 0
    for name, p in m.named_parameters():
        if po[name].grad is not None:
            assert rel_err(p.grad, po[name].grad) < 5 * TOL, name
            n += 1
    assert n > 100
    import vae_npvc_b200.trainer.basic as TB
    monkeypatch.setattr(TB, 'Pm', prims_ref)
    cfg = dict(arch, device='cpu', cuda_graph=False, precision='fp32', learning_rate=1e-3, max_grad_norm=10)
    tr = TB.Trainer(cfg)
    tr.model.load_state_dict(sd)
    it, d1 = tr.train_step(xs + [y], 0)
    assert abs(d1['Total'] - olosses['Total']) <= TOL * abs(olosses['Total'])
    for _ in range(3):
        it, d2 = tr.train_step(xs + [y], it)
    assert it == 4 and d2['Total'] < d1['Total']


def test_weight_bank_segments_equal_whole(cpu_prims):
    """bank.py segments: materialising / back-propagating the encoder and the decoder layers by separate calls
    (the trainer issues the decoder's on a side stream) gives the same operand planes and (dv, dg) as one call."""
    import vae_npvc_b200.trainer.basic as TB
    old = TB.Pm
    TB.Pm = prims_ref
    try:
        arch = synth.small_config()
        cfg = dict(arch, model_type='vae_npvc_b200.model.vqvae', device='cpu', cuda_graph=False, precision='fp32')
        tr = TB.Trainer(cfg)
        tr.model.load_state_dict(synth.synthetic_state_dict(arch, 777))
        bank = tr.bank
        assert len(bank.segments) == 3 and bank.segment_kinds == [0, 1, 2]     # encoder | decoder | conditioning convs
        for sa, sb in zip(bank.segments[:-1], bank.segments[1:]):               # contiguous, cover every layer
            assert sa['layers'][1] == sb['layers'][0] and sa['rows'][1] == sb['rows'][0] and sa['tiles'][1] == sb['tiles'][0]
        assert bank.segments[0]['layers'][0] == 0 and bank.segments[-1]['layers'][1] == bank.n_layers
        assert bank.segments[-1]['rows'][1] == bank.row2layer.numel()
        assert bank.segments[-1]['tiles'][1] == bank.tile2layer.numel()
        bank.prepare()
        whole_w, whole_w32 = bank.wbuf.clone(), bank.w32buf.clone()
        bank.wbuf.zero_()
        bank.w32buf.zero_()
        bank.prepare(1)
        assert not torch.equal(bank.wbuf, whole_w)                               # only the decoder so far
        bank.prepare(2)
        bank.prepare(0)
        assert torch.equal(bank.wbuf, whole_w) and torch.equal(bank.w32buf, whole_w32)
        g = torch.Generator().manual_seed(3)
        dw = torch.randn(tr._flat_g.shape, generator=g)
        tr._flat_g.copy_(dw)
        bank.finish_backward()
        whole_g = tr._flat_g.clone()
        tr._flat_g.copy_(dw)
        bank.finish_backward(1)
        bank.finish_backward(0)
        bank.finish_backward(2)
        assert torch.equal(tr._flat_g, whole_g)
    finally:
        TB.Pm = old
