"""-m gpu: the drop-in Model / Trainer on the B200 (through the C ABI) against the goldens recorded
from the reference and against the CPU oracle.

Tolerances (north_star): codebook indices bit-exact wherever the top-2 distance gap exceeds fp32
round-off; fp32 path (precision='fp32', hi/mid/lo operand planes) <= 1e-5 on forward values and
losses, 5e-5 on small-model gradients; bf16 path <= 1e-3 on losses.  Full-size gradients: 5e-3
because one LeakyReLU-kink flip moves an early-encoder gradient by ~1e-3 even between two fp32
evaluations of the reference itself (make_golden.py)."""
import numpy as np
import pytest
import torch

from conftest import load_npz, rel_err
from oracle import synth
from oracle import vqvae_oracle as O

pytestmark = pytest.mark.gpu
TOL = 2e-5


@pytest.fixture()
def fp32_mode():
    import vae_npvc_b200.functional as Fn
    from vae_npvc_b200 import _C
    _C.require_cuda()
    old = Fn.get_precision()
    Fn.set_precision('fp32')
    yield Fn
    Fn.set_precision(old)


def _model(arch, sd):
    from vae_npvc_b200.model.vqvae import Model
    m = Model(arch)
    m.load_state_dict({k: v.clone() for k, v in sd.items()})
    return m.cuda()


@pytest.mark.parametrize('name', ['small_vqnorm', 'small_vqplain', 'small_2layer_dil', 'small_jitter'])
def test_small_model_matches_reference_golden(fp32_mode, golden_meta, name):
    case = golden_meta['cases'][name]
    arch = case['arch']
    x, y = synth.synthetic_batch(arch, case['B'], case['T'], 777)
    m = _model(arch, synth.synthetic_state_dict(arch, 777)).train()
    if 'jitter_src' in case:
        src = case['jitter_src']
        m.jitter.sources = lambda length: np.asarray(src)
    xhat, loss, losses = m([x.cuda(), y.cuda()])
    loss.backward()
    gold = load_npz(name)
    assert xhat.is_cuda and rel_err(xhat, gold['xhat']) < TOL
    with torch.no_grad():
        assert np.array_equal(m.encode(x.cuda()).cpu().numpy().reshape(-1), gold['idx'])
    for k, v in case['losses'].items():
        assert isinstance(losses[k], float) and abs(losses[k] - v) <= TOL * max(1.0, abs(v)), k
    params = dict(m.named_parameters())
    for key, gv in gold.items():
        if key.startswith('grad:'):
            assert rel_err(params[key[5:]].grad, gv) < 5 * TOL, key


def test_convert_paths(fp32_mode, golden_meta):
    arch = golden_meta['cases']['small_convert']['arch']
    x, y = synth.synthetic_batch(arch, 3, 20, 777)
    gold = load_npz('small_convert')
    m = _model(arch, synth.synthetic_state_dict(arch, 777)).eval()
    with torch.no_grad():
        ids = m.encode(x.cuda())
        assert ids.dtype == torch.int64 and np.array_equal(ids.cpu().numpy(), gold['ids'])
        assert rel_err(m.decode((ids, y.cuda())), gold['decode']) < TOL
        assert rel_err(m.infer((x.cuda(), y.cuda())), gold['infer']) < TOL


@pytest.mark.parametrize('name', ['vcc2020_vqnorm', 'vcc2020_vqplain'])
def test_full_vcc2020_shape_fp32(fp32_mode, golden_meta, name):
    """BASELINE.json configs[0]/[1] shape: 16 x 80 x 128, C512, codebook 512 x 64."""
    case = golden_meta['cases'][name]
    arch = case['arch']
    x, y = synth.synthetic_batch(arch, 16, 128, 777)
    m = _model(arch, synth.synthetic_state_dict(arch, 777)).train()
    xhat, loss, losses = m([x.cuda(), y.cuda()])
    loss.backward()
    for k, v in case['losses'].items():
        assert abs(losses[k] - v) <= 5e-5 * max(1.0, abs(v)), (k, losses[k], v)
    assert abs(float(xhat.double().norm()) - case['xhat_norm']) < 5e-5 * case['xhat_norm']
    for n, p in m.named_parameters():
        gn = case['grad_norms'][n]
        assert abs(float(p.grad.double().norm()) - gn) <= 5e-3 * gn + 1e-12, n
    # indices vs the CPU oracle on the same weights: bit-exact
    po = O.params_from_state_dict(synth.synthetic_state_dict(arch, 777), requires_grad=False)
    with torch.no_grad():
        assert torch.equal(m.encode(x.cuda()).cpu(), O.model_encode(po, arch, x))


def test_full_vcc2020_shape_bf16(golden_meta):
    """bf16 operand planes (the fast path): losses within 1e-3 relative of the reference."""
    import vae_npvc_b200.functional as Fn
    old = Fn.get_precision()
    Fn.set_precision('bf16')
    try:
        case = golden_meta['cases']['vcc2020_vqplain']
        arch = case['arch']
        x, y = synth.synthetic_batch(arch, 16, 128, 777)
        m = _model(arch, synth.synthetic_state_dict(arch, 777)).train()
        xhat, loss, losses = m([x.cuda(), y.cuda()])
        loss.backward()
        for k in ('Total', 'X like'):
            assert abs(losses[k] - case['losses'][k]) <= 1e-3 * abs(case['losses'][k]), (k, losses[k])
        worst = max(abs(float(p.grad.double().norm()) - case['grad_norms'][n]) / case['grad_norms'][n]
                    for n, p in m.named_parameters())
        assert worst < 0.1                      # sanity only: per-tensor bf16 gradients are ~1e-2
    finally:
        Fn.set_precision(old)


def test_ema_model_matches_oracle(fp32_mode, golden_meta):
    from test_composition_cpu import _ema_preset
    arch = golden_meta['cases']['small_ema']['arch']
    x, y = synth.synthetic_batch(arch, 3, 20, 777)
    sd = _ema_preset(arch, synth.synthetic_state_dict(arch, 777), x)
    m = _model(arch, sd).train()
    xhat, loss, losses = m([x.cuda(), y.cuda()])
    loss.backward()
    po = O.params_from_state_dict(sd)
    oxhat, oloss, olosses, _ = O.model_forward(po, arch, x, y, True)
    oloss.backward()
    assert rel_err(xhat, oxhat) < TOL
    for k, v in olosses.items():
        assert abs(losses[k] - v) <= TOL * max(1.0, abs(v)), k
    for n, p in m.named_parameters():
        if po[n].grad is not None:
            assert rel_err(p.grad, po[n].grad) < 5 * TOL, n
    for b in ('emb_sum', 'emb_elem', 'embeddings'):
        assert rel_err(m.state_dict()['quantizer.' + b], po['quantizer.' + b]) < TOL, b


@pytest.mark.parametrize('use_graph', [False, True])
def test_trainer_steps_match_reference(fp32_mode, golden_meta, use_graph):
    """3 clipped Adam steps against parameters recorded from the reference; then 3 more steps on
    the captured CUDA graph against the CPU oracle continuing from the same state."""
    from vae_npvc_b200.trainer.basic import Trainer
    case = golden_meta['cases']['small_train3']
    arch = case['arch']
    cfg = dict(arch)
    cfg.update(model_type='vae_npvc_b200.model.vqvae', precision='fp32', cuda_graph=use_graph, graph_warmup=3,
               learning_rate=case['lr'], max_grad_norm=case['max_grad_norm'], lr_scheduler='StepLR',
               lr_param={'step_size': case['step_size'], 'gamma': case['gamma'], 'last_epoch': -1})
    tr = Trainer(cfg)
    sd = synth.synthetic_state_dict(arch, 777)
    tr.model.load_state_dict(sd)
    params = O.params_from_state_dict(sd)
    ost = O.AdamState(O.trainable_names(params), lr=case['lr'], step_size=case['step_size'], gamma=case['gamma'])
    it = 0
    for i in range(6):
        seed = case['batch_seeds'][i] if i < 3 else 200 + i
        x, y = synth.synthetic_batch(arch, case['B'], case['T'], seed)
        it, detail = tr.train_step([x, y], it)
        ol, _, _ = O.train_step(params, arch, ost, x, y, max_grad_norm=case['max_grad_norm'])
        assert tr.last_step_on_graph == (use_graph and i >= 3)
        for k, v in ol.items():
            assert abs(detail[k] - v) <= 5 * TOL * max(1.0, abs(v)), (i, k, detail[k], v)
        if i == 2:
            gold = load_npz('small_train3')
            mp = dict(tr.model.named_parameters())
            for key, p in gold.items():
                assert rel_err(mp[key[6:]], p) < TOL, key
    for n, p in tr.model.named_parameters():
        assert rel_err(p, params[n]) < 5 * TOL, n
    assert it == 6
