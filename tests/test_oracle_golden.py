"""The oracle restatement vs the goldens recorded from the imported reference
(oracle/make_golden.py).  Runs anywhere - does NOT need /root/reference."""
import numpy as np
import pytest
import torch

from conftest import load_npz, rel_err
from oracle import synth
from oracle import vqvae_oracle as O

SMALL = ['small_vqnorm', 'small_vqplain', 'small_ema', 'small_ema_bigK', 'small_2layer_dil', 'small_jitter']
TOL = 2e-5        # fp32 round-off between two evaluation orders of the same math


def _run(meta, name, dtype=torch.float32):
    case = meta['cases'][name]
    arch = case['arch']
    sd = synth.synthetic_state_dict(arch, 777)
    x, y = synth.synthetic_batch(arch, case['B'], case['T'], 777)
    params = O.params_from_state_dict(sd, dtype=dtype)
    torch.manual_seed(1234)
    xhat, loss, losses, idx = O.model_forward(params, arch, x.to(dtype), y, True, case.get('jitter_src'))
    loss.backward()
    return case, params, xhat, losses, idx


@pytest.mark.parametrize('name', SMALL)
def test_small_cases_match_reference_tensors(golden_meta, name):
    case, params, xhat, losses, idx = _run(golden_meta, name)
    gold = load_npz(name)
    assert rel_err(xhat, gold['xhat']) < TOL
    assert np.array_equal(idx.numpy(), gold['idx'])                    # indices bit-exact
    for k, v in case['losses'].items():
        assert abs(losses[k] - v) <= TOL * max(1.0, abs(v)), k
    n_grads = 0
    for key, g in gold.items():
        if key.startswith('grad:'):
            n_grads += 1
            assert rel_err(params[key[5:]].grad, g) < 5 * TOL, key
        if key.startswith('buf:'):
            assert rel_err(params['quantizer.' + key[4:]].float(), g.astype(np.float32)) < TOL, key
    assert n_grads == len(case['grad_norms']) > 0


def test_ema_no_ste_bug_compat(golden_meta):
    """with use_ema the reconstruction loss never reaches the encoder through
    z_vq; the only encoder gradient is beta * z_enc_loss."""
    arch = golden_meta['cases']['small_ema']['arch']
    sd = synth.synthetic_state_dict(arch, 777)
    x, y = synth.synthetic_batch(arch, 3, 20, 777)
    params = O.params_from_state_dict(sd)
    torch.manual_seed(1234)
    z = O.encoder_forward(params, arch['encoder'], x)
    buf = {k: params['quantizer.' + k] for k in O.EMA_BUFFERS}
    z_vq, lq, le, det, idx = O.ema_vq_forward(buf, z)
    assert not z_vq.requires_grad and lq == 0.0 and le.requires_grad


def test_convert_paths(golden_meta):
    arch = golden_meta['cases']['small_convert']['arch']
    sd = synth.synthetic_state_dict(arch, 777)
    x, y = synth.synthetic_batch(arch, 3, 20, 777)
    gold = load_npz('small_convert')
    params = O.params_from_state_dict(sd, requires_grad=False)
    with torch.no_grad():
        ids = O.model_encode(params, arch, x)
        assert ids.dtype == torch.int64 and np.array_equal(ids.numpy(), gold['ids'])
        assert rel_err(O.model_decode(params, arch, ids, y), gold['decode']) < TOL
        assert rel_err(O.model_infer(params, arch, x, y), gold['infer']) < TOL


@pytest.mark.parametrize('normalize', [False, True])
@pytest.mark.parametrize('red', ['sum', 'mean', 'batch_mean', 'frame_mean', 'none'])
def test_vq_layer_all_reductions(normalize, red):
    gold = load_npz('vq_layer')
    key = f'vq_n{int(normalize)}_{red}'
    e = torch.tensor(gold['E']).requires_grad_(True)
    z = torch.tensor(gold['z']).requires_grad_(True)
    zv, lq, le, perp, idx = O.vq_forward(e, z, normalize, red)
    (zv.sum() * 0.5 + lq.sum() + 0.3 * le.sum()).backward()
    assert rel_err(zv, gold[key + ':z_vq']) < 1e-6
    assert rel_err(lq, gold[key + ':lq']) < 1e-6 and rel_err(le, gold[key + ':le']) < 1e-6
    assert rel_err(z.grad, gold[key + ':dz']) < 1e-5 and rel_err(e.grad, gold[key + ':dE']) < 1e-5
    assert rel_err(e.detach(), gold[key + ':E_after']) < 1e-6          # in-place embed_norm
    assert abs(float(perp) - float(gold[key + ':entropy'])) < 1e-4


def test_three_clipped_adam_steps(golden_meta):
    case = golden_meta['cases']['small_train3']
    arch = case['arch']
    params = O.params_from_state_dict(synth.synthetic_state_dict(arch, 777))
    opt = O.AdamState(O.trainable_names(params), lr=case['lr'], step_size=case['step_size'], gamma=case['gamma'])
    for it, seed in enumerate(case['batch_seeds']):
        x, y = synth.synthetic_batch(arch, case['B'], case['T'], seed)
        losses, _, _ = O.train_step(params, arch, opt, x, y, max_grad_norm=case['max_grad_norm'])
        for k, v in case['losses'][it].items():
            assert abs(losses[k] - v) <= TOL * max(1.0, abs(v)), (it, k)
    gold = load_npz('small_train3')
    for key, p in gold.items():
        assert rel_err(params[key[6:]], p) < TOL, key


@pytest.mark.parametrize('name', ['vcc2020_vqnorm', 'vcc2020_ema'])
def test_full_vcc2020_shape_scalars(golden_meta, name):
    """BASELINE.json configs[0]: 16x80x128, C512, codebook 512x64, fp32 fwd+bwd on CPU.
    Gradient norms: 5e-3 because single LeakyReLU-kink flips between two fp32 evaluation
    orders move early-encoder gradients by ~1e-3 (see make_golden.py / DESIGN.md)."""
    case, params, xhat, losses, idx = _run(golden_meta, name)
    for k, v in case['losses'].items():
        assert abs(losses[k] - v) <= 5e-5 * max(1.0, abs(v)), k
    assert abs(float(xhat.double().norm()) - case['xhat_norm']) < 5e-5 * case['xhat_norm']
    for n, gn in case['grad_norms'].items():
        assert abs(float(params[n].grad.double().norm()) - gn) <= 5e-3 * gn + 1e-12, n


def test_jitter_sources_follow_reference_draw_order():
    src = O.jitter_sources(20, 0.3, np.random.RandomState(5))
    assert len(src) == 20 and all(abs(int(s) - t) <= 1 for t, s in enumerate(src))
    assert src[0] in (0, 1) and src[-1] in (18, 19)
    # Reference quirk kept bug-compatibly: `[True, False][choice([1, 0], p=[p, 1-p])]`
    # (layers_vq.py:364) replaces when the draw is 0, i.e. with probability 1-p, not p.
    assert np.array_equal(O.jitter_sources(9, 1.0, np.random.RandomState(0)), np.arange(9))
    assert all(s != t for t, s in enumerate(O.jitter_sources(9, 0.0, np.random.RandomState(0))))
