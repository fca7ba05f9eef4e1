"""Hierarchical VQ-VAE2 + style tokens on CPU: the oracle against the golden taken from the
reference's own vqvae2.Model (oracle/make_golden_vqvae2.py), the new primitives' restatements against autograd, and the
REAL vae_npvc_b200.model.vqvae2 modules + autograd Functions on the torch-CPU contract (oracle/prims_ref.py) against the
same golden - everything above the C ABI without a GPU."""
import numpy as np
import pytest
import torch

from conftest import load_npz, rel_err
from oracle import prims_ref, synth
from oracle import vqvae2_oracle as O2
from oracle import vqvae_oracle as O

B, T = 3, 128


def _check_grads(get_grad, gold, tol):
    grads = {k[5:]: v for k, v in gold.items() if k.startswith('grad/')}
    gscale = max(float(np.linalg.norm(v)) for v in grads.values())
    assert len(grads) == 166
    for k, g in grads.items():
        got = get_grad(k)
        assert got is not None, k
        if float(np.linalg.norm(g)) < 1e-6 * gscale:
            # mathematically zero: the key bias shifts all token scores of a head alike, softmax cancels it
            assert float(got.norm()) < 1e-5 * gscale, k
        else:
            assert rel_err(got, g) < tol, (k, rel_err(got, g))


def test_oracle_matches_reference_golden():
    gold = load_npz('vqvae2_small')
    arch = synth.vqvae2_small_config()
    x, y = synth.vqvae2_batch(arch, B, T, 777)
    assert np.array_equal(x.numpy(), gold['x']) and np.array_equal(y.numpy(), gold['y'])
    params = O.params_from_state_dict(synth.vqvae2_state_dict(arch, 777))
    xhat, loss, losses, ids = O2.model_forward(params, arch, x, y)
    loss.backward()
    assert rel_err(xhat, gold['xhat']) < 2e-5
    assert abs(float(loss) - float(gold['loss'])) < 2e-5 * abs(float(gold['loss']))
    for k in ('Total', 'VQ loss', 'X like', 'entropy.0', 'entropy.1', 'quanti_err.0', 'quanti_err.1'):
        assert abs(losses[k] - float(gold['loss/' + k])) <= 2e-5 * max(1.0, abs(float(gold['loss/' + k]))), k
    for i, idx in enumerate(ids):
        assert np.array_equal(idx.numpy(), gold['ids/%d' % i])
    _check_grads(lambda k: params[k].grad, gold, 1e-4)


def test_upsample_follows_the_reference_definition():
    z = torch.randn(2, 4, 5)
    for target in (5, 10, 13, 19):
        rep = target // 5
        want = z.unsqueeze(-1).repeat(1, 1, 1, rep).flatten(-2, -1)          # vqvae2.py:134-137
        want = want[..., :target] if want.size(-1) >= target else torch.nn.functional.pad(want, (0, target - want.size(-1)), 'replicate')
        assert torch.equal(O2.upsample(z, target), want)
        assert torch.equal(prims_ref.upsample_cat_fwd([z], target), want)


def test_new_primitive_restatements_against_autograd():
    torch.manual_seed(0)
    a, b = torch.randn(2, 4, 3, dtype=torch.float64, requires_grad=True), torch.randn(2, 8, 1, dtype=torch.float64, requires_grad=True)
    out = prims_ref.upsample_cat_fwd([a, b], 7)
    g = torch.randn_like(out)
    ga, gb = torch.autograd.grad(out, (a, b), g)
    da, db = prims_ref.upsample_cat_bwd(g, [(4, 3), (8, 1)])
    assert torch.allclose(da, ga) and torch.allclose(db, gb)
    x = torch.randn(2, 4, 5, dtype=torch.float64, requires_grad=True)
    gm = torch.randn(2, 4, dtype=torch.float64)
    assert torch.allclose(prims_ref.time_mean_bwd(gm, 5), torch.autograd.grad(prims_ref.time_mean_fwd(x), x, gm)[0])
    q, K, V = (torch.randn(3, 8, dtype=torch.float64, requires_grad=True), torch.randn(5, 8, dtype=torch.float64, requires_grad=True),
               torch.randn(5, 8, dtype=torch.float64, requires_grad=True))
    attn, ctx = prims_ref.gst_attn_fwd(q, K, V, 2)
    gc = torch.randn_like(ctx)
    want = torch.autograd.grad(ctx, (q, K, V), gc)
    got = prims_ref.gst_attn_bwd(gc, q.detach(), K.detach(), V.detach(), attn.detach())
    for w_, g_ in zip(want, got):
        assert torch.allclose(w_, g_)


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


def test_strided_conv_relayout_equals_torch_strided_conv(cpu_prims):
    """functional.strided_conv1d (3 taps over rows of s frames) against F.conv1d(stride=s, kernel 2s, padding s/2) with
    weight-norm, forward and all gradients (vqvae2.py:204-212)."""
    Fn = cpu_prims
    for s in (2, 4):
        torch.manual_seed(s)
        Bc, C, Co, Tc = 2, 8, 8, 8 * s
        x = torch.randn(Bc, C, Tc, requires_grad=True)
        v = torch.randn(Co, C, 2 * s, requires_grad=True)
        g = (torch.rand(Co, 1, 1) + 0.5).requires_grad_(True)
        bias = torch.randn(Co, requires_grad=True)
        y = Fn.strided_conv1d(x, v, g, bias, s, act=Fn.ACT_LRELU)
        w = O.wn_weight(g, v)
        want = O2.strided_conv1d(O.leaky_relu(x), w, bias, s)
        assert y.shape == want.shape == (Bc, Co, Tc // s)
        assert rel_err(y, want) < 2e-5
        gy = torch.randn_like(want)
        got = torch.autograd.grad(y, (x, v, g, bias), gy)
        ref = torch.autograd.grad(want, (x, v, g, bias), gy)
        for a_, b_ in zip(got, ref):
            assert rel_err(a_, b_) < 1e-4


def test_model_forward_backward_matches_reference(cpu_prims):
    from vae_npvc_b200.model.vqvae2 import Model
    gold = load_npz('vqvae2_small')
    arch = synth.vqvae2_small_config()
    m = Model(arch)
    sd = synth.vqvae2_state_dict(arch, 777)
    assert list(m.state_dict().keys()) == list(sd.keys())
    m.load_state_dict({k: v.clone() for k, v in sd.items()})
    m.train()
    x, y = synth.vqvae2_batch(arch, B, T, 777)
    xhat, loss, losses = m([x, y])
    loss.backward()
    assert rel_err(xhat, gold['xhat']) < 2e-5
    for k in ('Total', 'VQ loss', 'X like', 'entropy.0', 'entropy.1', 'quanti_err.0', 'quanti_err.1'):
        assert abs(float(losses[k]) - float(gold['loss/' + k])) <= 2e-5 * max(1.0, abs(float(gold['loss/' + k]))), k
    params = dict(m.named_parameters())
    _check_grads(lambda k: params[k].grad, gold, 2e-4)
