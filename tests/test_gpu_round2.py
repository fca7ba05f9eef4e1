"""-m gpu: parity holes named by the round-1 review ("Next round" item 1c / Missing 4-6).

  * the SHIPPED recipe's quantiser at full size: use_ema true (egs/vcc20/vae1/conf/train_pytorch_vqvae.yaml:33-35)
    against the committed ``vcc2020_ema`` golden recorded from the reference
  * Trainer.valid_step / valid (trainer/basic.py:82-103) on the GPU, also after CUDA-graph replays
  * VectorQuantizer's five reductions (layers_vq.py:123-140) against tests/golden/vq_layer.npz
  * codebook sizes 4096 and 16384 (BASELINE.json configs[4]) through the tcgen05 lookup
  * conv shapes of the other named workloads: Cin = 513 (WORLD SP, pitch 520), Cin = 40 (MCC), Cin = 160 / S = 256
  * the 2-plane operand mode ('bf16x2') of every GEMM against float64
"""
import numpy as np
import pytest
import torch

from conftest import load_npz, rel_err
from oracle import prims_ref as R
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
    _C.set_backend(_C.BACKEND_AUTO)


def g(*shape, seed=0, scale=1.0):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed)) * scale


def _model(arch, sd):
    from vae_npvc_b200.model.vqvae import Model
    m = Model(arch)
    m.load_state_dict({k: v.clone() for k, v in sd.items()})
    return m.cuda()


def test_full_vcc2020_ema_matches_reference_golden(fp32_mode, golden_meta):
    """The reference's first training call initialises the EMA codebook with K random rows of the batch
    (layers_vq.py:192-201: _tile + randperm under torch.manual_seed(1234) in oracle/make_golden.py).  The draw is a
    HOST RNG stream, so the test reproduces it on the host from the oracle's encoder output and hands the CUDA model
    the initialised buffers; everything after the draw - lookup with the fresh codebook, EMA update, losses,
    statistics, gradients - is then the reference's first step and is held against its recorded numbers."""
    case = golden_meta['cases']['vcc2020_ema']
    arch = case['arch']
    sd = synth.synthetic_state_dict(arch, 777)
    x, y = synth.synthetic_batch(arch, 16, 128, 777)
    po = O.params_from_state_dict(sd, requires_grad=False)
    with torch.no_grad():
        z = O.encoder_forward(po, arch['encoder'], x)
    zf = z.transpose(1, 2).reshape(-1, z.shape[1])
    torch.manual_seed(1234)
    rows = zf[torch.randperm(zf.shape[0])][:arch['z_num']].clone()
    sd = dict(sd)
    sd['quantizer.emb_init'] = torch.tensor(True)
    sd['quantizer.embeddings'] = rows
    sd['quantizer.emb_sum'] = rows.clone()
    sd['quantizer.emb_elem'] = torch.ones(arch['z_num'])
    m = _model(arch, sd).train()
    xhat, loss, losses = m([x.cuda(), y.cuda()])
    loss.backward()
    for k, v in case['losses'].items():
        assert abs(losses[k] - v) <= 5e-5 * max(1.0, abs(v)), (k, losses[k], v)
    assert abs(float(xhat.double().norm()) - case['xhat_norm']) < 5e-5 * case['xhat_norm']
    seen = 0
    for n, p in m.named_parameters():
        if n in case['grad_norms']:
            gn = case['grad_norms'][n]
            assert abs(float(p.grad.double().norm()) - gn) <= 5e-3 * gn + 1e-12, n
            seen += 1
    assert seen == len(case['grad_norms'])
    # F6 (bug-compatible): the reconstruction loss never reaches the encoder - only beta * z_enc_loss does
    assert float(m.encoder.encode[0].weight_v.grad.abs().max()) > 0


@pytest.mark.parametrize('use_graph', [False, True])
def test_trainer_valid_step_uses_current_parameters(fp32_mode, use_graph):
    """trainer/basic.py:82-103.  valid_step after optimiser steps (eager, and replayed from the captured graph, where
    no Python runs) must evaluate the CURRENT parameters: compared with the oracle's eval-mode forward on the
    parameters read back from the trainer (stale operand planes after graph replays)."""
    from vae_npvc_b200.trainer.basic import Trainer
    arch = synth.small_config()
    cfg = dict(arch)
    cfg.update(model_type='vae_npvc_b200.model.vqvae', precision='fp32', cuda_graph=use_graph, graph_warmup=2,
               learning_rate=5e-3, max_grad_norm=10)
    tr = Trainer(cfg)
    tr.model.load_state_dict(synth.synthetic_state_dict(arch, 777))
    xv, yv = synth.synthetic_batch(arch, 3, 20, 99)
    it = 0
    for i in range(5):
        x, y = synth.synthetic_batch(arch, 3, 20, 500 + i)
        it, _ = tr.train_step([x, y], it)
        if i in (1, 4):
            assert tr.last_step_on_graph == (use_graph and i >= 2)
            detail = tr.valid_step([xv, yv])
            assert tr.model.training
            po = O.params_from_state_dict({k: v.detach().cpu() for k, v in tr.model.state_dict().items()},
                                          requires_grad=False)
            with torch.no_grad():
                _, _, want, _ = O.model_forward(po, arch, xv, yv, False)
            for k, v in want.items():
                assert abs(detail[k] - float(v)) <= 5 * TOL * max(1.0, abs(float(v))), (i, k, detail[k], float(v))
    out = tr.valid([[xv, yv], [xv, yv]])
    assert len(out['Total']) == 2 and all(isinstance(v, float) for v in out['Total'])
    assert abs(out['Total'][0] - out['Total'][1]) < 1e-6 * abs(out['Total'][0])


@pytest.mark.parametrize('normalize', [False, True])
@pytest.mark.parametrize('red', ['sum', 'mean', 'batch_mean', 'frame_mean', 'none'])
def test_vq_layer_all_reductions_gpu(fp32_mode, normalize, red):
    """VectorQuantizer.forward for every reduction (layers_vq.py:123-140) against outputs recorded from the
    reference class (tests/golden/vq_layer.npz)."""
    from vae_npvc_b200.model.layers_vq import VectorQuantizer
    gold = load_npz('vq_layer')
    key = f'vq_n{int(normalize)}_{red}'
    q = VectorQuantizer(16, 8, normalize=normalize, reduction=red).cuda()
    with torch.no_grad():
        q.embeddings.copy_(torch.tensor(gold['E']))
    z = torch.tensor(gold['z']).cuda().requires_grad_(True)
    zv, lq, le, det = q(z)
    (zv.sum() * 0.5 + lq.sum() + 0.3 * le.sum()).backward()
    assert rel_err(zv, gold[key + ':z_vq']) < 1e-6
    assert rel_err(lq, gold[key + ':lq']) < 1e-5 and rel_err(le, gold[key + ':le']) < 1e-5
    assert rel_err(z.grad, gold[key + ':dz']) < 2e-5 and rel_err(q.embeddings.grad, gold[key + ':dE']) < 2e-5
    assert rel_err(q.embeddings.detach(), gold[key + ':E_after']) < 1e-6
    assert abs(float(det['entropy']) - float(gold[key + ':entropy'])) < 1e-4


@pytest.mark.parametrize('K,D,normalize', [(4096, 64, False), (4096, 128, True), (16384, 64, False), (16384, 256, False)])
def test_vq_lookup_large_codebooks(fp32_mode, K, D, normalize):
    """BASELINE.json configs[4]: codebook sweep up to 16384 codes.  Indices equal an fp64 argmin wherever the top-2
    gap exceeds fp32 rounding; a frame equal to a code maps to that code."""
    from vae_npvc_b200 import _C, prims as Pm
    B, T = 8, 128
    z, E = g(B, D, T, seed=1).cuda(), g(K, D, seed=2).cuda()
    e_hat, e_pl, e_sq = Pm.vq_prep_codebook(E.clone(), normalize, False)
    zn, zpl, z_sq, _ = Pm.vq_prep_z(z, normalize)
    _C.set_backend(_C.BACKEND_TC)
    idx = Pm.vq_lookup(zn, zpl, z_sq, e_hat, e_pl, e_sq)
    torch.cuda.synchronize()
    assert _C.call('debug_tc_timeouts') == 0
    zf = zn.double().transpose(1, 2).reshape(B * T, D)
    d = (zf.pow(2).sum(1, keepdim=True) + e_hat.double().pow(2).sum(1).view(1, -1)) - 2 * zf @ e_hat.double().t()
    top2 = d.topk(2, dim=1, largest=False)
    clear = (top2.values[:, 1] - top2.values[:, 0]) > 1e-5 * d.abs().max()
    assert clear.float().mean() > 0.9
    assert torch.equal(idx[clear], top2.indices[clear, 0])
    pick = torch.randint(0, K, (B * T,), generator=torch.Generator().manual_seed(6)).cuda()
    zx = e_hat[pick].view(B, T, D).transpose(1, 2).contiguous()
    zn2, zpl2, z_sq2, _ = Pm.vq_prep_z(zx, False)
    got = Pm.vq_lookup(zn2, zpl2, z_sq2, e_hat, e_pl, e_sq)
    # duplicates of a code cannot occur with random rows; exact self-match has distance 0
    assert torch.equal(got, pick)


WORKLOAD_CONV_SHAPES = [
    # B, Cin, Cout, T, taps   (the first / last convs of the other named workloads)
    (2, 513, 512, 128, 3),       # CDVAE: WORLD spectrum encoder input, channel pitch 520
    (2, 128, 513, 128, 1),       # ... and its decoder's final 1x1
    (2, 40, 512, 128, 3),        # CDVAE: MCC encoder input
    (2, 128, 40, 128, 1),
    (2, 160, 512, 128, 3),       # AISHELL-3: 160-mel input
    (2, 512, 768, 128, 1),       # AISHELL-3: res/skip 1x1 with skip_channels 256
    (2, 256, 160, 128, 1),       # AISHELL-3: final 1x1
]


@pytest.mark.parametrize('P', [1, 2, 3])
@pytest.mark.parametrize('B,Cin,Cout,T,k', WORKLOAD_CONV_SHAPES)
def test_workload_conv_shapes_fwd_dgrad_wgrad(fp32_mode, P, B, Cin, Cout, T, k):
    from vae_npvc_b200 import _C, prims as Pm
    _C.set_backend(_C.BACKEND_TC)
    x = g(B, Cin, T, seed=1)
    w = g(Cout, Cin, k, seed=2, scale=1.0 / (Cin * k) ** 0.5)
    dy = g(B, Cout, T, seed=3)
    xpl = Pm.make_planes(x.cuda(), P)[0]
    dypl = Pm.make_planes(dy.cuda(), P)[0]
    W = Pm.wn_prepare(w.cuda(), None, False, P)
    out = Pm.conv_fwd(xpl, W.wf, k, 1, Cout)
    dx = Pm.conv_fwd(dypl, W.wb, k, 1, Cin)
    dw = Pm.conv_wgrad(dypl, xpl, k, 1, False)
    torch.cuda.synchronize()
    assert _C.call('debug_tc_timeouts') == 0
    xd = x.double().requires_grad_(True)
    wd = w.double().requires_grad_(True)
    ref = torch.nn.functional.conv1d(xd, wd, padding=(k - 1) // 2)
    ref.backward(dy.double())
    tol = {1: 1e-2, 2: 1e-4, 3: 2e-5}[P]
    assert rel_err(out, ref) < tol, rel_err(out, ref)
    assert rel_err(dx, xd.grad) < tol, rel_err(dx, xd.grad)
    assert rel_err(dw, wd.grad) < tol, rel_err(dw, wd.grad)


@pytest.mark.parametrize('name', ['small_vqnorm', 'small_2layer_dil'])
def test_small_model_two_plane_mode(golden_meta, name):
    """'bf16x2' on the small golden models: every tensor within 1e-4 of the reference's outputs."""
    import vae_npvc_b200.functional as Fn
    from vae_npvc_b200 import _C
    _C.require_cuda()
    old = Fn.get_precision()
    Fn.set_precision('bf16x2')
    try:
        case = golden_meta['cases'][name]
        arch = case['arch']
        x, y = synth.synthetic_batch(arch, case['B'], case['T'], 777)
        m = _model(arch, synth.synthetic_state_dict(arch, 777)).train()
        xhat, loss, losses = m([x.cuda(), y.cuda()])
        loss.backward()
        gold = load_npz(name)
        assert rel_err(xhat, gold['xhat']) < 1e-4
        with torch.no_grad():
            assert np.array_equal(m.encode(x.cuda()).cpu().numpy().reshape(-1), gold['idx'])
        for k, v in case['losses'].items():
            assert abs(losses[k] - v) <= 1e-4 * max(1.0, abs(v)), k
        params = dict(m.named_parameters())
        for key, gv in gold.items():
            if key.startswith('grad:'):
                assert rel_err(params[key[5:]].grad, gv) < 3e-4, (key, rel_err(params[key[5:]].grad, gv))
    finally:
        Fn.set_precision(old)


def test_aishell3_shape_matches_oracle(fp32_mode):
    """BASELINE configs[3] / egs/aishell3/vc2/conf/train_pytorch_vqvae.yaml:30-38,53-54: 160-mel, skip 256, codebook
    128 x 128, jitter_p 0.12 (pinned source table, as the reference's np.random stream cannot be reproduced on a GPU)."""
    arch = synth.arch_config(X=160, C=512, D=128, K=128, Y=128, S=256, y_num=218, stacks=10, use_ema=False, jitter_p=0.12)
    B, T = 4, 128
    sd = synth.synthetic_state_dict(arch, 777)
    x, y = synth.synthetic_batch(arch, B, T, 777)
    src = O.jitter_sources(T, 0.12, np.random.RandomState(7))
    m = _model(arch, sd).train()
    m.jitter.sources = lambda length: np.asarray(src)
    xhat, loss, losses = m([x.cuda(), y.cuda()])
    loss.backward()
    po = O.params_from_state_dict(sd)
    oxhat, oloss, olosses, oidx = O.model_forward(po, arch, x, y, True, src)
    oloss.backward()
    assert rel_err(xhat, oxhat) < 5e-5
    for k, v in olosses.items():
        assert abs(losses[k] - v) <= 5e-5 * max(1.0, abs(v)), (k, losses[k], v)
    errs = sorted(rel_err(p.grad, po[n].grad) for n, p in m.named_parameters() if po[n].grad is not None)
    # LeakyReLU-kink floor of full-width layers (oracle/make_golden.py): with only 512 frames ONE slope flip moves an
    # early-encoder gradient by several 1e-3; the bulk of the tensors is at fp32 round-off
    assert errs[len(errs) // 2] < 3e-3 and errs[-1] < 5e-2, (errs[len(errs) // 2], errs[-1])
    with torch.no_grad():
        assert torch.equal(m.encode(x.cuda()).cpu(), O.model_encode(O.params_from_state_dict(sd, requires_grad=False), arch, x))


def test_device_jitter_draws_distribution_and_graph(fp32_mode):
    """layers_vq.py:353-379 on the device: src[t] in {t-1, t, t+1}, boundaries clamp inwards, a frame is replaced with
    probability 1 - p (the reference's inverted flag), left / right equally likely, and every call draws afresh."""
    from vae_npvc_b200 import prims as Pm
    T, p = 4096, 0.12
    state = torch.tensor([12345, 0], dtype=torch.int64, device='cuda')
    a = Pm.jitter_sources(T, p, state)
    b = Pm.jitter_sources(T, p, state)
    assert int(state[1]) == 2 and not torch.equal(a, b)
    t = torch.arange(T, device='cuda')
    for s in (a, b):
        d = s - t
        assert int(d.abs().max()) <= 1 and int(s.min()) >= 0 and int(s.max()) < T
        rep = (d != 0).float().mean().item()
        assert abs(rep - (1 - p)) < 0.03, rep
        right = (d[1:-1] == 1).float().sum().item() / max(1.0, (d[1:-1] != 0).float().sum().item())
        assert abs(right - 0.5) < 0.05, right
        assert int(s[0]) in (0, 1) and int(s[-1]) in (T - 2, T - 1)
    # the Trainer keeps the captured graph with jitter on and draws a new table on every replay
    from vae_npvc_b200.trainer.basic import Trainer
    arch = synth.small_config(jitter_p=0.3)
    cfg = dict(arch, model_type='vae_npvc_b200.model.vqvae', precision='fp32', cuda_graph=True, graph_warmup=2)
    tr = Trainer(cfg)
    tr.model.load_state_dict(synth.synthetic_state_dict(arch, 777))
    x, y = synth.synthetic_batch(arch, 3, 20, 777)
    it, seen = 0, []
    for i in range(5):
        it, detail = tr.train_step([x, y], it)
        seen.append(int(tr.model.jitter._dev_state[1]))
    assert tr.last_step_on_graph and seen == [1, 2, 3, 4, 5]


def test_cdvae_shape_on_gpu_matches_oracle_parts(fp32_mode):
    """BASELINE configs[2]: 513-dim spectrum + 40-dim MCC through the same kernels (channel pitch 520 / 40), two blocks
    per stack to keep the CPU oracle short."""
    from oracle import cdvae_oracle as CO
    from vae_npvc_b200.model.cdvae import Model
    arch = synth.cdvae_config(stacks=2)
    sd = synth.cdvae_state_dict(arch, 777)
    xs, y = synth.cdvae_batch(arch, 2, 128, 777)
    m = Model(arch)
    m.load_state_dict({k: v.clone() for k, v in sd.items()})
    m = m.cuda().train()
    outs, loss, losses = m([t.cuda() for t in xs] + [y.cuda()])
    loss.backward()
    po = O.params_from_state_dict(sd)
    oouts, oloss, olosses, _ = CO.cdvae_forward(po, arch, xs, y)
    oloss.backward()
    for a, b in zip(outs, oouts):
        assert rel_err(a, b) < 5e-5
    for k, v in olosses.items():
        assert abs(losses[k] - v) <= 5e-5 * max(1.0, abs(v)), (k, losses[k], v)
    errs = sorted(rel_err(p.grad, po[n].grad) for n, p in m.named_parameters() if po[n].grad is not None)
    # 2 x 2 x 128 frames: slope flips of near-zero pre-activations dominate (tests/parity.py perturbed_sensitivity)
    assert errs[len(errs) // 2] < 3e-3 and errs[-1] < 5e-2, (errs[len(errs) // 2], errs[-1])


def test_decoder_only_conversion_batch(fp32_mode):
    """BASELINE configs[4]: Model.decode((ids (B,T) int64, speaker ids)) -> (B,X,T), the decoder-only batched conversion
    (vqvae.py:55-60), at T = 512."""
    arch = synth.arch_config(stacks=3)
    sd = synth.synthetic_state_dict(arch, 777)
    B, T = 3, 512
    gen = torch.Generator().manual_seed(5)
    ids = torch.randint(0, arch['z_num'], (B, T), generator=gen)
    y = torch.randint(0, arch['y_num'], (B, 1), generator=gen)
    m = _model(arch, sd).eval()
    with torch.no_grad():
        got = m.decode((ids.cuda(), y.cuda()))
    want = O.model_decode(O.params_from_state_dict(sd, requires_grad=False), arch, ids, y)
    assert tuple(got.shape) == (B, 80, T) and got.is_contiguous()
    assert rel_err(got, want) < 2e-5


def test_batched_conditioning_linear_kernels(fp32_mode):
    """vnb_linear_fwd_batch / vnb_linear_bwd_batch (all decoder blocks' conditioning convs in one launch each way)
    against the per-layer contract."""
    from vae_npvc_b200 import prims as Pm
    L, B, I, O = 10, 16, 128, 1024
    y = g(B, I, seed=1)
    Ws = [g(O, I, seed=10 + l, scale=0.1) for l in range(L)]
    b1 = [g(O, seed=30 + l) for l in range(L)]
    b2 = [g(O, seed=50 + l) if l % 2 else None for l in range(L)]
    dp = g(L, B, O, seed=3)
    want_p = R.linear_fwd_batch(y, Ws, b1, b2)
    cW = [w.cuda() for w in Ws]
    got_p = Pm.linear_fwd_batch(y.cuda(), cW, [b.cuda() for b in b1], [b.cuda() if b is not None else None for b in b2])
    assert rel_err(got_p, want_p) < 1e-5
    rdW = [torch.zeros(O, I) for _ in range(L)]
    rdW[3] = None
    want_dy = R.linear_bwd_batch(dp, y, Ws, rdW)
    cdW = [torch.zeros(O, I, device='cuda') if w is not None else None for w in rdW]
    got_dy = Pm.linear_bwd_batch(dp.cuda(), y.cuda(), cW, cdW)
    assert rel_err(got_dy, want_dy) < 1e-5
    for a, b in zip(cdW, rdW):
        if b is not None:
            assert rel_err(a, b) < 1e-5
    assert Pm.linear_bwd_batch(dp.cuda(), y.cuda(), cW, cdW, want_dy=False) is None
    # odd sizes: B not a multiple of 16, O not a multiple of 64, I not a multiple of 32
    L, B, I, O = 3, 5, 24, 100
    y, dp = g(B, I, seed=4), g(L, B, O, seed=5)
    Ws = [g(O, I, seed=60 + l) for l in range(L)]
    rdW = [torch.zeros(O, I) for _ in range(L)]
    want_dy = R.linear_bwd_batch(dp, y, Ws, rdW)
    cdW = [torch.zeros(O, I, device='cuda') for _ in range(L)]
    got_dy = Pm.linear_bwd_batch(dp.cuda(), y.cuda(), [w.cuda() for w in Ws], cdW)
    assert rel_err(got_dy, want_dy) < 1e-5 and all(rel_err(a, b) < 1e-5 for a, b in zip(cdW, rdW))
    assert rel_err(Pm.linear_fwd_batch(y.cuda(), [w.cuda() for w in Ws], [None] * L, [None] * L),
                   R.linear_fwd_batch(y, Ws, [None] * L, [None] * L)) < 1e-5


@pytest.mark.parametrize('P', [1, 2, 3])
def test_consumers_finalise_groupnorm_statistics(fp32_mode, P):
    """vnb_enc_out_s / vnb_dec_gate_s (statistics finalised inside the consumer) == vnb_gn_finalize + vnb_enc_out /
    vnb_dec_gate, bit for bit, and the stored (mean, rstd) are the finalised ones."""
    from vae_npvc_b200 import prims as Pm
    B, C, T = 5, 256, 72
    x = g(B, C, T, seed=1)
    w1 = g(C, C, 3, seed=2, scale=0.05)
    w2 = g(2 * C, C, 3, seed=3, scale=0.05)
    gamma1, beta1 = g(C, seed=4) + 1.0, g(C, seed=5)
    gamma2, beta2 = g(2 * C, seed=6) + 1.0, g(2 * C, seed=7)
    sk = g(B, C, T, seed=8)
    xpl = Pm.make_planes(x.cuda(), P)[0]
    W1, W2 = Pm.wn_prepare(w1.cuda(), None, False, P), Pm.wn_prepare(w2.cuda(), None, False, P)
    for groups, W, Cout in ((1, W1, C), (2, W2, 2 * C)):
        h_a, mean_a, rstd_a = Pm.conv_fwd(xpl, W.wf, 3, 1, Cout, gn_groups=groups)
        h_b, mean_b, rstd_b = Pm.conv_fwd(xpl, W.wf, 3, 1, Cout, gn_groups=groups, defer_stats=True)
        assert hasattr(mean_b, '_vnb_stats') and torch.equal(h_a, h_b)
        if groups == 1:
            ya = Pm.enc_out(h_a, sk.cuda(), mean_a, rstd_a, gamma1.cuda(), beta1.cuda(), P)
            yb = Pm.enc_out(h_b, sk.cuda(), mean_b, rstd_b, gamma1.cuda(), beta1.cuda(), P)
            assert torch.equal(ya[0], yb[0]) and torch.equal(ya[1].data, yb[1].data) and torch.equal(ya[2].data, yb[2].data)
        else:
            ga = Pm.dec_gate(h_a, mean_a, rstd_a, gamma2.cuda(), beta2.cuda(), P)
            gb = Pm.dec_gate(h_b, mean_b, rstd_b, gamma2.cuda(), beta2.cuda(), P)
            assert torch.equal(ga.data, gb.data)
        assert not hasattr(mean_b, '_vnb_stats')
        assert torch.equal(mean_a, mean_b) and torch.equal(rstd_a, rstd_b)


def test_row_sum_kernels_publish_group_sums(fp32_mode):
    """vnb_gn_bwd_rows_s / vnb_gate_bwd_rows_s: same R as the plain kernels, and S == the group sums
    vnb_gn_bwd_finalize derives from R (so the apply kernels no longer wait for the finalisation)."""
    from vae_npvc_b200 import prims as Pm
    B, C, T = 6, 128, 50
    dn, h = g(B, C, T, seed=1).cuda(), g(B, C, T, seed=2).cuda()
    gamma = (g(C, seed=3) + 1.0).cuda()
    for groups in (1, 2, 4):
        mean, rstd = Pm.gn_stats(h, groups)
        R0 = Pm.gn_bwd_rows(dn, h, mean, rstd, groups)
        R1, S1 = Pm.gn_bwd_rows(dn, h, mean, rstd, groups, gamma)
        assert S1 is not None and torch.equal(R0, R1)
        _, _, S0, _, _ = Pm.gn_bwd_finalize(R0, gamma, rstd, T, groups)
        assert rel_err(S1, S0) < 1e-6
        want = R.gn_bwd_rows(dn.cpu(), h.cpu(), mean.cpu(), rstd.cpu(), groups, gamma.cpu())[1]
        assert rel_err(S1, want) < 1e-5
    Cg = 64
    a, dg = g(B, 2 * Cg, T, seed=4).cuda(), g(B, Cg, T, seed=5).cuda()
    gamma2, beta2 = (g(2 * Cg, seed=6) + 1.0).cuda(), g(2 * Cg, seed=7).cuda()
    mean, rstd = Pm.gn_stats(a, 2)
    for P in (1, 3):
        R0 = Pm.gate_bwd_rows(dg, a, mean, rstd, gamma2, beta2, P)
        R1, S1 = Pm.gate_bwd_rows(dg, a, mean, rstd, gamma2, beta2, P, want_S=True)
        assert S1 is not None and torch.equal(R0, R1)
        _, _, S0, _, _ = Pm.gn_bwd_finalize(R0, gamma2, rstd, T, 2)
        assert rel_err(S1, S0) < 1e-6
    # shapes whose groups are not multiples of 32 channels fall back (S None)
    assert Pm.gn_bwd_rows(dn[:, :48].contiguous(), h[:, :48].contiguous(), *Pm.gn_stats(h[:, :48].contiguous(), 1), 1,
                          gamma[:48].contiguous())[1] is None


@pytest.mark.parametrize('normalize', [False, True])
@pytest.mark.parametrize('B,D,T,K', [(16, 64, 128, 512), (3, 128, 77, 300), (2, 16, 50, 40), (5, 32, 128, 4096),
                                     (4, 64, 100, 1024), (2, 64, 33, 8192)])
def test_vq_fused_lookup_gather(fp32_mode, normalize, B, D, T, K):
    """vnb_vq_fused (three bf16 plane products on the tensor cores as a shortlist + rigorous window + exact fp32 re-rank,
    frames read once, rows gathered in the same kernel) against the contract of prep_z + lookup + gather: indices bit-exact wherever the fp64 top-2
    gap exceeds fp32 rounding (north_star), gathered rows, loss sums, histogram, normalised frames."""
    from vae_npvc_b200 import _C, prims as Pm
    z, E = g(B, D, T, seed=1, scale=1.5), g(K, D, seed=2)
    e_hat, e_pl, e_sq = Pm.vq_prep_codebook(E.clone().cuda(), normalize, False)
    got = Pm.vq_fused(Pm.as_cl(z.cuda()), normalize, e_hat, e_pl, e_sq)
    assert got is not None
    zn, z_norm, idx, q, sums, counts = got
    torch.cuda.synchronize()
    assert _C.call('debug_tc_timeouts') == 0
    rzn, rzpl, rz_sq, rz_norm = R.vq_prep_z(z, normalize)
    assert rel_err(zn, rzn) < 1e-6
    if normalize:
        assert rel_err(z_norm, rz_norm) < 1e-6
    zf = rzn.double().transpose(1, 2).reshape(B * T, D)
    eh = e_hat.double().cpu()
    d = (zf.pow(2).sum(1, keepdim=True) + eh.pow(2).sum(1).view(1, -1)) - 2 * zf @ eh.t()
    top2 = d.topk(2, dim=1, largest=False)
    clear = (top2.values[:, 1] - top2.values[:, 0]) > 1e-5 * d.abs().max()
    assert clear.float().mean() > 0.9
    assert torch.equal(idx.cpu()[clear], top2.indices[clear, 0])
    # whatever a near-tie resolved to, the chosen code is within rounding of the minimum
    chosen = d.gather(1, idx.cpu().view(-1, 1)).squeeze(1)
    assert float((chosen - top2.values[:, 0]).max()) <= 2e-5 * float(d.abs().max())
    # downstream values on the kernel's own indices
    rq, rsums, rcounts = R.vq_gather(idx.cpu(), e_hat.cpu(), rzn, z if normalize else None)
    assert rel_err(q, rq) < 1e-6 and torch.equal(counts.cpu(), rcounts)
    assert abs(float(sums[0]) - float(rsums[0])) <= 1e-5 * abs(float(rsums[0]))
    if normalize:
        assert abs(float(sums[1]) - float(rsums[1])) <= 1e-4 * abs(float(rsums[1])) + 1e-6
    # a frame equal to a code maps to that code (distance 0 to itself)
    pick = torch.randint(0, K, (B * T,), generator=torch.Generator().manual_seed(6)).cuda()
    zx = e_hat[pick].view(B, T, D).transpose(1, 2).contiguous()
    got2 = Pm.vq_fused(Pm.as_cl(zx), False, e_hat, e_pl, e_sq, want_q=False, want_stats=False)
    assert torch.equal(got2[2], pick)


@pytest.mark.parametrize('copies', [2, 3, 5])
def test_vq_fused_duplicate_codes_take_the_lowest_index(fp32_mode, copies):
    """Ties: torch.argmin returns the lowest index (layers_vq.py:107).  Two equal codes go through the exact two-candidate
    re-rank, three or more put a third score inside the window and take the exhaustive fp32 search."""
    from vae_npvc_b200 import _C, prims as Pm
    K, D, B, T = 300, 64, 3, 100
    E = g(K, D, seed=3)
    where = [7, 70, 133, 200, 290][:copies]
    for w in where[1:]:
        E[w] = E[where[0]]
    e_hat, e_pl, e_sq = Pm.vq_prep_codebook(E.clone().cuda(), False, False)
    z = g(B, D, T, seed=4)
    # a third of the frames sit next to the duplicated code
    zz = z.transpose(1, 2).reshape(B * T, D)
    zz[::3] = E[where[0]] + 1e-3 * zz[::3]
    z = zz.view(B, T, D).transpose(1, 2).contiguous()
    got = Pm.vq_fused(Pm.as_cl(z.cuda()), False, e_hat, e_pl, e_sq)
    torch.cuda.synchronize()
    assert _C.call('debug_tc_timeouts') == 0
    idx = got[2].cpu()
    assert bool((idx[::3] == where[0]).all())
    d = (zz.double().pow(2).sum(1, keepdim=True) + E.double().pow(2).sum(1).view(1, -1)) - 2 * zz.double() @ E.double().t()
    top2 = d.topk(2, dim=1, largest=False)
    clear = (top2.values[:, 1] - top2.values[:, 0]) > 1e-5 * d.abs().max()
    assert torch.equal(idx[clear], top2.indices[clear, 0])


def test_vq_fused_leaves_huge_codebooks_to_the_three_kernel_path(fp32_mode):
    from vae_npvc_b200 import prims as Pm
    e_hat, e_pl, e_sq = Pm.vq_prep_codebook(g(16384, 64, seed=2).cuda(), False, False)
    assert Pm.vq_fused(Pm.as_cl(g(2, 64, 33, seed=1).cuda()), False, e_hat, e_pl, e_sq) is None


@pytest.mark.parametrize('P', [1, 2, 3])
@pytest.mark.parametrize('B,Cin,Cout,T,taps', [(16, 512, 512, 128, 3), (4, 128, 80, 128, 1), (3, 24, 20, 37, 3), (16, 1024, 512, 128, 3)])
def test_conv_with_activation_backward_in_the_epilogue(fp32_mode, P, B, Cin, Cout, T, taps):
    """vnb_conv_fwd_mask = vnb_conv_fwd followed by vnb_act_bwd_add (full tiles, ragged tiles, cluster split-K shapes),
    with and without a residual, on both backends."""
    from vae_npvc_b200 import _C, prims as Pm
    x = g(B, Cin, T, seed=1).cuda()
    v = (g(Cout, Cin, taps, seed=2) / (Cin * taps) ** 0.5).cuda()
    W = Pm.wn_prepare(v, None, False, P)
    xpl = Pm.make_planes(Pm.as_cl(x), P)[0]
    m = Pm.as_cl(g(B, Cout, T, seed=3).cuda())
    res = Pm.as_cl(g(B, Cout, T, seed=4).cuda())
    for backend in (_C.BACKEND_AUTO, _C.BACKEND_SIMT):
        _C.set_backend(backend)
        for act, scale, r in [(Pm.ACT_LRELU, 1.0, None), (Pm.ACT_RELU, 0.5, res), (Pm.ACT_LRELU, 1.0, res)]:
            want = Pm.act_bwd_add(Pm.conv_fwd(xpl, W.wf, taps, 1, Cout), m, r, act, scale)
            got = Pm.conv_dgrad_act(xpl, W.wf, taps, 1, Cout, m, act, scale, r)
            torch.cuda.synchronize()
            assert rel_err(got, want) < 1e-6, (backend, act, scale, r is not None)
    assert _C.call('debug_tc_timeouts') == 0


@pytest.mark.parametrize('P', [1, 2, 3])
@pytest.mark.parametrize('Ca,Cb', [(512, 128), (16, 12), (10, 6)])
def test_concat_planes_with_sums_and_missing_first_operand(fp32_mode, P, Ca, Cb):
    """vnb_concat_planes: the channel-sum path (register accumulation) against the restatement, and a null first operand
    (a gradient that does not exist) against explicit zeros."""
    from vae_npvc_b200 import prims as Pm
    B, T = 16, 128
    xa, xb = g(B, Ca, T, seed=1), g(B, Cb, T, seed=2)
    cs = torch.zeros(Ca + Cb, device='cuda')
    got = Pm.concat_planes(Pm.as_cl(xa.cuda()), Pm.as_cl(xb.cuda()), P, csum=cs)
    rcs = torch.zeros(Ca + Cb)
    want = R.concat_planes(xa, xb, P, csum=rcs)
    torch.cuda.synchronize()
    assert torch.equal(got.data.cpu().view(torch.int16), want.data.view(torch.int16))
    assert rel_err(cs, rcs) < 1e-5
    cs0, cs1 = torch.zeros(Ca + Cb, device='cuda'), torch.zeros(Ca + Cb, device='cuda')
    a = Pm.concat_planes(None, Pm.as_cl(xb.cuda()), P, csum=cs0, Ca=Ca)
    b = Pm.concat_planes(Pm.as_cl(torch.zeros(B, Ca, T).cuda()), Pm.as_cl(xb.cuda()), P, csum=cs1)
    assert torch.equal(a.data.view(torch.int16), b.data.view(torch.int16)) and rel_err(cs0, cs1) < 1e-6
    c = Pm.concat_planes(None, Pm.as_cl(xb.cuda()), P, Ca=Ca)            # without sums: the first-generation kernel
    assert torch.equal(c.data.view(torch.int16), b.data.view(torch.int16))
