"""-m gpu: the second-generation streaming kernels (csrc/ew_v2.cuh: make_planes, dec_gate, gn_bwd_apply,
gate_bwd_apply, enc_out behind option VNB_OPT_EW_V2, approximate gate math behind VNB_OPT_FAST_GATE) against the contract
restatement in oracle/prims_ref.py and against the first-generation kernels, for both operand modes, on shapes that
take one loop iteration (the VCC2020 step) and several (more rows than the grid covers), channel counts whose
vector count does and does not divide the CTA size, and a shape the scheme must hand back to the old kernels.

Tolerances: P == 3 (fp32-faithful planes) 5e-5 as for the first generation; P == 1 compares bf16-rounded values,
where an fp32 last-bit difference can flip a rounding: 2e-3 norm-wise; with the approximate gate math
(tanh.approx.f32, 2^-11 absolute) 1e-2 - half the 2e-2 the bf16 mode allows per gradient tensor."""
import pytest
import torch

from conftest import rel_err
from oracle import prims_ref as R

pytestmark = pytest.mark.gpu

# (B, C, T): one iteration; several iterations (rows * C/4 > 1184 * 256 * U); Cv = 20 (grid must be a multiple of 5);
# C % 4 != 0 (falls back to the first generation)
SHAPES = [(2, 512, 128), (40, 512, 300), (3, 80, 37), (2, 24, 20), (2, 14, 9)]


@pytest.fixture(scope='module')
def Pm():
    from vae_npvc_b200 import _C, prims
    _C.require_cuda()
    return prims


@pytest.fixture()
def opts():
    from vae_npvc_b200 import _C
    old = (_C.get_option(_C.OPT_EW_V2), _C.get_option(_C.OPT_FAST_GATE))

    def set_(v2, fast=0):
        _C.set_option(_C.OPT_EW_V2, v2)
        _C.set_option(_C.OPT_FAST_GATE, fast)
    yield set_
    _C.set_option(_C.OPT_EW_V2, old[0])
    _C.set_option(_C.OPT_FAST_GATE, old[1])


def g(*shape, seed=0, scale=1.0):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed)) * scale


def val(pl):
    d = pl.data.float().cpu()
    return (d.sum(0) if pl.P == 3 else d[0])[..., :pl.C].transpose(1, 2)


def tol(P, fast=False):
    return 5e-5 if P == 3 else (1e-2 if fast else 2e-3)


def test_options_round_trip(opts):
    from vae_npvc_b200 import _C
    opts(1, 1)
    assert _C.get_option(_C.OPT_EW_V2) == 1 and _C.get_option(_C.OPT_FAST_GATE) == 1
    opts(0, 0)
    assert _C.get_option(_C.OPT_EW_V2) == 0 and _C.get_option(_C.OPT_FAST_GATE) == 0


@pytest.mark.parametrize('P', [1, 3])
@pytest.mark.parametrize('B,C,T', SHAPES)
def test_make_planes_v2(Pm, opts, P, B, C, T):
    x = g(B, C, T, seed=1, scale=3.0)
    xc = Pm.as_cl(x.cuda())
    for act in (R.ACT_NONE, R.ACT_LRELU, R.ACT_RELU):
        opts(0)
        a1, b1 = Pm.make_planes(xc, P, 0.7, act, True, True)
        opts(1)
        a2, b2 = Pm.make_planes(xc, P, 0.7, act, True, True)
        assert torch.equal(a1.data, a2.data) and torch.equal(b1.data, b2.data)      # same arithmetic: same bits
        only = Pm.make_planes(xc, P, 0.7, act, False, True)[1]
        assert torch.equal(only.data, b1.data)
    ra, rb = R.make_planes(x, P, 0.7, R.ACT_RELU, True, True)
    assert rel_err(val(a2), val(ra)) < tol(P) and rel_err(val(b2), val(rb)) < tol(P)


@pytest.mark.parametrize('P', [1, 3])
@pytest.mark.parametrize('B,C,T', SHAPES)
def test_enc_out_v2(Pm, opts, P, B, C, T):
    h, sk = g(B, C, T, seed=1, scale=2.0) + 0.3, g(B, C, T, seed=2)
    gamma, beta = g(C, seed=3) * 0.3 + 1.0, g(C, seed=4) * 0.5
    rm, rr = R.gn_stats(h, 1)
    hc, sc = Pm.as_cl(h.cuda()), Pm.as_cl(sk.cuda())
    args = (rm.cuda(), rr.cuda(), gamma.cuda(), beta.cuda())
    for skip, rs in ((sc, sk), (None, None)):
        ry, rypl, rupl = R.enc_out(h, rs, rm, rr, gamma, beta, P)
        out = {}
        for v2 in (0, 1):
            opts(v2)
            out[v2] = Pm.enc_out(hc, skip, *args, P, True, True, True)
            assert rel_err(out[v2][0], ry) < 5e-5
            assert rel_err(val(out[v2][1]), val(rypl)) < tol(P) and rel_err(val(out[v2][2]), val(rupl)) < tol(P)
        assert rel_err(out[1][0], out[0][0]) < 1e-6
        # outputs are individually optional
        opts(1)
        y, ypl, upl = Pm.enc_out(hc, skip, *args, P, False, False, True)
        assert y is None and ypl is None and rel_err(val(upl), val(out[1][2])) < 1e-6


@pytest.mark.parametrize('P,fast', [(1, 0), (1, 1), (3, 0), (3, 1)])
@pytest.mark.parametrize('B,C,T', SHAPES)
def test_gate_kernels_v2(Pm, opts, P, fast, B, C, T):
    a, dg = g(B, 2 * C, T, seed=1, scale=1.5) + 0.2, g(B, C, T, seed=2)
    gamma, beta = g(2 * C, seed=3) * 0.3 + 1.0, g(2 * C, seed=4) * 0.5
    rm, rr = R.gn_stats(a, 2)
    rR = R.gate_bwd_rows(dg, a, rm, rr, gamma, beta)
    S = R.gn_bwd_finalize(rR, gamma, rr, T, 2, True)[2]
    rg = R.dec_gate(a, rm, rr, gamma, beta, P)
    rda = R.gate_bwd_apply(dg, a, rm, rr, gamma, beta, S, P)
    ac, dgc = Pm.as_cl(a.cuda()), Pm.as_cl(dg.cuda())
    args = (rm.cuda(), rr.cuda(), gamma.cuda(), beta.cuda())
    opts(1, fast)
    cg = Pm.dec_gate(ac, *args, P)
    cda = Pm.gate_bwd_apply(dgc, ac, *args, S.cuda(), P)
    approx = bool(fast) and P == 1                  # P == 3 never uses the approximate functions
    assert rel_err(val(cg), val(rg)) < tol(P, approx)
    assert rel_err(val(cda), val(rda)) < tol(P, approx)
    opts(0)
    og = Pm.dec_gate(ac, *args, P)
    oda = Pm.gate_bwd_apply(dgc, ac, *args, S.cuda(), P)
    assert rel_err(val(cg), val(og)) < tol(P, approx) and rel_err(val(cda), val(oda)) < tol(P, approx)
    if not approx and P == 3:
        assert rel_err(val(cg), val(og)) < 1e-6 and rel_err(val(cda), val(oda)) < 1e-6


@pytest.mark.parametrize('P', [1, 3])
@pytest.mark.parametrize('groups', [1, 2])
@pytest.mark.parametrize('B,C,T', SHAPES)
def test_gn_bwd_apply_v2(Pm, opts, P, groups, B, C, T):
    h, dn = g(B, C, T, seed=1, scale=2.0) + 0.3, g(B, C, T, seed=2)
    gamma = g(C, seed=3) * 0.3 + 1.0
    rm, rr = R.gn_stats(h, groups)
    S = R.gn_bwd_finalize(R.gn_bwd_rows(dn, h, rm, rr, groups), gamma, rr, T, groups, True)[2]
    rdp, rdn = R.gn_bwd_apply(dn, h, rm, rr, gamma, S, groups, P, True)
    hc, dnc = Pm.as_cl(h.cuda()), Pm.as_cl(dn.cuda())
    out = {}
    for v2 in (0, 1):
        opts(v2)
        out[v2] = Pm.gn_bwd_apply(dnc, hc, rm.cuda(), rr.cuda(), gamma.cuda(), S.cuda(), groups, P, True)
        assert rel_err(val(out[v2][0]), val(rdp)) < tol(P) and rel_err(val(out[v2][1]), val(rdn)) < tol(P)
        assert Pm.gn_bwd_apply(dnc, hc, rm.cuda(), rr.cuda(), gamma.cuda(), S.cuda(), groups, P, False)[1] is None
    assert torch.equal(out[0][1].data, out[1][1].data)                    # planes of dn: a pure conversion
    assert rel_err(val(out[1][0]), val(out[0][0])) < (1e-6 if P == 3 else tol(P))


@pytest.mark.parametrize('precision', ['bf16', 'fp32'])
def test_train_step_v2_matches_v1(Pm, opts, precision):
    """One optimiser step of the small model with the second-generation kernels (+ approximate gate) against the same
    step on the first generation: the step is the same function of the same inputs."""
    from oracle import synth
    from vae_npvc_b200 import functional as Fn
    from vae_npvc_b200.trainer.basic import Trainer
    old = Fn.get_precision()
    try:
        arch = synth.small_config(C=32, S=16)
        cfg = dict(arch, model_type='vae_npvc_b200.model.vqvae', precision=precision, cuda_graph=False,
                   learning_rate=1e-3, max_grad_norm=10)
        x, y = synth.synthetic_batch(arch, 4, 24, 777)
        res = []
        for v2 in (0, 1):
            opts(v2, v2)
            tr = Trainer(cfg)
            tr.model.load_state_dict(synth.synthetic_state_dict(arch, 777))
            _, detail = tr.train_step([x, y], 0)
            res.append((detail, {n: p.detach().clone() for n, p in tr.model.named_parameters()}))
        t = 1e-5 if precision == 'fp32' else 2e-3
        for k, v in res[0][0].items():
            assert abs(res[1][0][k] - v) <= t * max(1.0, abs(v)), k
        worst = max(rel_err(res[1][1][n], p) for n, p in res[0][1].items())
        assert worst < (1e-5 if precision == 'fp32' else 5e-3)
    finally:
        Fn.set_precision(old)
