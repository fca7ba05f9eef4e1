"""-m gpu: every CUDA primitive (through the C ABI) against the torch-CPU contract restatement
of the same name in oracle/prims_ref.py, on seeded inputs incl. ragged shapes (T not a multiple
of 8, channel counts that do not fill a tile)."""
import pytest
import torch

from conftest import rel_err
from oracle import prims_ref as R

pytestmark = pytest.mark.gpu
TOL = 1e-5          # fp32 path (north_star: <= 1e-5 for fp32)


@pytest.fixture(scope='module')
def Pm():
    from vae_npvc_b200 import _C, prims
    _C.require_cuda()
    return prims


def g(*shape, seed=0, scale=1.0):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed)) * scale


def cu(t):
    return t.cuda() if t is not None else None


def planes_val(pl):
    d = pl.data.float().cpu()                                     # (P, B, T, Cp) channels-last
    return (d.sum(0) if pl.P == 3 else d[0])[..., :pl.C].transpose(1, 2)


def to_cuda_planes(Pm, rp):
    return Pm.Planes(rp.data.cuda(), rp.P, rp.B, rp.C, rp.T, rp.Cp)


SHAPES = [(2, 8, 16, 20, 3, 1), (3, 80, 48, 37, 3, 1), (2, 24, 70, 64, 1, 1), (2, 16, 16, 24, 3, 2), (1, 130, 66, 9, 1, 1)]


@pytest.mark.parametrize('P', [1, 3])
@pytest.mark.parametrize('B,Cin,Cout,T,k,dil', SHAPES)
@pytest.mark.parametrize('transposed', [False, True])
def test_wn_prepare_conv_fwd_wgrad(Pm, P, B, Cin, Cout, T, k, dil, transposed):
    v = g(Cin, Cout, k, seed=1) if transposed else g(Cout, Cin, k, seed=1)
    gg = g(v.shape[0], 1, 1, seed=2).abs() + 0.5
    x, bias, res = g(B, Cin, T, seed=3), g(Cout, seed=4), g(B, Cout, T, seed=5)
    rw = R.wn_prepare(v, gg, transposed, P, need_w32=True)
    cw = Pm.wn_prepare(cu(v), cu(gg), transposed, P, need_w32=True)
    assert torch.equal(cw.wf.cpu().float(), rw.wf.float()) or rel_err(cw.wf.float(), rw.wf.float()) < 1e-2
    assert rel_err(cw.w32, rw.w32) < 1e-6
    # same operand planes into both: the conv contraction itself
    rx = R.Planes.of(x, P)
    cx = to_cuda_planes(Pm, rx)
    ro = R.conv_fwd(rx, rw.wf, k, dil, Cout, bias, 0, res)
    co = Pm.conv_fwd(cx, rw.wf.cuda(), k, dil, Cout, cu(bias), 0, cu(res))
    assert rel_err(co, ro) < TOL
    # data-gradient operand (flipped taps, transposed channels)
    dy = g(B, Cout, T, seed=6)
    rdy = R.Planes.of(dy, P)
    rdx = R.conv_fwd(rdy, rw.wb, k, dil, Cin)
    cdx = Pm.conv_fwd(to_cuda_planes(Pm, rdy), cw.wb, k, dil, Cin)
    assert rel_err(cdx, rdx) < (TOL if P == 3 else 2e-2)
    # accumulate + per-sample bias
    b2 = g(B, Cout, seed=7)
    base = g(B, Cout, T, seed=8)
    ro2 = R.conv_fwd(rx, rw.wf, k, dil, Cout, b2, Cout, None, base.clone(), True)
    co2 = Pm.conv_fwd(cx, rw.wf.cuda(), k, dil, Cout, cu(b2), Cout, None, base.clone().cuda(), True)
    assert rel_err(co2, ro2) < TOL
    # weight gradient in the parameter layout
    rdw = R.conv_wgrad(rdy, rx, k, dil, transposed)
    cdw = Pm.conv_wgrad(to_cuda_planes(Pm, rdy), cx, k, dil, transposed)
    assert tuple(cdw.shape) == tuple(v.shape) and rel_err(cdw, rdw) < TOL
    rdv, rdg = R.wn_backward(rdw, v, gg)
    cdv, cdg = Pm.wn_backward(cdw, cu(v), cu(gg))
    assert rel_err(cdv, rdv) < TOL and rel_err(cdg, rdg) < TOL


def test_conv_matches_autograd_end_to_end(Pm):
    """P=3 conv fwd / dgrad / wgrad against torch autograd of F.conv1d / conv_transpose1d in fp64."""
    import torch.nn.functional as F
    for transposed in (False, True):
        B, Cin, Cout, T, k = 2, 20, 12, 29, 3
        w = (g(Cin, Cout, k, seed=1) if transposed else g(Cout, Cin, k, seed=1)).double().requires_grad_(True)
        x = g(B, Cin, T, seed=2).double().requires_grad_(True)
        y = F.conv_transpose1d(x, w, padding=1) if transposed else F.conv1d(x, w, padding=1)
        dy = g(B, Cout, T, seed=3).double()
        y.backward(dy)
        cw = Pm.wn_prepare(w.detach().float().cuda(), None, transposed, 3)
        xp = Pm.make_planes(x.detach().float().cuda(), 3)[0]
        dyp = Pm.make_planes(dy.float().cuda(), 3)[0]
        assert rel_err(Pm.conv_fwd(xp, cw.wf, k, 1, Cout), y) < TOL
        assert rel_err(Pm.conv_fwd(dyp, cw.wb, k, 1, Cin), x.grad) < TOL
        assert rel_err(Pm.conv_wgrad(dyp, xp, k, 1, transposed), w.grad) < TOL


@pytest.mark.parametrize('P', [1, 3])
def test_planes(Pm, P):
    x = g(3, 7, 21, seed=1, scale=3.0)
    for act in (R.ACT_LRELU, R.ACT_RELU):
        rid, ract = R.make_planes(x, P, 0.7, act, True, True)
        cid, cact = Pm.make_planes(cu(x), P, 0.7, act, True, True)
        assert torch.equal(cid.data.cpu()[..., :7], rid.data[..., :7])
        assert torch.equal(cact.data.cpu()[..., :7], ract.data[..., :7])
    if P == 3:
        assert rel_err(planes_val(cid), x * 0.7) < 1e-7          # hi+mid+lo reproduces fp32
    xb = g(3, 5, 21, seed=2)
    assert torch.equal(Pm.concat_planes(cu(x), cu(xb), P).data.cpu(), R.concat_planes(x, xb, P).data)
    # fused per-channel sums (bias gradients), vector (C % 4 == 0) and scalar channel counts
    for C1, C2 in ((7, 5), (8, 12), (512, 128)):
        xa, xc = g(4, C1, 33, seed=3), g(4, C2, 33, seed=4)
        cs = torch.zeros(C1, device='cuda')
        Pm.make_planes(cu(xa), P, 0.5, R.ACT_NONE, True, False, csum=cs)
        assert rel_err(cs, 0.5 * xa.sum(dim=(0, 2))) < 1e-5
        cs2 = torch.zeros(C1 + C2, device='cuda')
        Pm.concat_planes(cu(xa), cu(xc), P, csum=cs2)
        assert rel_err(cs2, torch.cat([xa, xc], 1).sum(dim=(0, 2))) < 1e-5


@pytest.mark.parametrize('B,C,T', [(2, 8, 20), (3, 64, 37), (2, 512, 128)])
def test_encoder_block_elementwise(Pm, B, C, T):
    P = 3
    h, s, dn = g(B, C, T, seed=1, scale=2.0) + 0.3, g(B, C, T, seed=2), g(B, C, T, seed=3)
    gamma, beta = g(C, seed=4) + 1.0, g(C, seed=5)
    rm, rr = R.gn_stats(h, 1)
    cm, cr = Pm.gn_stats(cu(h), 1)
    assert rel_err(cm, rm) < TOL and rel_err(cr, rr) < TOL
    ry, rypl, rupl = R.enc_out(h, s, rm, rr, gamma, beta, P)
    cy, cypl, cupl = Pm.enc_out(cu(h), cu(s), cu(rm), cu(rr), cu(gamma), cu(beta), P)
    assert rel_err(cy, ry) < TOL and rel_err(planes_val(cypl), ry) < TOL
    assert rel_err(planes_val(cupl), planes_val(rupl)) < TOL
    rR = R.gn_bwd_rows(dn, h, rm, rr, 1)
    cR = Pm.gn_bwd_rows(cu(dn), cu(h), cu(rm), cu(rr), 1)
    assert rel_err(cR, rR) < TOL
    rf = R.gn_bwd_finalize(rR, gamma, rr, T, 1, True)
    cf = Pm.gn_bwd_finalize(cu(rR), cu(gamma), cu(rr), T, 1, True)
    for a, b in zip(cf, rf):
        assert rel_err(a, b) < 5 * TOL
    rdp, rdnp = R.gn_bwd_apply(dn, h, rm, rr, gamma, rf[2], 1, P, True)
    cdp, cdnp = Pm.gn_bwd_apply(cu(dn), cu(h), cu(rm), cu(rr), cu(gamma), cu(rf[2]), 1, P, True)
    assert rel_err(planes_val(cdp), planes_val(rdp)) < 5 * TOL and rel_err(planes_val(cdnp), dn) < 1e-6
    # GroupNorm backward against autograd
    hh = h.double().requires_grad_(True)
    torch.nn.functional.group_norm(hh, 1, gamma.double(), beta.double(), 1e-5).backward(dn.double())
    assert rel_err(planes_val(cdp), hh.grad) < 5 * TOL
    assert rel_err(Pm.act_bwd_add(cu(dn), cu(h), cu(s), R.ACT_LRELU, 1.0), R.act_bwd_add(dn, h, s, R.ACT_LRELU)) < TOL
    assert rel_err(Pm.act_bwd_add(cu(dn), cu(h), None, R.ACT_RELU, 0.3), R.act_bwd_add(dn, h, None, R.ACT_RELU, 0.3)) < TOL
    assert rel_err(Pm.channel_sum(cu(dn)), R.channel_sum(dn)) < TOL


@pytest.mark.parametrize('B,C,S,T', [(2, 8, 5, 20), (3, 64, 24, 37), (2, 512, 128, 128)])
def test_decoder_block_elementwise(Pm, B, C, S, T):
    P = 3
    a, dg = g(B, 2 * C, T, seed=1, scale=1.5), g(B, C, T, seed=2)
    gamma, beta = g(2 * C, seed=3) * 0.3 + 1.0, g(2 * C, seed=4) * 0.3
    rm, rr = R.gn_stats(a, 2)
    cm, cr = Pm.gn_stats(cu(a), 2)
    assert rel_err(cm, rm) < TOL and rel_err(cr, rr) < TOL
    rg = R.dec_gate(a, rm, rr, gamma, beta, P)
    cg = Pm.dec_gate(cu(a), cu(rm), cu(rr), cu(gamma), cu(beta), P)
    assert rel_err(planes_val(cg), planes_val(rg)) < TOL
    rR = R.gate_bwd_rows(dg, a, rm, rr, gamma, beta)
    cR = Pm.gate_bwd_rows(cu(dg), cu(a), cu(rm), cu(rr), cu(gamma), cu(beta))
    assert rel_err(cR, rR) < 5 * TOL
    rf = R.gn_bwd_finalize(rR, gamma, rr, T, 2, True)
    cf = Pm.gn_bwd_finalize(cu(rR), cu(gamma), cu(rr), T, 2, True)
    for x1, x2 in zip(cf, rf):
        assert rel_err(x1, x2) < 5 * TOL
    rda = R.gate_bwd_apply(dg, a, rm, rr, gamma, beta, rf[2], P)
    cda = Pm.gate_bwd_apply(cu(dg), cu(a), cu(rm), cu(rr), cu(gamma), cu(beta), cu(rf[2]), P)
    assert rel_err(planes_val(cda), planes_val(rda)) < 5 * TOL
    # gate + GroupNorm(2) backward against autograd in fp64
    aa = a.double().requires_grad_(True)
    hn = torch.nn.functional.group_norm(aa, 2, gamma.double(), beta.double(), 1e-5)
    (torch.tanh(hn[:, :C]) * torch.sigmoid(hn[:, C:])).backward(dg.double())
    assert rel_err(planes_val(cda), aa.grad) < 5 * TOL
    r, x = g(B, C + S, T, seed=5), g(B, C, T, seed=6)
    rxo, rxp, rsk = R.dec_out(r, x, S, P)
    cxo, cxp, csk = Pm.dec_out(cu(r), cu(x), S, P)
    assert rel_err(cxo, rxo) < TOL and rel_err(csk, rsk) < TOL and rel_err(planes_val(cxp), rxo) < TOL
    r2 = g(B, C + S, T, seed=7)
    R.dec_out(r2, x, S, P, rsk, False)
    Pm.dec_out(cu(r2), cu(x), S, P, csk, False)
    assert rel_err(csk, rsk) < TOL
