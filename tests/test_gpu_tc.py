"""-m gpu: the tcgen05/TMEM/TMA GEMM backend against the SIMT backend and the CPU contract on the SAME
operand planes.  bf16 planes (P=1): both backends contract identical bf16 operands in fp32, so they
must agree to fp32 round-off.  P=3 (hi/mid/lo): fp32-faithful, tolerance stated per assert."""
import pytest
import torch

from conftest import rel_err
from oracle import prims_ref as R

pytestmark = pytest.mark.gpu


@pytest.fixture()
def backends():
    from vae_npvc_b200 import _C, prims
    _C.require_cuda()
    yield _C, prims
    _C.set_backend(_C.BACKEND_AUTO)


def g(*shape, seed=0, scale=1.0):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed)) * scale


CONV_SHAPES = [
    # B, Cin, Cout, T, taps, dil
    (2, 64, 128, 128, 3, 1),
    (2, 80, 512, 128, 3, 1),         # first encoder conv: K = 80 does not fill a 64-wide stage
    (3, 512, 1024, 128, 3, 1),       # decoder conv_in
    (2, 512, 640, 128, 1, 1),        # res/skip 1x1, Cout not a multiple of 128
    (2, 128, 80, 64, 1, 1),          # final layer, BN = 64 tile
    (2, 64, 64, 100, 3, 2),          # ragged T (multiple of 4), dilation 2
    (1, 72, 96, 37, 3, 1),           # T not a multiple of 4 -> scalar stores; pitch padding
    (2, 16, 16, 24, 3, 1),           # tiny (the small golden model)
    (1, 512, 512, 317, 3, 1),        # conversion-length utterance, 3 time tiles
]


@pytest.mark.parametrize('P', [1, 2, 3])
@pytest.mark.parametrize('B,Cin,Cout,T,k,dil', CONV_SHAPES)
def test_conv_fwd_tc_vs_simt(backends, P, B, Cin, Cout, T, k, dil):
    _C, Pm = backends
    x = g(B, Cin, T, seed=1)
    w = g(Cout, Cin, k, seed=2, scale=1.0 / (Cin * k) ** 0.5)
    bias, res = g(B, Cout, seed=3), g(B, Cout, T, seed=4)
    xpl = Pm.make_planes(x.cuda(), P)[0]
    W = Pm.wn_prepare(w.cuda(), None, False, P)
    outs = {}
    for name, be in (('simt', _C.BACKEND_SIMT), ('tc', _C.BACKEND_TC)):
        _C.set_backend(be)
        o1 = Pm.conv_fwd(xpl, W.wf, k, dil, Cout)
        o2 = Pm.conv_fwd(xpl, W.wf, k, dil, Cout, bias.cuda(), Cout, res.cuda(), o1.clone(), True)
        outs[name] = (o1, o2)
    torch.cuda.synchronize()
    tol = 1e-5 if P == 1 else 2e-5
    assert rel_err(outs['tc'][0], outs['simt'][0]) < tol, rel_err(outs['tc'][0], outs['simt'][0])
    assert rel_err(outs['tc'][1], outs['simt'][1]) < tol
    # and against the exact value in float64 (P=3 reproduces fp32 operands; P=1 rounds them to bf16)
    ref = torch.nn.functional.conv1d(x.double(), w.double(), padding=(k - 1) // 2 * dil, dilation=dil)
    err = rel_err(outs['tc'][0], ref)
    assert err < {1: 1e-2, 2: 1e-4, 3: 2e-5}[P], err


def test_conv_tc_data_gradient_operand(backends):
    """dgrad = the same kernel on the wb operand (taps flipped, channels transposed)."""
    _C, Pm = backends
    B, Cin, Cout, T, k = 2, 512, 1024, 128, 3
    wt = g(Cin, Cout, k, seed=1, scale=0.03)          # ConvTranspose1d layout
    dy = g(B, Cout, T, seed=2)
    W = Pm.wn_prepare(wt.cuda(), None, True, 1)
    dypl = Pm.make_planes(dy.cuda(), 1)[0]
    _C.set_backend(_C.BACKEND_SIMT)
    a = Pm.conv_fwd(dypl, W.wb, k, 1, Cin)
    _C.set_backend(_C.BACKEND_TC)
    b = Pm.conv_fwd(dypl, W.wb, k, 1, Cin)
    assert rel_err(b, a) < 1e-5          # K = 3072 products: fp32 summation-order differences only


WGRAD_SHAPES = [
    # B, Cin, Cout, T, taps, dil
    (2, 64, 128, 128, 3, 1),
    (16, 512, 512, 128, 3, 1),       # encoder block k3 (k-split over samples, atomic reduce)
    (4, 512, 1024, 128, 3, 1),       # decoder conv_in: dy has 2C channels
    (3, 512, 640, 128, 1, 1),        # res/skip 1x1: Cout not a multiple of 128
    (2, 80, 512, 100, 3, 1),         # first conv: Cin = 80, ragged T
    (2, 128, 80, 64, 1, 1),          # final layer
    (1, 72, 96, 37, 3, 2),           # ragged everything, dilation 2
    (2, 16, 16, 24, 3, 1),           # the small golden model
    (3, 64, 192, 50, 3, 2),          # tap-fused tiles (columns % 64 == 0) with dilation 2 and ragged T
    (2, 128, 64, 77, 3, 4),          # tap-fused, rows < 128 when transposed, dilation 4
]


@pytest.mark.parametrize('fuse', [0, 1])
@pytest.mark.parametrize('P', [1, 2, 3])
@pytest.mark.parametrize('transposed', [False, True])
@pytest.mark.parametrize('B,Cin,Cout,T,k,dil', WGRAD_SHAPES)
def test_conv_wgrad_tc_vs_simt(backends, fuse, P, transposed, B, Cin, Cout, T, k, dil):
    _C, Pm = backends
    if fuse and (k != 3 or (Cout if transposed else Cin) % 64 != 0):
        pytest.skip('tap-fused tiles need k = 3 and a column count divisible by 64')
    _C.set_option(_C.OPT_WG_FUSE, fuse)
    dy, u = g(B, Cout, T, seed=5), g(B, Cin, T, seed=6)
    dypl, upl = Pm.make_planes(dy.cuda(), P)[0], Pm.make_planes(u.cuda(), P)[0]
    _C.set_backend(_C.BACKEND_SIMT)
    a = Pm.conv_wgrad(dypl, upl, k, dil, transposed)
    _C.set_backend(_C.BACKEND_TC)
    b = Pm.conv_wgrad(dypl, upl, k, dil, transposed)
    torch.cuda.synchronize()
    _C.set_option(_C.OPT_WG_FUSE, 0)
    assert _C.call('debug_tc_timeouts') == 0
    assert a.shape == b.shape == ((Cin, Cout, k) if transposed else (Cout, Cin, k))
    assert rel_err(b, a) < 2e-5, rel_err(b, a)
    pad = (k - 1) // 2 * dil
    up = torch.nn.functional.pad(u.double(), (pad, pad))
    ref = torch.stack([torch.einsum('bot,bit->oi', dy.double(), up[:, :, j * dil:j * dil + T]) for j in range(k)], dim=2)
    if transposed:
        ref = ref.transpose(0, 1).flip(2)
    err = rel_err(b, ref)
    assert err < {1: 1e-2, 2: 1e-4, 3: 2e-5}[P], err


@pytest.mark.parametrize('B,D,T,K,normalize', [(16, 64, 128, 512, True), (16, 64, 128, 512, False), (3, 128, 77, 300, True),
                                               (2, 24, 50, 40, False), (8, 256, 128, 1024, False)])
def test_vq_lookup_tc_vs_simt_and_fp64(backends, B, D, T, K, normalize):
    """Indices of the tcgen05 distance+argmin kernel: identical to the fp32 SIMT kernel and to an fp64
    argmin wherever the top-2 distance gap exceeds fp32 rounding (north_star's criterion)."""
    _C, Pm = backends
    z, E = g(B, D, T, seed=1).cuda(), g(K, D, seed=2).cuda()
    e_hat, e_pl, e_sq = Pm.vq_prep_codebook(E.clone(), normalize, False)
    zn, zpl, z_sq, _ = Pm.vq_prep_z(z, normalize)
    _C.set_backend(_C.BACKEND_SIMT)
    i_simt = Pm.vq_lookup(zn, zpl, z_sq, e_hat, e_pl, e_sq)
    _C.set_backend(_C.BACKEND_TC)
    i_tc = Pm.vq_lookup(zn, zpl, z_sq, e_hat, e_pl, e_sq)
    torch.cuda.synchronize()
    assert _C.call('debug_tc_timeouts') == 0
    zf = zn.double().transpose(1, 2).reshape(B * T, D)
    d = (zf.pow(2).sum(1, keepdim=True) + e_hat.double().pow(2).sum(1).view(1, -1)) - 2 * zf @ e_hat.double().t()
    top2 = d.topk(2, dim=1, largest=False)
    clear = (top2.values[:, 1] - top2.values[:, 0]) > 1e-5 * d.abs().max()
    assert clear.float().mean() > 0.9
    assert torch.equal(i_tc[clear], top2.indices[clear, 0])
    assert torch.equal(i_simt[clear], top2.indices[clear, 0])
    assert (i_tc != i_simt).float().mean() < 1e-3


@pytest.mark.parametrize('be', ['simt', 'tc'])
@pytest.mark.parametrize('B,Cin,Cout,T,k,groups', [(4, 512, 1024, 128, 3, 2), (3, 64, 64, 100, 3, 1), (2, 512, 512, 317, 3, 1)])
def test_conv_fused_groupnorm_statistics(backends, be, B, Cin, Cout, T, k, groups):
    """GroupNorm mean / rstd accumulated in the GEMM epilogue == torch statistics of the conv output."""
    _C, Pm = backends
    _C.set_backend(_C.BACKEND_SIMT if be == 'simt' else _C.BACKEND_TC)
    x = g(B, Cin, T, seed=1) + 0.5
    w = g(Cout, Cin, k, seed=2, scale=1.0 / (Cin * k) ** 0.5)
    bias = g(B, Cout, seed=3)
    xpl = Pm.make_planes(x.cuda(), 3)[0]
    W = Pm.wn_prepare(w.cuda(), None, False, 3)
    out, mean, rstd = Pm.conv_fwd(xpl, W.wf, k, 1, Cout, bias.cuda(), Cout, gn_groups=groups)
    og = out.double().cpu().reshape(B, groups, -1) if out.is_contiguous() else \
        Pm.to_nct(out).double().cpu().reshape(B, groups, -1)
    m = og.mean(2)
    rs = 1.0 / (og.var(2, unbiased=False) + 1e-5).sqrt()
    assert rel_err(mean, m) < 1e-5 and rel_err(rstd, rs) < 1e-5
