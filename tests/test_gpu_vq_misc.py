"""-m gpu: vector-quantiser, conditioning, loss and optimiser primitives against
oracle/prims_ref.py, plus the domain's size-independent properties at full size."""
import pytest
import torch

from conftest import rel_err
from oracle import prims_ref as R

pytestmark = pytest.mark.gpu
TOL = 1e-5


@pytest.fixture(scope='module')
def Pm():
    from vae_npvc_b200 import _C, prims
    _C.require_cuda()
    return prims


def g(*shape, seed=0, scale=1.0):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed)) * scale


def cu(t):
    return t.cuda() if t is not None else None


def _gap_mask(zn, e_hat, z_sq, e_sq, min_gap=1e-5):
    """frames whose top-2 distance gap exceeds fp32 round-off (north_star: indices bit-exact there)"""
    B, D, T = zn.shape
    zf = zn.double().transpose(1, 2).reshape(-1, D)
    d = (zf.pow(2).sum(1, keepdim=True) + e_hat.double().pow(2).sum(1)) - 2 * zf @ e_hat.double().t()
    top2 = d.topk(2, dim=1, largest=False).values
    return (top2[:, 1] - top2[:, 0]) > min_gap, d.argmin(1)


@pytest.mark.parametrize('normalize', [False, True])
@pytest.mark.parametrize('B,D,T,K', [(2, 8, 11, 16), (3, 64, 37, 512), (16, 64, 128, 512), (2, 128, 50, 1000)])
def test_vq_forward_backward(Pm, normalize, B, D, T, K):
    z, E = g(B, D, T, seed=1, scale=2.0), g(K, D, seed=2)
    Er, Ec = E.clone(), E.clone().cuda()
    re_hat, re_pl, re_sq = R.vq_prep_codebook(Er, normalize, True)
    ce_hat, ce_pl, ce_sq = Pm.vq_prep_codebook(Ec, normalize, True)
    assert rel_err(Ec, Er) < 1e-6 and rel_err(ce_hat, re_hat) < 1e-6 and rel_err(ce_sq, re_sq) < 1e-6
    assert rel_err(ce_pl.float().sum(0)[:, :D], re_hat) < 1e-6
    rzn, rzpl, rz_sq, rz_norm = R.vq_prep_z(z, normalize)
    czn, czpl, cz_sq, cz_norm = Pm.vq_prep_z(cu(z), normalize)
    assert rel_err(czn, rzn) < 1e-6 and rel_err(cz_sq, rz_sq) < 1e-6 and rel_err(cz_norm, rz_norm) < 1e-6
    ridx = R.vq_lookup(rzn, rzpl, rz_sq, re_hat, re_pl, re_sq)
    cidx = Pm.vq_lookup(czn, czpl, cz_sq, ce_hat, ce_pl, ce_sq)
    assert cidx.dtype == torch.int64
    safe, idx64 = _gap_mask(rzn, re_hat, rz_sq, re_sq)
    assert torch.equal(cidx.cpu()[safe], ridx[safe]) and torch.equal(cidx.cpu()[safe], idx64[safe])
    assert float(safe.float().mean()) > 0.99
    # downstream on the reference's indices so a near-tie cannot poison the comparison
    rq, rsums, rcounts = R.vq_gather(ridx, re_hat, rzn, z if normalize else None)
    cq, csums, ccounts = Pm.vq_gather(cu(ridx), ce_hat, czn, cu(z) if normalize else None)
    assert rel_err(cq, rq) < 1e-6 and rel_err(csums, rsums) < TOL and torch.equal(ccounts.cpu(), rcounts)
    scale = 1.0 / (B * T)
    rs = R.vq_finalize(rsums, rcounts, B * T, scale, normalize, 1e-10)
    cs = Pm.vq_finalize(cu(rsums), cu(rcounts), B * T, scale, normalize, 1e-10)
    assert rel_err(cs, rs) < TOL
    g_out, g_lq, g_le = g(B, D, T, seed=3), torch.tensor(0.7), torch.tensor(0.01)
    rdz, rdEh = R.vq_bwd(g_out, g_lq, g_le, ridx, z, rzn, rz_norm, re_hat, Er, scale, normalize)
    cdz, cdEh = Pm.vq_bwd(cu(g_out), cu(g_lq), cu(g_le), cu(ridx), cu(z), czn, cz_norm, ce_hat, Ec, scale, normalize)
    assert rel_err(cdz, rdz) < 5 * TOL and rel_err(cdEh, rdEh) < 5 * TOL
    assert rel_err(Pm.vq_bwd_codebook(cu(rdEh), Ec, normalize), R.vq_bwd_codebook(rdEh, Er, normalize)) < 5 * TOL


def test_vq_lookup_properties_full_size(Pm):
    """BASELINE config 5 corner of the sweep at N = 2^17 frames: (1) a frame equal to a code maps to
    that code; (2) idempotence: looking up the gathered rows returns the same indices; (3) the
    histogram sums to N."""
    K, D, B, T = 512, 64, 512, 256
    E = g(K, D, seed=5).cuda()
    e_hat, e_pl, e_sq = Pm.vq_prep_codebook(E, False, False)
    pick = torch.randint(0, K, (B * T,), generator=torch.Generator().manual_seed(6)).cuda()
    zexact = E[pick].view(B, T, D).transpose(1, 2).contiguous()
    zn, zpl, z_sq, _ = Pm.vq_prep_z(zexact, False)
    assert torch.equal(Pm.vq_lookup(zn, zpl, z_sq, e_hat, e_pl, e_sq), pick)
    z = g(B, D, T, seed=7).cuda()
    zn, zpl, z_sq, _ = Pm.vq_prep_z(z, False)
    idx = Pm.vq_lookup(zn, zpl, z_sq, e_hat, e_pl, e_sq)
    q, sums, counts = Pm.vq_gather(idx, e_hat, zn, None)
    qn, qpl, q_sq, _ = Pm.vq_prep_z(q, False)
    assert torch.equal(Pm.vq_lookup(qn, qpl, q_sq, e_hat, e_pl, e_sq), idx)
    assert int(counts.sum()) == B * T and int(idx.min()) >= 0 and int(idx.max()) < K
    # minimality against a brute-force fp32 distance matrix on a slice
    zf = z[:4].transpose(1, 2).reshape(-1, D)
    d = (zf.pow(2).sum(1, keepdim=True) + E.pow(2).sum(1)) - 2 * zf @ E.t()
    chosen = d.gather(1, idx[:4 * T].view(-1, 1)).squeeze(1)
    assert float((chosen - d.min(1).values).abs().max()) < 1e-4


def test_ema_update_and_jitter(Pm):
    B, D, T, K = 3, 8, 20, 32
    z = g(B, D, T, seed=1)
    idx = torch.randint(0, K - 4, (B * T,), generator=torch.Generator().manual_seed(2))   # last 4 codes unused
    rz, rc = R.vq_ema_scatter(idx, z, K)
    cz, cc = Pm.vq_ema_scatter(cu(idx), cu(z), K)
    assert rel_err(cz, rz) < TOL and torch.equal(cc.cpu(), rc)
    rows = g(K, D, seed=3)
    bufs = [g(K, D, seed=4), torch.full((K,), 1.05), g(K, D, seed=5)]
    rb = [b.clone() for b in bufs]
    cb = [b.clone().cuda() for b in bufs]
    rs = R.vq_ema_update(rz, rc, B * T, 0.9, 1.0, rows, *rb)
    cs = Pm.vq_ema_update(cu(rz), cu(rc), B * T, 0.9, 1.0, cu(rows), *cb)
    assert rel_err(cs, rs) < TOL
    for a, b in zip(cb, rb):
        assert rel_err(a, b) < TOL
    assert float(rs[2]) < K                      # some codes died -> restart rows were used
    perm = torch.randperm(B * T * 2, generator=torch.Generator().manual_seed(6))[:K]
    assert rel_err(Pm.vq_frame_rows(cu(z), cu(perm), K), R.vq_frame_rows(z, perm, K)) < 1e-7
    src = torch.tensor([1, 1, 2, 2, 5, 5] + list(range(6, T)))
    assert torch.equal(Pm.jitter_fwd(cu(z), cu(src)).cpu(), R.jitter_fwd(z, src))
    assert torch.equal(Pm.jitter_bwd(cu(z), cu(src)).cpu(), R.jitter_bwd(z, src))


def test_conditioning_loss_optimiser(Pm):
    table, idx = g(7, 12, seed=1), torch.tensor([[3], [0], [3], [6]])
    assert torch.equal(Pm.embed_fwd(cu(table), cu(idx)).cpu(), R.embed_fwd(table, idx))
    dout = g(4, 1, 12, seed=2)
    assert rel_err(Pm.embed_bwd(cu(dout), cu(idx), 7), R.embed_bwd(dout, idx, 7)) < 1e-6
    y, W, b1, b2, dp = g(4, 12, seed=3), g(30, 12, seed=4), g(30, seed=5), g(30, seed=6), g(4, 30, seed=7)
    assert rel_err(Pm.linear_fwd(cu(y), cu(W), cu(b1), cu(b2)), R.linear_fwd(y, W, b1, b2)) < TOL
    rdW, rdy = R.linear_bwd(dp, y, W)
    cdW, cdy = Pm.linear_bwd(cu(dp), cu(y), cu(W))
    assert rel_err(cdW, rdW) < TOL and rel_err(cdy, rdy) < TOL
    acc = g(4, 12, seed=8)
    _, rdy2 = R.linear_bwd(dp, y, W, acc.clone())
    _, cdy2 = Pm.linear_bwd(cu(dp), cu(y), cu(W), acc.clone().cuda())
    assert rel_err(cdy2, rdy2) < TOL
    xh, x = g(16, 80, 128, seed=9), g(16, 80, 128, seed=10)
    sc = 1.0 / (16 * 128)
    assert rel_err(Pm.logloss_fwd(cu(xh), cu(x), sc), R.logloss_fwd(xh, x, sc)) < TOL
    gsc = torch.tensor(0.5)
    assert rel_err(Pm.logloss_bwd(cu(xh), cu(x), cu(gsc), sc), R.logloss_bwd(xh, x, gsc, sc)) < 1e-6
    # clip + Adam + StepLR over a flat buffer, 5 steps, step_size 2
    n = 100003
    p, m, v = g(n, seed=11), torch.zeros(n), torch.zeros(n)
    rp, rm, rv, rst = p.clone(), m.clone(), v.clone(), torch.zeros(2)
    cp, cm, cv, cst = p.clone().cuda(), m.clone().cuda(), v.clone().cuda(), torch.zeros(2).cuda()
    for it in range(5):
        gr = g(n, seed=20 + it, scale=0.01 * (it + 1))
        rs, cs = torch.zeros(1), torch.zeros(1).cuda()
        R.sqnorm(gr, rs)
        Pm.sqnorm(cu(gr), cs)
        assert rel_err(cs, rs) < TOL
        R.adam_step(rp, gr, rm, rv, rs, 0.5, 1e-3, 0.5, 0.999, 1e-8, rst, 2, 0.5)
        Pm.adam_step(cp, cu(gr), cm, cv, cs, 0.5, 1e-3, 0.5, 0.999, 1e-8, cst, 2, 0.5)
    assert rel_err(cp, rp) < 1e-6 and rel_err(cm, rm) < TOL and rel_err(cv, rv) < TOL
    assert cst.tolist() == [5.0, 5.0]


def test_no_cpu_path_and_errors(Pm):
    from vae_npvc_b200 import _C
    with pytest.raises(RuntimeError, match='no CPU path'):
        Pm.gn_stats(torch.zeros(1, 2, 3), 1)
    with pytest.raises(RuntimeError, match='bad'):
        _C.call('gn_stats', torch.zeros(1, 3, 3).cuda(), 1, 3, 3, 2, 1e-5, torch.zeros(256).cuda(), torch.zeros(2).cuda(),
                 torch.zeros(2).cuda())
    assert 'gn_stats' in _C.last_error()
