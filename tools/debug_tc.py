"""GPU debug helper: tcgen05 conv kernel vs the SIMT backend on a list of shapes (not a test)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from vae_npvc_b200 import _C, prims as Pm

def rel(a, b):
    return float((a.double() - b.double()).norm() / (b.double().norm() + 1e-30))

shapes = [(2, 64, 128, 128, 3, 1, 1), (2, 16, 16, 24, 3, 1, 1), (3, 512, 1024, 128, 3, 1, 1), (2, 80, 512, 128, 3, 1, 3),
          (1, 72, 96, 37, 3, 1, 1)]
if len(sys.argv) > 1:
    shapes = shapes[:int(sys.argv[1])]
for (B, Cin, Cout, T, k, dil, P) in shapes:
    g = torch.Generator().manual_seed(1)
    x = torch.randn(B, Cin, T, generator=g).cuda()
    w = (torch.randn(Cout, Cin, k, generator=g) / (Cin * k) ** 0.5).cuda()
    xpl = Pm.make_planes(x, P)[0]
    W = Pm.wn_prepare(w, None, False, P)
    _C.set_backend(_C.BACKEND_SIMT)
    a = Pm.conv_fwd(xpl, W.wf, k, dil, Cout)
    torch.cuda.synchronize()
    _C.set_backend(_C.BACKEND_TC)
    try:
        b = Pm.conv_fwd(xpl, W.wf, k, dil, Cout)
        torch.cuda.synchronize()
        print((B, Cin, Cout, T, k, dil, P), 'rel err tc vs simt %.3e' % rel(b, a), 'timeouts', _C.call('debug_tc_timeouts'), flush=True)
    except Exception as e:
        print((B, Cin, Cout, T, k, dil, P), 'FAILED', str(e)[:300], flush=True)
        break

# ---- weight gradient ----
for (B, Cin, Cout, T, k, dil, P) in [(2, 64, 128, 128, 3, 1, 1), (16, 512, 512, 128, 3, 1, 1), (2, 16, 16, 24, 3, 1, 3)]:
    g = torch.Generator().manual_seed(2)
    dy, u = torch.randn(B, Cout, T, generator=g).cuda(), torch.randn(B, Cin, T, generator=g).cuda()
    dypl, upl = Pm.make_planes(dy, P)[0], Pm.make_planes(u, P)[0]
    _C.set_backend(_C.BACKEND_SIMT)
    a = Pm.conv_wgrad(dypl, upl, k, dil, False)
    torch.cuda.synchronize()
    _C.set_backend(_C.BACKEND_TC)
    try:
        b = Pm.conv_wgrad(dypl, upl, k, dil, False)
        torch.cuda.synchronize()
        print('wgrad', (B, Cin, Cout, T, k, dil, P), 'rel err tc vs simt %.3e' % rel(b, a), 'timeouts', _C.call('debug_tc_timeouts'), flush=True)
    except Exception as e:
        print('wgrad', (B, Cin, Cout, T, k, dil, P), 'FAILED', str(e)[:300], flush=True)
        break

# ---- VQ lookup ----
for (B, D, T, K) in [(16, 64, 128, 512), (2, 24, 50, 40)]:
    g = torch.Generator().manual_seed(3)
    z, E = torch.randn(B, D, T, generator=g).cuda(), torch.randn(K, D, generator=g).cuda()
    e_hat, e_pl, e_sq = Pm.vq_prep_codebook(E, False, False)
    zn, zpl, z_sq, _ = Pm.vq_prep_z(z, False)
    _C.set_backend(_C.BACKEND_SIMT)
    i0 = Pm.vq_lookup(zn, zpl, z_sq, e_hat, e_pl, e_sq)
    torch.cuda.synchronize()
    _C.set_backend(_C.BACKEND_TC)
    try:
        i1 = Pm.vq_lookup(zn, zpl, z_sq, e_hat, e_pl, e_sq)
        torch.cuda.synchronize()
        print('vq', (B, D, T, K), 'mismatch frac %.2e' % float((i0 != i1).float().mean()), 'timeouts', _C.call('debug_tc_timeouts'), flush=True)
    except Exception as e:
        print('vq', (B, D, T, K), 'FAILED', str(e)[:300], flush=True)
        break
