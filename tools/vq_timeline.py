"""In-kernel clock64 timeline of vq_fused_kernel (VNB_TC_DBG=9): where CTA 0's time goes."""
import ctypes
import os
import sys

os.environ.setdefault('VNB_TC_DBG', '9')
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from vae_npvc_b200 import _C  # noqa: E402
from vae_npvc_b200 import prims as Pm  # noqa: E402

lib = _C.load()
lib.vnb_debug_tc_timeline.argtypes = [ctypes.c_void_p]
N = 1 << 17
for K, D, which in [(512, 64, 9), (1024, 64, 9), (1024, 64, 10), (1024, 128, 9), (1024, 128, 10)]:
    _C.set_option(_C.OPT_TC_DBG, which)
    torch.manual_seed(0)
    z = Pm.cl_empty(N // 128, D, 128, 'cuda').normal_()
    E = torch.randn(K, D, device='cuda')
    e_hat, e_pl, e_sq = Pm.vq_prep_codebook(E, False)
    for rep in range(3):
        r = Pm.vq_fused(z, False, e_hat, e_pl, e_sq)
    torch.cuda.synchronize()
    buf = (ctypes.c_longlong * 128)()
    assert lib.vnb_debug_tc_timeline(buf) == 0
    t = [buf[64 + i] - buf[64] for i in range(64)]
    print('K %d D %d, %s CTA: prologue done %d' % (K, D, 'first' if which == 9 else 'last', t[1]))
    for ch in range(min(4, (K + 127) // 128)):
        print('  chunk %d  MMA: tmem free %d, mid tile landed %d, hi tile landed %d, all issued %d | epilogue: es staged %d, '
              'accumulator ready %d, scanned %d' % (ch, t[2 + 4 * ch], t[3 + 4 * ch], t[4 + 4 * ch], t[5 + 4 * ch], t[20 + 4 * ch],
                                                    t[21 + 4 * ch], t[22 + 4 * ch]))
    print('  scan done %d, re-rank done %d, gather done / end %d' % (t[40], t[41], t[43]))

# wall time of the same launch against the number of CTA waves, with and without the statistics atomics
def _time(fn, reps=5):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps


_C.set_option(_C.OPT_TC_DBG, 0)
K, D = 1024, 64
E = torch.randn(K, D, device='cuda')
e_hat, e_pl, e_sq = Pm.vq_prep_codebook(E, False)
for ctas in (148, 296, 444, 592, 740, 1024):
    n = ctas * 128
    z = Pm.cl_empty(ctas, D, 128, 'cuda').normal_()
    q = Pm.cl_empty(ctas, D, 128, 'cuda')
    idx = torch.empty(n, dtype=torch.int64, device='cuda')
    sums = torch.zeros(2, device='cuda')
    counts = torch.zeros(K, dtype=torch.int32, device='cuda')
    for label, s_, c_, q_ in (('all outputs', sums, counts, q),):
        dt = _time(lambda: _C.call('vq_fused', z.transpose(1, 2), ctas, D, 128, 0, e_hat, e_pl, e_sq, K, None, None, idx,
                                   q_.transpose(1, 2) if q_ is not None else None, s_, c_))
        print('K %d D %d  %4d CTAs  %-15s %8.1f us' % (K, D, ctas, label, dt))
