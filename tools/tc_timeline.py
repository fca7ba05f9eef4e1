"""In-kernel clock64 timeline of conv_fwd_tc_kernel (VNB_TC_DBG=9): where CTA (0,0,0)'s time goes."""
import ctypes
import os
import sys

os.environ.setdefault('VNB_TC_DBG', '9')
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from vae_npvc_b200 import _C  # noqa: E402
from vae_npvc_b200 import prims as Pm  # noqa: E402

B, T, P = 16, 128, 1
lib = _C.load()
lib.vnb_debug_tc_timeline.argtypes = [ctypes.c_void_p]
NAMES = ['start', 'setup done', 'first stage landed', 'accumulator complete', 'staged in smem', 'cluster barrier',
         'reduced + stored', 'end']
for (Cin, Cout, taps) in [(512, 512, 1), (512, 512, 3), (1024, 512, 3), (512, 1024, 3), (512, 640, 1)]:
    x = torch.randn(B, Cin, T, device='cuda')
    v = torch.randn(Cout, Cin, taps, device='cuda') * 0.02
    W = Pm.wn_prepare(v, None, False, P)
    xpl = Pm.make_planes(x, P)[0]
    out = Pm.cl_empty(B, Cout, T, 'cuda')
    for rep in range(3):
        Pm.conv_fwd(xpl, W.wf, taps, 1, Cout, None, 0, None, out)
    torch.cuda.synchronize()
    buf = (ctypes.c_longlong * 128)()
    assert lib.vnb_debug_tc_timeline(buf) == 0
    t0 = buf[0]
    print('Cin %d Cout %d k%d (BN %s, ksplit %s):' % (Cin, Cout, taps, os.environ.get('VNB_TC_BN', 'auto'),
                                                      os.environ.get('VNB_TC_KSPLIT', 'auto')),
          ' | '.join('%s %d' % (NAMES[i], buf[i] - t0) for i in range(1, 8)), '| rerun reduce', buf[9] - buf[8])
