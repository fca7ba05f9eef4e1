"""Graph-timed sweep of the weight-gradient GEMM (vnb_conv_wgrad) over tile width / cluster split-K (VNB_WG_BN,
VNB_WG_KSPLIT are read once per process: one process per setting) with the clock64 timeline of CTA (0,0,0)."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from vae_npvc_b200 import _C  # noqa: E402
from vae_npvc_b200 import prims as Pm  # noqa: E402

B, T = 16, 128
P = int(os.environ.get('VNB_BENCH_P', 2))
SHAPES = [('dec conv_in wgrad 512->1024 k3 (transposed)', 512, 1024, 3, True), ('enc k3 wgrad 512->512 k3', 512, 512, 3, False),
          ('dec res/skip wgrad 512->640 k1', 512, 640, 1, False), ('enc skip wgrad 512->512 k1', 512, 512, 1, False)]
NAMES = {1: 'setup', 2: 'first stage', 8: 'last load issued', 9: 'last mma issued', 3: 'acc complete', 4: 'staged', 5: 'partials in', 6: 'stored', 7: 'end'}
lib = _C.load()
lib.vnb_debug_tc_timeline.argtypes = [ctypes.c_void_p]
flush = torch.empty(256 * 1024 * 1024 // 4, device='cuda')
tag = 'BN %s ksplit %s' % (os.environ.get('VNB_WG_BN', 'auto'), os.environ.get('VNB_WG_KSPLIT', 'auto'))
for name, Cin, Cout, taps, transposed in SHAPES:
    u = Pm.make_planes(Pm.cl_empty(B, Cin, T, 'cuda').normal_(), P)[0]
    dy = Pm.make_planes(Pm.cl_empty(B, Cout, T, 'cuda').normal_(), P)[0]

    def fn():
        return Pm.conv_wgrad(dy, u, taps, 1, transposed)
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(10):
            fn()
    ts = []
    for _ in range(10):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1) * 100)
    ts.sort()
    _C.set_option(_C.OPT_TC_DBG, int(os.environ.get('VNB_DBG_VAL', 9)))
    fn()
    torch.cuda.synchronize()
    _C.set_option(_C.OPT_TC_DBG, 0)
    buf = (ctypes.c_longlong * 128)()
    lib.vnb_debug_tc_timeline(buf)
    tl = ' | '.join('%s %d' % (NAMES[i], buf[i] - buf[0]) for i in (1, 2, 8, 9, 3, 4, 5, 6, 7) if buf[i] > buf[0])
    fl = 2.0 * B * T * Cin * Cout * taps
    print('P%d %-44s %-22s %7.2f us %6.1f TF/s (algorithmic) | %s' % (P, name, tag, ts[5], fl / ts[5] / 1e6, tl), flush=True)
