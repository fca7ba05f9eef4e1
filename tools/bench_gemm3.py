"""Graph-timed sweep of the conv GEMM kernels over tile / pair / stage settings, with the in-kernel timeline.

    python tools/bench_gemm3.py [tag]

Each configuration: `reps` launches captured in ONE CUDA graph (no host launch cost between kernels - what the captured
training step sees), L2 flushed before every replay, CUDA events around the replay.  Also prints the clock64 timeline of
CTA (0,0,0) (VNB_OPT_TC_DBG = 9) of one eager launch and, for comparison, torch.matmul (cuBLAS) at the same GEMM shape
timed the same way.
"""
import ctypes
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from vae_npvc_b200 import _C  # noqa: E402
from vae_npvc_b200 import prims as Pm  # noqa: E402

B, T = 16, 128
SHAPES = [  # (name, Cin, Cout, taps, transposed, gn_groups)
    ('dec conv_in fwd 512->1024 k3 +GN', 512, 1024, 3, True, 2),
    ('dec conv_in dgrad 1024->512 k3', 1024, 512, 3, False, 0),
    ('enc k3 fwd 512->512 k3 +GN', 512, 512, 3, False, 1),
    ('enc k1 skip 512->512 k1', 512, 512, 1, False, 0),
    ('dec res/skip 512->640 k1', 512, 640, 1, False, 0),
]
TS_NAMES = {1: 'setup', 2: 'first stage', 8: 'last load issued', 9: 'last mma issued', 3: 'acc complete', 4: 'staged',
            6: 'stored', 7: 'end'}


def graph_time(fn, flush, reps=10, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3 / reps)
    ts.sort()
    return ts[len(ts) // 2]


def timeline(fn):
    lib = _C.load()
    lib.vnb_debug_tc_timeline.argtypes = [ctypes.c_void_p]
    _C.set_option(_C.OPT_TC_DBG, 9)
    fn()
    torch.cuda.synchronize()
    _C.set_option(_C.OPT_TC_DBG, 0)
    buf = (ctypes.c_longlong * 128)()
    lib.vnb_debug_tc_timeline(buf)
    t0 = buf[0]
    return ' | '.join('%s %d' % (TS_NAMES[i], buf[i] - t0) for i in (1, 2, 8, 9, 3, 4, 6, 7) if buf[i] > t0)


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else 'r2'
    global PS
    PS = [int(v) for v in sys.argv[2].split(',')] if len(sys.argv) > 2 else [1, 2, 3]
    _C.require_cuda()
    flush = torch.empty(256 * 1024 * 1024 // 4, device='cuda')
    rows = []
    for P in PS:
        for name, Cin, Cout, taps, transposed, gn in SHAPES:
            x = torch.randn(B, Cin, T, device='cuda')
            v = torch.randn((Cin, Cout, taps) if transposed else (Cout, Cin, taps), device='cuda') * 0.02
            W = Pm.wn_prepare(v, None, transposed, P)
            xpl = Pm.make_planes(x, P)[0]
            out = Pm.cl_empty(B, Cout, T, 'cuda')
            sums = torch.zeros((B, max(gn, 1), 2), dtype=torch.float64, device='cuda')

            def fwd():
                _C.call('conv_fwd', xpl.data, P, B, Cin, T, W.wf, taps, 1, Cout, None, 0, None, out.transpose(1, 2), 0,
                        sums if gn else None, gn)

            fl = 2.0 * B * T * Cin * Cout * taps
            configs = [('v2 (round 1)', 0, 0, -1, 0), ('v3 auto', 2, 0, -1, 0)]
            if P == 1 or 'conv_in fwd' in name or 'enc k3' in name or os.environ.get('VNB_SWEEP_ALL'):
                for bn in (64, 128, 256):
                    for pair in (0, 1):
                        configs.append(('v3 BN%d %s' % (bn, 'pair' if pair else 'single'), 2, bn, pair, 0))
            if P == 1 or os.environ.get('VNB_SWEEP_ALL'):
                configs += [('v3 BN128 single 100KB', 2, 128, 0, 100), ('v3 BN64 single 100KB', 2, 64, 0, 100),
                            ('v3 BN128 pair 100KB', 2, 128, 1, 100), ('v3 BN64 pair 100KB', 2, 64, 1, 100)]
            for cname, mode, bn, pair, st in configs:
                _C.set_option(_C.OPT_GEMM_V3, mode)
                _C.set_option(_C.OPT_TC3_BN, bn)
                _C.set_option(_C.OPT_TC3_PAIR, pair)
                _C.set_option(_C.OPT_TC3_SMEM_KB, st)
                try:
                    us = graph_time(fwd, flush)
                    tl = timeline(fwd)
                except RuntimeError as e:
                    print('P%d %-34s %-18s failed: %s' % (P, name, cname, str(e)[:80]))
                    continue
                rows.append(dict(tag=tag, P=P, shape=name, config=cname, us=round(us, 2), tflops=round(fl / us / 1e6, 1)))
                print('P%d %-34s %-18s %7.2f us  %6.1f TF/s (algorithmic)  | %s' % (P, name, cname, us, fl / us / 1e6, tl), flush=True)
            _C.set_option(_C.OPT_GEMM_V3, 1)
            _C.set_option(_C.OPT_TC3_BN, 0)
            _C.set_option(_C.OPT_TC3_PAIR, -1)
            _C.set_option(_C.OPT_TC3_SMEM_KB, 0)
            if P == 1:
                M, N, K = B * T, Cout, Cin * taps
                a_ = torch.randn(M, K, device='cuda', dtype=torch.bfloat16)
                b_ = torch.randn(K, N, device='cuda', dtype=torch.bfloat16)
                o_ = torch.empty(M, N, device='cuda', dtype=torch.bfloat16)
                us = graph_time(lambda: torch.matmul(a_, b_, out=o_), flush)
                rows.append(dict(tag=tag, P=1, shape=name, config='cuBLAS bf16 %dx%dx%d (bf16 out, no epilogue)' % (M, N, K),
                                 us=round(us, 2), tflops=round(fl / us / 1e6, 1)))
                print('P1 %-34s %-18s %7.2f us  %6.1f TF/s' % (name, 'cuBLAS %dx%dx%d' % (M, N, K), us, fl / us / 1e6), flush=True)
    # calibration of the per-launch cost inside a graph: a chain of tiny dependent kernels (one 1 KB transpose each)
    xs = torch.randn(1, 16, 16, device='cuda')
    os_ = torch.empty(1, 16, 16, device='cuda')
    us = graph_time(lambda: _C.call('transpose', xs, 1, 16, 16, os_), flush, reps=50)
    print('chain of tiny kernels (PDL %s): %.2f us per launch' % (os.environ.get('VNB_PDL', '1'), us))
    big = torch.randn(16, 512, 128, device='cuda')
    us = graph_time(lambda: Pm.make_planes(big, 1), flush, reps=20)
    print('chain of make_planes 16x512x128: %.2f us per launch' % us)
    os.makedirs(os.path.join(ROOT, 'gpurun_out'), exist_ok=True)
    with open(os.path.join(ROOT, 'gpurun_out', 'bench_gemm3_%s.json' % tag), 'w') as f:
        json.dump(rows, f, indent=1)


if __name__ == '__main__':
    main()
