"""CUDA-event timing of the conv GEMM kernels (vnb_conv_fwd / vnb_conv_wgrad) on the step's shapes.

    python tools/bench_gemm.py [tag [shape_index]]

Kernel variants are selected with the library's env switches (VNB_TC_BN, VNB_TC_SHALLOW, ...).
Two timings per shape: 'cold' flushes L2 with a 256 MB write before every launch (one event pair per
launch); 'warm' is 20 back-to-back launches on resident operands (what the training step sees, since
one step's activations + weights of a layer sit in the 126 MB L2).
"""
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
    ('dec conv_in fwd   512->1024 k3 +GN', 512, 1024, 3, True, 2),
    ('dec conv_in dgrad 1024->512 k3', 1024, 512, 3, False, 0),
    ('enc k3 fwd        512->512  k3 +GN', 512, 512, 3, False, 1),
    ('enc k1 skip       512->512  k1', 512, 512, 1, False, 0),
    ('dec res/skip fwd  512->640  k1', 512, 640, 1, False, 0),
    ('dec res/skip dgrad 640->512 k1', 640, 512, 1, False, 0),
]


def timeit(fn, flush=None, iters=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    if flush is None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(iters):
            fn()
        e1.record()
        e1.synchronize()
        return e0.elapsed_time(e1) * 1e3 / iters
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else 'default'
    only = int(sys.argv[2]) if len(sys.argv) > 2 else None          # index into SHAPES (for ncu captures)
    _C.require_cuda()
    P = int(os.environ.get('VNB_BENCH_P', 1))
    flush = torch.empty(256 * 1024 * 1024 // 4, device='cuda')
    rows = []
    for name, Cin, Cout, taps, transposed, gn in (SHAPES if only is None else SHAPES[only:only + 1]):
        x = torch.randn(B, Cin, T, device='cuda')
        v = torch.randn((Cin, Cout, taps) if transposed else (Cout, Cin, taps), device='cuda') * 0.02
        W = Pm.wn_prepare(v, None, transposed, P)
        xpl = Pm.make_planes(x, P)[0]
        out = Pm.cl_empty(B, Cout, T, 'cuda')
        sums = torch.zeros((B, max(gn, 1), 2), dtype=torch.float64, device='cuda')

        def fwd():
            _C.call('conv_fwd', xpl.data, P, B, Cin, T, W.wf, taps, 1, Cout, None, 0, None, out.transpose(1, 2), 0,
                    sums if gn else None, gn)

        dy = torch.randn(B, Cout, T, device='cuda')
        dypl = Pm.make_planes(dy, P)[0]
        dw = torch.zeros_like(v)

        def wgrad():
            _C.call('conv_wgrad', dypl.data, xpl.data, P, B, Cout, Cin, T, taps, 1, int(transposed), dw)

        fl = 2.0 * B * T * Cin * Cout * taps          # algorithmic (P = 2 / 3 execute 3x / 6x)
        for kind, fn in (('conv_fwd', fwd), ('wgrad', wgrad)):
            cold, warm = timeit(fn, flush), timeit(fn)
            rows.append(dict(tag=tag, shape=name, kind=kind, cold_us=round(cold, 2), warm_us=round(warm, 2),
                             warm_tflops=round(fl / warm / 1e6, 1), cold_tflops=round(fl / cold / 1e6, 1)))
            print('%-10s %-38s %-8s cold %7.2f us (%6.1f TF/s)  warm %7.2f us (%6.1f TF/s)'
                  % (tag, name, kind, cold, fl / cold / 1e6, warm, fl / warm / 1e6), flush=True)
    if only is None:
        # what the library GEMM gets at the same one-wave shapes (torch.matmul bf16 -> cuBLAS): the forward of the
        # decoder conv_in is M = 2048 frames, N = 1024, K = 3 * 512; its weight gradient M = 512, N = 3 * 1024, K = 2048
        for name, M, N, K in (('cuBLAS bf16 2048x1024x1536 (dec conv_in fwd)', 2048, 1024, 1536),
                              ('cuBLAS bf16 2048x512x3072  (dec conv_in dgrad)', 2048, 512, 3072),
                              ('cuBLAS bf16 512x3072x2048  (dec conv_in wgrad)', 512, 3072, 2048),
                              ('cuBLAS bf16 2048x512x1536  (enc k3 fwd)', 2048, 512, 1536),
                              ('cuBLAS bf16 2048x640x512   (dec res/skip fwd)', 2048, 640, 512)):
            a_ = torch.randn(M, K, device='cuda', dtype=torch.bfloat16)
            b_ = torch.randn(K, N, device='cuda', dtype=torch.bfloat16)
            o_ = torch.empty(M, N, device='cuda', dtype=torch.bfloat16)

            def mm():
                torch.matmul(a_, b_, out=o_)
            fl = 2.0 * M * N * K
            cold, warm = timeit(mm, flush), timeit(mm)
            rows.append(dict(tag=tag, shape=name, kind='cublas', cold_us=round(cold, 2), warm_us=round(warm, 2),
                             warm_tflops=round(fl / warm / 1e6, 1), cold_tflops=round(fl / cold / 1e6, 1)))
            print('%-10s %-38s %-8s cold %7.2f us (%6.1f TF/s)  warm %7.2f us (%6.1f TF/s)'
                  % (tag, name, 'cublas', cold, fl / cold / 1e6, warm, fl / warm / 1e6), flush=True)
    os.makedirs(os.path.join(ROOT, 'gpurun_out'), exist_ok=True)
    with open(os.path.join(ROOT, 'gpurun_out', 'bench_gemm_%s.json' % tag), 'w') as f:
        json.dump(rows, f, indent=1)


if __name__ == '__main__':
    main()
