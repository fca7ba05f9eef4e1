"""The batched speaker-conditioning linears (vnb_linear_fwd_batch / vnb_linear_bwd_batch) at the VCC2020 shape:
L = 10 blocks, B = 16 utterances, I = 128, O = 1024.  CUDA-event time of 20 launches in one graph."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from vae_npvc_b200 import prims as Pm  # noqa: E402

L, B, I, O = 10, 16, 128, 1024
y = torch.randn(B, I, device='cuda')
Ws = [torch.randn(O, I, device='cuda') * 0.05 for _ in range(L)]
b1 = [torch.randn(O, device='cuda') for _ in range(L)]
b2 = [torch.randn(O, device='cuda') for _ in range(L)]
dWs = [torch.empty(O, I, device='cuda') for _ in range(L)]
dp = torch.randn(L, B, O, device='cuda')


def timeit(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            fn()
    ts = []
    for _ in range(5):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        e1.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3 / reps)
    return sorted(ts)[2]


p = Pm.linear_fwd_batch(y, Ws, b1, b2)
ref = torch.stack([y @ W.t() + a + b for W, a, b in zip(Ws, b1, b2)])
print('fwd max err %.2e' % float((p - ref).abs().max()))
dy = Pm.linear_bwd_batch(dp, y, Ws, dWs)
print('bwd dy err %.2e  dW err %.2e' % (float((dy - sum(dp[l] @ Ws[l] for l in range(L))).abs().max()),
                                        float((dWs[3] - dp[3].t() @ y).abs().max())))
print('linear_fwd_batch %.2f us   linear_bwd_batch (+ zeroing dy) %.2f us'
      % (timeit(lambda: Pm.linear_fwd_batch(y, Ws, b1, b2)), timeit(lambda: Pm.linear_bwd_batch(dp, y, Ws, dWs))))
