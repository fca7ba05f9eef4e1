"""Per-kernel roofline microbenchmarks (BASELINE config 5 and SURVEY.md section 8d).

    python tools/bench_kernels.py [--quick] [--mem-only] [--tag NAME]    # writes gpurun_out/bench_kernels[_NAME].json

The streaming-kernel generation / gate math are whatever the options say (VNB_EW_V2, VNB_FAST_GATE in the
environment); they are recorded in the output.

* VQ lookup sweep: K in {256,512,1024,4096,16384} x D in {64,128,256}, N = 2^17 frames, z = randn, E = randn
  (the reference's own codebook init, layers_vq.py:18).  Reports frames/s, executed TFLOP/s (6 bf16 plane-pair
  products per fp32-faithful dot) against the measured bf16 peak, and algorithmic bytes (2*D*4+8 per frame)
  against the measured HBM peak.
* Memory-bound fused kernels at a size that cannot hide in L2 (2^17 frames x 512 channels = 256 MB per fp32
  tensor): achieved GB/s = algorithmic bytes (each tensor read once / written once) / time.
* Parameter-sized passes at the VCC2020 model size (31.1 M weights).
Timing: CUDA events around 5 back-to-back launches after 2 warm-ups; inputs are larger than the 126 MB L2 for
every memory-bound case (so no explicit flush is needed); the small-K VQ cases flush L2 with a 256 MB write.
"""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from vae_npvc_b200 import _C  # noqa: E402
from vae_npvc_b200 import prims as Pm  # noqa: E402


def peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        d = json.load(open(p))
        return d['hbm_gbs'], d['bf16_tflops'], 'measured'
    return 6650.0, 1590.0, 'fallback'


def timeit(fn, reps=5, flush=None):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    if flush is not None:
        flush.zero_()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    e1.synchronize()
    return e0.elapsed_time(e1) * 1e-3 / reps


def main():
    quick = '--quick' in sys.argv
    mem_only = '--mem-only' in sys.argv
    tag = sys.argv[sys.argv.index('--tag') + 1] if '--tag' in sys.argv else ''
    _C.require_cuda()
    hbm, tf, src = peaks()
    out = {'peaks': {'hbm_gbs': hbm, 'bf16_tflops': tf, 'source': src}, 'vq_lookup': [], 'memory_bound': [],
           'options': {'ew_v2': _C.get_option(_C.OPT_EW_V2), 'fast_gate': _C.get_option(_C.OPT_FAST_GATE)}}
    print('options:', out['options'], flush=True)
    flush = torch.empty(256 * 1024 * 1024 // 4, device='cuda')
    # ---------------- VQ lookup sweep ----------------
    N = 1 << 17
    Bv, Tv = N // 128, 128
    for D in ([] if mem_only else [64] if quick else [64, 128, 256]):
        for K in ([512, 4096] if quick else [256, 512, 1024, 4096, 16384]):
            g = torch.Generator(device='cuda').manual_seed(0)
            z = torch.randn(Bv, D, Tv, device='cuda', generator=g)
            E = torch.randn(K, D, device='cuda', generator=g)
            e_hat, e_pl, e_sq = Pm.vq_prep_codebook(E, False, False)
            zn, zpl, z_sq, _ = Pm.vq_prep_z(Pm.as_cl(z), False)
            idx = torch.empty((N,), dtype=torch.int64, device='cuda')

            def run():
                _C.call('vq_lookup', zn.transpose(1, 2), zpl.data, z_sq, e_hat, e_pl, e_sq, Bv, D, Tv, K, idx)

            dt = timeit(run, flush=flush)
            # check against torch on a slice (argmin of the fp32 distance)
            zz = zn.transpose(1, 2).reshape(N, D)[:4096]
            dist = (zz * zz).sum(1, keepdim=True) + (E * E).sum(1)[None] - 2 * zz @ E.t()
            agree = float((dist.argmin(1) == idx[:4096]).float().mean())
            ex_tf = 6 * 2.0 * N * K * D / dt / 1e12
            bytes_alg = N * (2 * D * 4 + 8)
            row = dict(K=K, D=D, N=N, us=dt * 1e6, frames_per_s=N / dt, executed_tflops=ex_tf, tensor_frac=ex_tf / tf,
                       algorithmic_tflops=2.0 * N * K * D / dt / 1e12, hbm_gbs=bytes_alg / dt / 1e9,
                       hbm_frac=bytes_alg / dt / 1e9 / hbm, argmin_agreement_vs_torch_fp32=agree)
            out['vq_lookup'].append(row)
            print('vq_lookup K=%5d D=%3d: %8.1f us  %7.2f Mframes/s  executed %6.1f TF/s (%.2f of %s peak)  '
                  'alg %.0f GB/s (%.2f)  argmin agreement %.4f' % (K, D, dt * 1e6, N / dt / 1e6, ex_tf, ex_tf / tf, src,
                                                                  bytes_alg / dt / 1e9, bytes_alg / dt / 1e9 / hbm, agree),
                  flush=True)
            # round 2: the fused kernel (frames read once, three bf16 plane products + exact fp32 re-rank, rows gathered and written):
            # read z + write q + write idx = the algorithmic bytes of the WHOLE lookup
            zc = Pm.as_cl(z)
            fused = Pm.vq_fused(zc, False, e_hat, e_pl, e_sq)
            if fused is not None:
                qf = Pm.cl_empty(Bv, D, Tv, 'cuda')
                sums = torch.zeros(2, device='cuda')
                counts = torch.zeros(K, dtype=torch.int32, device='cuda')
                idx2 = torch.empty((N,), dtype=torch.int64, device='cuda')

                def run_f():
                    _C.call('vq_fused', zc.transpose(1, 2), Bv, D, Tv, 0, e_hat, e_pl, e_sq, K, None, None, idx2, qf.transpose(1, 2),
                            sums, counts)

                dtf = timeit(run_f, flush=flush)
                agree_f = float((dist.argmin(1) == idx2[:4096]).float().mean())
                same = float((idx2 == idx).float().mean())
                alg_tf = 3 * 2.0 * N * K * D / dtf / 1e12
                rowf = dict(K=K, D=D, N=N, kernel='vq_fused', us=dtf * 1e6, frames_per_s=N / dtf, executed_tflops=alg_tf,
                            tensor_frac=alg_tf / tf, algorithmic_tflops=alg_tf / 3, hbm_gbs=bytes_alg / dtf / 1e9, hbm_frac=bytes_alg / dtf / 1e9 / hbm,
                            argmin_agreement_vs_torch_fp32=agree_f, same_as_six_product_kernel=same)
                out['vq_lookup'].append(rowf)
                print('vq_fused  K=%5d D=%3d: %8.1f us  %7.2f Mframes/s  executed %6.1f TF/s (%.2f of %s peak)  '
                      'alg %.0f GB/s (%.2f)  argmin agreement %.4f  same as 6-product kernel %.6f'
                      % (K, D, dtf * 1e6, N / dtf / 1e6, alg_tf, alg_tf / tf, src, bytes_alg / dtf / 1e9,
                         bytes_alg / dtf / 1e9 / hbm, agree_f, same), flush=True)
            del z, E, zn, zpl, idx
    # ---------------- memory-bound fused kernels ----------------
    B, T, C, S, P = (256 if quick else 1024), 128, 512, 128, 1
    rows = B * T
    dev = 'cuda'

    def rec(name, dt, nbytes):
        r = dict(kernel=name, us=dt * 1e6, algorithmic_bytes=nbytes, gbs=nbytes / dt / 1e9, hbm_frac=nbytes / dt / 1e9 / hbm)
        out['memory_bound'].append(r)
        print('%-34s %9.1f us  %8.1f MB  %7.0f GB/s  (%.2f of %s HBM peak)' % (name, dt * 1e6, nbytes / 1e6, r['gbs'],
                                                                               r['hbm_frac'], src), flush=True)

    h = Pm.cl_empty(B, C, T, dev).normal_()
    s = Pm.cl_empty(B, C, T, dev).normal_()
    mean, rstd = torch.zeros(B, 1, device=dev), torch.ones(B, 1, device=dev)
    gamma, beta = torch.ones(C, device=dev), torch.zeros(C, device=dev)
    dt = timeit(lambda: Pm.enc_out(h, s, mean, rstd, gamma, beta, P, True, True, True))
    rec('enc_out (GN apply+skip add -> y, 2 planes)', dt, rows * C * (4 + 4 + 4 + 2 + 2))
    dt = timeit(lambda: Pm.make_planes(h, P))
    rec('make_planes (fp32 -> bf16 plane)', dt, rows * C * (4 + 2))
    dn = s
    R = Pm.gn_bwd_rows(dn, h, mean, rstd, 1)
    _, _, Ssum, _, _ = Pm.gn_bwd_finalize(R, gamma, rstd, T, 1, True)
    dt = timeit(lambda: Pm.gn_bwd_rows(dn, h, mean, rstd, 1))
    rec('gn_bwd_rows (2 reads, row sums)', dt, rows * C * 8)
    dt = timeit(lambda: Pm.gn_bwd_apply(dn, h, mean, rstd, gamma, Ssum, 1, P, True))
    rec('gn_bwd_apply (2 reads -> 2 planes)', dt, rows * C * (8 + 4))
    dt = timeit(lambda: Pm.act_bwd_add(dn, h, s, 1))
    rec('act_bwd_add (3 reads, 1 write)', dt, rows * C * 16)
    dt = timeit(lambda: Pm.channel_sum(h))
    rec('channel_sum (1 read)', dt, rows * C * 4)
    del s, dn
    a = Pm.cl_empty(B, 2 * C, T, dev).normal_()
    mean2, rstd2 = torch.zeros(B, 2, device=dev), torch.ones(B, 2, device=dev)
    gamma2, beta2 = torch.ones(2 * C, device=dev), torch.zeros(2 * C, device=dev)
    dt = timeit(lambda: Pm.dec_gate(a, mean2, rstd2, gamma2, beta2, P))
    rec('dec_gate (GN2 + tanh*sigmoid -> plane)', dt, rows * (2 * C * 4 + C * 2))
    dg = h
    R2 = Pm.gate_bwd_rows(dg, a, mean2, rstd2, gamma2, beta2)
    _, _, S2, _, _ = Pm.gn_bwd_finalize(R2, gamma2, rstd2, T, 2, True)
    dt = timeit(lambda: Pm.gate_bwd_rows(dg, a, mean2, rstd2, gamma2, beta2))
    rec('gate_bwd_rows (3C reads, row sums)', dt, rows * 3 * C * 4)
    dt = timeit(lambda: Pm.gate_bwd_apply(dg, a, mean2, rstd2, gamma2, beta2, S2, P))
    rec('gate_bwd_apply (3C reads -> 2C plane)', dt, rows * (3 * C * 4 + 2 * C * 2))
    del a
    r = Pm.cl_empty(B, C + S, T, dev).normal_()
    skip = Pm.cl_empty(B, S, T, dev)
    dt = timeit(lambda: Pm.dec_out(r, h, S, P, skip, False))
    rec('dec_out (res add, skip accumulate)', dt, rows * ((C + S) * 4 + C * 4 + C * 4 + C * 2 + S * 8))
    del r, skip, h
    # VQ gather / backward at D=64
    D, K = 64, 512
    z = Pm.cl_empty(B, D, T, dev).normal_()
    E = torch.randn(K, D, device=dev)
    e_hat, e_pl, e_sq = Pm.vq_prep_codebook(E, False, False)
    dt = timeit(lambda: Pm.vq_prep_z(z, False))
    rec('vq_prep_z (z -> zn, 3 planes, |z|^2)', dt, rows * (D * 4 * 2 + D * 2 * 3 + 8))
    idx = torch.randint(0, K, (rows,), device=dev)
    dt = timeit(lambda: Pm.vq_gather(idx, e_hat, z, None))
    rec('vq_gather (idx, z -> q, losses, hist)', dt, rows * (8 + D * 4 * 2))
    # ---------------- parameter-sized passes (VCC2020 model: 31.1 M parameters) ----------------
    n = 31_100_000 // 4 * 4
    p, g, m, v = (torch.randn(n, device=dev) * 0.01 for _ in range(4))
    v.abs_()
    sqn = torch.zeros(1, device=dev)
    state = torch.zeros(2, device=dev)
    dt = timeit(lambda: Pm.adam_step(p, g, m, v, sqn, 10.0, 1e-3, 0.5, 0.999, 1e-8, state, 100000, 0.5))
    rec('adam_step (4 reads, 3 writes)', dt, n * 4 * 7)
    dt = timeit(lambda: Pm.sqnorm(g, sqn))
    rec('sqnorm (1 read)', dt, n * 4)
    # ---------------- batched weight-norm passes over the VCC2020 model's flat buffers ----------------
    del p, g, m, v
    from vae_npvc_b200 import synth
    from vae_npvc_b200.trainer.basic import Trainer
    arch = synth.arch_config()
    tr = Trainer(dict(arch, model_type='vae_npvc_b200.model.vqvae', precision='bf16', cuda_graph=False))
    pw = sum(e.v.numel() for e in tr.bank.by_ptr.values())
    out['options']['wn_v2'] = _C.get_option(_C.OPT_WN_V2)
    dt = timeit(lambda: tr.bank.prepare(), flush=flush)
    rec('wn_prepare_batch (%.1f M weights -> 2 bf16 operands)' % (pw / 1e6), dt, pw * (4 + 2 + 2))
    tr._flat_g.normal_()
    dt = timeit(lambda: tr.bank.finish_backward(), flush=flush)
    rec('wn_backward_batch (dw, v -> dv, dg)', dt, pw * 12)
    os.makedirs(os.path.join(ROOT, 'gpurun_out'), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, 'gpurun_out', 'bench_kernels%s.json' % ('_' + tag if tag else '')), 'w'), indent=1)


if __name__ == '__main__':
    main()
