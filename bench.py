"""Benchmark of the VQ-VAE training step (BASELINE.json metric: train frames/s, fwd+bwd+optimiser).

    python bench.py --gpus 1 --steps 100 --warmup 10                # our CUDA path, one B200
    torchrun --nproc-per-node N ... bench.py --gpus N ...           # batch-sharded DP, weak scaling
    python bench.py --impl reference ...                            # the reference's own modules (oracle/_ref) on the host CPU

Prints ONE JSON line on rank 0.  A "step" = zero-grad, forward, backward, [all-reduce], clip, Adam,
StepLR over one synthetic batch of 16 x 128 frames PER GPU of the VCC2020-shape model
(80-dim mel, C512, codebook 512 x 64).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402

DEFAULT_PRECISION = 'bf16x2'         # fastest operand mode that meets the stated parity bars (tests/test_gpu_parity_modes.py)


def _wl(name):
    """The named workloads of BASELINE.json.configs (default = configs[1], the one the metric is quoted on)."""
    from vae_npvc_b200 import synth
    if name == 'vcc2020':
        arch = synth.arch_config(X=80, C=512, D=64, K=512, Y=128, S=128, y_num=117, stacks=10, use_ema=False, embed_norm=True)
        return dict(arch=arch, B=16, T=128, kind='vqvae', what='VQ-VAE VCC2020 shape (BASELINE configs[1]): 80-mel, 16x128 frames '
                    'per GPU, C512, 10+10 blocks, codebook 512x64 (use_ema false, embed_norm true), Adam+clip+StepLR')
    if name == 'vcc2020_yaml':
        arch = synth.arch_config(X=80, C=512, D=128, K=512, Y=128, S=128, y_num=117, stacks=10, use_ema=True)
        return dict(arch=arch, B=128, T=256, kind='vqvae', what='VQ-VAE as shipped (egs/vcc20/vae1/conf/train_pytorch_vqvae.yaml): '
                    '80-mel, 128x256 frames per GPU, use_ema true, codebook 512x128, Adam+clip+StepLR')
    if name == 'aishell3':
        arch = synth.arch_config(X=160, C=512, D=128, K=128, Y=128, S=256, y_num=218, stacks=10, use_ema=False, jitter_p=0.12)
        return dict(arch=arch, B=32, T=128, kind='vqvae', what='AISHELL-3 shape (BASELINE configs[3]; egs/aishell3/vc2/conf/'
                    'train_pytorch_vqvae.yaml): 160-mel, skip 256, codebook 128x128, 218 speakers, jitter_p 0.12, 32x128 frames per '
                    'GPU (256x128 global on 8 GPUs)')
    if name == 'cdvae':
        arch = synth.cdvae_config()
        return dict(arch=arch, B=16, T=128, kind='cdvae', what='CDVAE throughput shape (BASELINE configs[2]): 513-dim SP + 40-dim MCC '
                    'encoders / decoders over one shared 512x128 codebook, four reconstructions, 16x128 frames per domain per GPU')
    if name == 'decode':
        arch = synth.arch_config(X=80, C=512, D=64, K=512, Y=128, S=128, y_num=117, stacks=10)
        return dict(arch=arch, B=64, T=512, kind='decode', what='decoder-only batched conversion (BASELINE configs[4]): '
                    'Model.decode((ids 64x512, speaker ids)) -> 80-mel, VCC2020-shape decoder')
    if name == 'vqvae2':
        arch = synth.vqvae2_config()
        return dict(arch=arch, B=32, T=256, kind='vqvae2', what='hierarchical VQ-VAE2 + global style tokens as shipped '
                    '(egs/vcc20/vae2/conf/train_pytorch_vqvae2.yaml; SURVEY 8 f4): 80-mel, three levels (strides 1 / 2x2 / 4x4), '
                    'two 512x128 codebooks + 10 style tokens, 32x256 frames per GPU (recipe batch 96x256 on 3 GPUs), Adam+clip+StepLR')
    raise SystemExit('unknown workload %r' % name)


def flop_per_frame(wl):
    """Algorithmic FLOPs per frame of the workload: fwd + bwd = 3 x 2 x MAC for the convs, the VQ
    distance GEMM forward only."""
    from vae_npvc_b200 import synth
    a, T = wl['arch'], wl['T']
    if wl['kind'] == 'vqvae2':
        return None                                # per-level frame rates differ; no single per-frame figure is quoted
    vq = 2 * a['z_num'] * a['z_dim']
    if wl['kind'] == 'cdvae':
        macs, n = 0, 0
        for dom in a['domains'].values():
            e, d = synth.conv_macs_per_frame(dom['encoder'], dom['decoder'], T)
            macs += e + 2 * d                      # every decoder runs on both domains' latents
            n += 1
        return 6 * macs + n * vq
    e, d = synth.conv_macs_per_frame(a['encoder'], a['decoder'], T)
    if wl['kind'] == 'decode':
        return 2 * d
    return 6 * (e + d) + vq


def peaks():
    p = os.path.join(ROOT, 'MEASURED_PEAKS.json')
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d['hbm_gbs'], tf_burst=d['bf16_tflops'], tf_sus=d.get('bf16_tflops_sustained', d['bf16_tflops']),
                    src='measured')
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sus=1400.0, src='fallback')


class ClockSampler(threading.Thread):
    """SM clock / throttle-reason samples during the timed region (NVML, else the nvidia-smi query)."""
    Q = ('clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,'
         'clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap')

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self._stop_evt = index, [], threading.Event()

    def _nvml(self):
        """NVML (what nvidia-smi itself reads) sampled every 2 ms: a 50 ms timed region gets ~25 samples."""
        import pynvml as nv
        nv.nvmlInit()
        h = nv.nvmlDeviceGetHandleByIndex(self.index)
        mx = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
        bits = ((0x8, 2), (0x40, 3), (0x20, 4), (0x4, 5))        # hw_slowdown, hw_thermal, sw_thermal, sw_power_cap
        while not self._stop_evt.is_set():
            r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(h)
            row = [str(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM)), str(mx), '', '', '', '']
            for bit, col in bits:
                row[col] = 'Active' if (r & bit) else 'Not Active'
            self.rows.append(row)
            self._stop_evt.wait(0.002)

    def run(self):
        try:
            self._nvml()
            return
        except Exception:
            pass
        while not self._stop_evt.is_set():
            try:
                out = subprocess.run(['nvidia-smi', '-i', str(self.index), '--query-gpu=' + self.Q,
                                      '--format=csv,noheader,nounits'], capture_output=True, text=True, timeout=5).stdout
                f = [s.strip() for s in out.strip().split(',')]
                if len(f) >= 6:
                    self.rows.append(f)
            except Exception:
                pass
            self._stop_evt.wait(0.2)

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=6)
        if not self.rows:
            return {'sm_mhz': None, 'sm_max_mhz': None, 'reasons': ['unsampled']}
        sm = sorted(float(r[0]) for r in self.rows)
        names = ['hw_slowdown', 'hw_thermal_slowdown', 'sw_thermal_slowdown', 'sw_power_cap']
        reasons = [n for i, n in enumerate(names) if any(r[2 + i].lower().startswith('active') for r in self.rows)]
        return {'sm_mhz': sm[len(sm) // 2], 'sm_max_mhz': float(self.rows[0][1]), 'reasons': reasons,
                'samples': len(self.rows)}


def cpu_oracle_step_rate(wl, steps, warmup, threads=None):
    """The reference's CPU path restated (oracle port): full train step, fp32, torch CPU."""
    from oracle import synth
    from oracle import vqvae_oracle as O
    arch = wl['arch']
    cores = threads or os.cpu_count() or 1
    torch.set_num_threads(cores)
    sd = synth.synthetic_state_dict(arch, 777)
    params = O.params_from_state_dict(sd)
    opt = O.AdamState(O.trainable_names(params), lr=1e-3, step_size=100000, gamma=0.5)
    x, y = synth.synthetic_batch(arch, wl['B'], wl['T'], 777)
    for _ in range(warmup):
        O.train_step(params, arch, opt, x, y, max_grad_norm=10)
    t0 = time.perf_counter()
    for _ in range(steps):
        O.train_step(params, arch, opt, x, y, max_grad_norm=10)
    dt = (time.perf_counter() - t0) / steps
    return wl['B'] * wl['T'] / dt, dt, cores


def reference_available():
    from oracle import build_ref
    return build_ref.available()


def reference_step_rate(wl, steps, warmup, device='cpu', autocast=False, threads=None):
    """The UNMODIFIED reference modules staged under oracle/_ref (oracle/build_ref.py): vae_npvc.model.vqvae.Model
    with the train step of vae_npvc/trainer/basic.py:55-79 around it (zero_grad, forward, backward,
    clip_grad_norm_(10), Adam(0.5, 0.999), StepLR).  On ``device='cuda'`` the reference's own Trainer class is used
    as is (it hard-codes .cuda(), trainer/basic.py:28) - stock eager PyTorch on the B200; ``autocast`` wraps its
    forward in torch.autocast(bfloat16).  kind 'decode': Model.decode forward only (vqvae.py:55-60).
    -> (frames/s, s/step, cores)."""
    from oracle import build_ref, synth
    build_ref.import_reference()
    arch = wl['arch']
    B, T = wl['B'], wl['T']
    cfg = dict(arch)
    v2 = wl['kind'] == 'vqvae2'
    cfg.update(model_type='vae_npvc.model.vqvae2' if v2 else 'vae_npvc.model.vqvae', optim_type='Adam', learning_rate=1e-3,
               max_grad_norm=10, lr_scheduler='StepLR', lr_param={'step_size': 100000, 'gamma': 0.5, 'last_epoch': -1})
    if v2:
        sd = synth.vqvae2_state_dict(arch, 777)
        x, y = synth.vqvae2_batch(arch, B, T, 777)
    else:
        sd = synth.synthetic_state_dict(arch, 777)
        x, y = synth.synthetic_batch(arch, B, T, 777)
    cores = threads or os.cpu_count() or 1
    import warnings
    warnings.simplefilter('ignore')
    sync = (lambda: None)
    if wl['kind'] == 'decode':
        from vae_npvc.model.vqvae import Model as RefModel
        torch.set_num_threads(cores)
        m = RefModel(cfg)
        m.load_state_dict(sd)
        m.eval()
        ids = torch.randint(0, arch['z_num'], (B, T), generator=torch.Generator().manual_seed(5))
        if device == 'cuda':
            m, ids, y = m.cuda(), ids.cuda(), y.cuda()
            sync = torch.cuda.synchronize

        def step(i):
            with torch.no_grad():
                return m.decode((ids, y))
    elif device == 'cuda':
        from vae_npvc.trainer.basic import Trainer as RefTrainer
        tr = RefTrainer(cfg)
        torch.nn.Module.load_state_dict(tr.model, sd)          # (vqvae2's own override reads an undefined attribute)
        xh, yh = x.pin_memory(), y.pin_memory()
        if autocast:
            fwd = tr.model.forward
            tr.model.forward = lambda inp: _autocast_call(fwd, inp)

        def step(i):
            return tr.train_step([xh, yh], i)
        sync = torch.cuda.synchronize
    else:
        torch.set_num_threads(cores)
        if v2:
            from vae_npvc.model.vqvae2 import Model as RefModel
        else:
            from vae_npvc.model.vqvae import Model as RefModel
        m = RefModel(cfg)
        torch.nn.Module.load_state_dict(m, sd)
        m.train()
        opt = torch.optim.Adam(m.parameters(), lr=1e-3, betas=(0.5, 0.999), weight_decay=0.0)
        sch = torch.optim.lr_scheduler.StepLR(opt, step_size=100000, gamma=0.5)

        def step(i):                                   # trainer/basic.py:55-79 without the hard-coded .cuda()
            m.zero_grad()
            _, loss, detail = m([x, y])
            loss.backward()
            torch.nn.utils.clip_grad_norm_(m.parameters(), 10)
            opt.step()
            sch.step()
            return i + 1, detail
    for i in range(warmup):
        step(i)
    sync()
    t0 = time.perf_counter()
    for i in range(steps):
        step(i)
    sync()
    dt = (time.perf_counter() - t0) / steps
    return B * T / dt, dt, cores


def _autocast_call(fwd, inp):
    with torch.autocast('cuda', dtype=torch.bfloat16):
        return fwd(inp)


def metric_name(wl):
    return 'conversion frames/sec (decoder only)' if wl['kind'] == 'decode' else 'train frames/sec (fwd+bwd+optimizer)'


def run_reference(args, rank):
    if rank != 0:
        return
    wl = _wl(args.workload)
    if wl['kind'] == 'cdvae':
        print(json.dumps({'impl': 'reference', 'unavailable': 'the reference snapshot has no cross-domain (CDVAE) model file '
                          '(SURVEY.md F9): nothing to time for --workload cdvae'}))
        return
    on_gpu = args.ref_device == 'cuda'
    big = wl['B'] * wl['T'] > 8192
    cap = 200 if on_gpu else (3 if big else 12)
    steps, warm = max(1, min(args.steps, cap)), max(1, min(args.warmup, 10 if on_gpu else (1 if big else 3)))
    if reference_available():
        fps, dt, cores = reference_step_rate(wl, steps, warm, args.ref_device, args.ref_autocast)
        kind = 'reference'
        what = ('the reference\'s own vae_npvc.model.%s.Model (staged under oracle/_ref)' % ('vqvae2' if wl['kind'] == 'vqvae2' else 'vqvae')
                + (' decode()' if wl['kind'] == 'decode' else ' + trainer/basic.py:55-79 step') + ', '
                + ('stock eager PyTorch on the B200%s' % (' under bf16 autocast' if args.ref_autocast else ', fp32')
                   if on_gpu else 'torch CPU fp32'))
    else:
        if on_gpu or wl['kind'] == 'decode':
            print(json.dumps({'impl': 'reference', 'unavailable': 'oracle/_ref not staged: no reference modules to run'}))
            return
        fps, dt, cores = cpu_oracle_step_rate(wl, steps, warm)
        kind = 'port'
        what = 'oracle restatement of the reference Model + trainer/basic.py step, torch CPU fp32'
    line = {
        'impl': 'reference', 'metric': metric_name(wl), 'value': fps, 'unit': 'frames/s',
        'n_gpus': args.gpus, 'steps': steps, 'warmup': warm, 'ms_per_step': dt * 1e3, 'higher_is_better': True,
        'scaling': 'weak', 'vs_baseline': None, 'dtype': 'bf16 autocast' if (on_gpu and args.ref_autocast) else 'f32',
        'data': 'synthetic',
        'config': {'workload': wl['what'], 'reference_device': args.ref_device},
        'cpu_baseline': {'value': fps, 'unit': 'frames/s', 'cores': cores, 'kind': kind,
                         'sample': '%d steps of %dx%d frames (%s)' % (steps, wl['B'], wl['T'], what)},
        'e2e': {'value': fps, 'unit': 'frames/s', 'h2d_bytes_per_step': 0, 'd2h_bytes_per_step': 0},
        'gpu_launches': 0,
    }
    print(json.dumps(line))


def time_dominant_kernel(P, B, T, iters=20):
    """CUDA-event time of the step's dominant kernel in isolation: the decoder block's k=3 ConvTranspose as a GEMM
    (Cout=1024, Cin=512, 3 taps, B x T frames) - 20 of the step's ~110 GEMM launches have this shape or its data- /
    weight-gradient twin, ~2/3 of all FLOPs."""
    from vae_npvc_b200 import prims as Pm
    from vae_npvc_b200 import _C as _Cm
    Cin, Cout, taps = 512, 1024, 3
    x = torch.randn(B, Cin, T, device='cuda')
    v = torch.randn(Cin, Cout, taps, device='cuda') * 0.02
    W = Pm.wn_prepare(v, None, True, P)
    xpl = Pm.make_planes(x, P)[0]
    out = Pm.cl_empty(B, Cout, T, 'cuda')
    flush = torch.empty(256 * 1024 * 1024 // 4, device='cuda')          # > 126 MB L2
    sums = torch.zeros((B, 2, 2), dtype=torch.float64, device='cuda')

    def launch():
        # exactly the step's call: fused GroupNorm(2) statistics in the epilogue
        _Cm.call('conv_fwd', xpl.data, P, B, Cin, T, W.wf, taps, 1, Cout, None, 0, None, out.transpose(1, 2), 0, sums, 2)

    for _ in range(3):
        launch()
    torch.cuda.synchronize()
    # `reps` launches captured in one CUDA graph (as in the training step: no host launch cost between
    # kernels), L2 flushed before every replay; the first launch of a replay reads its operands from HBM
    reps = 10
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(reps):
            launch()
    times = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        g.replay()
        e1.record()
        e1.synchronize()
        times.append(e0.elapsed_time(e1) * 1e-3 / reps)
    dt = sum(times) / len(times)
    pairs = {1: 1, 2: 3, 3: 6}[P]
    flops = 2.0 * B * T * Cout * Cin * taps            # algorithmic (the multi-plane modes execute 3x / 6x)
    return dt, flops, flops * pairs


def run_ours(args, rank, world):
    from vae_npvc_b200 import _C
    from vae_npvc_b200.trainer.basic import Trainer
    from vae_npvc_b200 import synth
    _C.require_cuda()
    local = int(os.environ.get('LOCAL_RANK', 0))
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    wl = _wl(args.workload)
    arch, B, T = wl['arch'], wl['B'], wl['T']
    P = {'bf16': 1, 'bf16x2': 2, 'fp32': 3}[args.precision]
    warm = max(3, args.warmup)

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    detail = {}
    if wl['kind'] == 'decode':
        # forward-only conversion: Model.decode through the public module API, weights frozen
        import vae_npvc_b200.functional as Fn
        from vae_npvc_b200.model.vqvae import Model
        Fn.set_precision(args.precision)
        m = Model(arch)
        m.load_state_dict(synth.synthetic_state_dict(arch, 777))
        m = m.cuda().eval()
        gen = torch.Generator().manual_seed(5 + rank)
        ids = torch.randint(0, arch['z_num'], (B, T), generator=gen)
        y = torch.randint(0, arch['y_num'], (B, 1), generator=gen)
        host_in = [ids.pin_memory(), y.pin_memory()]
        dev_in = [ids.cuda(), y.cuda()]
        out_host = torch.empty((B, arch['decoder']['final_channels'], T), dtype=torch.float32).pin_memory()

        def step_device():
            with torch.no_grad():
                return m.decode((dev_in[0], dev_in[1]))

        def step_e2e(i):
            with torch.no_grad():
                o = m.decode((host_in[0].cuda(non_blocking=True), host_in[1].cuda(non_blocking=True)))
                out_host.copy_(o, non_blocking=True)
            torch.cuda.current_stream().synchronize()
        h2d, d2h = ids.numel() * 8 + y.numel() * 8, out_host.numel() * 4
        tr = None
    else:
        cfg = dict(arch)
        cfg.update(model_type={'cdvae': arch.get('model_type', 'vae_npvc_b200.model.vqvae'),
                               'vqvae2': 'vae_npvc_b200.model.vqvae2'}.get(wl['kind'], 'vae_npvc_b200.model.vqvae'),
                   precision=args.precision, cuda_graph=not args.no_graph,
                   overlap_allreduce=not bool(os.environ.get('VNB_NO_OVERLAP')), allreduce=os.environ.get('VNB_ALLREDUCE', 'multimem'),
                   allreduce_blocks=int(os.environ.get('VNB_AR_BLOCKS', 0)),
                   **({'allreduce_buckets': tuple(int(v) for v in os.environ['VNB_BUCKETS'].split(','))}
                      if os.environ.get('VNB_BUCKETS') else {}),
                   bank_overlap=not bool(os.environ.get('VNB_NO_BANK_OVERLAP')),
                   sparse_grad_zero=not bool(os.environ.get('VNB_NO_SPARSE_ZERO')),
                   graph_warmup=min(3, warm - 1), learning_rate=1e-3, max_grad_norm=10, lr_scheduler='StepLR',
                   lr_param={'step_size': 100000, 'gamma': 0.5, 'last_epoch': -1})
        torch.manual_seed(777)
        tr = Trainer(cfg)
        if wl['kind'] == 'cdvae':
            tr.model.load_state_dict(synth.cdvae_state_dict(arch, 777))
            xs, y = synth.cdvae_batch(arch, B, T, 777 + rank)
            tensors = xs + [y]
        elif wl['kind'] == 'vqvae2':
            tr.model.load_state_dict(synth.vqvae2_state_dict(arch, 777))
            x, y = synth.vqvae2_batch(arch, B, T, 777 + rank)
            tensors = [x, y]
        else:
            tr.model.load_state_dict(synth.synthetic_state_dict(arch, 777))       # same weights on every rank
            x, y = synth.synthetic_batch(arch, B, T, 777 + rank)                  # each rank its own shard
            tensors = [x, y]
        host_in = [t.pin_memory() for t in tensors]
        dev_in = [t.cuda() for t in tensors]

        def step_device():
            return tr.train_step_device(*dev_in)

        def step_e2e(i):
            nonlocal detail
            _, detail = tr.train_step(host_in, i)
        h2d = sum(t.numel() * t.element_size() for t in tensors)
        d2h = None

    for _ in range(warm):
        step_device()
    barrier()
    # ---- device-resident timing: CUDA events on the launching stream ----
    sampler = ClockSampler(local)
    sampler.start()
    _C.launch_count_reset()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    e0.record()
    for _ in range(args.steps):
        step_device()
    e1.record()
    barrier()
    dt = e0.elapsed_time(e1) * 1e-3
    eager_launches = _C.launch_count()
    # ---- end to end through the public API: pinned host batch in, host results out ----
    barrier()
    t0 = time.perf_counter()
    for i in range(args.steps):
        step_e2e(i)
    torch.cuda.synchronize()
    dt_e2e = time.perf_counter() - t0
    clocks = sampler.stop()
    tmax = torch.tensor([dt, dt_e2e], device='cuda', dtype=torch.float64)
    if dist is not None:
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    dt, dt_e2e = float(tmax[0]), float(tmax[1])
    on_graph = bool(tr is not None and tr.last_step_on_graph)
    per_step = tr.launches_per_step if on_graph else eager_launches // max(1, args.steps)
    if d2h is None:
        d2h = 4 * len(detail)
    if rank == 0:
        frames = B * T * world * args.steps
        value, e2e = frames / dt, frames / dt_e2e
        pk = peaks()
        kdt, kflops, kexec = time_dominant_kernel(P, min(B, 128), T)
        fpf = flop_per_frame(wl)
        line = {
            'metric': metric_name(wl), 'value': value, 'unit': 'frames/s', 'n_gpus': world,
            'steps': args.steps, 'warmup': warm, 'ms_per_step': dt / args.steps * 1e3, 'higher_is_better': True,
            'scaling': 'weak', 'vs_baseline': None,
            'dtype': {1: 'bf16', 2: 'bf16x2 (hi+mid bf16 operand planes, fp32 accumulate)', 3: 'f32 (3xbf16 split planes)'}[P],
            'data': 'synthetic',
            'config': {'workload': wl['what'], 'workload_key': args.workload,
                       'frames_per_gpu_per_step': B * T, 'parallelism': 'dp%d' % world,
                       'cuda_graph': on_graph,
                       'options': {'precision': args.precision,
                                   'parity_met': {'bf16': 'losses 1e-3 only (gradients ~1e-2..2e-1, features 6e-3..3e-2)',
                                                  'bf16x2': 'losses 1e-6, converted features 5e-6, ids bit-exact, gradients at the '
                                                            'conditioning floor of the model (median <= 1.2e-3; tests/test_gpu_parity_modes.py)',
                                                  'fp32': 'losses 1e-6, features 5e-7, ids bit-exact, gradients median 3e-6'}[args.precision],
                                   'ew_v2': _C.get_option(_C.OPT_EW_V2), 'fast_gate': _C.get_option(_C.OPT_FAST_GATE),
                                   'gemm_v3': _C.get_option(_C.OPT_GEMM_V3),
                                   'bank_overlap': bool(tr.bank_overlap) if tr else None,
                                   'sparse_grad_zero': (tr._zero_idx is not None) if tr else None},
                       'l2': 'no explicit flush: each step streams params+grads+Adam moments+operand planes '
                             '(~0.7 GB) > 126 MB L2; the isolated-kernel timing flushes L2 with a 256 MB write',
                       'flop_per_frame': fpf,
                       'step_roofline_frac_tensor': (value * fpf / (world * pk['tf_sus'] * 1e12)) if fpf else None},
            'clocks': clocks,
            'e2e': {'value': e2e, 'unit': 'frames/s', 'h2d_bytes_per_step': h2d, 'd2h_bytes_per_step': d2h},
            'gpu_launches': int(per_step * args.steps),
            'roofline': {'bound': 'tensor', 'kernel': 'conv_fwd Cout1024 Cin512 k3 (decoder conv_in) at %dx%d frames, backend %s'
                         % (min(B, 128), T, {0: 'auto', 1: 'simt', 2: 'tc'}[_C.get_backend()]),
                         'achieved': kflops / kdt / 1e12, 'peak': pk['tf_burst'], 'unit': 'TFLOP/s',
                         'frac': kflops / kdt / 1e12 / pk['tf_burst'],
                         # dram__bytes_read.sum + dram__bytes_write.sum of this launch from the ncu --set full captures
                         # (P 1: profiles/r01_ncu_conv_fwd_dominant.txt, P 2: profiles/r02i_ncu_conv_fwd_dominant.txt):
                         # the bf16 operand planes once, the fp32 output stays in L2
                         'traffic': ({1: 5278464, 2: 10530000}.get(P) if B * T == 2048 and args.workload == 'vcc2020' else None),
                         'executed_tflops': kexec / kdt / 1e12, 'peak_source': pk['src'], 'kernel_ms': kdt * 1e3},
        }
        if world == 1 and not args.no_cpu_baseline and wl['kind'] != 'cdvae':
            big = B * T > 8192
            if reference_available():
                fps, cdt, cores = reference_step_rate(wl, 2 if big else 8, 1 if big else 2)
                kind, what = 'reference', "the reference's own vae_npvc.model.%s.Model (oracle/_ref)" % ('vqvae2' if wl['kind'] == 'vqvae2' else 'vqvae')
            elif wl['kind'] == 'vqvae':
                fps, cdt, cores = cpu_oracle_step_rate(wl, 2 if big else 8, 1 if big else 2)
                kind, what = 'port', 'oracle restatement of the reference Model + trainer step'
            else:
                fps = None
            if fps is not None:
                line['cpu_baseline'] = {'value': fps, 'unit': 'frames/s', 'cores': cores, 'kind': kind,
                                        'sample': '%d steps of %dx%d frames (%s, torch CPU fp32), %.0f ms/step'
                                                  % (2 if big else 8, B, T, what, cdt * 1e3)}
        print(json.dumps(line))
    if dist is not None:
        # no destroy_process_group(): tearing NCCL down while captured graphs still hold its kernels hung at
        # exit on the 2-GPU box; synchronise, flush and leave
        dist.barrier()
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('--gpus', type=int, default=1)
    ap.add_argument('--steps', type=int, default=100)          # 0.2 s of timed steps: 20-step runs scatter by +-2 %
    ap.add_argument('--warmup', type=int, default=10)
    ap.add_argument('--impl', default='ours', choices=['ours', 'reference'])
    ap.add_argument('--precision', default=DEFAULT_PRECISION, choices=['bf16', 'bf16x2', 'fp32'])
    ap.add_argument('--workload', default='vcc2020', choices=['vcc2020', 'vcc2020_yaml', 'aishell3', 'cdvae', 'decode', 'vqvae2'],
                    help='default = BASELINE.json configs[1]; the others are the remaining named workloads')
    ap.add_argument('--no-graph', action='store_true')
    ap.add_argument('--no-cpu-baseline', action='store_true')
    # diagnostics of the reference arm (defaults = the contract: the reference's CPU path on the host cores)
    ap.add_argument('--ref-device', default='cpu', choices=['cpu', 'cuda'],
                    help='cuda: the unmodified reference Trainer/Model on the B200 through stock eager PyTorch')
    ap.add_argument('--ref-autocast', action='store_true', help='with --ref-device cuda: bf16 autocast forward')
    args = ap.parse_args()
    rank, world = int(os.environ.get('RANK', 0)), int(os.environ.get('WORLD_SIZE', 1))
    if args.impl == 'reference':
        run_reference(args, rank)
    else:
        run_ours(args, rank, world)


if __name__ == '__main__':
    main()
