"""Drop-in mirror of vae_npvc/trainer/basic.py: ``Trainer``.

Select it with ``trainer_type: vae_npvc_b200.trainer.basic`` (vae_npvc/bin/train.py:33-34,49-51).
Same constructor keys, ``train_step(batch, iteration) -> (iteration, dict[str, float])``,
``valid``, ``valid_step``, ``save_checkpoint`` / ``load_checkpoint`` (keys model / optimizer /
iteration; the optimizer entry is in torch.optim.Adam's state_dict format) and
``get_model_info`` as the reference.

B200-native differences (all behind the same API):
  * parameters, gradients and Adam moments live in four flat fp32 buffers; clip_grad_norm_ +
    Adam(betas=(0.5, 0.999)) + StepLR (trainer/basic.py:38-46, 62-72) are two kernels over them
  * after ``graph_warmup`` eager steps the whole step (zero-grad, forward, backward, clip, Adam)
    is captured once into a CUDA graph and replayed; the step's scalars come back in ONE
    device->host copy instead of the reference's 3-7 ``.item()`` syncs
  * under torch.distributed (one process per GPU) the flat gradient buffer is all-reduced
    (averaged) over NCCL and the EMA codebook statistics are all-reduced inside the model

New optional config keys (defaults keep the reference behaviour): ``cuda_graph`` (True),
``graph_warmup`` (3), ``precision`` ('bf16' | 'bf16x2' | 'fp32'), ``device`` ('cuda'), ``wgrad_stream`` (True),
``bank_overlap`` (True), ``sparse_grad_zero`` (True), ``jitter_draws`` ('device' on CUDA | 'host').
"""
import os
import contextlib
from importlib import import_module

import numpy as np
import torch

from .. import functional as Fn
from .. import prims as Pm
from ..bank import WeightBank
from ..model import layers_vq as _lvq
from ..model import vqvae as _model_mod


class FlatAdam(object):
    """Adam state over flat buffers, serialised in torch.optim.Adam's state_dict layout so that
    reference checkpoints (trainer/basic.py:109-121) load and save unchanged."""

    def __init__(self, params, flat_p, flat_g, lr, betas=(0.5, 0.999), eps=1e-8, offsets=None):
        self.params = params
        if offsets is None:
            offsets, off = [], 0
            for p in params:
                offsets.append(off)
                off += p.numel()
        self.offsets = list(offsets)        # element offset of every parameter in the flat buffers (16-byte aligned)
        self.flat_p, self.flat_g = flat_p, flat_g
        self.m = torch.zeros_like(flat_p)
        self.v = torch.zeros_like(flat_p)
        self.state = torch.zeros(2, dtype=torch.float32, device=flat_p.device)   # [adam step, sched step]
        self.sqn = torch.zeros(1, dtype=torch.float32, device=flat_p.device)
        self.lr, self.betas, self.eps = lr, betas, eps
        self.lr0 = lr                     # the config value ('initial_lr' of torch's StepLR)
        self.step_size, self.gamma = 0, 1.0
        self.param_groups = [{'lr': lr, 'betas': betas, 'eps': eps, 'weight_decay': 0.0, 'amsgrad': False,
                              'params': list(range(len(params)))}]

    def step(self, max_grad_norm, step_size, gamma):
        self.step_size, self.gamma = int(step_size), float(gamma)
        self.sqn.zero_()
        if max_grad_norm > 0:
            Pm.sqnorm(self.flat_g, self.sqn)
        Pm.adam_step(self.flat_p, self.flat_g, self.m, self.v, self.sqn, float(max_grad_norm), self.lr,
                     self.betas[0], self.betas[1], self.eps, self.state, int(step_size), float(gamma))

    def current_lr(self):
        """The learning rate StepLR has decayed to: lr * gamma ** floor(scheduler steps / step_size) - what
        torch's ``optimizer.param_groups[0]['lr']`` holds at this point of the reference's run."""
        if not self.step_size or self.gamma == 1.0:
            return float(self.lr)
        return float(self.lr) * float(self.gamma) ** int(float(self.state[1].item()) // self.step_size)

    def state_dict(self):
        step = float(self.state[0].item())
        st = {}
        for i, p in enumerate(self.params):
            n, off = p.numel(), self.offsets[i]
            if step > 0:
                st[i] = {'step': torch.tensor(step), 'exp_avg': self.m[off:off + n].view_as(p).clone(),
                         'exp_avg_sq': self.v[off:off + n].view_as(p).clone()}
        g = dict(self.param_groups[0])
        # torch.optim.Adam + StepLR save the DECAYED lr in param_groups and the undecayed one as 'initial_lr'
        # (trainer/basic.py:109-115 stores optimizer.state_dict() as is)
        g['lr'] = self.current_lr()
        if self.step_size:
            g['initial_lr'] = float(self.lr0)
        return {'state': st, 'param_groups': [g]}

    def load_state_dict(self, sd):
        step = 0.0
        for i, p in enumerate(self.params):
            n, off = p.numel(), self.offsets[i]
            s = sd['state'].get(i)
            if s is not None:
                self.m[off:off + n].copy_(s['exp_avg'].reshape(-1))
                self.v[off:off + n].copy_(s['exp_avg_sq'].reshape(-1))
                step = float(s['step'])
        self.state[0] = step
        # Resume semantics of the reference (trainer/basic.py:117-121): the Trainer - and its StepLR - is built
        # BEFORE load_checkpoint; optimizer.load_state_dict then replaces param_groups, so training continues at
        # the checkpoint's (decayed) lr, while the scheduler's own counter restarts at 0 (its state is never
        # saved) and keeps chaining gamma off group['lr'] every step_size steps from here.
        groups = sd.get('param_groups') or [{}]
        if 'lr' in groups[0]:
            self.lr = float(groups[0]['lr'])
            self.param_groups[0]['lr'] = self.lr
        self.state[1] = 0.0


def _maybe_init_distributed(config):
    """One process per GPU under torchrun WITHOUT touching the reference's launcher: vae_npvc/bin/train.py knows nothing
    about ranks, so when the torchrun environment is present (WORLD_SIZE > 1) and no process group exists
    yet, the Trainer binds this process to cuda:LOCAL_RANK and opens the NCCL group itself.  Returns the device string
    to use (config 'device' wins when given)."""
    import os
    import torch.distributed as dist
    world = int(os.environ.get('WORLD_SIZE', '1'))
    dev = config.get('device', 'cuda')
    if world <= 1 or not str(dev).startswith('cuda') or not dist.is_available() or not bool(config.get('ddp', True)):
        return dev
    local = int(os.environ.get('LOCAL_RANK', os.environ.get('RANK', '0')))
    if torch.cuda.is_available():
        torch.cuda.set_device(local)
    if not dist.is_initialized():
        os.environ.setdefault('MASTER_ADDR', '127.0.0.1')
        os.environ.setdefault('MASTER_PORT', '29500')
        dist.init_process_group('nccl', device_id=torch.device('cuda', local))
    return 'cuda:%d' % local if 'device' not in config else dev


def _dist():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        return dist
    return None


class Trainer(object):
    def __init__(self, config):
        model_type = config.get('model_type', 'vae_npvc_b200.model.vqvae:Model').split(':')
        optim_type = config.get('optim_type', 'Adam')
        learning_rate = config.get('learning_rate', 1e-3)
        max_grad_norm = config.get('max_grad_norm', 5)
        lr_scheduler = config.get('lr_scheduler', None)
        lr_param = config.get('lr_param', {'step_size': 100000, 'gamma': 0.5, 'last_epoch': -1})
        if optim_type.upper() != 'ADAM':
            # the reference's RAdam uses add_(scalar, tensor) overloads removed from torch 2.x
            raise NotImplementedError('vae_npvc_b200.trainer: only optim_type Adam is supported (every shipped '
                                      'recipe uses it; the reference RAdam does not run on this torch)')
        if 'precision' in config:
            Fn.set_precision(config['precision'])
        module = import_module(model_type[0], package=None)
        model_name = 'Model' if len(model_type) < 2 else model_type[1]
        model = getattr(module, model_name)(config)

        self.device = torch.device(_maybe_init_distributed(config))
        if self.device.type == 'cuda':
            from .. import _C
            _C.require_cuda()
        self.model = model.to(self.device)
        self.learning_rate = learning_rate
        self.max_grad_norm = max_grad_norm
        self._allreduce_kind = str(config.get('allreduce', 'multimem'))      # 'multimem' (NVSwitch, own kernel) | 'nccl'
        self._allreduce_blocks = int(config.get('allreduce_blocks', 0))     # 0 = 16 CTAs at 2 ranks, 8 beyond (measured)
        self._flatten()
        self.optimizer = FlatAdam(self._params, self._flat_p, self._flat_g, learning_rate, offsets=self._offsets)
        # all conv weights materialised / back-propagated through weight-norm by one kernel each per step
        self.bank = WeightBank(self.model, self._flat_p, self._flat_g, Fn.get_precision(), Pm)
        self._zero_idx = self._plan_grad_zero() if bool(config.get('sparse_grad_zero', True)) else None
        if lr_scheduler is not None:
            self._step_size, self._gamma = int(lr_param.get('step_size', 100000)), float(lr_param.get('gamma', 0.5))
            self.scheduler = 'StepLR'
        else:
            self._step_size, self._gamma = 0, 1.0
            self.scheduler = None
        self.optimizer.step_size, self.optimizer.gamma = self._step_size, self._gamma
        self.iteration = 0
        self.model.train()
        # Jitter (layers_vq.py:353-379) draws on the host in the reference; 'device' (default here) draws the same
        # distribution in a kernel so that recipes with jitter_p != 0 (AISHELL-3) keep the captured graph
        jit = getattr(self.model, 'jitter', None)
        if jit is not None and hasattr(jit, 'draws'):
            jit.draws = str(config.get('jitter_draws', 'device' if self.device.type == 'cuda' else 'host'))
        # bucketed all-reduce: trigger name (model/vqvae.py set_grad_ready_callback) -> slice of the flat buffers
        # its CTAs are placed ahead of the queued GEMM CTAs (stream priority, kept by graph capture): the collective's
        # progress beside backward sets the exposed tail (profiles/r02g_dp_sweep.txt)
        self._comm = (torch.cuda.Stream(device=self.device, priority=int(config.get('comm_priority',
                                                                                  os.environ.get('VNB_COMM_PRIO', -1))))
                      if self._side_wanted() else None)
        # gradient slices all-reduced while backward still runs: (decoder, encoder) counts; what is left after the
        # last one (the lowest encoder group, codebook, speaker table, conditioning convs) follows backward
        self._n_buckets = tuple(config.get('allreduce_buckets', (6, 6)))     # measured best at 2 and 8 GPUs (profiles/r02g_*)
        # diagnostics (tools/diag_ddp_timeline.py): GPU-timer stamps at the step's milestones, readable after a replay
        self._stamps = (torch.zeros(64, dtype=torch.int64, device=self.device)
                        if bool(config.get('debug_stamps', False)) and self.device.type == 'cuda' else None)
        self._stamp_next = 10
        self._early_segments = set()
        self.config_early_cond = bool(int(config.get('early_cond_bucket', os.environ.get('VNB_EARLY_COND', 1))))
        self._buckets = self._plan_buckets()
        self._reduced = []
        self.overlap_allreduce = bool(config.get('overlap_allreduce', True))

        self.use_graph = bool(config.get('cuda_graph', True)) and self.device.type == 'cuda'
        # weight-gradient GEMMs on a second stream, beside the data-gradient chain (functional.set_side_stream)
        self._side = (torch.cuda.Stream(device=self.device)
                      if self.device.type == 'cuda' and bool(config.get('wgrad_stream', True)) else None)
        self.graph_warmup = int(config.get('graph_warmup', 3))
        # the decoder's share of the two parameter-sized weight-norm passes on the side stream, beside the encoder's
        # forward / backward (bank.py segments)
        self.bank_overlap = bool(config.get('bank_overlap', True))
        self._early_finished = False
        self._graphs = {}                 # (B, X, T) -> captured step
        self._cap_stream, self._cap_stream_prio = None, None
        self._main_prio = int(config.get('main_priority', os.environ.get('VNB_MAIN_PRIO', 0)))
        self._eager_steps = 0
        self.last_step_on_graph = False
        self.launches_per_step = 0

    def _side_wanted(self):
        return self.device.type == 'cuda'

    def _plan_buckets(self):
        """Slices of the flat gradient buffer that are complete at known points of the backward pass, keyed by the
        model's gradient-ready notification (model/vqvae.py set_grad_ready_callback): the upper half of the
        decoder's blocks (+ final layers) when the gradient reaches the middle of the decoder, the rest of the
        decoder at the decoder's input, the upper half of the encoder in the middle of the encoder.  What remains
        (lower encoder half, codebook, speaker table) is reduced after backward.  {} when the parameter order does
        not allow it (sub-module slices not contiguous)."""
        spans, order, off = {}, [], 0
        offs = {id(p): o for p, o in zip(self._params, self._offsets)}
        flat_order = sorted(((offs[id(p)], n, p) for n, p in self.model.named_parameters() if p.requires_grad),
                            key=lambda t: t[0])
        for off, n, p in flat_order:
            parts = n.split('.')
            key = '.'.join(parts[:3]) if parts[0] in ('encoder', 'decoder') else parts[0]
            if '.conv_cond.' in n:
                key = 'cond'                                # trailing group, reduced after backward
            if key not in spans:
                if order and spans[order[-1]][1] != off:
                    return {}
                spans[key] = [off, off]
                order.append(key)
            elif spans[key][1] != off:
                return {}                                   # a sub-module's parameters are not contiguous
            spans[key][1] = off + (p.numel() + 3) // 4 * 4
            off = spans[key][1]
        dec = [k for k in order if k.startswith('decoder.layers.')]
        enc = [k for k in order if k.startswith('encoder.encode.')]
        dec_all = [k for k in order if k.startswith('decoder.')]
        plan = {}
        for k in spans:                     # 16-byte aligned bucket starts (the multimem kernel moves float4): a
            spans[k][0] = (spans[k][0] + 3) // 4 * 4    # start moved UP leaves those elements to the later bucket

        def _end(key):                      # aligned end of a slice = aligned start of the next one (or the buffer end)
            i = order.index(key)
            return spans[order[i + 1]][0] if i + 1 < len(order) else (off + 3) // 4 * 4

        def _cuts(n, parts):                # split n modules into `parts` groups: start indices of groups 1..parts-1
            return sorted({(n * j) // parts for j in range(1, parts)} - {0})

        self._dec_buckets_complete = False
        contiguous = lambda ks: bool(ks) and order.index(ks[-1]) - order.index(ks[0]) == len(ks) - 1
        if len(dec) >= 4 and contiguous(dec_all):
            # decoder: `nb` slices; slice j (modules dec[c_j] ..) is complete when the gradient reaches the OUTPUT of
            # dec[c_j - 1]; the first slice (from the decoder's first module) at the decoder's input
            cuts = _cuts(len(dec), int(self._n_buckets[0]))
            ends = cuts[1:] + [None]
            for c, e in zip(cuts, ends):
                plan[dec[c - 1]] = (spans[dec[c]][0], spans[dec[e]][0] if e is not None else _end(dec_all[-1]))
            plan['decoder'] = (spans[dec_all[0]][0], spans[dec[cuts[0]]][0] if cuts else _end(dec_all[-1]))
            self._dec_buckets_complete = True
        elif contiguous(dec_all):
            plan['decoder'] = (spans[dec_all[0]][0], _end(dec_all[-1]))
            self._dec_buckets_complete = True
        if len(enc) >= 4 and contiguous(enc):
            cuts = _cuts(len(enc), int(self._n_buckets[1]) + 1)      # the lowest group stays for after backward
            ends = cuts[1:] + [None]
            for c, e in zip(cuts, ends):
                plan[enc[c - 1]] = (spans[enc[c]][0], spans[enc[e]][0] if e is not None else _end(enc[-1]))
        if 'cond' in spans and bool(self.config_early_cond):
            # the conditioning convs' gradients are final when the gradient reaches the speaker embedding the decoder was
            # given (model/vqvae.py marks it 'cond': functional.CondAllFn's backward has run) - mid-backward, not at its end
            plan['cond'] = (spans['cond'][0], _end('cond'))
        return plan

    def _plan_grad_zero(self):
        """zero_grad (trainer/basic.py:57) without the 125 MB memset: the weight-gradient GEMMs WRITE dW into the v
        slots of the flat gradient buffer (99.6 % of it), so only the slots that are accumulated into - biases,
        weight_g, GroupNorm affine, codebook, speaker table - need zeroing.  -> int64 index tensor of those elements
        (None: nothing to gain).  Every conv of the model mirrors is used in every forward, which is what makes the
        v slots fully rewritten each step."""
        if not self.bank.by_ptr:
            return None
        total = self._flat_g.numel()
        keep = torch.ones(total, dtype=torch.bool)
        base = self._flat_p.data_ptr()
        for e in self.bank.by_ptr.values():
            off = (e.v.data_ptr() - base) // 4
            keep[off:off + e.v.numel()] = False
        idx = keep.nonzero().view(-1)
        if idx.numel() * 8 > total:
            return None
        return idx.to(self.device)

    # ---- flat parameter / gradient buffers -------------------------------------------------------
    def _flatten(self):
        """Parameters, gradients (and, in FlatAdam, both moments) as views of flat fp32 buffers.  Every parameter starts
        on a 16-byte boundary (the kernels read biases / GroupNorm affine / weight rows as float4): a 513-entry bias
        must not shift everything behind it.  The padding elements stay zero in all four buffers."""
        named = [(n, p) for n, p in self.model.named_parameters() if p.requires_grad]
        params = [p for _, p in named]                  # reference order = torch.optim's state indices
        # flat order: the conditioning convs of the decoder blocks go LAST - their weight gradients are produced by
        # one batched kernel after the whole decoder has been back-propagated (functional.CondAllFn), so they must
        # not sit inside the decoder slices that the data-parallel trainer all-reduces while backward still runs
        order = sorted(range(len(named)), key=lambda i: ('.conv_cond.' in named[i][0], i))
        offsets, total = [0] * len(named), 0
        for i in order:
            offsets[i] = total
            total += (params[i].numel() + 3) // 4 * 4
        flat_p = torch.zeros(total, dtype=torch.float32, device=self.device)
        flat_g = self._symmetric_grad_buffer(total)
        if flat_g is None:
            flat_g = torch.zeros(total, dtype=torch.float32, device=self.device)
        for p, off in zip(params, offsets):
            n = p.numel()
            flat_p[off:off + n].copy_(p.data.reshape(-1))
            p.data = flat_p[off:off + n].view_as(p)
            p.grad = flat_g[off:off + n].view_as(p)
        self._params, self._flat_p, self._flat_g, self._offsets = params, flat_p, flat_g, offsets
        self._grad_len = total                          # a multiple of 4 by construction

    def _symmetric_grad_buffer(self, total):
        """Data parallel on NVSwitch: the flat gradient buffer is allocated in symmetric memory with a multicast
        mapping so that csrc/allreduce.cu can have the switch sum and broadcast it (multimem.ld_reduce / .st).
        The rendezvous is plumbing (torch symmetric memory); the collective itself is our kernel.  Returns None -
        the NCCL all-reduce is used then - when there is one rank, no CUDA, ``allreduce: nccl`` is configured or
        the rendezvous / multicast mapping is not available."""
        self._mc = None
        dist = _dist()
        if dist is None or self.device.type != 'cuda' or self._allreduce_kind != 'multimem':
            return None
        try:
            import torch.distributed._symmetric_memory as symm
            world, rank = dist.get_world_size(), dist.get_rank()
            group = dist.group.WORLD
            padded = (total + 3) // 4 * 4
            buf = symm.empty(padded, dtype=torch.float32, device=self.device)
            buf.zero_()
            hb = symm.rendezvous(buf, group)
            nblocks = int(self._allreduce_blocks) or (16 if world <= 2 else 8)
            flags = symm.empty(128 * world, dtype=torch.int32, device=self.device)
            flags.zero_()
            hf = symm.rendezvous(flags, group)
            mc = int(hb.multicast_ptr)
            ok = torch.tensor([1 if mc != 0 else 0], device=self.device)
            dist.all_reduce(ok, op=dist.ReduceOp.MIN)         # also orders the zero-fills before any peer's first signal
            torch.cuda.synchronize()
            if int(ok) == 0:
                return None
            self._mc = dict(mc=mc, buf=buf, flags=flags, handles=(hb, hf), rank=rank, world=world, nblocks=nblocks,
                            flag_ptrs=torch.tensor([int(p) for p in hf.buffer_ptrs], dtype=torch.int64, device=self.device))
            return buf[:total]
        except Exception as e:                                   # no NVSwitch multicast on this machine / torch build
            import warnings
            warnings.warn('multimem all-reduce unavailable (%s: %s); using NCCL' % (type(e).__name__, e))
            self._mc = None
            return None

    def _stamp(self, slot):
        if self._stamps is not None:
            from .. import _C
            _C.call('debug_stamp', self._stamps, slot)

    def _all_reduce_mean(self, lo, hi):
        """Mean over ranks of flat_g[lo:hi] on the current stream (lo, hi multiples of 4 on the multimem path)."""
        if self._stamps is not None and self._stamp_next < 62:
            k = self._stamp_next
            self._stamp_next += 2
            self._stamp(k)
            self._all_reduce_mean_(lo, hi)
            self._stamp(k + 1)
            return
        self._all_reduce_mean_(lo, hi)

    def _all_reduce_mean_(self, lo, hi):
        import torch.distributed as dist
        if self._mc is not None:
            from .. import _C
            m = self._mc
            _C.call('allreduce_multimem', m['mc'], lo, hi - lo, m['flag_ptrs'], m['rank'], m['world'], 1.0 / m['world'],
                    m['nblocks'])
        else:
            dist.all_reduce(self._flat_g[lo:hi])

    # ---- one optimisation step on device-resident inputs -----------------------------------------
    def _step_body(self, *inputs, jitter_src=None):
        """zero_grad -> forward -> backward -> [all-reduce] -> clip + Adam + StepLR
        (trainer/basic.py:57-72).  Returns (loss-detail keys, stacked 0-d device scalars)."""
        if self._zero_idx is not None:
            self._flat_g.index_fill_(0, self._zero_idx, 0.0)
        else:
            self._flat_g.zero_()
        _lvq.defer_items(True)
        Fn.set_bank(self.bank)
        Fn.set_side_stream(self._side)
        if hasattr(Pm, 'stats_arena_begin') and self.device.type == 'cuda':
            Pm.stats_arena_begin(self.device)
        try:
            split = self.bank_overlap and self._side is not None and len(self.bank.segments) > 1
            if split:
                # encoder weights on this stream; the rest (decoder: 2/3 of the parameters) on the side stream while
                # the encoder's forward runs - the first decoder conv waits for it (bank.weights)
                ev0 = torch.cuda.Event()
                ev0.record()
                self._side.wait_event(ev0)
                self.bank.prepare(0)
                with torch.cuda.stream(self._side):
                    ev1 = torch.cuda.Event()
                    for sidx in range(1, len(self.bank.segments)):
                        self.bank.prepare(sidx, event=ev1)
                    ev1.record()
                self.bank.fresh = True
            else:
                self.bank.prepare()
            if jitter_src is not None:
                self._saved_sources = self.model.jitter.__dict__.get('sources')
                self.model.jitter.sources = lambda length: jitter_src
            dist = _dist()
            self._reduced = []
            overlap = (dist is not None and self._allreduce_kind != 'none' and self._side is not None and self._comm is not None and bool(self._buckets)
                       and self.overlap_allreduce)
            self._early_finished = False
            self._early_segments = set()
            if overlap:
                _model_mod.set_grad_ready_callback(self._grad_ready)                 # hooks registered during forward
            elif split and dist is None:
                _model_mod.set_grad_ready_callback(self._decoder_done)
            self._stamp_next = 10
            self._stamp(0)
            output, loss, detail = self.model(list(inputs))
            for sidx in range(len(self.bank.segments)):
                self.bank.wait_ready(sidx)          # joins the side stream even if a segment was never used
            self._stamp(1)
            loss.backward()
            self._stamp(2)
            _model_mod.set_grad_ready_callback(None)
            if self._side is not None:
                torch.cuda.current_stream().wait_stream(self._side)
            # data parallel: sum the RAW weight gradients (weight-norm's backward is linear in dW, so it commutes
            # with the sum); the slices that completed during backward are already in flight on the comm stream
            if dist is not None and self._allreduce_kind != 'none':       # 'none': independent replicas (diagnostic)
                # every collective of the step goes through the communication stream (one ordered sequence on
                # every rank; the multimem kernel's flag words are per sequence)
                comm = self._comm if (self._comm is not None and self.device.type == 'cuda') else None
                if comm is not None:
                    comm.wait_stream(torch.cuda.current_stream())
                with (torch.cuda.stream(comm) if comm is not None else contextlib.nullcontext()):
                    pos, total = 0, self._grad_len
                    for lo, hi in sorted(self._reduced):
                        if lo > pos:
                            self._all_reduce_mean(pos, lo)
                        pos = max(pos, hi)
                    if pos < total:
                        self._all_reduce_mean(pos, total)
                if comm is not None:
                    torch.cuda.current_stream().wait_stream(comm)
            self._stamp(3)
            for sidx in range(len(self.bank.segments)):
                if sidx not in self._early_segments:      # (the decoder's went through on a side stream already)
                    self.bank.finish_backward(sidx)
            self._stamp(4)
            if dist is not None and self._mc is None and self._allreduce_kind != 'none':
                self._flat_g.mul_(1.0 / dist.get_world_size())      # the multimem kernel already scaled
        finally:
            if jitter_src is not None:                    # the pinned table is for THIS step only
                if self._saved_sources is None:
                    self.model.jitter.__dict__.pop('sources', None)
                else:
                    self.model.jitter.sources = self._saved_sources
            _model_mod.set_grad_ready_callback(None)
            if hasattr(Pm, 'stats_arena_end'):
                Pm.stats_arena_end()
            Fn.set_side_stream(None)
            Fn.set_bank(None)
            _lvq.defer_items(False)
        self.optimizer.step(self.max_grad_norm, self._step_size, self._gamma)
        self._stamp(5)
        self.bank.fresh = False
        keys = list(detail.keys())
        scal = torch.stack([detail[k].t.reshape(()).float() for k in keys])
        return keys, scal

    def _decoder_done(self, name):
        """Single-GPU twin of _grad_ready: when the gradient reaches the decoder's input every decoder weight gradient
        is final (its GEMMs are queued on the side stream), so the decoder's weight-norm backward runs there while the
        encoder is back-propagated."""
        if name != 'decoder' or self._early_finished:
            return
        segs = self._decoder_segments()
        self._side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(self._side):
            for sidx in segs:
                self.bank.finish_backward(sidx)
        self._early_segments.update(segs)
        self._early_finished = True

    def _decoder_segments(self):
        """Bank segments whose weight gradients are final when the gradient reaches the decoder's input: the decoder's
        convs WITHOUT the conditioning convs (their gradients come from one batched kernel later, functional.CondAllFn)."""
        return [i for i, kind in enumerate(self.bank.segment_kinds) if kind == 1]

    def _grad_ready(self, name):
        """Called from a tensor hook during backward (model/vqvae.py _mark): the modules after ``name`` have been
        back-propagated and their weight-gradient GEMMs are queued on the side stream, so their slice of the flat
        gradient buffer is all-reduced on the communication stream while the rest of backward runs."""
        rng = self._buckets.get(name)
        if rng is None or rng in self._reduced:
            return
        self._comm.wait_stream(torch.cuda.current_stream())
        self._comm.wait_stream(self._side)
        with torch.cuda.stream(self._comm):
            self._all_reduce_mean(rng[0], rng[1])
            if name == 'decoder' and self._dec_buckets_complete and not self._early_finished:
                # every decoder slice has been averaged: the decoder's weight-norm backward follows on the
                # communication stream instead of waiting behind the last bucket after backward
                segs = self._decoder_segments()
                for sidx in segs:
                    self.bank.finish_backward(sidx)
                self._early_segments.update(segs)
                self._early_finished = True
        self._reduced.append(rng)

    def _capture(self, *inputs):
        from .. import _C
        static = [t.clone() for t in inputs]
        g = torch.cuda.CUDAGraph()
        n0 = _C.launch_count()
        # `main_priority` (default 0): priority of the capture stream, i.e. of the dependent chain's kernel nodes relative to
        # the weight-gradient side stream (0) and the communication stream (-1).  Measured: -1 / -2 change nothing at
        # 16 x 128 frames (3.04 vs 3.05 ms) - every GEMM fills the SMs on its own, so the streams time-slice whatever the
        # order; kept as a knob.
        prio = self._main_prio
        if self._cap_stream is None or self._cap_stream_prio != prio:
            self._cap_stream = torch.cuda.Stream(device=self.device, priority=prio)
            self._cap_stream_prio = prio
        with (torch.cuda.graph(g, stream=self._cap_stream) if os.environ.get('VNB_CAP_STREAM', '1') != '0' else torch.cuda.graph(g)):
            keys, scal = self._step_body(*static)
        self.launches_per_step = int(_C.launch_count() - n0)      # our kernels inside one replay
        return dict(graph=g, inputs=static, keys=keys, scal=scal)

    def train_step_device(self, *inputs):
        """The step on inputs already resident on the device ((x, y) for vqvae.Model; (x_a, x_b, y) for the
        cross-domain model).  Returns (keys, scalars tensor) without synchronising."""
        jit = getattr(self.model, 'jitter', None)
        needs_host_draws = (jit is not None and jit._probability != 0.0
                            and (getattr(jit, 'draws', 'host') != 'device' or 'sources' in jit.__dict__))
        q = getattr(self.model, 'quantizer', None)
        # an EMA codebook initialises itself from the first training batch (layers_vq.py:192-201): that step must
        # run eagerly, a captured graph would repeat the initialisation on every replay (and .item() cannot be
        # captured) - resolved here, on the host, before any capture
        needs_init = q is not None and hasattr(q, '_is_init') and not q._is_init()
        if (self.use_graph and not needs_host_draws and not needs_init and self._eager_steps >= self.graph_warmup):
            key = tuple(tuple(t.shape) for t in inputs)
            ent = self._graphs.get(key)
            if ent is None:
                torch.cuda.synchronize()
                ent = self._capture(*inputs)
                self._graphs[key] = ent
            for dst, src in zip(ent['inputs'], inputs):
                dst.copy_(src, non_blocking=True)
            ent['graph'].replay()
            self.last_step_on_graph = True
            return ent['keys'], ent['scal']
        self._eager_steps += 1
        self.last_step_on_graph = False
        return self._step_body(*inputs)

    def train_step(self, input, iteration=None):
        assert self.model.training
        keys, scal = self.train_step_device(*[t.to(self.device, non_blocking=True) for t in input])
        vals = scal.tolist()                                   # the step's only device->host sync
        loss_detail = dict(zip(keys, vals))
        self.iteration += 1                                    # (reference :74-77, with its inverted
        return self.iteration, loss_detail                     #  branch resolved the way train.py uses it)

    def valid(self, data_loader):
        loss_detail = dict()
        for i, batch in enumerate(data_loader):
            step_detail = self.valid_step(batch)
            for key, val in step_detail.items():
                loss_detail.setdefault(key, list()).append(val)
        return loss_detail

    def valid_step(self, input):
        self.model.eval()
        Fn.set_bank(self.bank)
        try:
            with torch.no_grad():
                # the operand planes in the bank belong to the parameters BEFORE the last optimiser step (they are
                # materialised at the start of a step; a graph replay never runs python, so no flag can tell):
                # always re-derive them from the current parameters
                self.bank.prepare()
                input = [t.to(self.device) for t in input]
                output, loss, loss_detail = self.model(input)
        finally:
            Fn.set_bank(None)
        self.model.train()
        return loss_detail

    def get_model_info(self):
        return self.model

    def save_checkpoint(self, checkpoint_file):
        dist = _dist()
        if dist is not None and dist.get_rank() != 0:
            return                         # replicas are bit-identical: rank 0 writes (train.py calls this on every rank)
        torch.save({'model': self.model.state_dict(), 'optimizer': self.optimizer.state_dict(),
                    'iteration': self.iteration}, checkpoint_file)
        print('Saved state dict. to {}'.format(checkpoint_file))

    def load_checkpoint(self, checkpoint_file):
        checkpoint_data = torch.load(checkpoint_file, map_location='cpu')
        # copy INTO the flat-buffer views (keeps parameter storage and captured graphs valid)
        self.model.load_state_dict(checkpoint_data['model'])
        self.bank.fresh = False
        self.optimizer.load_state_dict(checkpoint_data['optimizer'])
        self._graphs = {}                  # captured steps hold the old learning rate as a kernel argument
        return checkpoint_data['iteration']
