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
``graph_warmup`` (3), ``precision`` ('bf16' | 'fp32'), ``device`` ('cuda'), ``wgrad_stream`` (True).
"""
from importlib import import_module

import numpy as np
import torch

from .. import functional as Fn
from .. import prims as Pm
from ..bank import WeightBank
from ..model import layers_vq as _lvq


class FlatAdam(object):
    """Adam state over flat buffers, serialised in torch.optim.Adam's state_dict layout so that
    reference checkpoints (trainer/basic.py:109-121) load and save unchanged."""

    def __init__(self, params, flat_p, flat_g, lr, betas=(0.5, 0.999), eps=1e-8):
        self.params = params
        self.flat_p, self.flat_g = flat_p, flat_g
        self.m = torch.zeros_like(flat_p)
        self.v = torch.zeros_like(flat_p)
        self.state = torch.zeros(2, dtype=torch.float32, device=flat_p.device)   # [adam step, sched step]
        self.sqn = torch.zeros(1, dtype=torch.float32, device=flat_p.device)
        self.lr, self.betas, self.eps = lr, betas, eps
        self.param_groups = [{'lr': lr, 'betas': betas, 'eps': eps, 'weight_decay': 0.0, 'amsgrad': False,
                              'params': list(range(len(params)))}]

    def step(self, max_grad_norm, step_size, gamma):
        self.sqn.zero_()
        if max_grad_norm > 0:
            Pm.sqnorm(self.flat_g, self.sqn)
        Pm.adam_step(self.flat_p, self.flat_g, self.m, self.v, self.sqn, float(max_grad_norm), self.lr,
                     self.betas[0], self.betas[1], self.eps, self.state, int(step_size), float(gamma))

    def state_dict(self):
        step = float(self.state[0].item())
        st, off = {}, 0
        for i, p in enumerate(self.params):
            n = p.numel()
            if step > 0:
                st[i] = {'step': torch.tensor(step), 'exp_avg': self.m[off:off + n].view_as(p).clone(),
                         'exp_avg_sq': self.v[off:off + n].view_as(p).clone()}
            off += n
        return {'state': st, 'param_groups': [dict(g) for g in self.param_groups]}

    def load_state_dict(self, sd):
        off, step = 0, 0.0
        for i, p in enumerate(self.params):
            n = p.numel()
            s = sd['state'].get(i)
            if s is not None:
                self.m[off:off + n].copy_(s['exp_avg'].reshape(-1))
                self.v[off:off + n].copy_(s['exp_avg_sq'].reshape(-1))
                step = float(s['step'])
            off += n
        self.state[0] = step
        # The LR is NOT taken from the checkpoint: the reference builds a fresh StepLR on resume,
        # which resets lr to 'initial_lr' (= the config value) and restarts the decay
        # (scheduler state is never saved, trainer/basic.py:109-115); state[1] stays 0.


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

        self.device = torch.device(config.get('device', 'cuda'))
        if self.device.type == 'cuda':
            from .. import _C
            _C.require_cuda()
        self.model = model.to(self.device)
        self.learning_rate = learning_rate
        self.max_grad_norm = max_grad_norm
        self._flatten()
        self.optimizer = FlatAdam(self._params, self._flat_p, self._flat_g, learning_rate)
        # all conv weights materialised / back-propagated through weight-norm by one kernel each per step
        self.bank = WeightBank(self.model, self._flat_p, self._flat_g, Fn.get_precision(), Pm)
        if lr_scheduler is not None:
            self._step_size, self._gamma = int(lr_param.get('step_size', 100000)), float(lr_param.get('gamma', 0.5))
            self.scheduler = 'StepLR'
        else:
            self._step_size, self._gamma = 0, 1.0
            self.scheduler = None
        self.iteration = 0
        self.model.train()
        # contiguous slice of the flat buffers that holds the decoder's parameters (bucketed all-reduce)
        self._comm = torch.cuda.Stream(device=self.device) if self._side_wanted() else None
        self._dec_range = self._param_range('decoder.')
        self._dec_reduced = False
        self.overlap_allreduce = bool(config.get('overlap_allreduce', True))

        self.use_graph = bool(config.get('cuda_graph', True)) and self.device.type == 'cuda'
        # weight-gradient GEMMs on a second stream, beside the data-gradient chain (functional.set_side_stream)
        self._side = (torch.cuda.Stream(device=self.device)
                      if self.device.type == 'cuda' and bool(config.get('wgrad_stream', True)) else None)
        self.graph_warmup = int(config.get('graph_warmup', 3))
        self._graphs = {}                 # (B, X, T) -> captured step
        self._eager_steps = 0
        self.last_step_on_graph = False
        self.launches_per_step = 0

    def _side_wanted(self):
        return self.device.type == 'cuda'

    def _param_range(self, prefix):
        """(lo, hi) element range of the parameters whose name starts with ``prefix`` when they are contiguous in
        the flat buffers, else None."""
        off, lo, hi, seen_after = 0, None, None, False
        for n, p in self.model.named_parameters():
            if not p.requires_grad:
                continue
            if n.startswith(prefix):
                if seen_after:
                    return None
                if lo is None:
                    lo = off
                hi = off + p.numel()
            elif lo is not None:
                seen_after = True
            off += p.numel()
        return (lo, hi) if lo is not None else None

    # ---- flat parameter / gradient buffers -------------------------------------------------------
    def _flatten(self):
        params = [p for p in self.model.parameters() if p.requires_grad]
        total = sum(p.numel() for p in params)
        flat_p = torch.empty(total, dtype=torch.float32, device=self.device)
        flat_g = torch.zeros(total, dtype=torch.float32, device=self.device)
        off = 0
        for p in params:
            n = p.numel()
            flat_p[off:off + n].copy_(p.data.reshape(-1))
            p.data = flat_p[off:off + n].view_as(p)
            p.grad = flat_g[off:off + n].view_as(p)
            off += n
        self._params, self._flat_p, self._flat_g = params, flat_p, flat_g

    # ---- one optimisation step on device-resident inputs -----------------------------------------
    def _step_body(self, x, y, jitter_src=None):
        """zero_grad -> forward -> backward -> [all-reduce] -> clip + Adam + StepLR
        (trainer/basic.py:57-72).  Returns (loss-detail keys, stacked 0-d device scalars)."""
        self._flat_g.zero_()
        _lvq.defer_items(True)
        Fn.set_bank(self.bank)
        Fn.set_side_stream(self._side)
        if hasattr(Pm, 'stats_arena_begin') and self.device.type == 'cuda':
            Pm.stats_arena_begin(self.device)
        try:
            self.bank.prepare()
            if jitter_src is not None:
                self.model.jitter.sources = lambda length: jitter_src
            dist = _dist()
            self._dec_reduced = False
            if (dist is not None and self._side is not None and self._dec_range is not None and self.overlap_allreduce):
                self.model._decoder_backward_done = self._reduce_decoder_grads       # registered during forward
            output, loss, detail = self.model([x, y])
            loss.backward()
            self.model._decoder_backward_done = None
            if self._side is not None:
                torch.cuda.current_stream().wait_stream(self._side)
            # data parallel: sum the RAW weight gradients (weight-norm's backward is linear in dW, so it commutes
            # with the sum); the decoder's two thirds of the buffer were already started behind its backward pass
            if dist is not None:
                if self._dec_reduced:
                    lo, hi = self._dec_range
                    if lo > 0:
                        dist.all_reduce(self._flat_g[:lo])
                    if hi < self._flat_g.numel():
                        dist.all_reduce(self._flat_g[hi:])
                    torch.cuda.current_stream().wait_stream(self._comm)
                else:
                    dist.all_reduce(self._flat_g)
            self.bank.finish_backward()
            if dist is not None:
                self._flat_g.mul_(1.0 / dist.get_world_size())
        finally:
            self.model._decoder_backward_done = None
            if hasattr(Pm, 'stats_arena_end'):
                Pm.stats_arena_end()
            Fn.set_side_stream(None)
            Fn.set_bank(None)
            _lvq.defer_items(False)
        self.optimizer.step(self.max_grad_norm, self._step_size, self._gamma)
        self.bank.fresh = False
        keys = list(detail.keys())
        scal = torch.stack([detail[k].t.reshape(()).float() for k in keys])
        return keys, scal

    def _reduce_decoder_grads(self, grad):
        """Tensor hook on the decoder input's gradient: every decoder layer has been back-propagated (its weight
        gradients are queued on the side stream), so their slice of the flat gradient buffer is all-reduced on a
        communication stream while the quantiser and the encoder are still being back-propagated."""
        import torch.distributed as dist
        main = torch.cuda.current_stream()
        self._comm.wait_stream(main)
        self._comm.wait_stream(self._side)
        lo, hi = self._dec_range
        with torch.cuda.stream(self._comm):
            dist.all_reduce(self._flat_g[lo:hi])
        self._dec_reduced = True
        return None

    def _capture(self, x, y):
        from .. import _C
        static_x, static_y = x.clone(), y.clone()
        g = torch.cuda.CUDAGraph()
        n0 = _C.launch_count()
        with torch.cuda.graph(g):
            keys, scal = self._step_body(static_x, static_y)
        self.launches_per_step = int(_C.launch_count() - n0)      # our kernels inside one replay
        return dict(graph=g, x=static_x, y=static_y, keys=keys, scal=scal)

    def train_step_device(self, x, y):
        """The step on inputs already resident on the device.  Returns (keys, scalars tensor)
        without synchronising."""
        jit = getattr(self.model, 'jitter', None)
        needs_host_draws = jit is not None and jit._probability != 0.0
        if (self.use_graph and not needs_host_draws and self._eager_steps >= self.graph_warmup):
            key = tuple(x.shape)
            ent = self._graphs.get(key)
            if ent is None:
                torch.cuda.synchronize()
                ent = self._capture(x, y)
                self._graphs[key] = ent
            ent['x'].copy_(x, non_blocking=True)
            ent['y'].copy_(y, non_blocking=True)
            ent['graph'].replay()
            self.last_step_on_graph = True
            return ent['keys'], ent['scal']
        self._eager_steps += 1
        self.last_step_on_graph = False
        return self._step_body(x, y)

    def train_step(self, input, iteration=None):
        assert self.model.training
        x, y = [t.to(self.device, non_blocking=True) for t in input]
        keys, scal = self.train_step_device(x, y)
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
                if not self.bank.fresh:
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
        torch.save({'model': self.model.state_dict(), 'optimizer': self.optimizer.state_dict(),
                    'iteration': self.iteration}, checkpoint_file)
        print('Saved state dict. to {}'.format(checkpoint_file))

    def load_checkpoint(self, checkpoint_file):
        checkpoint_data = torch.load(checkpoint_file, map_location='cpu')
        # copy INTO the flat-buffer views (keeps parameter storage and captured graphs valid)
        self.model.load_state_dict(checkpoint_data['model'])
        self.bank.fresh = False
        self.optimizer.load_state_dict(checkpoint_data['optimizer'])
        return checkpoint_data['iteration']
