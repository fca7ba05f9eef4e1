"""Drop-in mirror of vae_npvc/model/layers_vq.py (VectorQuantizer, EMAVectorQuantizer, Jitter).

Same constructor signatures, attributes, buffers and return tuples as the reference; the
lookup / gather / loss / backward / EMA update run as CUDA kernels (functional.py).
"""
import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from .. import functional as Fn
from .. import prims as Pm


def _to_time_last(z, time_last):
    return z if time_last else z.transpose(1, 2)


class _Scalar:
    """Deferred ``.item()``: holds a 0-d device tensor; ``float()`` syncs.  Used when a caller
    (the CUDA-graph trainer) asked the model not to synchronise inside the step."""
    __slots__ = ('t',)

    def __init__(self, t):
        self.t = t

    def __float__(self):
        return float(self.t.item())

    def item(self):
        return float(self.t.item())


_DEFER = {'on': False}


def defer_items(on):
    """When on, the ``update_detail`` / ``losses`` dict values are _Scalar wrappers instead of
    python floats (no device sync inside forward)."""
    _DEFER['on'] = bool(on)


def host_value(t):
    return _Scalar(t) if _DEFER['on'] else t.item()


class VectorQuantizer(nn.Module):
    """Reference layers_vq.py:9-163."""

    def __init__(self, z_num, z_dim, normalize=False, reduction='frame_mean'):
        super(VectorQuantizer, self).__init__()
        self.target_norm = 1.0 if normalize else None
        self.embeddings = nn.Parameter(torch.randn(z_num, z_dim, requires_grad=True))
        self.embed_norm()
        self.z_num = z_num
        self.z_dim = z_dim
        self.normalize = normalize
        self.reduction = reduction
        self.quantize = True

    def embed_norm(self):
        if self.target_norm:
            with torch.no_grad():
                self.embeddings.mul_(self.target_norm / self.embeddings.norm(dim=1, keepdim=True))

    def encode(self, z, time_last=True):
        z = _to_time_last(z, time_last)
        B, D, T = z.shape
        with torch.no_grad():
            idx = Fn.VQLookupFn.apply(z, self.embeddings, bool(self.target_norm))
        return idx.view(B, T)

    def decode(self, z_id, time_last=True):
        B, T = z_id.shape
        with torch.no_grad():
            e_hat, _, _ = Pm.vq_prep_codebook(self.embeddings.detach(), bool(self.target_norm), False)
            q, _, _ = Pm.vq_gather(z_id.contiguous().view(-1), e_hat, None, None, True, False, shape=(B, T))
        return q if time_last else q.transpose(1, 2).contiguous()

    def forward(self, z, time_last=True):
        if not self.quantize:
            tensor_0 = torch.tensor(0.0, dtype=torch.float, device=z.device)
            return z, tensor_0, tensor_0, tensor_0
        zt = _to_time_last(z, time_last)
        B, D, T = zt.shape
        norm = bool(self.target_norm)
        if self.reduction == 'none':
            return self._forward_none(zt, time_last)
        z_vq, lq, le, perp, _ = Fn.VQFn.apply(zt, self.embeddings, norm, self.reduction)
        update_detail = {'entropy': host_value(perp)}
        if not time_last:
            z_vq = z_vq.transpose(1, 2).contiguous()
        return z_vq, lq, le, update_detail

    def _forward_none(self, zt, time_last):
        """reduction='none' (never used by Model): element-wise losses.  The lookup runs on the
        CUDA kernel; the element-wise loss algebra of layers_vq.py:119-142 is plain torch."""
        B, D, T = zt.shape
        zf = zt.transpose(1, 2).contiguous().view(-1, D)
        if self.target_norm:
            z_n = self.target_norm * zf / zf.norm(dim=1, keepdim=True)
            self.embed_norm()
            emb = self.target_norm * self.embeddings / self.embeddings.norm(dim=1, keepdim=True)
        else:
            z_n, emb = zf, self.embeddings
        with torch.no_grad():
            idx = Fn.VQLookupFn.apply(zt, self.embeddings, bool(self.target_norm))
            counts = torch.bincount(idx, minlength=self.z_num).float()
            p = counts / idx.numel()
            perp = torch.exp(-torch.sum(p * torch.log(p + 1e-10)))
        q = emb.index_select(0, idx)
        lq = F.mse_loss(q, z_n.detach(), reduction='none')
        le = F.mse_loss(q.detach(), z_n, reduction='none')
        if self.target_norm:
            le = le + F.mse_loss(z_n, zf, reduction='none')
        lq, le = lq.view(B, T, D), le.view(B, T, D)
        out = (z_n + (q - z_n).detach()).view(B, T, D)
        if time_last:
            lq, le, out = lq.transpose(1, 2), le.transpose(1, 2), out.transpose(1, 2).contiguous()
        return out, lq, le, {'entropy': host_value(perp)}

    def sparsity(self):
        s = torch.mm(self.embeddings, self.embeddings.t())
        return F.cross_entropy(s, torch.arange(s.size(0), device=s.device))

    def extra_repr(self):
        s = '{z_num}, {z_dim}'
        if self.normalize is not False:
            s += ', normalize=True'
        return s.format(**self.__dict__)


def _dist_world():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        return dist
    return None


class EMAVectorQuantizer(nn.Module):
    """Reference layers_vq.py:166-334.  Buffers and update order are the reference's; the lookup
    uses the OLD codebook, the EMA update happens before the loss, the loss uses rows gathered
    from the old codebook, and (bug-compatibly) the output has no gradient for
    the scalar reductions.  Under torch.distributed the per-code sums / counts are all-reduced
    and the restart rows broadcast from rank 0 so every rank keeps the same codebook."""

    def __init__(self, z_num, z_dim, mu, threshold=1.0, reduction='frame_mean'):
        super(EMAVectorQuantizer, self).__init__()
        self.register_buffer('emb_init', torch.tensor(0).bool())
        self.register_buffer('emb_sum', torch.zeros(z_num, z_dim))
        self.register_buffer('emb_elem', torch.ones(z_num))
        self.register_buffer('embeddings', torch.zeros(z_num, z_dim))
        self.mu = mu
        self.z_num = z_num
        self.z_dim = z_dim
        self.threshold = threshold
        self.reduction = reduction
        self.quantize = True
        self.update = True
        self._init_host = None          # host-side cache of emb_init (avoids a sync per step)

    def _load_from_state_dict(self, *args, **kwargs):
        self._init_host = None
        return super()._load_from_state_dict(*args, **kwargs)

    def _is_init(self):
        if self._init_host is None:
            self._init_host = bool(self.emb_init.item())
        return self._init_host

    def _random_rows(self, z):
        """K rows of the (tiled, noised when N < K) frame matrix in random order
        (_tile + randperm, layers_vq.py:183-190, 197, 213).  rand().topk() gives K distinct
        uniformly random positions and is CUDA-graph capturable."""
        B, D, T = z.shape
        N, K = B * T, self.z_num
        reps = 1 if N >= K else (K + N - 1) // N
        perm = torch.rand(reps * N, device=z.device).topk(K).indices
        rows = Pm.vq_frame_rows(z, perm, K)
        if N < K:
            rows = rows + torch.randn_like(rows) * (0.01 / np.sqrt(D))
        dist = _dist_world()
        if dist is not None:
            dist.broadcast(rows, 0)
        return rows

    def init_emb(self, z, time_last=True):
        z = _to_time_last(z, time_last).contiguous()
        with torch.no_grad():
            self.emb_init.fill_(True)
            self._init_host = True
            rows = self._random_rows(z)
            self.embeddings.copy_(rows)
            self.emb_sum.copy_(rows)
            self.emb_elem.fill_(1.0)

    def update_emb(self, z, z_idx):
        """z (B,D,T) fp32, z_idx (B*T,) int64."""
        with torch.no_grad():
            B, D, T = z.shape
            zsum, counts = Pm.vq_ema_scatter(z_idx, z, self.z_num)
            n = B * T
            dist = _dist_world()
            if dist is not None:
                dist.all_reduce(zsum)
                dist.all_reduce(counts)
                n *= dist.get_world_size()
            rows = self._random_rows(z)
            stats = Pm.vq_ema_update(zsum, counts, n, self.mu, self.threshold, rows, self.emb_sum, self.emb_elem,
                                     self.embeddings)
        return {'entropy': host_value(stats[0]), 'used_curr': host_value(stats[1]),
                'usage': host_value(stats[2]), 'diff_emb': host_value(stats[3])}

    def encode(self, z, time_last=True):
        z = _to_time_last(z, time_last)
        B, D, T = z.shape
        with torch.no_grad():
            idx = Fn.VQLookupFn.apply(z, self.embeddings, False)
        return idx.view(B, T)

    def decode(self, z_id, time_last=True):
        B, T = z_id.shape
        with torch.no_grad():
            q, _, _ = Pm.vq_gather(z_id.contiguous().view(-1), self.embeddings, None, None, True, False, shape=(B, T))
        return q if time_last else q.transpose(1, 2).contiguous()

    def forward(self, z, time_last=True):
        if not self.quantize:
            tensor_0 = torch.tensor(0.0, dtype=torch.float, device=z.device)
            return z, tensor_0, tensor_0, tensor_0
        zt = _to_time_last(z, time_last).contiguous()
        B, D, T = zt.shape
        if self.training and not self._is_init():
            self.init_emb(zt.detach())
        if self.reduction == 'none':
            return self._forward_none(zt, time_last)
        # clone: the lookup/loss must see the OLD codebook while update_emb overwrites the buffer
        z_vq, le, idx = Fn.EMAVQFn.apply(zt, self.embeddings, self.reduction)
        if self.update and self.training:
            update_detail = self.update_emb(zt.detach(), idx)
        else:
            update_detail = dict()
        if not time_last:
            z_vq = z_vq.transpose(1, 2).contiguous()
        return z_vq, 0.0, le, update_detail

    def _forward_none(self, zt, time_last):
        """reduction='none' (never used by Model): the one mode where the reference DOES apply
        the straight-through estimator (layers_vq.py:310-315).  Lookup on the CUDA kernel,
        element-wise algebra in torch."""
        B, D, T = zt.shape
        zf = zt.transpose(1, 2).contiguous().view(-1, D)
        with torch.no_grad():
            idx = Fn.VQLookupFn.apply(zt, self.embeddings, False)
            q = self.embeddings.index_select(0, idx)
        detail = self.update_emb(zt.detach(), idx) if (self.update and self.training) else dict()
        le = F.mse_loss(q.detach(), zf, reduction='none').view(B, T, D)
        out = (zf + (q - zf).detach()).view(B, T, D)
        if time_last:
            le, out = le.transpose(1, 2), out.transpose(1, 2).contiguous()
        return out, 0.0, le, detail

    def extra_repr(self):
        s = '{z_num}, {z_dim}, mu={mu}, threshold={threshold}'
        return s.format(**self.__dict__)


class Jitter(nn.Module):
    """Reference layers_vq.py:337-383.  The per-step host draws (np.random, same order as the
    reference) are made once per call and turned into a source-column table; the column copy and
    its zero-gradient mask run as one kernel each instead of T tiny in-place copies.
    Bug-compatible: a column is replaced with probability 1 - p (the reference indexes
    [True, False] with a draw that is 1 with probability p, layers_vq.py:364)."""

    def __init__(self, probability=0.12):
        super(Jitter, self).__init__()
        self._probability = probability
        # 'host': np.random draws in the reference's order (the module-level default).  'device': the table is drawn
        # by a kernel from (seed, call counter) - the Trainer switches to it so that the step stays inside its CUDA
        # graph; same distribution, different stream.
        self.draws = 'host'
        self._dev_state = None

    def device_sources(self, length, device):
        if self._dev_state is None or self._dev_state.device != device:
            seed = int(np.random.randint(0, 2 ** 31 - 1))          # derived from the stream the reference would use
            self._dev_state = torch.tensor([seed, 0], dtype=torch.int64, device=device)
        return Pm.jitter_sources(length, self._probability, self._dev_state)

    def sources(self, length):
        p = self._probability
        src = np.arange(length)
        for i in range(length):
            replace = [True, False][np.random.choice([1, 0], p=[p, 1 - p])]
            if replace:
                if i == 0:
                    src[i] = i + 1
                elif i == length - 1:
                    src[i] = i - 1
                else:
                    src[i] = i + np.random.choice([-1, 1], p=[0.5, 0.5])
        return src

    def forward(self, quantized, src=None):
        if self._probability == 0.0 or not self.training:
            return quantized
        if src is None and self.draws == 'device' and 'sources' not in self.__dict__:
            return Fn.JitterFn.apply(quantized, self.device_sources(quantized.size(2), quantized.device))
        if src is None:
            src = self.sources(quantized.size(2))
        src_t = torch.as_tensor(np.asarray(src), dtype=torch.int64).to(quantized.device, non_blocking=True)
        return Fn.JitterFn.apply(quantized, src_t)

    def extra_repr(self):
        return 'jitter_prob={_probability}'.format(**self.__dict__)
