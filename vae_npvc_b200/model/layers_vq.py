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
            q, _, _ = Pm.vq_gather(z_id.contiguous().view(-1), e_hat, _dummy_zn(B, e_hat.shape[1], T, e_hat), None,
                                   True, False)
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


def _dummy_zn(B, D, T, like):
    # vq_gather needs a zn pointer even when only the gathered rows are wanted
    return torch.empty((B, D, T), dtype=torch.float32, device=like.device)
