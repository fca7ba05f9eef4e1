"""Drop-in mirror of vae_npvc/model/layers.py for the hot-path layers.

Same class names, constructor signatures, attribute names (= state_dict prefixes) and return
values as the reference; the torch modules inside are PARAMETER HOLDERS only - every forward
runs the CUDA op-groups in vae_npvc_b200.functional.  There is no CPU path: calling a layer
with CPU tensors raises.
"""
import math

import torch
import torch.nn as nn

from .. import functional as Fn
from ..prims import ACT_LRELU, ACT_NONE, ACT_RELU  # noqa: F401

EPSILON = 1e-6
PI = math.pi
LOG_2PI = math.log(2.0 * PI)


def conv_params(m):
    """(v, g, bias) of a Conv1d/ConvTranspose1d holder; g is None once weight-norm is removed."""
    if hasattr(m, 'weight_g'):
        return m.weight_v, m.weight_g, m.bias
    return m.weight, None, m.bias


def run_conv(m, x, act=ACT_NONE, scale=1.0):
    """Stride-1 'same' convolution of holder ``m`` on act(x*scale) through the CUDA path."""
    v, g, bias = conv_params(m)
    transposed = isinstance(m, nn.ConvTranspose1d)
    k, dil = m.kernel_size[0], m.dilation[0]
    if m.stride[0] != 1 or m.padding[0] != (k - 1) // 2 * dil or k % 2 == 0:
        raise NotImplementedError('vae_npvc_b200: only stride-1 "same" odd-kernel convolutions are on the hot path '
                                  '(strided vqvae2 stages are out of scope)')
    return Fn.conv1d(x, v, g, bias, transposed, dil, act, scale)


class Conditions(nn.Module):
    """Speaker-embedding table (reference layers.py:12-60)."""

    def __init__(self, cond_num, cond_dim, normalize=False):
        super(Conditions, self).__init__()
        self._embedding = nn.Embedding(cond_num, cond_dim)
        self.target_norm = 1.0 if normalize else None
        self.embed_norm()
        self.cond_num = cond_num
        self.normalize = normalize

    def embed_norm(self):
        if self.target_norm:
            with torch.no_grad():
                w = self._embedding.weight
                w.mul_(self.target_norm / w.norm(dim=1, keepdim=True))

    def forward(self, input, pre_norm=True):
        if self.target_norm:
            # normalised tables are never built by Model (normalize=False, vqvae.py:34-38);
            # the row normalisation is a (cond_num x cond_dim) torch op, the gather is ours
            if pre_norm:
                self.embed_norm()
            w = self._embedding.weight
            table = (self.target_norm * w / w.norm(dim=1, keepdim=True)).contiguous()
        else:
            table = self._embedding.weight
        return Fn.EmbedFn.apply(table, input)

    def sparsity(self):
        w = self._embedding.weight
        s = torch.mm(w, w.t())
        return nn.functional.cross_entropy(s, torch.arange(s.size(0), device=s.device))


class Conv1d_Layernorm_LRelu_Residual(nn.Module):
    """Encoder residual block (reference layers.py:129-178)."""

    def __init__(self, kernel_size=3, channels=128, layers=2, dilation=1, use_causal_conv=False):
        super(Conv1d_Layernorm_LRelu_Residual, self).__init__()
        assert not use_causal_conv, 'causal convolutions are not on the B200 hot path'
        assert (kernel_size - 1) % 2 == 0, 'Not support even number kernel size.'
        self.use_causal_conv = use_causal_conv
        self.total_padding = None
        stack = []
        for j in range(layers):
            d = dilation if j == 0 else 1
            stack += [
                nn.LeakyReLU(negative_slope=0.2),
                nn.Conv1d(channels, channels, kernel_size, dilation=d, padding=(kernel_size - 1) // 2 * d, bias=True),
                nn.GroupNorm(num_groups=1, num_channels=channels, eps=1e-5, affine=True),
            ]
        self.stack = nn.Sequential(*stack)
        self.skip_layer = nn.Conv1d(channels, channels, 1, bias=True)
        self.n_layers = layers
        self.dilation = dilation

    def _params(self):
        ps = []
        for j in range(self.n_layers):
            conv, gn = self.stack[3 * j + 1], self.stack[3 * j + 2]
            ps += list(conv_params(conv)) + [gn.weight, gn.bias]
        return ps + list(conv_params(self.skip_layer))

    def forward(self, c):
        return Fn.enc_block(c, self._params(), self.n_layers, self.dilation)


class DeConv1d_Layernorm_GLU_ResSkip(nn.Module):
    """Gated decoder block (reference layers.py:181-249)."""

    def __init__(self, kernel_size=3, in_channels=128, cond_channels=128, skip_channels=80, dilation=0,
                 dropout=0.0, use_causal_conv=False):
        super(DeConv1d_Layernorm_GLU_ResSkip, self).__init__()
        assert not use_causal_conv, 'causal convolutions are not on the B200 hot path'
        assert (kernel_size - 1) % 2 == 0, 'Not support even number kernel size.'
        assert dropout == 0.0, 'dropout is 0.0 in every recipe; not on the B200 hot path'
        self.use_causal_conv = use_causal_conv
        self.dropout = dropout
        padding = (kernel_size - 1) // 2 * dilation
        self.conv_in = nn.ConvTranspose1d(in_channels, in_channels * 2, kernel_size, padding=padding,
                                          dilation=dilation, bias=True)
        self.norm_layer = nn.GroupNorm(num_groups=2, num_channels=in_channels * 2, eps=1e-5, affine=True)
        if cond_channels is not None and cond_channels > 0:
            self.conv_cond = nn.Conv1d(cond_channels, in_channels * 2, 1, bias=True)
        else:
            self.conv_cond = None
        self.res_skip_layers = nn.Conv1d(in_channels, in_channels + skip_channels, 1, bias=True)
        self.in_channels = in_channels
        self.dilation = dilation

    def _params(self):
        pc = list(conv_params(self.conv_cond)) if self.conv_cond is not None else [None, None, None]
        return (list(conv_params(self.conv_in)) + [self.norm_layer.weight, self.norm_layer.bias] + pc
                + list(conv_params(self.res_skip_layers)))

    def forward(self, x, c):
        """x (B,C,T); c (B,Y,T), or time-invariant as (B,Y,1) / (B,Y).  Returns (x, x_skip)."""
        x, skip = self.forward_acc(x, c, None)
        return x, skip.transpose(1, 2)

    def forward_acc(self, x, c, skip_acc, cond_slot=None):
        """As forward, but the skip output is ADDED IN PLACE into ``skip_acc``, the Decoder's running
        skip sum in its physical channels-last (B, T, S) form (None: a fresh one is returned).
        ``cond_slot`` = (l, shared, is_first): ``c`` is the Decoder's table of conditioning biases of all blocks
        (functional.cond_all) and this block takes row l."""
        if cond_slot is None and c is not None and c.dim() == 3 and c.size(2) == 1:
            c = c.squeeze(2)
        return Fn.dec_block(x, c, self._params(), skip_acc, self.dilation, cond_slot=cond_slot)


def log_loss(x, mu, reduction='frame_mean'):
    """Gaussian NLL with unit variance (reference layers.py:283-296)."""
    B, D, T = x.shape
    if reduction == 'none':
        return 0.5 * (LOG_2PI + (x - mu).pow(2))          # cold API corner: plain torch elementwise
    scale = {'sum': 1.0, 'mean': 1.0 / (B * D * T), 'batch_mean': 1.0 / B, 'frame_mean': 1.0 / (B * T)}[reduction]
    return Fn.LogLossFn.apply(x, mu, scale)


# Unused-by-Model Gaussian helpers kept for API completeness (reference layers.py:252-281)
def GaussianSampler(z_mu, z_lv):
    return torch.randn_like(z_mu) * torch.exp(0.5 * z_lv) + z_mu


def kl_loss(mu, lv):
    return 0.5 * (torch.exp(lv) + mu.pow(2) - lv - 1.0).sum()
