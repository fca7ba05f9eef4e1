"""Drop-in mirror of vae_npvc/model/vqvae2.py: the hierarchical VQ-VAE2 with a global-style-token top level
(recipe egs/vcc20/vae2/conf/train_pytorch_vqvae2.yaml).

Select it with ``model_type: vae_npvc_b200.model.vqvae2``.  Constructor keys (``levels``, ``encoder.i``, ``decoder.i``,
``quantizer.i``, ``use_gst``, ``use_ema``, ``y_num``, ``y_dim``, ``jitter_p``, ``beta``), sub-module names (= state_dict
keys: ``encoders.i.encode.*``, ``encoders.i.z_proj``, ``quantizers.i.*``, ``decoders.i.layers.*`` ...) and the return
tuple of ``forward`` are the reference's.  What is new against the single-level model and where it runs:

* down-sampling stages (kernel 2s, stride s, vqvae2.py:204-212): the existing tcgen05 conv on rows of s frames
  (functional.strided_conv1d);
* ``Encoder.forward`` returns (z_proj(h), h) and h feeds the next level (vqvae2.py:86-91, 244): the trailing LeakyReLU is
  folded into both consumers' operand conversion instead of being materialised;
* ``upsample`` + channel concatenation (vqvae2.py:104-114, 130-143): vnb_upsample_cat_fwd / bwd;
* time-varying conditioning of the middle decoder (vqvae2.py:110): a second 1x1 conv accumulated into the first
  (functional.DecBlockFn's general path); conditioning that is constant over time (the speaker embedding, the style
  embedding) keeps the one-launch batched path;
* the style-token level (vqvae2.py:95-96, layers_gst.py): vnb_time_mean_fwd, four small linears, vnb_gst_attn_fwd.

Differences from the reference, all where the reference cannot run: ``encode`` / ``decode`` / ``infer`` reference
attributes that do not exist (vqvae2.py:54-60) and are not provided; ``load_state_dict`` is the
plain nn.Module one (the reference's override reads an undefined ``self.use_ema``, vqvae2.py:160-172).
"""
import math

import torch
import torch.nn as nn

from .. import functional as Fn
from ..prims import ACT_LRELU, ACT_NONE, ACT_RELU
from .layers import (Conditions, Conv1d_Layernorm_LRelu_Residual, DeConv1d_Layernorm_GLU_ResSkip, conv_params,
                     log_loss, run_conv)
from .layers_gst import StyleTokenLayer
from .layers_vq import EMAVectorQuantizer, Jitter, VectorQuantizer, host_value
from .vqvae import _WeightNormMixin


class Encoder(nn.Module, _WeightNormMixin):
    """Reference vqvae2.py:175-258: stages of [conv (strided when ds != 1), residual blocks, LeakyReLU] + ``z_proj``."""

    def __init__(self, in_channels=[513, 1024, 512, 256], out_channels=[1024, 512, 256, 128],
                 downsample_scales=[1, 1, 1, 1], kernel_size=3, z_channels=128, dilation=True, stack_kernel_size=3,
                 stack_layers=2, stacks=[3, 3, 3, 3], use_weight_norm=True, use_causal_conv=False):
        super(Encoder, self).__init__()
        assert not use_causal_conv, 'Not supported yet.'
        layers = []
        for cin, cout, ds, stack in zip(in_channels, out_channels, downsample_scales, stacks):
            if ds == 1:
                k, pad, stride = kernel_size, (kernel_size - 1) // 2, 1
            else:
                k, pad, stride = ds * 2, ds // 2 + ds % 2, ds
            conv = nn.Conv1d(cin, cout, k, stride=stride, padding=pad)
            conv._vnb_no_bank = stride != 1          # its operand is a re-laid view of the parameter (bank.py skips it)
            layers += [conv]
            for j in range(stack):
                layers += [Conv1d_Layernorm_LRelu_Residual(kernel_size=stack_kernel_size, channels=cout,
                                                           layers=stack_layers, dilation=2 ** j if dilation else 1,
                                                           use_causal_conv=use_causal_conv)]
            layers += [nn.LeakyReLU(negative_slope=0.2)]
        self.encode = nn.Sequential(*layers)
        self.z_proj = nn.Conv1d(out_channels[-1], z_channels, 1)
        if use_weight_norm:
            self.apply_weight_norm()
        self.reset_parameters()

    def forward_fused(self, x, pending=ACT_NONE):
        """-> (z, h_pre): h_pre is the hidden sequence BEFORE the trailing LeakyReLU, which the consumers (``z_proj``
        here, the next level's first conv) apply while converting their operand; ``pending``: the same for ``x``."""
        h = x
        for m in self.encode:
            if isinstance(m, nn.LeakyReLU):
                assert pending == ACT_NONE
                pending = ACT_LRELU
            elif isinstance(m, nn.Conv1d):
                if m.stride[0] != 1:
                    v, g, bias = conv_params(m)
                    h = Fn.strided_conv1d(h, v, g, bias, m.stride[0], act=pending)
                else:
                    h = run_conv(m, h, act=pending)
                pending = ACT_NONE
            else:
                assert pending == ACT_NONE
                h = m(h)
        assert pending == ACT_LRELU
        return run_conv(self.z_proj, h, act=ACT_LRELU), h

    def forward(self, input):
        """Reference semantics: (z_proj(h), h) with h after the LeakyReLU (materialised with a torch op here; the Model
        uses forward_fused)."""
        z, h = self.forward_fused(input)
        return z, nn.functional.leaky_relu(h, 0.2)


class Decoder(nn.Module, _WeightNormMixin):
    """Reference vqvae2.py:261-396.  Unlike vqvae.Decoder the conditioning is used at every frame (vqvae2.py:358)."""

    def __init__(self, in_channels=[128, 256, 512, 1024], out_channels=[256, 512, 1024, 513],
                 upsample_scales=[1, 1, 1, 1], cond_channels=128, skip_channels=80, final_channels=80, kernel_size=5,
                 dilation=True, stack_kernel_size=3, stacks=[3, 3, 3, 3], use_weight_norm=True, use_causal_conv=False):
        super(Decoder, self).__init__()
        assert not use_causal_conv, 'Not supported yet.'
        layers = nn.ModuleList()
        for cin, cout, us, stack in zip(in_channels, out_channels, upsample_scales, stacks):
            if us != 1:
                raise NotImplementedError('up-sampling ConvTranspose stages are not used by the shipped vqvae2 recipe '
                                          '(upsample_scales [1]) and are not on the B200 hot path')
            layers += [nn.ConvTranspose1d(cin, cout, kernel_size, stride=1, padding=(kernel_size - 1) // 2, output_padding=0)]
            for j in range(stack):
                layers += [DeConv1d_Layernorm_GLU_ResSkip(kernel_size=stack_kernel_size, in_channels=cout,
                                                          cond_channels=cond_channels, skip_channels=skip_channels,
                                                          dilation=2 ** j if dilation else 1, dropout=0.0,
                                                          use_causal_conv=use_causal_conv)]
        self.layers = layers
        self.batch_cond = True
        self.final_layer = nn.Sequential(nn.ReLU(), nn.Conv1d(skip_channels, skip_channels, 1), nn.ReLU(),
                                         nn.Conv1d(skip_channels, final_channels, 1))
        if use_weight_norm:
            self.apply_weight_norm()
        self.reset_parameters()

    def forward(self, input):
        """input = (x (B, in, T), c (B, cond, T) or, constant over time, (B, cond, 1))."""
        x, c = input
        gated = [m for m in self.layers if isinstance(m, DeConv1d_Layernorm_GLU_ResSkip)]
        table = shared = None
        if c.size(2) == 1:
            # constant over time: each block's 1x1 conditioning conv is a per-utterance bias; all blocks in one launch
            c = c[:, :, 0].contiguous()
            if len(gated) > 1 and all(m.conv_cond is not None for m in gated) and self.batch_cond:
                params = []
                for m in gated:
                    params += list(conv_params(m.conv_cond)) + [m.conv_in.bias]
                table, shared = Fn.cond_all(c, params)
        else:
            # every block reads the same conditioning sequence: convert it to operand planes once
            c = Fn.Pm.as_cl(c)
            Fn.attach_planes(c, 'id', Fn.Pm.make_planes(c, Fn.get_precision())[0])
        skip = None
        gi = 0
        for layer in self.layers:
            if isinstance(layer, DeConv1d_Layernorm_GLU_ResSkip):
                if table is not None:
                    x, skip = layer.forward_acc(x, table, skip, cond_slot=(gi, shared, gi == 0))
                else:
                    x, skip = layer.forward_acc(x, c, skip)
                gi += 1
            else:
                x = run_conv(layer, x)
        scale = math.sqrt(1.0 / len(self.layers))
        o = run_conv(self.final_layer[1], skip.transpose(1, 2), act=ACT_RELU, scale=scale)
        return run_conv(self.final_layer[3], o, act=ACT_RELU)


class Model(nn.Module):
    """Reference vqvae2.py:11-172.  ``arch`` is the whole recipe YAML dict."""

    def __init__(self, arch):
        super(Model, self).__init__()
        levels = arch.get('levels', 3)
        use_gst = arch.get('use_gst', True)
        use_ema = arch.get('use_ema', True)
        encoders, quantizers, decoders = [], [], []
        for i in range(levels):
            encoders.append(Encoder(**arch[f'encoder.{i}']))
            decoders.append(Decoder(**arch[f'decoder.{i}']))
            if use_gst and i == levels - 1:
                quantizers.append(StyleTokenLayer(**arch[f'quantizer.{i}']))
            elif use_ema:
                quantizers.append(EMAVectorQuantizer(**arch[f'quantizer.{i}']))
            else:
                quantizers.append(VectorQuantizer(**arch[f'quantizer.{i}']))
        self.encoders = nn.ModuleList(encoders)
        self.quantizers = nn.ModuleList(quantizers)
        self.decoders = nn.ModuleList(decoders)
        self.embeds = Conditions(arch.get('y_num', 10), arch.get('y_dim', 128), normalize=False)
        self.jitter = Jitter(probability=arch.get('jitter_p', 0.0))
        self.beta = arch.get('beta', 0.01)
        self.levels = levels
        self.use_gst = use_gst
        if 'precision' in arch:
            Fn.set_precision(arch['precision'])

    def upsample(self, z, target_len):
        """Upsample the last dim of z by frame repetition (vqvae2.py:130-143)."""
        return Fn.upsample_cat([z], target_len)

    def forward(self, input):
        x, y_idx = input                                            # (B, x_dim, T), (B, 1)
        y = self.embeds(y_idx[..., :1]).transpose(1, 2)             # (B, y_dim, 1): constant over time
        # hierarchical encode (vqvae2.py:86-91)
        z_levels, time_levels = [], [x.size(-1)]
        h, pending = x, ACT_NONE
        for i in range(self.levels):
            z_, h = self.encoders[i].forward_fused(h, pending)
            pending = ACT_LRELU
            z_levels.append(z_)
            time_levels.append(z_.size(-1))
        # hierarchical quantise + decode, top level first (vqvae2.py:93-111).  quantised[k] = (z_vq, constant over time?)
        quantised, l_q, l_e, details = [], [], [], []
        z_ = z_levels.pop()
        for i in reversed(range(self.levels)):
            if self.use_gst and i == self.levels - 1:
                z_vq = self.quantizers[i](Fn.TimeMeanFn.apply(z_)).unsqueeze(-1)             # (B, token_dim, 1)
            else:
                z_vq, z_qut_loss, z_enc_loss, vq_detail = self.quantizers[i](z_)
                l_q.append(z_qut_loss)
                l_e.append(z_enc_loss)
                vq_detail = dict(vq_detail)
                vq_detail['quanti_err'] = host_value(z_enc_loss.detach())
                details.append(vq_detail)
                z_vq = self.jitter(z_vq)
            quantised.append(z_vq)
            if i > 0:
                z_ = z_levels.pop()
                z_ = self.decoders[i]((z_, self._cond(quantised, time_levels[i])))
        z_cat = Fn.upsample_cat(quantised, time_levels[0])
        xhat = self.decoders[0]((z_cat, y))                         # upsample(y, T) is constant over time
        z_qut_loss, z_enc_loss = sum(l_q), sum(l_e)
        x_loss = log_loss(xhat, x)
        loss = x_loss + z_qut_loss + self.beta * z_enc_loss
        losses = {'Total': host_value(loss.detach()), 'VQ loss': host_value(z_enc_loss.detach()),
                  'X like': host_value(x_loss.detach())}
        for i, d in enumerate(details):
            losses.update({key + f'.{i}': val for key, val in d.items()})
        return xhat, loss, losses

    @staticmethod
    def _cond(quantised, T):
        """torch.cat([upsample(z_vq, T) ...], dim=1) of the levels quantised so far (vqvae2.py:104,110); a single
        sequence that is constant over time stays (B, C, 1) so the decoder takes its per-utterance-bias path."""
        if len(quantised) == 1 and quantised[0].size(-1) == 1:
            return quantised[0]
        return Fn.upsample_cat(quantised, T)

    def remove_weight_norm(self):
        def _remove(m):
            try:
                torch.nn.utils.remove_weight_norm(m)
            except ValueError:
                return
        self.apply(_remove)
