"""Drop-in mirror of vae_npvc/model/vqvae.py: ``Model``, ``Encoder``, ``Decoder``.

Select it from a recipe YAML with ``model_type: vae_npvc_b200.model.vqvae`` (the reference's
plugin mechanism, vae_npvc/trainer/basic.py:24-26).  Constructor keys, sub-module names
(= state_dict prefixes, 207 keys for the VCC2020 model), return tuples and the nn.Module
protocol are the reference's; a reference checkpoint loads unchanged.  All arithmetic runs on
the CUDA op-groups; calling the model on CPU tensors raises.

Optional NEW config key (reference-preserving default): ``precision: 'bf16' | 'bf16x2' | 'fp32'`` - GEMM
operand planes (see functional.set_precision).
"""
import math

import torch
import torch.nn as nn

from .. import functional as Fn
from ..prims import ACT_LRELU, ACT_NONE, ACT_RELU
from .layers import (Conditions, Conv1d_Layernorm_LRelu_Residual, DeConv1d_Layernorm_GLU_ResSkip, conv_params,
                     log_loss, run_conv)
from .layers_vq import EMAVectorQuantizer, Jitter, VectorQuantizer, host_value


class _WeightNormMixin:
    """apply/remove weight-norm and the reference's (ineffective) reset, vqvae.py:194-217.

    reset_parameters() in the reference runs AFTER apply_weight_norm() and writes into the
    derived ``weight`` attribute, which the next forward recomputes from weight_g/weight_v
     - so it only consumes RNG draws.  It is reproduced so that constructing the
    model under the same torch seed yields bit-identical parameters to the reference."""

    def remove_weight_norm(self):
        def _remove(m):
            try:
                nn.utils.remove_weight_norm(m)
            except ValueError:
                return
        self.apply(_remove)

    def apply_weight_norm(self):
        def _apply(m):
            if isinstance(m, (nn.Conv1d, nn.ConvTranspose1d)):
                nn.utils.weight_norm(m)
        import warnings
        with warnings.catch_warnings():
            warnings.simplefilter('ignore')
            self.apply(_apply)

    def reset_parameters(self):
        def _reset(m):
            if isinstance(m, (nn.Conv1d, nn.ConvTranspose1d)):
                nn.init.kaiming_normal_(m.weight.data, nonlinearity='relu')
        self.apply(_reset)


class Encoder(nn.Module, _WeightNormMixin):
    """Reference vqvae.py:122-217.  ``encode`` is an nn.Sequential with the reference's indices
    (conv, residual blocks, LeakyReLU, ..., final 1x1 conv)."""

    def __init__(self, in_channels=[513, 1024, 512, 256], out_channels=[1024, 512, 256, 128],
                 downsample_scales=[1, 1, 1, 1], kernel_size=3, z_channels=128, dilation=True,
                 stack_kernel_size=3, stack_layers=2, stacks=[3, 3, 3, 3], use_weight_norm=True,
                 use_causal_conv=False):
        super(Encoder, self).__init__()
        assert not use_causal_conv, 'Not supported yet.'
        layers = []
        for cin, cout, ds, stack in zip(in_channels, out_channels, downsample_scales, stacks):
            if ds != 1:
                raise NotImplementedError('strided (vqvae2) stages are outside the B200 hot path')
            layers += [nn.Conv1d(cin, cout, kernel_size, stride=1, padding=(kernel_size - 1) // 2)]
            for j in range(stack):
                layers += [Conv1d_Layernorm_LRelu_Residual(kernel_size=stack_kernel_size, channels=cout,
                                                           layers=stack_layers, dilation=2 ** j if dilation else 1,
                                                           use_causal_conv=use_causal_conv)]
            layers += [nn.LeakyReLU(negative_slope=0.2)]
        layers += [nn.Conv1d(out_channels[-1], z_channels, 1)]
        self.encode = nn.Sequential(*layers)
        if use_weight_norm:
            self.apply_weight_norm()
        self.reset_parameters()

    def forward(self, input):
        h = input
        pending = ACT_NONE                  # a LeakyReLU entry is fused into the next conv's operand
        for i, m in enumerate(self.encode):
            if isinstance(m, nn.LeakyReLU):
                pending = ACT_LRELU
                continue
            elif isinstance(m, nn.Conv1d):
                h = run_conv(m, h, act=pending)
                pending = ACT_NONE
            else:
                assert pending == ACT_NONE
                h = m(h)
            h = _mark(h, 'encoder.encode.%d' % i)
        assert pending == ACT_NONE
        return h


# ---- gradient-ready notifications for the data-parallel trainer ---------------------------------
_GRAD_READY = None


def set_grad_ready_callback(cb):
    """``cb(name)`` is called during backward when the gradient wrt the OUTPUT of sub-module ``name`` exists, i.e.
    when every module after ``name`` in forward order has been back-propagated ('decoder' = the decoder's input).
    The data-parallel trainer starts all-reducing the finished part of the flat gradient buffer from there
    (trainer/basic.py); None switches the hooks off."""
    global _GRAD_READY
    _GRAD_READY = cb


def _mark(t, name):
    cb = _GRAD_READY
    if cb is not None and torch.is_tensor(t) and t.requires_grad:
        def hook(grad, cb=cb, name=name):
            cb(name)
            return None
        t.register_hook(hook)
    return t


class Decoder(nn.Module, _WeightNormMixin):
    """Reference vqvae.py:220-343."""

    def __init__(self, in_channels=[128, 256, 512, 1024], out_channels=[256, 512, 1024, 513],
                 upsample_scales=[1, 1, 1, 1], cond_channels=128, skip_channels=80, final_channels=80,
                 kernel_size=5, dilation=True, stack_kernel_size=3, stacks=[3, 3, 3, 3], use_weight_norm=True,
                 use_causal_conv=False):
        super(Decoder, self).__init__()
        assert not use_causal_conv, 'Not supported yet.'
        layers = nn.ModuleList()
        for cin, cout, us, stack in zip(in_channels, out_channels, upsample_scales, stacks):
            if us != 1:
                raise NotImplementedError('strided (vqvae2) stages are outside the B200 hot path')
            layers += [nn.ConvTranspose1d(cin, cout, kernel_size, stride=1, padding=(kernel_size - 1) // 2,
                                          output_padding=0)]
            for j in range(stack):
                layers += [DeConv1d_Layernorm_GLU_ResSkip(kernel_size=stack_kernel_size, in_channels=cout,
                                                          cond_channels=cond_channels, skip_channels=skip_channels,
                                                          dilation=2 ** j if dilation else 1, dropout=0.0,
                                                          use_causal_conv=use_causal_conv)]
        self.layers = layers
        self.batch_cond = True            # one launch for the conditioning convs of all blocks (functional.CondAllFn)
        self.final_layer = nn.Sequential(nn.ReLU(), nn.Conv1d(skip_channels, skip_channels, 1), nn.ReLU(),
                                         nn.Conv1d(skip_channels, final_channels, 1))
        if use_weight_norm:
            self.apply_weight_norm()
        self.reset_parameters()

    def forward(self, input):
        """input = (x (B,in,T), c (B,cond,>=1)); only c[:,:,:1] is used (vqvae.py:309) - it is
        time-invariant, so each block's 1x1 cond conv collapses to a per-utterance bias."""
        x, c = input
        c = _mark(c[:, :, 0].contiguous(), 'cond')   # its gradient exists once every conditioning conv has been back-propagated
        skip = None
        # the conditioning 1x1 convs of all gated blocks see the same (B, Y) embedding: one batched launch for all of
        # them (and one for their backward) instead of one small linear per block
        gated = [m for m in self.layers if isinstance(m, DeConv1d_Layernorm_GLU_ResSkip)]
        table = shared = None
        if len(gated) > 1 and all(m.conv_cond is not None for m in gated) and self.batch_cond:
            params = []
            for m in gated:
                params += list(conv_params(m.conv_cond)) + [m.conv_in.bias]
            table, shared = Fn.cond_all(c, params)
        gi = 0
        for i, layer in enumerate(self.layers):
            if isinstance(layer, DeConv1d_Layernorm_GLU_ResSkip):
                if table is not None:
                    x, skip = layer.forward_acc(x, table, skip, cond_slot=(gi, shared, gi == 0))
                else:
                    x, skip = layer.forward_acc(x, c, skip)
                gi += 1
            else:
                x = run_conv(layer, x)
            x = _mark(x, 'decoder.layers.%d' % i)
        # x_out * sqrt(1/len(layers)) then ReLU -> 1x1 -> ReLU -> 1x1 (vqvae.py:316-317)
        scale = math.sqrt(1.0 / len(self.layers))
        o = run_conv(self.final_layer[1], skip.transpose(1, 2), act=ACT_RELU, scale=scale)
        return run_conv(self.final_layer[3], o, act=ACT_RELU)


class Model(nn.Module):
    """Reference vqvae.py:11-119.  ``arch`` is the whole recipe YAML dict."""

    def __init__(self, arch):
        super(Model, self).__init__()
        self.encoder = Encoder(**arch['encoder'])
        self.decoder = Decoder(**arch['decoder'])
        self.use_ema = arch.get('use_ema', False)
        if self.use_ema:
            self.quantizer = EMAVectorQuantizer(arch.get('z_num', 512), arch.get('z_dim', 128), arch.get('mu', 0.9),
                                                reduction='frame_mean')
        else:
            self.quantizer = VectorQuantizer(arch.get('z_num', 512), arch.get('z_dim', 128),
                                             normalize=arch.get('embed_norm', True), reduction='frame_mean')
        self.embeds = Conditions(arch.get('y_num', 10), arch.get('y_dim', 128), normalize=False)
        self.jitter = Jitter(probability=arch.get('jitter_p', 0.0))
        self.beta = arch.get('beta', 0.01)
        if 'precision' in arch:
            Fn.set_precision(arch['precision'])

    def encode(self, input):
        x = input[0] if isinstance(input, (list, tuple)) else input
        z = self.encoder(x)
        return self.quantizer.encode(z)

    def decode(self, input):
        z_idx, y_idx = input
        y = self.embeds(y_idx).transpose(1, 2).contiguous()      # (B, y_dim, n)
        z_vq = self.quantizer.decode(z_idx)
        xhat = self.decoder((z_vq, y))
        # conversion output in the reference's (B, X, T)-contiguous layout (decoder/basic.py:73-75
        # hands it to numpy); forward() keeps the channels-last tensor, the loss consumes it directly
        return xhat.contiguous() if xhat.requires_grad else Fn.Pm.to_nct(xhat)

    def infer(self, input):
        x, y_idx = input
        z_idx = self.encode(x)
        return self.decode((z_idx, y_idx))

    def forward(self, input):
        x, y_idx = input                                          # (B, x_dim, T), (B, 1)
        y = self.embeds(y_idx).transpose(1, 2)                    # (B, y_dim, 1); Decoder reads column 0
        z = self.encoder(x)
        z_vq, z_qut_loss, z_enc_loss, vq_detail = self.quantizer(z)
        z_vq = self.jitter(z_vq)
        z_vq = _mark(z_vq, 'decoder')       # gradient wrt the decoder input: the whole decoder is back-propagated
        xhat = self.decoder((z_vq, y))
        x_loss = log_loss(xhat, x)
        loss = x_loss + z_qut_loss + self.beta * z_enc_loss
        losses = {'Total': host_value(loss.detach()), 'VQ loss': host_value(z_enc_loss.detach()),
                  'X like': host_value(x_loss.detach())}
        losses.update(vq_detail)
        return xhat, loss, losses

    def remove_weight_norm(self):
        def _remove(m):
            try:
                torch.nn.utils.remove_weight_norm(m)
            except ValueError:
                return
        self.apply(_remove)

    def load_state_dict(self, state_dict, *args, **kwargs):
        """Rebuilds the (non-EMA) quantiser when the checkpoint's codebook shape differs
        (reference vqvae.py:106-119)."""
        if not self.use_ema:
            ck, cur = state_dict['quantizer.embeddings'].shape, self.quantizer.embeddings.shape
            if ck != cur:
                print('Embedding size mismatch for model.quantizer: copying a param with shape {} from checkpoint, '
                      'resizing the param with shape {} in current model.'.format(ck, cur))
                dev = self.quantizer.embeddings.device
                self.quantizer = VectorQuantizer(ck[0], ck[1], normalize=self.quantizer.normalize,
                                                 reduction=self.quantizer.reduction).to(dev)
        return super(Model, self).load_state_dict(state_dict, *args, **kwargs)
