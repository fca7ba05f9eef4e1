"""Cross-domain VQ-VAE (BASELINE.json configs[2], "CDVAE"): two feature domains (513-dim WORLD spectrum + 40-dim MCC)
encoded into ONE shared quantised latent and decoded into both domains.

The reference snapshot ships NO model file for it: only the 513-channel constructor defaults
(vae_npvc/model/vqvae.py:124,223) and unused Gaussian helpers (layers.py:252-281) hint at it.  This composition is
therefore built from the in-tree parts only - ``Encoder`` / ``Decoder`` (vqvae.py:122-343), the shared
``VectorQuantizer`` (layers_vq.py:9-163), ``Conditions`` and ``log_loss`` - wired the way the single-domain ``Model``
wires them (vqvae.py:70-90), and serves as a THROUGHPUT shape (X = 513 and X = 40 through the same kernels), not as a
parity target against a reference model; its arithmetic is pinned against the same composition of the oracle's parts
(oracle/cdvae_oracle.py).

    forward([x_a (B,Xa,T), x_b (B,Xb,T), y_idx (B,1)]) -> ((xhat_a, xhat_b), loss, detail)

Both encoder outputs are quantised in ONE lookup (2B utterances, shared codebook) and every decoder runs ONCE on all
2B quantised sequences, so each domain is reconstructed from itself (self) and from the other domain (cross): four
reconstructions, as in the cross-domain VAE.  Each conv module is used once per step (which is what lets the trainer
write weight gradients straight into their slots).

    loss = sum_d log_loss(xhat_d, [x_d; x_d]) + z_qut_loss + beta * z_enc_loss          (frame_mean over 2B*T frames)

Select with ``model_type: vae_npvc_b200.model.cdvae`` and ``domains: {a: {encoder:..., decoder:...}, b: {...}}``.
"""
import torch
import torch.nn as nn

from .layers import Conditions, log_loss
from .layers_vq import VectorQuantizer, host_value
from .vqvae import Decoder, Encoder


def _cat_batch(tensors):
    """Concatenate logical (B,C,T) channels-last activations along the batch, keeping the channels-last layout."""
    return torch.cat([t.transpose(1, 2) for t in tensors], dim=0).transpose(1, 2)


class Model(nn.Module):
    def __init__(self, arch):
        super(Model, self).__init__()
        self.names = list(arch['domains'].keys())
        assert len(self.names) == 2, 'cross-domain model: exactly two domains'
        self.encoders = nn.ModuleDict({n: Encoder(**arch['domains'][n]['encoder']) for n in self.names})
        self.decoders = nn.ModuleDict({n: Decoder(**arch['domains'][n]['decoder']) for n in self.names})
        self.quantizer = VectorQuantizer(arch.get('z_num', 512), arch.get('z_dim', 128),
                                         normalize=arch.get('embed_norm', True), reduction='frame_mean')
        self.embeds = Conditions(arch.get('y_num', 10), arch.get('y_dim', 128), normalize=False)
        self.beta = arch.get('beta', 0.01)
        if 'precision' in arch:
            from .. import functional as Fn
            Fn.set_precision(arch['precision'])

    def encode(self, input):
        """input = (x, domain name) -> ids (B,T) int64"""
        x, name = input
        return self.quantizer.encode(self.encoders[name](x))

    def decode(self, input):
        """input = (ids (B,T), y_idx, domain name) -> features of that domain"""
        z_idx, y_idx, name = input
        y = self.embeds(y_idx).transpose(1, 2).contiguous()
        return self.decoders[name]((self.quantizer.decode(z_idx), y))

    def forward(self, input):
        xs, y_idx = list(input[:-1]), input[-1]
        y = self.embeds(y_idx).transpose(1, 2)                       # (B, Y, 1)
        z = _cat_batch([self.encoders[n](x) for n, x in zip(self.names, xs)])
        z_vq, z_qut_loss, z_enc_loss, vq_detail = self.quantizer(z)
        y2 = torch.cat([y, y], dim=0)
        loss = z_qut_loss + self.beta * z_enc_loss
        outs, losses = [], {}
        for n, x in zip(self.names, xs):
            xhat = self.decoders[n]((z_vq, y2))                      # rows [0,B): from domain a's latent, [B,2B): from b's
            x_loss = log_loss(xhat, _cat_batch([x, x]) if x.is_cuda else torch.cat([x, x], 0))
            loss = loss + x_loss
            losses['X like ' + n] = host_value(x_loss.detach())
            outs.append(xhat)
        losses = dict({'Total': host_value(loss.detach()), 'VQ loss': host_value(z_enc_loss.detach())}, **losses)
        losses.update(vq_detail)
        return tuple(outs), loss, losses
