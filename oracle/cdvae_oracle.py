"""CPU restatement of the cross-domain composition (vae_npvc_b200/model/cdvae.py) from the oracle's parts.

TEST INFRASTRUCTURE.  The reference has no CDVAE model file: "parity unpinned" for the WIRING - what is
pinned is every part (encoder / quantiser / decoder / log-loss restatements of oracle/vqvae_oracle.py, themselves pinned
against the imported reference classes by oracle/make_golden.py).  Parameters are keyed like the product model's
state_dict: encoders.<name>.encode.*, decoders.<name>.layers.* / final_layer.*, quantizer.embeddings,
embeds._embedding.weight."""
import torch

from . import vqvae_oracle as O


def cdvae_forward(params, arch, xs, y_idx):
    names = list(arch['domains'].keys())
    y = O.speaker_embed(params, y_idx)
    z = torch.cat([O.encoder_forward(params, arch['domains'][n]['encoder'], x, prefix=f'encoders.{n}.encode')
                   for n, x in zip(names, xs)], dim=0)
    z_vq, l_q, l_e, perp, idx = O.vq_forward(params['quantizer.embeddings'], z, arch.get('embed_norm', True), 'frame_mean')
    y2 = torch.cat([y, y], dim=0)
    loss = l_q + arch.get('beta', 0.01) * l_e
    outs, losses = [], {}
    for n, x in zip(names, xs):
        xhat = O.decoder_forward(params, arch['domains'][n]['decoder'], z_vq, y2, prefix=f'decoders.{n}')
        x_loss = O.log_loss(xhat, torch.cat([x, x], dim=0))
        loss = loss + x_loss
        losses['X like ' + n] = float(x_loss.detach())
        outs.append(xhat)
    losses = dict({'Total': float(loss.detach()), 'VQ loss': float(l_e.detach())}, **losses)
    losses['entropy'] = float(perp)
    return tuple(outs), loss, losses, idx
