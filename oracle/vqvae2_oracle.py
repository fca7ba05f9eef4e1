"""Functional CPU restatement of the hierarchical VQ-VAE2 + global-style-token model.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Pure torch-CPU, float32 or float64.  Re-uses the single-level parts of
oracle/vqvae_oracle.py (weight-norm, encoder / gated decoder blocks, quantiser, log-loss) and restates what is new in
/root/reference/vae_npvc/model/vqvae2.py and layers_gst.py: strided down-sampling convs, nearest-neighbour upsample,
time-varying conditioning, the 10-token multi-head style attention and the three-level encode / quantise / decode
order.  Parameters are a flat dict keyed exactly like the reference ``vqvae2.Model.state_dict()``.

Pinned by oracle/make_golden_vqvae2.py against the imported reference (forward, losses, every gradient) - see
tests/golden/vqvae2_small.npz.
"""
import math

import torch
import torch.nn.functional as F

from . import vqvae_oracle as O


def strided_conv1d(x, w, b, s):
    """Encoder stage entry with ds_scale != 1 (vqvae2.py:204-212): kernel 2s, stride s, padding s//2 + s%2."""
    return F.conv1d(x, w, b, stride=s, padding=s // 2 + s % 2)


def upsample(z, target_len):
    """Model.upsample (vqvae2.py:130-143): every frame repeated target_len // z_len times, then cut to target_len or
    padded with copies of the last frame."""
    z_len = z.shape[-1]
    rep = target_len // z_len
    idx = torch.clamp(torch.arange(target_len) // max(rep, 1), max=z_len - 1)
    assert rep >= 1, 'upsample to a shorter length is undefined in the reference (empty repeat)'
    return z[..., idx]


def encoder_forward(params, enc, x, prefix):
    """vqvae2.Encoder.forward (vqvae2.py:186-258): -> (z_proj(h), h) with h AFTER the last LeakyReLU."""
    idx = 0
    h = x
    for si, s in enumerate(enc.get('downsample_scales', [1] * len(enc['in_channels']))):
        w = O.layer_weight(params, f'{prefix}.encode.{idx}')
        b = params[f'{prefix}.encode.{idx}.bias']
        h = O.conv1d(h, w, b) if s == 1 else strided_conv1d(h, w, b, s)
        idx += 1
        for j in range(enc['stacks'][si]):
            d = 2 ** j if enc.get('dilation', True) else 1
            h = O.enc_block(params, f'{prefix}.encode.{idx}', h, enc.get('stack_layers', 2), d)
            idx += 1
        h = O.leaky_relu(h)
        idx += 1
    z = O.conv1d(h, O.layer_weight(params, f'{prefix}.z_proj'), params[f'{prefix}.z_proj.bias'])
    return z, h


def decoder_forward(params, dec, x, c, prefix):
    """vqvae2.Decoder.forward (vqvae2.py:349-366): the conditioning c (B, cond, T) is used at every frame (no
    c[:,:,:1] as in vqvae.py:309); skip sum scaled by sqrt(1 / len(layers))."""
    assert all(s == 1 for s in dec.get('upsample_scales', [1])), 'the shipped vqvae2 recipe only uses upsample_scales 1'
    idx = 0
    skip = 0.0
    for si, _ in enumerate(dec['in_channels']):
        x = O.conv_transpose1d_s1(x, O.layer_weight(params, f'{prefix}.layers.{idx}'), params[f'{prefix}.layers.{idx}.bias'])
        idx += 1
        for j in range(dec['stacks'][si]):
            d = 2 ** j if dec.get('dilation', True) else 1
            x, s = O.dec_block(params, f'{prefix}.layers.{idx}', x, c, d)
            skip = skip + s
            idx += 1
    o = skip * math.sqrt(1.0 / idx)
    o = O.conv1d(torch.relu(o), O.layer_weight(params, f'{prefix}.final_layer.1'), params[f'{prefix}.final_layer.1.bias'])
    return O.conv1d(torch.relu(o), O.layer_weight(params, f'{prefix}.final_layer.3'), params[f'{prefix}.final_layer.3.bias'])


def style_tokens(params, prefix, ref, heads):
    """StyleTokenLayer.forward + MultiHeadedAttention (layers_gst.py:46-60, 84-147), one query per utterance, no mask,
    dropout 0: ref (B, ref_dim) -> (B, n_feat)."""
    tok = torch.tanh(params[f'{prefix}.gst_embs'])                                   # (N, token_dim / heads)
    q = ref @ params[f'{prefix}.mha.linear_q.weight'].t() + params[f'{prefix}.mha.linear_q.bias']      # (B, F)
    k = tok @ params[f'{prefix}.mha.linear_k.weight'].t() + params[f'{prefix}.mha.linear_k.bias']      # (N, F)
    v = tok @ params[f'{prefix}.mha.linear_v.weight'].t() + params[f'{prefix}.mha.linear_v.bias']
    B, Fdim = q.shape
    N = k.shape[0]
    dk = Fdim // heads
    qh = q.view(B, heads, dk)
    kh = k.view(N, heads, dk)
    vh = v.view(N, heads, dk)
    scores = torch.einsum('bhd,nhd->bhn', qh, kh) / math.sqrt(dk)
    attn = torch.softmax(scores, dim=-1)
    ctx = torch.einsum('bhn,nhd->bhd', attn, vh).reshape(B, Fdim)
    return ctx @ params[f'{prefix}.mha.linear_out.weight'].t() + params[f'{prefix}.mha.linear_out.bias']


def model_forward(params, arch, x, y_idx):
    """vqvae2.Model.forward (vqvae2.py:74-127), jitter_p 0 (the shipped recipe), non-EMA quantisers.
    -> (xhat, loss, detail, [ids per quantised level, in the reference's order: top-down])"""
    levels = arch.get('levels', 3)
    use_gst = arch.get('use_gst', True)
    assert not arch.get('use_ema', True), 'the oracle restates the shipped recipe: non-EMA quantisers (use_ema: false)'
    y = params['embeds._embedding.weight'][y_idx[..., :1]].transpose(1, 2)          # (B, Y, 1)
    z_levels, time_levels = [], [x.shape[-1]]
    x_ = x
    for i in range(levels):
        z_, x_ = encoder_forward(params, arch[f'encoder.{i}'], x_, f'encoders.{i}')
        z_levels.append(z_)
        time_levels.append(z_.shape[-1])
    z_vq_levels, l_q, l_e, det, ids = [], [], [], [], []
    z_ = z_levels.pop()
    for i in reversed(range(levels)):
        if use_gst and i == levels - 1:
            q = arch[f'quantizer.{i}']
            z_vq = style_tokens(params, f'quantizers.{i}', z_.mean(dim=-1), q.get('gst_heads', 4)).unsqueeze(-1)
        else:
            q = arch[f'quantizer.{i}']
            z_vq, lq, le, perp, idx = O.vq_forward(params[f'quantizers.{i}.embeddings'], z_, q.get('normalize', False),
                                                   'frame_mean')
            l_q.append(lq)
            l_e.append(le)
            det.append({'entropy': float(perp), 'quanti_err': float(le.detach())})
            ids.append(idx)
        z_vq_levels.append([upsample(z_vq, t) for t in time_levels[:i + 1]])
        if i > 0:
            z_ = z_levels.pop()
            z_cat = torch.cat([zs[i] for zs in z_vq_levels], dim=1)
            z_ = decoder_forward(params, arch[f'decoder.{i}'], z_, z_cat, f'decoders.{i}')
    z_cat = torch.cat([zs[0] for zs in z_vq_levels], dim=1)
    xhat = decoder_forward(params, arch['decoder.0'], z_cat, upsample(y, time_levels[0]), 'decoders.0')
    z_qut_loss, z_enc_loss = sum(l_q), sum(l_e)
    x_loss = O.log_loss(xhat, x)
    loss = x_loss + z_qut_loss + arch.get('beta', 0.01) * z_enc_loss
    losses = {'Total': float(loss.detach()), 'VQ loss': float(z_enc_loss.detach()), 'X like': float(x_loss.detach())}
    for i, d in enumerate(det):
        losses.update({k + f'.{i}': v for k, v in d.items()})
    return xhat, loss, losses, ids
