"""Functional CPU restatement of the vae_npvc VQ-VAE train / convert step.

TEST INFRASTRUCTURE (see oracle/__init__.py).  Pure torch-CPU; works in float32 (the
reference's arithmetic) and float64 (to rank two fp32 implementations).

Every function cites the reference lines it restates (paths relative to
/root/reference).  The restatement is deliberately *explicit* - weight-norm, ConvTranspose
as a flipped Conv, GroupNorm, the straight-through estimator and the EMA update are all
written out - so that it documents the math the CUDA kernels must reproduce rather than
re-using the reference's nn.Module classes.  Parameters are a flat ``dict[str, Tensor]``
keyed exactly like the reference ``Model.state_dict()`` (207 keys for the VCC2020 model),
so a reference checkpoint feeds straight in.

Pinned by oracle/make_golden.py against the imported reference (fwd, losses, all grads,
EMA buffers, Adam step) - see tests/golden/.
"""
import math

import numpy as np
import torch
import torch.nn.functional as F

LOG_2PI = math.log(2.0 * math.pi)


# ----------------------------------------------------------------------------------------
# parameter helpers
# ----------------------------------------------------------------------------------------
def wn_weight(g, v):
    """Old-style ``nn.utils.weight_norm(dim=0)``: w[i] = g[i] * v[i] / ||v[i]||_F.

    vqvae.py:203-208 / 329-334 apply it to every Conv1d / ConvTranspose1d.  For a
    ConvTranspose1d the weight is (in, out, k), so the norm is per *input* channel
    (SURVEY.md F8).
    """
    norm = v.pow(2).sum(dim=(1, 2), keepdim=True).sqrt()
    return g * (v / norm)


def layer_weight(params, prefix):
    """Effective conv weight of ``prefix``: weight-normed if weight_g/weight_v exist."""
    if prefix + '.weight_g' in params:
        return wn_weight(params[prefix + '.weight_g'], params[prefix + '.weight_v'])
    return params[prefix + '.weight']


def conv1d(x, w, b, dilation=1):
    """Stride-1 'same' Conv1d with zero padding (layers.py:151-153, vqvae.py:155-157)."""
    k = w.shape[2]
    pad = (k - 1) // 2 * dilation
    return F.conv1d(x, w, b, stride=1, padding=pad, dilation=dilation)


def conv_transpose1d_s1(x, wt, b, dilation=1):
    """Stride-1 ConvTranspose1d (weight (in,out,k), padding=(k-1)//2*dilation) restated as a
    Conv1d with W'[o,i,j] = W[i,o,k-1-j]  (layers.py:199-200, vqvae.py:256-265;
    SURVEY.md section 9 B)."""
    w = wt.transpose(0, 1).flip(2)
    return conv1d(x, w, b, dilation)


def group_norm(x, groups, gamma, beta, eps=1e-5):
    """nn.GroupNorm(groups, C): per-sample, per-group mean / biased variance over
    (C/groups, T); per-channel affine (layers.py:154, 201)."""
    B, C, T = x.shape
    xg = x.reshape(B, groups, -1)
    mean = xg.mean(dim=2, keepdim=True)
    var = (xg - mean).pow(2).mean(dim=2, keepdim=True)
    xh = ((xg - mean) / torch.sqrt(var + eps)).reshape(B, C, T)
    return xh * gamma.view(1, C, 1) + beta.view(1, C, 1)


def leaky_relu(x, slope=0.2):
    return torch.where(x >= 0, x, x * slope)


# ----------------------------------------------------------------------------------------
# blocks
# ----------------------------------------------------------------------------------------
def enc_block(params, prefix, c, n_layers=1, dilation=1):
    """Conv1d_Layernorm_LRelu_Residual.forward (layers.py:168-176):
    stack = n_layers x [LeakyReLU(0.2) -> Conv1d(k) -> GroupNorm(1, C)]; out = stack(c) + skip(c).
    Only the first conv of the stack is dilated (layers.py:151-161)."""
    h = c
    for j in range(n_layers):
        w = layer_weight(params, f'{prefix}.stack.{3 * j + 1}')
        b = params[f'{prefix}.stack.{3 * j + 1}.bias']
        h = conv1d(leaky_relu(h), w, b, dilation if j == 0 else 1)
        h = group_norm(h, 1, params[f'{prefix}.stack.{3 * j + 2}.weight'],
                       params[f'{prefix}.stack.{3 * j + 2}.bias'])
    ws = layer_weight(params, f'{prefix}.skip_layer')
    return h + conv1d(c, ws, params[f'{prefix}.skip_layer.bias'])


def dec_block(params, prefix, x, c, dilation=1):
    """DeConv1d_Layernorm_GLU_ResSkip.forward (layers.py:218-249), non-causal, dropout 0."""
    C = x.shape[1]
    a = conv_transpose1d_s1(x, layer_weight(params, f'{prefix}.conv_in'),
                            params[f'{prefix}.conv_in.bias'], dilation)
    if f'{prefix}.conv_cond.bias' in params:
        a = a + conv1d(c, layer_weight(params, f'{prefix}.conv_cond'),
                       params[f'{prefix}.conv_cond.bias'])
    h = group_norm(a, 2, params[f'{prefix}.norm_layer.weight'], params[f'{prefix}.norm_layer.bias'])
    g = torch.tanh(h[:, :C]) * torch.sigmoid(h[:, C:])
    r = conv1d(g, layer_weight(params, f'{prefix}.res_skip_layers'),
               params[f'{prefix}.res_skip_layers.bias'])
    return r[:, :C] + x, r[:, C:]


def encoder_forward(params, enc, x, prefix='encoder.encode'):
    """Encoder.forward (vqvae.py:122-192).  Stride-1 stages only (all single-level recipes)."""
    assert all(s == 1 for s in enc.get('downsample_scales', [1])), 'strided stages are out of scope'
    idx = 0
    h = x
    stacks = enc['stacks']
    for si, _ in enumerate(enc['in_channels']):
        h = conv1d(h, layer_weight(params, f'{prefix}.{idx}'), params[f'{prefix}.{idx}.bias'])
        idx += 1
        for j in range(stacks[si]):
            d = 2 ** j if enc.get('dilation', True) else 1
            h = enc_block(params, f'{prefix}.{idx}', h, enc.get('stack_layers', 2), d)
            idx += 1
        h = leaky_relu(h)
        idx += 1
    return conv1d(h, layer_weight(params, f'{prefix}.{idx}'), params[f'{prefix}.{idx}.bias'])


def decoder_forward(params, dec, z, y, prefix='decoder'):
    """Decoder.forward (vqvae.py:298-318): only c[:,:,:1] is used, repeated over time;
    skip outputs summed and scaled by sqrt(1/len(layers)) (the count includes the
    ConvTranspose entries, vqvae.py:316)."""
    assert all(s == 1 for s in dec.get('upsample_scales', [1])), 'strided stages are out of scope'
    x = z
    c = y[:, :, :1]
    idx = 0
    skip = 0.0
    stacks = dec['stacks']
    for si, _ in enumerate(dec['in_channels']):
        x = conv_transpose1d_s1(x, layer_weight(params, f'{prefix}.layers.{idx}'),
                                params[f'{prefix}.layers.{idx}.bias'])
        idx += 1
        for j in range(stacks[si]):
            d = 2 ** j if dec.get('dilation', True) else 1
            x, s = dec_block(params, f'{prefix}.layers.{idx}', x, c.expand(-1, -1, x.shape[2]), d)
            skip = skip + s
            idx += 1
    o = skip * math.sqrt(1.0 / idx)
    o = conv1d(torch.relu(o), layer_weight(params, f'{prefix}.final_layer.1'),
               params[f'{prefix}.final_layer.1.bias'])
    return conv1d(torch.relu(o), layer_weight(params, f'{prefix}.final_layer.3'),
                  params[f'{prefix}.final_layer.3.bias'])


# ----------------------------------------------------------------------------------------
# vector quantisers
# ----------------------------------------------------------------------------------------
def _flatten(z, time_last=True):
    """layers_vq.py:85-90: (B,D,T) -> (B*T, D) rows in (b, t) order."""
    if time_last:
        B, D, T = z.shape
        return z.transpose(1, 2).contiguous().view(-1, D), (B, T, D)
    B, T, D = z.shape
    return z.contiguous().view(-1, D), (B, T, D)


def vq_distances(zf, emb):
    """layers_vq.py:103-105: ||z||^2 + ||e||^2 - 2 z e^T, materialised (N,K)."""
    return (zf.pow(2).sum(dim=1, keepdim=True) + emb.pow(2).sum(dim=1)
            - 2 * torch.matmul(zf, emb.t()))


def _reduce(loss, reduction, B, T, D, time_last):
    """layers_vq.py:123-140."""
    if reduction == 'sum':
        return loss.sum()
    if reduction == 'mean':
        return loss.mean()
    if reduction == 'batch_mean':
        return loss.sum() / B
    if reduction == 'frame_mean':
        return loss.sum() / (B * T)
    if reduction == 'none':
        loss = loss.view(B, T, D)
        return loss.transpose(1, 2) if time_last else loss
    raise ValueError(reduction)


def embed_norm_(emb):
    """VectorQuantizer.embed_norm (layers_vq.py:28-33): in place, outside autograd."""
    with torch.no_grad():
        emb.mul_(1.0 / emb.norm(dim=1, keepdim=True))


def vq_forward(emb, z, normalize=True, reduction='frame_mean', time_last=True):
    """VectorQuantizer.forward (layers_vq.py:79-150).

    ``emb`` is the codebook Parameter (K,D); when ``normalize`` it is re-normalised IN PLACE
    first (layers_vq.py:96), also in eval mode.  Returns (z_vq, z_qut_loss, z_enc_loss,
    perplexity 0-d tensor, idx (N,) int64).
    """
    zf, (B, T, D) = _flatten(z, time_last)
    if normalize:
        z_n = zf / zf.norm(dim=1, keepdim=True)
        embed_norm_(emb)
        e = emb / emb.norm(dim=1, keepdim=True)
    else:
        z_n, e = zf, emb
    idx = torch.argmin(vq_distances(z_n, e), dim=1)
    q = e.index_select(0, idx)
    counts = torch.bincount(idx, minlength=e.shape[0]).to(zf.dtype)
    p = counts / idx.numel()
    perplexity = torch.exp(-torch.sum(p * torch.log(p + 1e-10)))
    l_q = (q - z_n.detach()).pow(2)
    l_e = (q.detach() - z_n).pow(2)
    if normalize:
        l_e = l_e + (z_n - zf).pow(2)
    l_q = _reduce(l_q, reduction, B, T, D, time_last)
    l_e = _reduce(l_e, reduction, B, T, D, time_last)
    out = z_n + (q - z_n).detach()
    out = out.view(B, T, D)
    if time_last:
        out = out.transpose(1, 2).contiguous()
    return out, l_q, l_e, perplexity, idx


def vq_encode(emb, z, normalize=True, time_last=True):
    """VectorQuantizer.encode (layers_vq.py:35-58) - no in-place codebook renorm here."""
    zf, (B, T, D) = _flatten(z, time_last)
    if normalize:
        zf = zf / zf.norm(dim=1, keepdim=True)
        emb = emb / emb.norm(dim=1, keepdim=True)
    return torch.argmin(vq_distances(zf, emb), dim=1).view(B, T)


def vq_decode(emb, z_id, normalize=True, time_last=True):
    """VectorQuantizer.decode (layers_vq.py:61-76)."""
    B, T = z_id.shape
    if normalize:
        emb = emb / emb.norm(dim=1, keepdim=True)
    q = emb.index_select(0, z_id.flatten()).view(B, T, -1)
    return q.transpose(1, 2).contiguous() if time_last else q


def _ema_tile(z, z_num):
    """EMAVectorQuantizer._tile (layers_vq.py:183-190)."""
    n, d = z.shape
    if n < z_num:
        reps = (z_num + n - 1) // n
        z = z.repeat(reps, 1)
        z = z + torch.randn_like(z) * (0.01 / np.sqrt(d))
    return z


def ema_vq_forward(buf, z, mu=0.9, threshold=1.0, reduction='frame_mean', training=True,
                   update=True, time_last=True):
    """EMAVectorQuantizer.forward + init_emb + update_emb (layers_vq.py:192-233, 268-323).

    ``buf`` is a dict with 'emb_init' (0-d bool), 'emb_sum' (K,D), 'emb_elem' (K,),
    'embeddings' (K,D); it is updated in place like the reference's buffers.  Random draws
    (randn_like in _tile, randperm) are made in the reference's order so an identical
    torch seed gives identical results.  Bug-compatible: for every reduction except 'none'
    the output carries NO gradient (layers_vq.py:310-315).
    Returns (z_vq, 0.0, z_enc_loss, detail dict of 0-d tensors, idx).
    """
    zf, (B, T, D) = _flatten(z, time_last)
    K = buf['embeddings'].shape[0]
    if (not bool(buf['emb_init'])) and training:
        buf['emb_init'] = ~buf['emb_init']
        _z = _ema_tile(zf, K)
        buf['embeddings'] = _z[torch.randperm(_z.shape[0])][:K].detach().clone()
        buf['emb_sum'] = buf['embeddings'].clone()
        buf['emb_elem'] = torch.ones(K, dtype=zf.dtype)
    with torch.no_grad():
        idx = torch.argmin(vq_distances(zf, buf['embeddings']), dim=1)
        q = buf['embeddings'].index_select(0, idx)
    detail = {}
    if update and training:
        with torch.no_grad():
            s = torch.zeros(K, D, dtype=zf.dtype).index_add_(0, idx, zf)      # one-hot @ z (:207-210)
            cnt = torch.bincount(idx, minlength=K).to(zf.dtype)
            _z = _ema_tile(zf, K)
            rand = _z[torch.randperm(_z.shape[0])][:K]
            old = buf['embeddings'].clone()
            buf['emb_sum'] = mu * buf['emb_sum'] + (1.0 - mu) * s
            buf['emb_elem'] = mu * buf['emb_elem'] + (1.0 - mu) * cnt
            alive = (buf['emb_elem'].view(K, 1) >= threshold).to(zf.dtype)
            buf['embeddings'] = alive * (buf['emb_sum'] / buf['emb_elem'].view(K, 1)) + (1 - alive) * rand
            p = cnt / cnt.sum()
            detail = {
                'entropy': torch.exp(-torch.sum(p * torch.log(p + 1e-8))),
                'used_curr': (cnt >= threshold).sum(),
                'usage': alive.sum(),
                'diff_emb': torch.norm(buf['embeddings'] - old) / np.sqrt(K * D),
            }
    l_e = _reduce((q.detach() - zf).pow(2), reduction, B, T, D, time_last)
    if reduction == 'none':
        q = zf + (q - zf).detach()
    out = q.view(B, T, D)
    if time_last:
        out = out.transpose(1, 2).contiguous()
    return out, 0.0, l_e, detail, idx


def jitter_sources(T, p, rng=np.random):
    """Jitter.forward's host-side draws (layers_vq.py:361-378) restated as a source-column
    table: src[t] == t where the column is kept.  Draw order matches the reference (one
    Bernoulli per step, plus one direction draw for replaced interior columns).
    Bug-compatible: the reference indexes [True, False] with a draw that is 1 with
    probability p, so a column is REPLACED with probability 1-p (layers_vq.py:364)."""
    src = np.arange(T)
    for i in range(T):
        replace = [True, False][rng.choice([1, 0], p=[p, 1 - p])]
        if replace:
            if i == 0:
                src[i] = i + 1
            elif i == T - 1:
                src[i] = i - 1
            else:
                src[i] = i + rng.choice([-1, 1], p=[0.5, 0.5])
    return src


def jitter_apply(q, src):
    """Replaced columns come from a detached clone, so they carry no gradient
    (layers_vq.py:357, 377)."""
    src_t = torch.as_tensor(src, dtype=torch.long)
    keep = (src_t == torch.arange(q.shape[2])).view(1, 1, -1)
    return torch.where(keep, q, q.detach()[:, :, src_t])


def log_loss(x, mu, reduction='frame_mean'):
    """layers.py:283-296 (Gaussian NLL with unit variance)."""
    B, D, T = x.shape
    loss = 0.5 * (LOG_2PI + (x - mu).pow(2))
    if reduction == 'sum':
        return loss.sum()
    if reduction == 'mean':
        return loss.mean()
    if reduction == 'batch_mean':
        return loss.sum() / B
    if reduction == 'frame_mean':
        return loss.sum() / (B * T)
    return loss


# ----------------------------------------------------------------------------------------
# model + trainer
# ----------------------------------------------------------------------------------------
EMA_BUFFERS = ('emb_init', 'emb_sum', 'emb_elem', 'embeddings')


def speaker_embed(params, y_idx):
    """Conditions.forward with normalize=False (layers.py:42-54) + vqvae.py:73 transpose:
    (B,n) int64 -> (B, Y, n)."""
    return params['embeds._embedding.weight'][y_idx].transpose(1, 2).contiguous()


def model_forward(params, arch, x, y_idx, training=True, jitter_src=None):
    """Model.forward (vqvae.py:70-90).  ``params`` holds parameters and, for use_ema, the
    four quantiser buffers under 'quantizer.<name>' (updated in place).
    Returns (xhat, loss, detail dict of python floats, idx)."""
    y = speaker_embed(params, y_idx)
    z = encoder_forward(params, arch['encoder'], x)
    if arch.get('use_ema', False):
        buf = {k: params['quantizer.' + k] for k in EMA_BUFFERS}
        z_vq, l_q, l_e, det, idx = ema_vq_forward(buf, z, arch.get('mu', 0.9), 1.0, 'frame_mean', training)
        for k in EMA_BUFFERS:
            params['quantizer.' + k] = buf[k]
        det = {k: float(v) for k, v in det.items()}
    else:
        z_vq, l_q, l_e, perp, idx = vq_forward(params['quantizer.embeddings'], z,
                                               arch.get('embed_norm', True), 'frame_mean')
        det = {'entropy': float(perp)}
    p = arch.get('jitter_p', 0.0)
    if p != 0.0 and training:
        if jitter_src is None:
            jitter_src = jitter_sources(z_vq.shape[2], p)
        z_vq = jitter_apply(z_vq, jitter_src)
    xhat = decoder_forward(params, arch['decoder'], z_vq, y)
    x_loss = log_loss(xhat, x)
    loss = x_loss + l_q + arch.get('beta', 0.01) * l_e
    losses = {'Total': float(loss.detach()), 'VQ loss': float(l_e.detach()), 'X like': float(x_loss.detach())}
    losses.update(det)
    return xhat, loss, losses, idx


def model_encode(params, arch, x):
    """Model.encode (vqvae.py:45-52)."""
    z = encoder_forward(params, arch['encoder'], x)
    if arch.get('use_ema', False):
        zf, (B, T, _) = _flatten(z)
        return torch.argmin(vq_distances(zf, params['quantizer.embeddings']), dim=1).view(B, T)
    return vq_encode(params['quantizer.embeddings'], z, arch.get('embed_norm', True))


def model_decode(params, arch, z_idx, y_idx):
    """Model.decode (vqvae.py:55-60)."""
    y = speaker_embed(params, y_idx)
    norm = (not arch.get('use_ema', False)) and arch.get('embed_norm', True)
    z_vq = vq_decode(params['quantizer.embeddings'], z_idx, norm)
    return decoder_forward(params, arch['decoder'], z_vq, y)


def model_infer(params, arch, x, y_idx):
    """Model.infer (vqvae.py:63-67)."""
    return model_decode(params, arch, model_encode(params, arch, x), y_idx)


def trainable_names(params):
    return [k for k, v in params.items() if v.is_floating_point() and v.requires_grad]


class AdamState:
    """torch.optim.Adam(lr, betas=(0.5,0.999), eps=1e-8, wd=0) + StepLR restated
    (trainer/basic.py:38-46)."""

    def __init__(self, names, lr=1e-3, betas=(0.5, 0.999), eps=1e-8, step_size=100000, gamma=0.5):
        self.names = list(names)
        self.lr0, self.betas, self.eps = lr, betas, eps
        self.step_size, self.gamma = step_size, gamma
        self.t = 0
        self.m, self.v = {}, {}

    def lr(self):
        return self.lr0 * self.gamma ** (self.t // self.step_size) if self.step_size else self.lr0


def train_step(params, arch, opt, x, y_idx, max_grad_norm=10.0, jitter_src=None):
    """Trainer.train_step (trainer/basic.py:55-79): zero_grad -> forward -> backward ->
    clip_grad_norm_(max_grad_norm) -> Adam -> StepLR.  Returns (losses, grads dict after
    clipping, total grad norm before clipping)."""
    for n in opt.names:
        params[n].grad = None
    xhat, loss, losses, idx = model_forward(params, arch, x, y_idx, True, jitter_src)
    loss.backward()
    grads = {n: params[n].grad for n in opt.names if params[n].grad is not None}
    total = torch.sqrt(sum(g.pow(2).sum() for g in grads.values()))
    coef = 1.0
    if max_grad_norm > 0:
        coef = min(float(max_grad_norm / (total + 1e-6)), 1.0)       # clip_grad_norm_ semantics
    lr = opt.lr()
    opt.t += 1
    b1, b2 = opt.betas
    with torch.no_grad():
        for n, g in grads.items():
            g = g * coef
            grads[n] = g
            if n not in opt.m:
                opt.m[n] = torch.zeros_like(g)
                opt.v[n] = torch.zeros_like(g)
            opt.m[n].mul_(b1).add_(g, alpha=1 - b1)
            opt.v[n].mul_(b2).addcmul_(g, g, value=1 - b2)
            bc1 = 1 - b1 ** opt.t
            bc2 = 1 - b2 ** opt.t
            denom = (opt.v[n].sqrt() / math.sqrt(bc2)).add_(opt.eps)
            params[n].addcdiv_(opt.m[n], denom, value=-lr / bc1)
    return losses, grads, float(total)


def params_from_state_dict(sd, dtype=torch.float32, requires_grad=True):
    """Clone a reference state_dict into oracle params (floating leaves require grad;
    EMA buffers and the bool flag never do)."""
    out = {}
    for k, v in sd.items():
        v = v.detach().clone()
        is_buf = k.startswith('quantizer.emb_') or (k == 'quantizer.embeddings' and 'quantizer.emb_sum' in sd)
        if v.is_floating_point():
            v = v.to(dtype)
            if requires_grad and not is_buf:
                v.requires_grad_(True)
        out[k] = v
    return out
