"""Synthetic workload generator: recipe-shaped configs, seeded weights and mel-like feature batches.

BASELINE.json's north_star puts "a synthetic mel/SP feature generator of the recipe's shape" in the dataset's
place; this is it (with ``dataset.utt2mel_spk.SyntheticDataset`` as the item-level form).  It lives in the product
because ``bench.py`` and the conversion benchmark feed the CUDA path from it; the oracle and the tests use the SAME
functions (``oracle/synth.py`` re-exports them), so both sides of every parity check see identical inputs.

The reference's own initialisation cannot be re-derived without the reference (parity must copy a
state_dict), and a 31 M-parameter state_dict is too big to commit.  So goldens are made from weights produced HERE
by a seeded torch CPU generator - the same function feeds the imported reference (oracle/make_golden.py, build
container only), the oracle and the CUDA path (any box).  Shapes and key names follow the reference
``Model.state_dict()`` (vqvae.py:122-183, 220-296; layers.py:130-165, 182-215).
"""
import math

import torch


def arch_config(X=80, C=512, D=64, K=512, Y=128, S=128, y_num=117, stacks=10, use_ema=False,
                embed_norm=True, jitter_p=0.0, beta=0.01, mu=0.9, stack_layers=1, dilation=False):
    """A config dict with the keys of egs/vcc20/vae1/conf/train_pytorch_vqvae.yaml:28-61.
    Defaults = BASELINE.json configs[0]/[1] ("VCC2020 shape", codebook 512x64)."""
    return {
        'model_type': 'vae_npvc.model.vqvae',
        'y_dim': Y, 'y_num': y_num, 'z_dim': D, 'z_num': K,
        'use_ema': use_ema, 'embed_norm': embed_norm, 'beta': beta, 'mu': mu, 'jitter_p': jitter_p,
        'encoder': {
            'in_channels': [X], 'out_channels': [C], 'kernel_size': 3, 'downsample_scales': [1],
            'z_channels': D, 'dilation': dilation, 'stack_kernel_size': 3,
            'stack_layers': stack_layers, 'stacks': [stacks], 'use_weight_norm': True,
            'use_causal_conv': False,
        },
        'decoder': {
            'in_channels': [D], 'out_channels': [C], 'cond_channels': Y, 'skip_channels': S,
            'final_channels': X, 'kernel_size': 3, 'upsample_scales': [1], 'dilation': dilation,
            'stack_kernel_size': 3, 'stacks': [stacks], 'use_weight_norm': True,
            'use_causal_conv': False,
        },
    }


def small_config(**kw):
    """A tiny model whose full state and gradients fit in a committed fixture."""
    base = dict(X=8, C=16, D=8, K=32, Y=8, S=12, y_num=5, stacks=2)
    base.update(kw)
    return arch_config(**base)


def _conv(sd, gen, prefix, d0, d1, k, bias_n, fan_in):
    """weight_v (d0,d1,k) ~ U(+-1/sqrt(fan_in)); weight_g = ||v|| * U(0.8,1.2) so the
    normalisation is not the identity; bias ~ U(+-1/sqrt(fan_in))."""
    bound = 1.0 / math.sqrt(fan_in)
    v = (torch.rand(d0, d1, k, generator=gen) * 2 - 1) * bound
    g = v.pow(2).sum(dim=(1, 2), keepdim=True).sqrt() * (0.8 + 0.4 * torch.rand(d0, 1, 1, generator=gen))
    sd[prefix + '.bias'] = (torch.rand(bias_n, generator=gen) * 2 - 1) * bound
    sd[prefix + '.weight_g'] = g
    sd[prefix + '.weight_v'] = v


def _gn(sd, gen, prefix, c):
    sd[prefix + '.weight'] = 1.0 + 0.1 * torch.randn(c, generator=gen)
    sd[prefix + '.bias'] = 0.1 * torch.randn(c, generator=gen)


def synthetic_state_dict(arch, seed=777):
    """Reference-keyed state_dict (insertion order == reference ``state_dict()`` order)."""
    gen = torch.Generator().manual_seed(seed)
    sd = {}
    enc, dec = arch['encoder'], arch['decoder']
    ks = enc.get('stack_kernel_size', 3)
    idx = 0
    for si, cin in enumerate(enc['in_channels']):
        cout = enc['out_channels'][si]
        k = enc.get('kernel_size', 3)
        _conv(sd, gen, f'encoder.encode.{idx}', cout, cin, k, cout, cin * k)
        idx += 1
        for _ in range(enc['stacks'][si]):
            p = f'encoder.encode.{idx}'
            for j in range(enc.get('stack_layers', 2)):
                _conv(sd, gen, f'{p}.stack.{3 * j + 1}', cout, cout, ks, cout, cout * ks)
                _gn(sd, gen, f'{p}.stack.{3 * j + 2}', cout)
            _conv(sd, gen, f'{p}.skip_layer', cout, cout, 1, cout, cout)
            idx += 1
        idx += 1                                            # the LeakyReLU entry
    _conv(sd, gen, f'encoder.encode.{idx}', enc['z_channels'], enc['out_channels'][-1], 1,
          enc['z_channels'], enc['out_channels'][-1])
    idx = 0
    kd = dec.get('stack_kernel_size', 3)
    Y, S = dec['cond_channels'], dec['skip_channels']
    for si, cin in enumerate(dec['in_channels']):
        cout = dec['out_channels'][si]
        k = dec.get('kernel_size', 5)
        _conv(sd, gen, f'decoder.layers.{idx}', cin, cout, k, cout, cin * k)       # ConvT: (in,out,k)
        idx += 1
        for _ in range(dec['stacks'][si]):
            p = f'decoder.layers.{idx}'
            _conv(sd, gen, f'{p}.conv_in', cout, 2 * cout, kd, 2 * cout, cout * kd)   # ConvT (in,2in,k)
            _gn(sd, gen, f'{p}.norm_layer', 2 * cout)
            _conv(sd, gen, f'{p}.conv_cond', 2 * cout, Y, 1, 2 * cout, Y)
            _conv(sd, gen, f'{p}.res_skip_layers', cout + S, cout, 1, cout + S, cout)
            idx += 1
    _conv(sd, gen, 'decoder.final_layer.1', S, S, 1, S, S)
    _conv(sd, gen, 'decoder.final_layer.3', dec['final_channels'], S, 1, dec['final_channels'], S)
    K, D = arch['z_num'], arch['z_dim']
    if arch.get('use_ema', False):
        sd['quantizer.emb_init'] = torch.tensor(0).bool()
        sd['quantizer.emb_sum'] = torch.zeros(K, D)
        sd['quantizer.emb_elem'] = torch.ones(K)
        sd['quantizer.embeddings'] = torch.zeros(K, D)
    else:
        e = torch.randn(K, D, generator=gen)
        if arch.get('embed_norm', True):
            e = e / e.norm(dim=1, keepdim=True)              # ctor calls embed_norm() (layers_vq.py:20)
        sd['quantizer.embeddings'] = e
    sd['embeds._embedding.weight'] = torch.randn(arch['y_num'], arch['y_dim'], generator=gen)
    return sd


def synthetic_batch(arch, B, T, seed=777):
    """x ~ N(0,1) (B,X,T) float32 (CMVN-normalised features, egs/vcc20/vae1/run.sh:106-120);
    y ~ U{0..y_num-1} (B,1) int64 - the contract of dataset/utt2mel_spk.py:42-74 after
    default collation."""
    gen = torch.Generator().manual_seed(seed)
    x = torch.randn(B, arch['encoder']['in_channels'][0], T, generator=gen)
    y = torch.randint(0, arch['y_num'], (B, 1), generator=gen)
    return x, y


def cdvae_config(Xa=513, Xb=40, C=512, D=128, K=512, Y=128, S=128, y_num=117, stacks=10, embed_norm=True, beta=0.01):
    """BASELINE.json configs[2]: 513-dim WORLD spectrum + 40-dim MCC encoders / decoders over one shared latent
    (vae_npvc_b200/model/cdvae.py).  Each domain is a single-stage Encoder / Decoder of the VCC2020 recipe's shape."""
    doms = {}
    for name, X in (('sp', Xa), ('mcc', Xb)):
        a = arch_config(X=X, C=C, D=D, K=K, Y=Y, S=S, y_num=y_num, stacks=stacks)
        doms[name] = {'encoder': a['encoder'], 'decoder': a['decoder']}
    return {'model_type': 'vae_npvc_b200.model.cdvae', 'domains': doms, 'y_dim': Y, 'y_num': y_num, 'z_dim': D, 'z_num': K,
            'embed_norm': embed_norm, 'beta': beta}


def cdvae_state_dict(arch, seed=777):
    """Seeded weights keyed like cdvae.Model.state_dict()."""
    sd = {}
    for i, (name, dom) in enumerate(arch['domains'].items()):
        single = dict(arch, encoder=dom['encoder'], decoder=dom['decoder'], use_ema=False)
        part = synthetic_state_dict(single, seed + 31 * i)
        for k, v in part.items():
            if k.startswith('encoder.'):
                sd['encoders.%s.%s' % (name, k[len('encoder.'):])] = v
            elif k.startswith('decoder.'):
                sd['decoders.%s.%s' % (name, k[len('decoder.'):])] = v
            elif i == 0:
                sd[k] = v                                   # quantizer.embeddings, embeds._embedding.weight
    # nn.Module order: encoders.*, decoders.*, quantizer, embeds
    order = ([k for k in sd if k.startswith('encoders.')] + [k for k in sd if k.startswith('decoders.')]
             + [k for k in sd if k.startswith('quantizer.')] + [k for k in sd if k.startswith('embeds.')])
    return {k: sd[k] for k in order}


def cdvae_batch(arch, B, T, seed=777):
    gen = torch.Generator().manual_seed(seed)
    xs = [torch.randn(B, dom['encoder']['in_channels'][0], T, generator=gen) for dom in arch['domains'].values()]
    y = torch.randint(0, arch['y_num'], (B, 1), generator=gen)
    return xs, y


def conv_macs_per_frame(enc, dec, T):
    """Algorithmic conv multiply-accumulates per frame of one Encoder and one Decoder pass (the
    time-invariant conditioning conv counted once per utterance, i.e. / T per frame)."""
    e = 0
    for cin, cout, st in zip(enc['in_channels'], enc['out_channels'], enc['stacks']):
        e += cin * cout * enc.get('kernel_size', 3)
        e += st * enc.get('stack_layers', 2) * cout * cout * enc.get('stack_kernel_size', 3) + st * cout * cout
    e += enc['out_channels'][-1] * enc['z_channels']
    d = 0
    Y, S = dec['cond_channels'], dec['skip_channels']
    for cin, cout, st in zip(dec['in_channels'], dec['out_channels'], dec['stacks']):
        d += cin * cout * dec.get('kernel_size', 5)
        d += st * (cout * 2 * cout * dec.get('stack_kernel_size', 3) + cout * (cout + S)) + st * Y * 2 * cout / T
    d += S * S + S * dec['final_channels']
    return e, d


# ---- hierarchical VQ-VAE2 + style tokens (egs/vcc20/vae2/conf/train_pytorch_vqvae2.yaml:28-127) ---------------------
def vqvae2_config(X=80, C=512, D=128, K=512, Y=128, S=128, y_num=117, stacks=(6, 3, 3), dec_stacks=(10, 6, 6),
                  tokens=10, heads=4, scales=((1,), (2, 2), (4, 4))):
    """The shipped three-level recipe (defaults) or a scaled-down model of the same topology."""
    def enc(cin, n_stage, ds, st):
        return {'in_channels': [cin] + [C] * (n_stage - 1), 'out_channels': [C] * n_stage, 'kernel_size': 3,
                'downsample_scales': list(ds), 'z_channels': D, 'dilation': False, 'stack_kernel_size': 3,
                'stack_layers': 1, 'stacks': [st] * n_stage, 'use_weight_norm': True, 'use_causal_conv': False}

    def dec(cin, cond, final, st):
        return {'in_channels': [cin], 'out_channels': [C], 'cond_channels': cond, 'skip_channels': S,
                'final_channels': final, 'kernel_size': 3, 'upsample_scales': [1], 'dilation': False,
                'stack_kernel_size': 3, 'stacks': [st], 'use_weight_norm': True, 'use_causal_conv': False}
    return {
        'model_type': 'vae_npvc.model.vqvae2', 'levels': 3, 'y_dim': Y, 'y_num': y_num, 'beta': 0.01, 'use_gst': True,
        'use_ema': False, 'jitter_p': 0.0,
        'encoder.0': enc(X, len(scales[0]), scales[0], stacks[0]),
        'encoder.1': enc(C, len(scales[1]), scales[1], stacks[1]),
        'encoder.2': enc(C, len(scales[2]), scales[2], stacks[2]),
        'quantizer.0': {'z_dim': D, 'z_num': K, 'normalize': True},
        'quantizer.1': {'z_dim': D, 'z_num': K, 'normalize': True},
        'quantizer.2': {'ref_embed_dim': D, 'gst_tokens': tokens, 'gst_token_dim': D, 'gst_heads': heads},
        'decoder.0': dec(3 * D, Y, X, dec_stacks[0]),
        'decoder.1': dec(D, 2 * D, D, dec_stacks[1]),
        'decoder.2': dec(D, D, D, dec_stacks[2]),
    }


def vqvae2_small_config():
    return vqvae2_config(X=8, C=16, D=8, K=32, Y=8, S=12, y_num=5, stacks=(2, 1, 1), dec_stacks=(2, 2, 1), tokens=5, heads=2)


def vqvae2_state_dict(arch, seed=777):
    """Seeded weights keyed like ``vqvae2.Model.state_dict()`` (the key list and shapes come from the mirror model,
    whose sub-module names are the reference's)."""
    from .model.vqvae2 import Model
    shapes = {k: tuple(v.shape) for k, v in Model(arch).state_dict().items()}
    gen = torch.Generator().manual_seed(seed)
    sd = {}
    for k, shp in shapes.items():
        if k.endswith('.weight_g'):
            continue                                            # written with its weight_v
        if k.endswith('.weight_v'):
            pre = k[:-len('.weight_v')]
            transposed = '.conv_in' in k or (k.startswith('decoders.') and '.layers.' in k and k.count('.') == 4)
            fan_in = (shp[0] if transposed else shp[1]) * shp[2]
            bound = 1.0 / math.sqrt(fan_in)
            v = (torch.rand(*shp, generator=gen) * 2 - 1) * bound
            sd[pre + '.weight_g'] = v.pow(2).sum(dim=(1, 2), keepdim=True).sqrt() * (0.8 + 0.4 * torch.rand(shp[0], 1, 1, generator=gen))
            sd[k] = v
        elif k.endswith('.embeddings'):
            e = torch.randn(*shp, generator=gen)
            sd[k] = e / e.norm(dim=1, keepdim=True)
        elif k.endswith('gst_embs') or k.endswith('_embedding.weight'):
            sd[k] = torch.randn(*shp, generator=gen)
        elif '.mha.' in k:
            fan_in = shapes[k.rsplit('.', 1)[0] + '.weight'][1]
            sd[k] = (torch.rand(*shp, generator=gen) * 2 - 1) / math.sqrt(fan_in)
        elif k.endswith('.weight'):                             # GroupNorm scale
            sd[k] = 1.0 + 0.1 * torch.randn(*shp, generator=gen)
        elif k.endswith('.bias'):
            sd[k] = 0.1 * (torch.rand(*shp, generator=gen) * 2 - 1)
        else:
            raise KeyError('vqvae2_state_dict: unexpected key %s' % k)
    return {k: sd[k] for k in shapes}


def vqvae2_batch(arch, B, T, seed=777):
    gen = torch.Generator().manual_seed(seed)
    x = torch.randn(B, arch['encoder.0']['in_channels'][0], T, generator=gen)
    y = torch.randint(0, arch['y_num'], (B, 1), generator=gen)
    return x, y
