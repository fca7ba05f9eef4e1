"""torch.autograd.Functions that compose the CUDA primitives into the reference's op-groups.

One Function per op-group of SURVEY.md section 8(a): stride-1 conv (a2/a8 ends), encoder residual
block (a3), gated decoder block (a9), vector quantisers (a4/a6), jitter (a7), log-loss (a10),
speaker embedding (a1).  Forward and backward are both hand-sequenced primitive calls
(prims.py -> C ABI); autograd is only the glue that lets the reference-shaped nn.Modules
compose and lets ``loss.backward()`` work as in vae_npvc/trainer/basic.py:62.

Operand planes written by one op's epilogue are handed to the next op through a side channel
on the output tensor (``attach_planes`` / ``cached_planes``), so a chain of layers never
re-reads fp32 activations just to re-quantise them.
"""
import os

import torch
from torch.autograd import Function

from . import prims as Pm
from .prims import ACT_LRELU, ACT_NONE, ACT_RELU

_ACT_KIND = {ACT_NONE: 'id', ACT_LRELU: 'lrelu', ACT_RELU: 'relu'}
_PRECISION = {'P': 3 if os.environ.get('VNB_PRECISION', 'bf16').lower() in ('fp32', 'f32', '3') else 1}


def set_precision(mode):
    """'bf16' -> single bf16 operand plane (fast);  'fp32' -> hi/mid/lo split planes
    (fp32-faithful contraction on the same kernels)."""
    _PRECISION['P'] = {'bf16': 1, 'fp32': 3, 1: 1, 3: 3}[mode]


def get_precision():
    return _PRECISION['P']


def attach_planes(t, kind, planes):
    if planes is None:
        return
    d = getattr(t, '_vnb_planes', None)
    if d is None:
        d = {}
        t._vnb_planes = d
    d[kind] = (planes, t._version)


def cached_planes(t, kind, P):
    d = getattr(t, '_vnb_planes', None)
    if d is not None and kind in d:
        pl, ver = d[kind]
        if ver == t._version and pl.P == P and (pl.B, pl.C, pl.T) == tuple(t.shape):
            return pl
    return None


def _input_planes(x, P, act=ACT_NONE, scale=1.0):
    """Operand planes of act(x*scale): from the side channel when the producer left them."""
    if scale == 1.0:
        pl = cached_planes(x, _ACT_KIND[act], P)
        if pl is not None:
            return pl
    if act == ACT_NONE:
        return Pm.make_planes(x, P, scale, ACT_NONE, True, False)[0]
    return Pm.make_planes(x, P, scale, act, False, True)[1]


def _dw_to_param_grads(dw, v, g):
    """weight gradient (parameter layout) -> (dv, dg) through weight-norm, or (dw, None)."""
    if g is None:
        return dw, None
    dv, dg = Pm.wn_backward(dw, v, g)
    return dv, dg.view_as(g)


class ConvFn(Function):
    """y = Conv(act(x*scale)) for nn.Conv1d / stride-1 nn.ConvTranspose1d with optional
    weight-norm.  Reference call sites: vqvae.py:155-157,174 (encoder ends), :256-265 (decoder
    ConvTranspose), :281-286 (final ReLU-conv-ReLU-conv)."""

    @staticmethod
    def forward(ctx, x, v, g, bias, transposed, dil, act, scale, P, box):
        x = x.contiguous()
        xpl = _input_planes(x, P, act, scale)
        need_dx = ctx.needs_input_grad[0]
        W = Pm.wn_prepare(v, g, transposed, P, need_wb=need_dx)
        out = Pm.conv_fwd(xpl, W.wf, W.taps, dil, W.Cout, bias, 0)
        ctx.save_for_backward(x if (act != ACT_NONE or scale != 1.0) else None, v, g)
        ctx.xpl, ctx.wb = xpl, W.wb
        ctx.cfg = (transposed, dil, act, scale, P, W.taps, W.Cin, bias is not None)
        return out

    @staticmethod
    def backward(ctx, dy):
        x, v, g = ctx.saved_tensors
        transposed, dil, act, scale, P, taps, Cin, has_bias = ctx.cfg
        dy = dy.contiguous()
        dypl = Pm.make_planes(dy, P)[0]
        db = Pm.channel_sum(dy) if (has_bias and ctx.needs_input_grad[3]) else None
        dx = None
        if ctx.needs_input_grad[0]:
            dx = Pm.conv_fwd(dypl, ctx.wb, taps, dil, Cin)
            if x is not None:
                dx = Pm.act_bwd_add(dx, x, None, act, scale)
        dv = dg = None
        if ctx.needs_input_grad[1]:
            dw = Pm.conv_wgrad(dypl, ctx.xpl, taps, dil, transposed)
            dv, dg = _dw_to_param_grads(dw, v, g)
        return dx, dv, dg, db, None, None, None, None, None, None


def conv1d(x, v, g, bias, transposed=False, dil=1, act=ACT_NONE, scale=1.0, P=None):
    return ConvFn.apply(x, v, g, bias, transposed, dil, act, scale, P or get_precision(), None)


class EncBlockFn(Function):
    """Conv1d_Layernorm_LRelu_Residual.forward (vae_npvc/model/layers.py:168-176):
    out = stack(c) + skip(c),  stack = n_layers x [LeakyReLU(0.2) -> Conv1d(k) -> GroupNorm(1,C)].
    params = n_layers x (v, g, bias, gamma, beta) + (v_skip, g_skip, bias_skip).
    Math: SURVEY.md section 9 A."""

    @staticmethod
    def forward(ctx, c, P, dil, box, n_layers, *params):
        c = c.contiguous()
        cpl = cached_planes(c, 'id', P)
        upl = cached_planes(c, 'lrelu', P)
        if cpl is None or upl is None:
            a, b = Pm.make_planes(c, P, 1.0, ACT_LRELU, cpl is None, upl is None)
            cpl, upl = cpl or a, upl or b
        C = c.shape[1]
        layers = []
        inp = upl
        y = ypl = yupl = None
        for j in range(n_layers):
            v, g, bias, gamma, beta = params[5 * j:5 * j + 5]
            d = dil if j == 0 else 1
            W = Pm.wn_prepare(v, g, False, P)
            h = Pm.conv_fwd(inp, W.wf, W.taps, d, C, bias, 0)
            mean, rstd = Pm.gn_stats(h, 1)
            if j < n_layers - 1:
                yj, _, nxt = Pm.enc_out(h, None, mean, rstd, gamma, beta, P, True, False, True)
            else:
                vs, gs, bs = params[5 * n_layers:5 * n_layers + 3]
                Ws = Pm.wn_prepare(vs, gs, False, P)
                s = Pm.conv_fwd(cpl, Ws.wf, Ws.taps, 1, C, bs, 0)
                y, ypl, yupl = Pm.enc_out(h, s, mean, rstd, gamma, beta, P, True, True, True)
                yj, nxt = None, None
            layers.append(dict(inp=inp, h=h, mean=mean, rstd=rstd, wb=W.wb, taps=W.taps, d=d, yj=yj))
            inp = nxt
        if box is not None:
            box['id'], box['lrelu'] = ypl, yupl
        ctx.save_for_backward(c, *params)
        ctx.layers, ctx.cpl, ctx.ws_wb = layers, cpl, Ws.wb
        ctx.cfg = (P, n_layers, C)
        return y

    @staticmethod
    def backward(ctx, dy):
        c, *params = ctx.saved_tensors
        P, n_layers, C = ctx.cfg
        dy = dy.contiguous()
        T = dy.shape[2]
        grads = [None] * len(params)
        dn = dy
        dypl = None
        for j in range(n_layers - 1, -1, -1):
            L = ctx.layers[j]
            v, g, bias, gamma, beta = params[5 * j:5 * j + 5]
            last = j == n_layers - 1
            R = Pm.gn_bwd_rows(dn, L['h'], L['mean'], L['rstd'], 1)
            dgamma, dbeta, S, _, csum = Pm.gn_bwd_finalize(R, gamma, L['rstd'], T, 1, True)
            dpre, dnpl = Pm.gn_bwd_apply(dn, L['h'], L['mean'], L['rstd'], gamma, S, 1, P, last)
            if last:
                dypl = dnpl
                grads[5 * n_layers + 2] = dbeta.clone()               # skip bias: sum_{b,t} dy
            dw = Pm.conv_wgrad(dpre, L['inp'], L['taps'], L['d'], False)
            dv, dg = _dw_to_param_grads(dw, v, g)
            grads[5 * j:5 * j + 5] = [dv, dg, csum, dgamma, dbeta]
            du = Pm.conv_fwd(dpre, L['wb'], L['taps'], L['d'], C)
            if j > 0:
                dn = Pm.act_bwd_add(du, ctx.layers[j - 1]['yj'], None, ACT_LRELU)
        vs, gs, bs = params[5 * n_layers:5 * n_layers + 3]
        dws = Pm.conv_wgrad(dypl, ctx.cpl, 1, 1, False)
        grads[5 * n_layers], grads[5 * n_layers + 1] = _dw_to_param_grads(dws, vs, gs)
        dc = None
        if ctx.needs_input_grad[0]:
            dcs = Pm.conv_fwd(dypl, ctx.ws_wb, 1, 1, C)
            dc = Pm.act_bwd_add(du, c, dcs, ACT_LRELU)
        return (dc, None, None, None, None) + tuple(grads)


def enc_block(c, params, n_layers=1, dil=1, P=None):
    box = {}
    y = EncBlockFn.apply(c, P or get_precision(), dil, box, n_layers, *params)
    for k, pl in box.items():
        attach_planes(y, k, pl)
    return y
