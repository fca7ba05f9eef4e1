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
_MODES = {'bf16': 1, 'bf16x2': 2, 'fp32': 3, 'f32': 3, '1': 1, '2': 2, '3': 3, 1: 1, 2: 2, 3: 3}
_PRECISION = {'P': _MODES.get(os.environ.get('VNB_PRECISION', 'bf16').lower(), 1)}


def set_precision(mode):
    """GEMM operand mode (number of bf16 planes an fp32 operand is split into; accumulation is always fp32 in TMEM):
    'bf16'   -> 1 plane  (8 significant bits per operand, 1 product)             - fastest, ~1e-2 on gradients
    'bf16x2' -> 2 planes (hi + mid, 16 bits, products hi*hi + hi*mid + mid*hi)   - <= 1e-3 on losses, gradients, features
    'fp32'   -> 3 planes (hi + mid + lo, 24 bits, 6 products)                    - fp32-faithful, the strict parity mode."""
    _PRECISION['P'] = _MODES[mode]


def get_precision():
    return _PRECISION['P']


_BANK = {'bank': None}


def set_bank(bank):
    """Install (or remove, with None) a bank.WeightBank: conv operands then come from it and parameter
    gradients are written straight into the parameters' ``.grad`` slots (the Functions return None for
    them).  Used by the Trainer; plain module-level use of the layers never sets one."""
    _BANK['bank'] = bank


def get_bank():
    return _BANK['bank']


_SIDE = {'stream': None, 'aux': None}


def set_side_stream(stream):
    """Install (or remove, with None) a CUDA stream for the weight-gradient GEMMs.  Nothing in the backward
    chain consumes dW - only the optimiser does - so with a bank installed they are forked onto this stream
    (event after the producer of their operands) and run beside the data-gradient chain; the caller joins
    the stream before it turns dW into (dv, dg).  Works under CUDA-graph capture (fork / join pattern)."""
    _SIDE['stream'] = stream
    if stream is None:
        _SIDE['aux'] = None
    else:
        if _SIDE.get('aux_cached') is None or _SIDE['aux_cached'].device != stream.device:
            _SIDE['aux_cached'] = torch.cuda.Stream(device=stream.device, priority=int(os.environ.get('VNB_AUX_PRIO', 0)))
        _SIDE['aux'] = _SIDE['aux_cached']


class _Fork(object):
    """Run an independent op on the auxiliary stream between ``with _Fork(tensors...)`` and ``join()``:
    the encoder block's skip conv beside its k=3 conv chain, the skip-path data gradient beside the
    GroupNorm backward chain.  A no-op when no side stream is installed."""

    def __init__(self, *inputs):
        self.aux = _SIDE['aux']
        self.inputs = inputs

    def __enter__(self):
        if self.aux is not None:
            ev = torch.cuda.Event()
            ev.record()
            self.aux.wait_event(ev)
            self.ctx = torch.cuda.stream(self.aux)
            self.ctx.__enter__()
        return self

    def __exit__(self, *exc):
        if self.aux is not None:
            self.ctx.__exit__(*exc)
            for t in self.inputs:
                if t is not None:
                    t.record_stream(self.aux)
        return False

    def join(self, *outputs):
        if self.aux is not None:
            torch.cuda.current_stream().wait_stream(self.aux)
            for t in outputs:
                if t is not None:
                    t.record_stream(torch.cuda.current_stream())


def _slot(p):
    """The gradient slot of leaf parameter ``p`` when gradients are written directly, else None."""
    if _BANK['bank'] is None or p is None or not getattr(p, 'is_leaf', False):
        return None
    return p.grad


def _weights(v, g, transposed, P, need_wf=True, need_wb=True, need_w32=False):
    bank = _BANK['bank']
    if bank is not None and bank.P == P and bank.has(v):
        W = bank.weights(v)
        if not need_w32 or W.w32 is not None:
            return W
    return Pm.wn_prepare(v, g, transposed, P, need_wf=need_wf, need_wb=need_wb, need_w32=need_w32)


def _wgrad(dypl, upl, taps, dil, transposed, v, g):
    """Weight gradient of a conv -> (dv, dg) for autograd, or (None, None) after writing raw dW into the
    v gradient slot (the bank's finish_backward turns it into dv, dg)."""
    bank = _BANK['bank']
    if bank is not None and bank.has(v) and _slot(v) is not None:
        side = _SIDE['stream']
        if side is None:
            Pm.conv_wgrad(dypl, upl, taps, dil, transposed, out=v.grad)
            return None, None
        ev = torch.cuda.Event()
        ev.record()
        side.wait_event(ev)
        with torch.cuda.stream(side):
            Pm.conv_wgrad(dypl, upl, taps, dil, transposed, out=v.grad)
        dypl.data.record_stream(side)        # operands were allocated on the main stream
        upl.data.record_stream(side)
        return None, None
    return _dw_to_param_grads(Pm.conv_wgrad(dypl, upl, taps, dil, transposed), v, g)


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
        x = Pm.as_cl(x)
        xpl = _input_planes(x, P, act, scale)
        need_dx = ctx.needs_input_grad[0]
        W = _weights(v, g, transposed, P, need_wb=need_dx)
        out = Pm.conv_fwd(xpl, W.wf, W.taps, dil, W.Cout, bias, 0)
        ctx.save_for_backward(x if (act != ACT_NONE or scale != 1.0) else None, v, g, bias)
        ctx.xpl, ctx.wb = xpl, W.wb
        ctx.cfg = (transposed, dil, act, scale, P, W.taps, W.Cin, bias is not None)
        return out

    @staticmethod
    def backward(ctx, dy):
        x, v, g, bias = ctx.saved_tensors
        transposed, dil, act, scale, P, taps, Cin, has_bias = ctx.cfg
        dy = Pm.as_cl(dy)
        db = None
        if has_bias and ctx.needs_input_grad[3]:
            # bias gradient = channel sums of dy, accumulated by the kernel that quantises dy (the gradient slot is
            # zeroed by the trainer's flat-gradient memset)
            slot = _slot(bias)
            db = slot if slot is not None else torch.zeros_like(bias)
            dypl = Pm.make_planes(dy, P, csum=db)[0]
            if slot is not None:
                db = None
        else:
            dypl = Pm.make_planes(dy, P)[0]
        dx = None
        if ctx.needs_input_grad[0]:
            if x is not None:
                # the activation's backward rides in the data-gradient GEMM's epilogue
                dx = Pm.conv_dgrad_act(dypl, ctx.wb, taps, dil, Cin, x, act, scale)
            else:
                dx = Pm.conv_fwd(dypl, ctx.wb, taps, dil, Cin)
        dv = dg = None
        if ctx.needs_input_grad[1]:
            dv, dg = _wgrad(dypl, ctx.xpl, taps, dil, transposed, v, g)
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
        c = Pm.as_cl(c)
        cpl = cached_planes(c, 'id', P)
        upl = cached_planes(c, 'lrelu', P)
        if cpl is None or upl is None:
            a, b = Pm.make_planes(c, P, 1.0, ACT_LRELU, cpl is None, upl is None)
            cpl, upl = cpl or a, upl or b
        C = c.shape[1]
        layers = []
        inp = upl
        y = ypl = yupl = None
        # the 1x1 skip conv only depends on the block input: run it beside the k-tap conv chain
        vs, gs, bs = params[5 * n_layers:5 * n_layers + 3]
        Ws = _weights(vs, gs, False, P)
        with _Fork(cpl.data) as fk:
            s_skip = Pm.conv_fwd(cpl, Ws.wf, Ws.taps, 1, C, bs, 0)
        for j in range(n_layers):
            v, g, bias, gamma, beta = params[5 * j:5 * j + 5]
            d = dil if j == 0 else 1
            W = _weights(v, g, False, P)
            h, mean, rstd = Pm.conv_fwd(inp, W.wf, W.taps, d, C, bias, 0, gn_groups=1, defer_stats=True)
            if j < n_layers - 1:
                yj, _, nxt = Pm.enc_out(h, None, mean, rstd, gamma, beta, P, True, False, True)
            else:
                fk.join(s_skip)
                y, ypl, yupl = Pm.enc_out(h, s_skip, mean, rstd, gamma, beta, P, True, True, True)
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
        dy = Pm.as_cl(dy)
        T = dy.shape[2]
        grads = [None] * len(params)
        dn = dy
        dypl = None
        finals = []
        for j in range(n_layers - 1, -1, -1):
            L = ctx.layers[j]
            v, g, bias, gamma, beta = params[5 * j:5 * j + 5]
            last = j == n_layers - 1
            # the row-sum kernel also publishes the group sums the apply kernel needs; the finalisation (parameter
            # gradients only) then runs beside the rest of the chain
            R, S = Pm.gn_bwd_rows(dn, L['h'], L['mean'], L['rstd'], 1, gamma)
            bs = params[5 * n_layers + 2]
            slots = {'dgamma': _slot(gamma), 'dbeta': _slot(beta), 'csum': _slot(bias),
                     'dbeta2': _slot(bs) if last else None}          # skip bias grad = sum_{b,t} dy = dbeta
            if S is None:
                dgamma, dbeta, S, _, csum = Pm.gn_bwd_finalize(R, gamma, L['rstd'], T, 1, True, slots)
            else:
                with _Fork(R) as fk_fin:
                    dgamma, dbeta, _, _, csum = Pm.gn_bwd_finalize(R, gamma, L['rstd'], T, 1, True, slots)
                finals.append((fk_fin, dgamma, dbeta, csum))
            dpre, dnpl = Pm.gn_bwd_apply(dn, L['h'], L['mean'], L['rstd'], gamma, S, 1, P, last)
            if last:
                dypl = dnpl
                if ctx.needs_input_grad[0]:
                    with _Fork(dypl.data) as fk:       # skip-path data gradient, beside the rest of the chain
                        dcs = Pm.conv_fwd(dypl, ctx.ws_wb, 1, 1, C)
                if slots['dbeta2'] is None:
                    grads[5 * n_layers + 2] = (dbeta if dbeta is not None else beta.grad).clone()
            dv, dg = _wgrad(dpre, L['inp'], L['taps'], L['d'], False, v, g)
            grads[5 * j:5 * j + 5] = [dv, dg, csum, dgamma, dbeta]
            du = Pm.conv_fwd(dpre, L['wb'], L['taps'], L['d'], C)
            if j > 0:
                dn = Pm.act_bwd_add(du, ctx.layers[j - 1]['yj'], None, ACT_LRELU)
        vs, gs, bs = params[5 * n_layers:5 * n_layers + 3]
        grads[5 * n_layers], grads[5 * n_layers + 1] = _wgrad(dypl, ctx.cpl, 1, 1, False, vs, gs)
        dc = None
        if ctx.needs_input_grad[0]:
            fk.join(dcs)
            dc = Pm.act_bwd_add(du, c, dcs, ACT_LRELU)
        for fk_fin, a_, b_, c_ in finals:
            fk_fin.join(a_, b_, c_)
        return (dc, None, None, None, None) + tuple(grads)


def enc_block(c, params, n_layers=1, dil=1, P=None):
    box = {}
    y = EncBlockFn.apply(c, P or get_precision(), dil, box, n_layers, *params)
    for k, pl in box.items():
        attach_planes(y, k, pl)
    return y


class DecBlockFn(Function):
    """DeConv1d_Layernorm_GLU_ResSkip.forward (vae_npvc/model/layers.py:218-249), non-causal:
    a = ConvT_k(x) + Conv_1(cond); h = GroupNorm(2)(a); g = tanh(h[:C]) * sigmoid(h[C:]);
    r = Conv_1(g); x' = r[:C] + x; skip = r[C:].
    ``cond`` is (B,Y) when it is time-invariant (the Decoder's c.repeat, vqvae.py:309-312: the
    1x1 cond conv collapses to a per-utterance bias) or (B,Y,T) in general.
    ``skip_acc``: optional running skip sum updated IN PLACE (vqvae.py:313).  It is passed and
    returned in its PHYSICAL (B, T, S) form (a base tensor, not a transposed view: autograd does not
    allow in-place updates of view inputs of multi-output Functions); callers take ``.transpose(1, 2)``
    once at the end.
    Math: SURVEY.md section 9 B."""

    @staticmethod
    def forward(ctx, x, cond, skip_acc, P, dil, box, v_in, g_in, b_in, gamma, beta, v_c, g_c, b_c, v_rs, g_rs, b_rs):
        x = Pm.as_cl(x)
        B, C, T = x.shape
        S = v_rs.shape[0] - C
        xpl = _input_planes(x, P)
        Win = _weights(v_in, g_in, True, P, need_wb=ctx.needs_input_grad[0])
        # cond_slot (box['cond_slot'] = (l, shared, is_first)): ``cond`` is CondAllFn's (L, B, 2C) table of per-utterance
        # conditioning biases of ALL blocks (both conv biases folded in); this block uses row l
        cond_slot = box.get('cond_slot') if box is not None else None
        has_cond = v_c is not None and cond is not None and cond_slot is None
        bcast = has_cond and cond.dim() == 2
        cpl = wc32 = wc_wb = None
        if cond_slot is not None:
            a, mean, rstd = Pm.conv_fwd(xpl, Win.wf, Win.taps, dil, 2 * C, cond[cond_slot[0]], 2 * C, gn_groups=2, defer_stats=True)
        elif has_cond and bcast:
            cond = cond.contiguous()
            Wc = _weights(v_c, g_c, False, P, need_wf=False, need_wb=False, need_w32=True)
            wc32 = Wc.w32.view(2 * C, -1)
            p = Pm.linear_fwd(cond, wc32, b_c, b_in)
            a, mean, rstd = Pm.conv_fwd(xpl, Win.wf, Win.taps, dil, 2 * C, p, 2 * C, gn_groups=2, defer_stats=True)
        elif not has_cond:
            a, mean, rstd = Pm.conv_fwd(xpl, Win.wf, Win.taps, dil, 2 * C, b_in, 0, gn_groups=2, defer_stats=True)
        else:
            a = Pm.conv_fwd(xpl, Win.wf, Win.taps, dil, 2 * C, b_in, 0)
            if has_cond:
                cond = Pm.as_cl(cond)
                cpl = _input_planes(cond, P)
                Wc = _weights(v_c, g_c, False, P, need_wb=ctx.needs_input_grad[1])
                wc_wb = Wc.wb
                a, mean, rstd = Pm.conv_fwd(cpl, Wc.wf, 1, 1, 2 * C, b_c, 0, None, a, True, gn_groups=2, defer_stats=True)
        gpl = Pm.dec_gate(a, mean, rstd, gamma, beta, P)
        Wrs = _weights(v_rs, g_rs, False, P)
        r = Pm.conv_fwd(gpl, Wrs.wf, 1, 1, C + S, b_rs, 0)
        first = skip_acc is None
        if first:
            skip_acc = torch.empty((B, T, S), dtype=torch.float32, device=x.device)
        x_out, xopl, _ = Pm.dec_out(r, x, S, P, skip_acc.transpose(1, 2), first)
        skip = skip_acc
        if not first:
            ctx.mark_dirty(skip_acc)
        if box is not None:
            box['id'] = xopl
        ctx.save_for_backward(cond if has_cond else None, a, mean, rstd, gamma, beta, v_in, g_in, v_c, g_c, v_rs, g_rs,
                              b_in, b_c, b_rs)
        ctx.cond_slot = cond_slot
        ctx.xpl, ctx.gpl, ctx.cpl = xpl, gpl, cpl
        ctx.win_wb, ctx.wrs_wb, ctx.wc_wb, ctx.wc32 = Win.wb, Wrs.wb, wc_wb, wc32
        ctx.cfg = (P, dil, B, C, S, T, Win.taps, has_cond, bcast)
        ctx.has_acc = not first
        return x_out, skip

    @staticmethod
    def backward(ctx, dxo, dskip):
        cond, a, mean, rstd, gamma, beta, v_in, g_in, v_c, g_c, v_rs, g_rs, b_in, b_c, b_rs = ctx.saved_tensors
        P, dil, B, C, S, T, taps, has_cond, bcast = ctx.cfg
        dev = a.device
        # (the last block's x output is unused: its gradient does not exist and is never materialised as zeros)
        no_dxo = dxo is None
        dxo = None if no_dxo else (Pm.cl_zeros(B, C, T, dev) if dxo is None else Pm.as_cl(dxo))
        # dskip arrives in the physical (B, T, S) form of the skip accumulator
        dskip_phys = torch.zeros((B, T, S), device=dev) if dskip is None else dskip.contiguous()
        dskip = dskip_phys.transpose(1, 2)
        rs_slot = _slot(b_rs)                  # zeroed by the trainer's flat-gradient memset
        db_rs = rs_slot if rs_slot is not None else torch.zeros((C + S,), dtype=torch.float32, device=dev)
        drpl = Pm.concat_planes(dxo, dskip, P, csum=db_rs, Ca=C)      # res/skip bias gradient rides along
        if rs_slot is not None:
            db_rs = None
        dg = Pm.conv_fwd(drpl, ctx.wrs_wb, 1, 1, C)
        dv_rs, dg_rs = _wgrad(drpl, ctx.gpl, 1, 1, False, v_rs, g_rs)
        R, S_rows = Pm.gate_bwd_rows(dg, a, mean, rstd, gamma, beta, P, want_S=True)
        cond_slot = ctx.cond_slot
        any_cond = has_cond or (cond_slot is not None and b_c is not None)
        slots = {'dgamma': _slot(gamma), 'dbeta': _slot(beta), 'csum': _slot(b_in),
                 'csum2': _slot(b_c) if any_cond else None}          # both biases add to the GN input
        if cond_slot is not None:
            slots['bsum'] = cond_slot[1]['dP'][cond_slot[0]]          # d(cond bias) of this block -> CondAllFn's table
        fk_fin = None
        if S_rows is None:
            dgamma, dbeta, Ssum, bsum, csum = Pm.gn_bwd_finalize(R, gamma, rstd, T, 2, True, slots)
        else:
            # the group sums came with the row sums: the finalisation (parameter gradients, conditioning-bias gradient)
            # leaves the critical path
            with _Fork(R) as fk_fin:
                dgamma, dbeta, _, bsum, csum = Pm.gn_bwd_finalize(R, gamma, rstd, T, 2, True, slots)
            Ssum = S_rows
        dcond = dv_c = dg_c = db_c = None
        fk_lin = None
        if has_cond and bcast:
            # the conditioning path's backward only needs bsum: forked beside the gate backward + k=3 data gradient
            bank = get_bank()
            with _Fork(bsum) as fk_lin:
                if bank is not None and bank.has(v_c) and _slot(v_c) is not None:
                    _, dcond = Pm.linear_bwd(bsum, cond, ctx.wc32, None, v_c.grad.view(2 * C, -1))
                else:
                    dwc, dcond = Pm.linear_bwd(bsum, cond, ctx.wc32)
                    dv_c, dg_c = _dw_to_param_grads(dwc.view_as(v_c), v_c, g_c)
        dapl = Pm.gate_bwd_apply(dg, a, mean, rstd, gamma, beta, Ssum, P)
        dx = Pm.conv_fwd(dapl, ctx.win_wb, taps, dil, C, None, 0, dxo) if ctx.needs_input_grad[0] else None
        dv_in, dg_in = _wgrad(dapl, ctx.xpl, taps, dil, True, v_in, g_in)
        if fk_fin is not None:
            fk_fin.join(dgamma, dbeta, bsum, csum)
        if cond_slot is not None:
            if b_c is not None and slots['csum2'] is None:
                db_c = (csum if csum is not None else b_in.grad).clone()
            # the whole table of gradients is handed to autograd once, by the block whose backward runs last
            dcond = cond_slot[1]['dP'] if cond_slot[2] else None
        if has_cond:
            if slots['csum2'] is None:
                db_c = (csum if csum is not None else b_in.grad).clone()
            if bcast:
                fk_lin.join(dcond, dv_c, dg_c)
            else:
                dv_c, dg_c = _wgrad(dapl, ctx.cpl, 1, 1, False, v_c, g_c)
                if ctx.needs_input_grad[1]:
                    dcond = Pm.conv_fwd(dapl, ctx.wc_wb, 1, 1, cond.shape[1])
            if not ctx.needs_input_grad[1]:
                dcond = None
        return (dx, dcond, dskip_phys if ctx.has_acc else None, None, None, None, dv_in, dg_in, csum, dgamma, dbeta, dv_c, dg_c, db_c,
                dv_rs, dg_rs, db_rs)


class CondAllFn(Function):
    """The 1x1 conditioning convs of ALL gated decoder blocks on the same time-invariant speaker embedding
    (vqvae.py:309-312 feeds c.repeat(1,1,T) to every block; layers.py:231) in ONE launch:
        p[l] = conv_cond_l(c) + b_cond_l + b_in_l                       (L, B, 2C)
    and, in backward, ONE launch for every dW_l and dc = sum_l W_l^T dp_l (30 launches and 9 gradient adds per step
    before).  params = L x (v_c, g_c, b_c, b_in); the bias gradients are produced by the blocks (they are the channel
    sums of the GroupNorm input gradient), so they return None here.  ``shared['dP']``: the (L, B, 2C) buffer the
    blocks' backward write their rows into."""

    @staticmethod
    def forward(ctx, cond, shared, P, *params):
        cond = cond.contiguous()
        L = len(params) // 4
        Ws, b1, b2 = [], [], []
        for l in range(L):
            v_c, g_c, b_c, b_in = params[4 * l:4 * l + 4]
            W = _weights(v_c, g_c, False, P, need_wf=False, need_wb=False, need_w32=True)
            Ws.append(W.w32.view(v_c.shape[0], -1))
            b1.append(b_c)
            b2.append(b_in)
        p_all = Pm.linear_fwd_batch(cond, Ws, b1, b2)
        shared['dP'] = torch.empty_like(p_all)
        ctx.save_for_backward(cond, *params)
        ctx.Ws, ctx.L = Ws, L
        return p_all

    @staticmethod
    def backward(ctx, dP):
        cond, *params = ctx.saved_tensors
        L = ctx.L
        bank = _BANK['bank']
        targets, tmp = [], []
        for l in range(L):
            v_c, g_c = params[4 * l], params[4 * l + 1]
            if not ctx.needs_input_grad[3 + 4 * l]:
                targets.append(None)
            elif bank is not None and bank.has(v_c) and _slot(v_c) is not None:
                targets.append(v_c.grad.view(v_c.shape[0], -1))          # raw dW into the slot (bank.finish_backward)
            else:
                t = torch.empty((v_c.shape[0], v_c.shape[1] * v_c.shape[2]), dtype=torch.float32, device=cond.device)
                targets.append(t)
                tmp.append(l)
        dy = Pm.linear_bwd_batch(dP.contiguous(), cond, ctx.Ws, targets, ctx.needs_input_grad[0])
        grads = [None] * (4 * L)
        for l in tmp:
            v_c, g_c = params[4 * l], params[4 * l + 1]
            grads[4 * l], grads[4 * l + 1] = _dw_to_param_grads(targets[l].view_as(v_c), v_c, g_c)
        return (dy, None, None) + tuple(grads)


def cond_all(cond, params, P=None):
    """-> (p_all (L, B, 2C), shared) for dec_block(..., cond_slot=(l, shared, is_first))."""
    shared = {}
    return CondAllFn.apply(cond, shared, P or get_precision(), *params), shared


def dec_block(x, cond, params, skip_acc=None, dil=1, P=None, cond_slot=None):
    """params = (v_in, g_in, b_in, gamma, beta, v_c, g_c, b_c, v_rs, g_rs, b_rs) -> (x', skip_phys);
    skip_acc / skip_phys are physical (B, T, S) tensors (see DecBlockFn).  ``cond_slot`` = (l, shared, is_first):
    ``cond`` is cond_all's table and this block takes row l."""
    box = {}
    if cond_slot is not None:
        box['cond_slot'] = cond_slot
        params = tuple(params[:5]) + (None, None, params[7]) + tuple(params[8:])      # conv_cond weights live in CondAllFn
    x_out, skip = DecBlockFn.apply(x, cond, skip_acc, P or get_precision(), dil, box, *params)
    attach_planes(x_out, 'id', box.get('id'))
    return x_out, skip


def _vq_scale(reduction, B, T, D):
    """Scalar the summed squared error is multiplied by (layers_vq.py:123-134)."""
    return {'sum': 1.0, 'mean': 1.0 / (B * T * D), 'batch_mean': 1.0 / B, 'frame_mean': 1.0 / (B * T)}[reduction]


def _vq_lookup_gather(z, normalize, e_hat, e_pl, e_sq, want_q=True, want_stats=True):
    """-> (zn, z_norm, idx, q, sums, counts): one fused launch where the shape allows (prims.vq_fused), else the three
    separate kernels (frame preparation, six-product tensor-core lookup or SIMT lookup, gather)."""
    fused = Pm.vq_fused(z, normalize, e_hat, e_pl, e_sq, want_q, want_stats) if hasattr(Pm, 'vq_fused') else None
    if fused is not None:
        return fused
    zn, zpl, z_sq, z_norm = Pm.vq_prep_z(z, normalize)
    idx = Pm.vq_lookup(zn, zpl, z_sq, e_hat, e_pl, e_sq)
    q = sums = counts = None
    if want_q or want_stats:
        q, sums, counts = Pm.vq_gather(idx, e_hat, zn, z if normalize else None, want_q, want_stats)
    return zn, z_norm, idx, q, sums, counts


class VQFn(Function):
    """VectorQuantizer.forward for the scalar reductions (layers_vq.py:79-150): in-place codebook
    renormalisation, distance GEMM + argmin, gather, both losses, perplexity, straight-through
    output.  Returns (z_vq, z_qut_loss, z_enc_loss, perplexity, idx).  SURVEY.md section 9 C/D."""

    @staticmethod
    def forward(ctx, z, E, normalize, reduction):
        z = Pm.as_cl(z)
        B, D, T = z.shape
        e_hat, e_pl, e_sq = Pm.vq_prep_codebook(E.detach(), normalize, True)
        zn, z_norm, idx, q, sums, counts = _vq_lookup_gather(z, normalize, e_hat, e_pl, e_sq)
        scale = _vq_scale(reduction, B, T, D)
        scal = Pm.vq_finalize(sums, counts, B * T, scale, normalize, 1e-10)
        ctx.save_for_backward(z, zn, z_norm, e_hat, E, idx)
        ctx.cfg = (scale, normalize)
        lq, le, perp = scal[0], scal[1], scal[2]
        ctx.mark_non_differentiable(perp, idx)
        return q, lq, le, perp, idx

    @staticmethod
    def backward(ctx, g_out, g_lq, g_le, _gp, _gi):
        z, zn, z_norm, e_hat, E, idx = ctx.saved_tensors
        scale, normalize = ctx.cfg
        g_out = Pm.as_cl(g_out) if g_out is not None else None
        g_lq = g_lq.contiguous() if g_lq is not None else None
        g_le = g_le.contiguous() if g_le is not None else None
        want_dE = ctx.needs_input_grad[1]
        dz, dEh = Pm.vq_bwd(g_out, g_lq, g_le, idx, z, zn, z_norm, e_hat, E, scale, normalize, want_dE)
        dE = None
        if want_dE:
            slot = _slot(E)
            dE = Pm.vq_bwd_codebook(dEh, E.detach(), normalize, slot)
            if slot is not None:
                dE = None
        return dz, dE, None, None


class VQLookupFn(Function):
    """Lookup only (VectorQuantizer.encode, layers_vq.py:35-58; the EMA twin :236-252)."""

    @staticmethod
    def forward(ctx, z, E, normalize):
        z = Pm.as_cl(z)
        e_hat, e_pl, e_sq = Pm.vq_prep_codebook(E.detach(), normalize, False)
        idx = _vq_lookup_gather(z, normalize, e_hat, e_pl, e_sq, want_q=False, want_stats=False)[2]
        ctx.mark_non_differentiable(idx)
        return idx

    @staticmethod
    def backward(ctx, _g):
        return None, None, None


class EMAVQFn(Function):
    """Lookup + gather + commitment loss of EMAVectorQuantizer.forward (layers_vq.py:285-308).
    Bug-compatible with the reference: for the scalar reductions the quantised output carries
    NO gradient; only z_enc_loss reaches z.  Returns (z_vq, z_enc_loss, idx)."""

    @staticmethod
    def forward(ctx, z, embeddings, reduction):
        z = Pm.as_cl(z)
        B, D, T = z.shape
        e_hat, e_pl, e_sq = Pm.vq_prep_codebook(embeddings, False, False)
        _, _, idx, q, sums, counts = _vq_lookup_gather(z, False, e_hat, e_pl, e_sq)
        scale = _vq_scale(reduction, B, T, D)
        scal = Pm.vq_finalize(sums, counts, B * T, scale, False, 1e-8)
        ctx.save_for_backward(z, e_hat, idx)
        ctx.scale = scale
        ctx.mark_non_differentiable(q, idx)
        return q, scal[1], idx

    @staticmethod
    def backward(ctx, _gq, g_le, _gi):
        z, e_hat, idx = ctx.saved_tensors
        g_le = g_le.contiguous() if g_le is not None else None
        dz, _ = Pm.vq_bwd(None, None, g_le, idx, None, z, None, e_hat, e_hat, ctx.scale, False, False)
        return dz, None, None


class JitterFn(Function):
    """Jitter.forward (layers_vq.py:353-379) with the host draws given as a source table."""

    @staticmethod
    def forward(ctx, q, src):
        ctx.save_for_backward(src)
        return Pm.jitter_fwd(Pm.as_cl(q), src)

    @staticmethod
    def backward(ctx, dout):
        (src,) = ctx.saved_tensors
        return Pm.jitter_bwd(Pm.as_cl(dout), src), None


class LogLossFn(Function):
    """log_loss (layers.py:283-296) for the scalar reductions: scale * sum 0.5*(log 2pi + (x-mu)^2)."""

    @staticmethod
    def forward(ctx, x, mu, scale):
        x, mu = Pm.as_cl(x), Pm.as_cl(mu)
        ctx.save_for_backward(x, mu)
        ctx.scale = scale
        return Pm.logloss_fwd(x, mu, scale)[0]

    @staticmethod
    def backward(ctx, g):
        x, mu = ctx.saved_tensors
        g = g.contiguous()
        # d/dx = g*scale*(x-mu); d/dmu = -that
        dx = Pm.logloss_bwd(x, mu, g, ctx.scale) if ctx.needs_input_grad[0] else None
        dmu = Pm.logloss_bwd(mu, x, g, ctx.scale) if ctx.needs_input_grad[1] else None
        return dx, dmu, None


class EmbedFn(Function):
    """nn.Embedding row gather of Conditions (layers.py:42-54) and its scatter-add backward."""

    @staticmethod
    def forward(ctx, table, idx):
        ctx.save_for_backward(idx, table)
        ctx.rows = table.shape[0]
        return Pm.embed_fwd(table, idx)

    @staticmethod
    def backward(ctx, dout):
        idx, table = ctx.saved_tensors
        slot = _slot(table)                    # zeroed by the trainer's flat-gradient memset
        dt = Pm.embed_bwd(dout, idx, ctx.rows, slot)
        return (None if slot is not None else dt), None


# ---------------------------------------------------------------------------------------------------------------
# hierarchical VQ-VAE2 + global style tokens (vae_npvc/model/vqvae2.py, layers_gst.py)
# ---------------------------------------------------------------------------------------------------------------
def strided_conv1d(x, v, g, bias, s, act=ACT_NONE, P=None):
    """The down-sampling conv of a vqvae2 Encoder stage (vqvae2.py:204-212): kernel 2s, stride s, padding s/2 (s even).

    On channels-last frames no strided kernel is needed: the (B, T, C) buffer viewed as (B, T/s, s*C) - rows of s
    consecutive frames, a free view - turns the conv into a 3-TAP STRIDE-1 'same' conv with s*C input channels.  Output
    frame t reads input frames s*t - s/2 ... s*t + 3s/2 - 1, i.e. the upper half of row t-1, all of row t and the lower
    half of row t+1:
        v'[o, j*C + c, dr + 1] = v[o, c, s*dr + j + s/2]   where that tap exists, else 0        (dr in -1..1, j in 0..s-1)
    The zeros do not change the weight-norm (||v'|| = ||v||), the conv's own zero padding of one row each side is the
    reference's padding of s/2 frames, and the re-laid weight is a (differentiable) pad + permute of the parameter, so
    the existing tcgen05 conv, its data gradient and its weight-gradient GEMM serve (1.5x the FLOPs of the 2s taps).
    ``act`` is applied to x first (the previous stage's LeakyReLU, vqvae2.py:226)."""
    assert s >= 2 and s % 2 == 0, 'down-sampling scales of the shipped recipe are 2 and 4 (even)'
    B, C, T = x.shape
    if T % s != 0:
        raise NotImplementedError('vae_npvc_b200: strided conv needs a frame count divisible by the stride (got T=%d, s=%d)' % (T, s))
    Cout, Cin, k = v.shape
    assert k == 2 * s and Cin == C
    p = s // 2
    vp = torch.nn.functional.pad(v, (s - p, p))                                    # (Cout, C, 3s): tap index m + s - p
    v3 = vp.view(Cout, C, 3, s).permute(0, 3, 1, 2).reshape(Cout, s * C, 3)         # [o, j*C + c, dr + 1]
    xt = x.transpose(1, 2)                                                           # physical (B, T, C) of a channels-last x
    if not xt.is_contiguous():
        xt = xt.contiguous()                                                         # (autograd-aware copy of a foreign layout)
    rows = xt.reshape(B, T // s, s * C).transpose(1, 2)                              # logical (B, s*C, T/s), channels-last
    return conv1d(rows, v3, g, bias, False, 1, act, 1.0, P)


class UpsampleCatFn(Function):
    """torch.cat([upsample(z_i, T) for i], dim=1) (vqvae2.py:104-114, 130-143) in one pass per source."""

    @staticmethod
    def forward(ctx, T, *srcs):
        srcs = [Pm.as_cl(s) for s in srcs]
        ctx.shapes = [(s.shape[1], s.shape[2]) for s in srcs]
        return Pm.upsample_cat_fwd(srcs, T)

    @staticmethod
    def backward(ctx, dout):
        return (None,) + tuple(Pm.upsample_cat_bwd(Pm.as_cl(dout), ctx.shapes))


def upsample_cat(srcs, T):
    return UpsampleCatFn.apply(T, *srcs)


class TimeMeanFn(Function):
    """z.mean(dim=-1) (vqvae2.py:96): (B, C, T) -> (B, C)."""

    @staticmethod
    def forward(ctx, x):
        x = Pm.as_cl(x)
        ctx.T = x.shape[2]
        return Pm.time_mean_fwd(x)

    @staticmethod
    def backward(ctx, dy):
        return Pm.time_mean_bwd(dy.contiguous(), ctx.T)


class LinearFn(Function):
    """torch.nn.Linear on (rows, in) (layers_gst.py:94-96, 137)."""

    @staticmethod
    def forward(ctx, x, W, b):
        x = x.contiguous()
        ctx.save_for_backward(x, W)
        return Pm.linear_fwd(x, W, b)

    @staticmethod
    def backward(ctx, dp):
        x, W = ctx.saved_tensors
        dp = dp.contiguous()
        dW, dx = Pm.linear_bwd(dp, x, W)
        return dx, dW, Pm.channel_sum(dp.unsqueeze(0).transpose(1, 2))         # bias gradient: column sums of dp


class GstAttnFn(Function):
    """softmax(q K^T / sqrt(dk)) V per head, one query per utterance (layers_gst.py:128-147)."""

    @staticmethod
    def forward(ctx, q, K, V, H):
        q, K, V = q.contiguous(), K.contiguous(), V.contiguous()
        attn, out = Pm.gst_attn_fwd(q, K, V, H)
        ctx.save_for_backward(q, K, V, attn)
        return out

    @staticmethod
    def backward(ctx, dctx):
        q, K, V, attn = ctx.saved_tensors
        dq, dK, dV = Pm.gst_attn_bwd(dctx.contiguous(), q, K, V, attn)
        return dq, dK, dV, None
