"""Python-level primitives: one function per C-ABI entry point (include/vae_npvc_b200.h).

Each allocates its outputs with torch (the library itself allocates nothing) and launches on
the current CUDA stream.  This module is the single seam between the host-side mirror of the
reference classes (functional.py, model/*.py) and the CUDA kernels; there is no other
implementation behind it.
"""
import torch

from . import _C
from ._C import ACT_LRELU, ACT_NONE, ACT_RELU  # noqa: F401


def pitch(C):
    return (C + 7) // 8 * 8


# ---- channels-last helpers -------------------------------------------------------------------
# Activations keep the reference's LOGICAL shape (B, C, T) but live in memory as a contiguous
# (B, T, C) array (strides (T*C, 1, C)): frames are rows, channels contiguous.  Any torch op sees
# an ordinary (B, C, T) tensor; the kernels get the contiguous (B, T, C) view.
def cl_empty(B, C, T, device, dtype=torch.float32):
    return torch.empty((B, T, C), dtype=dtype, device=device).transpose(1, 2)


def cl_zeros(B, C, T, device, dtype=torch.float32):
    return torch.zeros((B, T, C), dtype=dtype, device=device).transpose(1, 2)


def is_cl(x):
    return x.dim() == 3 and x.transpose(1, 2).is_contiguous()


def as_cl(x):
    """Same logical (B,C,T) tensor, channels-last in memory (converted by vnb_transpose if needed)."""
    if is_cl(x):
        return x
    if x.dtype == torch.float32 and x.is_cuda:
        xc = x.contiguous()
        B, C, T = xc.shape
        out = cl_empty(B, C, T, x.device)
        _C.call('transpose', xc, B, C, T, out.transpose(1, 2))
        return out
    return x.transpose(1, 2).contiguous().transpose(1, 2)


def to_nct(x):
    """Logical (B,C,T) tensor -> reference-layout contiguous copy (for callers that need it)."""
    if x.is_contiguous():
        return x
    if is_cl(x) and x.dtype == torch.float32 and x.is_cuda:
        B, C, T = x.shape
        out = torch.empty((B, C, T), dtype=torch.float32, device=x.device)
        _C.call('transpose', x.transpose(1, 2), B, T, C, out)
        return out
    return x.contiguous()


def _r(x):
    """contiguous (B,T,C) storage view of a logical (B,C,T) activation (None passes through)."""
    return None if x is None else as_cl(x).transpose(1, 2)


class Planes:
    """bf16 GEMM-operand planes of an activation: data (P, B, T, Cp), channels-last."""
    __slots__ = ('data', 'P', 'B', 'C', 'T', 'Cp')

    def __init__(self, data, P, B, C, T, Cp):
        self.data, self.P, self.B, self.C, self.T, self.Cp = data, P, B, C, T, Cp

    @staticmethod
    def empty(P, B, C, T, device):
        Cp = pitch(C)
        # zero-filled so the pitch padding is finite (the kernels never read it as data)
        alloc = torch.zeros if Cp != C else torch.empty
        return Planes(alloc((P, B, T, Cp), dtype=torch.bfloat16, device=device), P, B, C, T, Cp)


class ConvWeights:
    """Materialised conv weights: wf (P,k,co,ci_p) forward operand, wb (P,k,ci,co_p) data-grad
    operand, w32 optional fp32 effective weight in the parameter layout."""
    __slots__ = ('wf', 'wb', 'w32', 'taps', 'Cout', 'Cin', 'transposed')

    def __init__(self, wf, wb, w32, taps, Cout, Cin, transposed):
        self.wf, self.wb, self.w32, self.taps, self.Cout, self.Cin, self.transposed = wf, wb, w32, taps, Cout, Cin, transposed


def _f32(shape, like):
    return torch.empty(shape, dtype=torch.float32, device=like.device)


# --------------------------------------------------------------------------------------------
# weight norm
# --------------------------------------------------------------------------------------------
def _wbuf(shape, inner, device):
    # pitch padding (if any) must be finite; otherwise skip the memset
    alloc = torch.zeros if shape[-1] != inner else torch.empty
    return alloc(shape, dtype=torch.bfloat16, device=device)


def wn_prepare(v, g, transposed, P, need_wf=True, need_wb=True, need_w32=False):
    d0, d1, k = v.shape
    co, ci = (d1, d0) if transposed else (d0, d1)
    wf = _wbuf((P, k, co, pitch(ci)), ci, v.device) if need_wf else None
    wb = _wbuf((P, k, ci, pitch(co)), co, v.device) if need_wb else None
    w32 = _f32((d0, d1, k), v) if need_w32 else None
    _C.call('wn_prepare', v, g, d0, d1, k, int(transposed), P, wf, wb, w32)
    return ConvWeights(wf, wb, w32, k, co, ci, bool(transposed))


def wn_prepare_batch(flat_p, desc, row2layer, P, wbuf, w32buf, tile2layer, max_k, row_scale):
    _C.call('wn_prepare_batch', flat_p, desc, row2layer, row2layer.numel(), tile2layer, tile2layer.numel(), max_k, P,
            row_scale, wbuf, w32buf)


def wn_backward_batch(flat_p, flat_g, desc, row2layer):
    _C.call('wn_backward_batch', flat_p, flat_g, desc, row2layer, row2layer.numel())


def wn_backward(dw, v, g):
    d0, d1, k = v.shape
    dv, dg = torch.empty_like(v), torch.empty_like(g)
    _C.call('wn_backward', dw, v, g, d0, d1, k, dv, dg)
    return dv, dg


# --------------------------------------------------------------------------------------------
# GEMM-shaped ops
# --------------------------------------------------------------------------------------------
# Zeroed scratch for the fused GroupNorm statistics: the trainer zeroes ONE arena per step (stats_arena_begin)
# and conv_fwd carves its (B, groups, 2) fp64 accumulators out of it, instead of one fill kernel per conv.
_ARENA = {'buf': None, 'off': -1}


def stats_arena_begin(device, n_doubles=1 << 16):
    buf = _ARENA['buf']
    if buf is None or buf.device != torch.device(device) or buf.numel() < n_doubles:
        buf = torch.empty(n_doubles, dtype=torch.float64, device=device)
        _ARENA['buf'] = buf
    buf.zero_()
    _ARENA['off'] = 0


def stats_arena_end():
    _ARENA['off'] = -1


def _stats_zeros(B, groups, device):
    n = B * groups * 2
    buf, off = _ARENA['buf'], _ARENA['off']
    if buf is not None and off >= 0 and buf.device == torch.device(device) and off + n <= buf.numel():
        _ARENA['off'] = off + n
        return buf[off:off + n].view(B, groups, 2)
    return torch.zeros((B, groups, 2), dtype=torch.float64, device=device)


def conv_fwd(x_pl, w_op, taps, dil, Cout, bias=None, bias_bstride=0, residual=None, out=None, accumulate=False,
             gn_groups=0, eps=1e-5, lengths=None, defer_stats=False):
    """x_pl: Planes (B,Cin,T); w_op: (P,taps,Cout,Cin_p) bf16 -> out fp32, logical (B,Cout,T).
    gn_groups > 0: also returns the GroupNorm (mean, rstd) (B, gn_groups) of the result, accumulated
    in the GEMM epilogue -> (out, mean, rstd).  ``lengths`` (B,) int32: packed variable-length batch - the
    statistics cover the first lengths[b] frames of each sample only."""
    user_out = None
    dev = x_pl.data.device
    if out is None:
        out = cl_empty(x_pl.B, Cout, x_pl.T, dev)
    elif not is_cl(out):                      # reference-layout buffer from a caller: stage through channels-last
        user_out, out = out, as_cl(out)
    sums = _stats_zeros(x_pl.B, gn_groups, dev) if gn_groups else None
    _C.call('conv_fwd', x_pl.data, x_pl.P, x_pl.B, x_pl.C, x_pl.T, w_op, taps, dil, Cout, bias,
            bias_bstride, _r(residual), out.transpose(1, 2), int(accumulate), sums, gn_groups)
    if user_out is not None:
        user_out.copy_(out)
        out = user_out
    if not gn_groups:
        return out
    mean, rstd = _f32((x_pl.B, gn_groups), out), _f32((x_pl.B, gn_groups), out)
    if defer_stats and lengths is None:
        # the consumer (enc_out / dec_gate) finalises the statistics itself and fills mean / rstd for the backward pass
        mean._vnb_stats = (sums, float(eps))
        return out, mean, rstd
    if lengths is not None:
        _C.call('gn_sums_unpad', out.transpose(1, 2), x_pl.B, Cout, x_pl.T, gn_groups, lengths, sums)
        _C.call('gn_finalize_masked', sums, x_pl.B, gn_groups, Cout // gn_groups, x_pl.T, lengths, eps, mean, rstd)
    else:
        _C.call('gn_finalize', sums, x_pl.B * gn_groups, float(Cout // gn_groups * x_pl.T), eps, mean, rstd)
    return out, mean, rstd


def conv_wgrad(dy_pl, u_pl, taps, dil, transposed, out=None):
    """-> dw in the parameter layout; ``out``: write it there (e.g. the parameter's gradient slot)."""
    Cout, Cin = dy_pl.C, u_pl.C
    shape = (Cin, Cout, taps) if transposed else (Cout, Cin, taps)
    dw = out if out is not None else torch.empty(shape, dtype=torch.float32, device=dy_pl.data.device)
    assert tuple(dw.shape) == shape
    _C.call('conv_wgrad', dy_pl.data, u_pl.data, dy_pl.P, dy_pl.B, Cout, Cin, dy_pl.T, taps, dil,
            int(transposed), dw)
    return dw


# --------------------------------------------------------------------------------------------
# operand planes
# --------------------------------------------------------------------------------------------
def make_planes(x, P, scale=1.0, act=ACT_NONE, want_id=True, want_act=False, csum=None):
    """``csum``: optional ZEROED (C,) tensor that also receives the per-channel sums of x*scale."""
    B, C, T = x.shape
    pid = Planes.empty(P, B, C, T, x.device) if want_id else None
    pact = Planes.empty(P, B, C, T, x.device) if want_act else None
    _C.call('make_planes', _r(x), B, C, T, P, scale, act, pid.data if pid else None,
            pact.data if pact else None, csum)
    return pid, pact


def concat_planes(xa, xb, P, csum=None, Ca=None):
    """``csum``: optional ZEROED (Ca+Cb,) tensor that also receives the per-channel sums.  ``xa`` None with ``Ca``: the
    first Ca channels are zeros (a gradient that does not exist) - no tensor of zeros is materialised."""
    B, Cb, T = xb.shape
    if xa is not None:
        Ca = xa.shape[1]
    out = Planes.empty(P, B, Ca + Cb, T, xb.device)
    _C.call('concat_planes', _r(xa), Ca, _r(xb), Cb, B, T, P, out.data, csum)
    return out


# --------------------------------------------------------------------------------------------
# forward elementwise / reductions
# --------------------------------------------------------------------------------------------
def gn_stats(h, groups, eps=1e-5):
    B, C, T = h.shape
    mean, rstd = _f32((B, groups), h), _f32((B, groups), h)
    ws = _f32((_C.call('gn_stats_ws_floats', B, groups),), h)
    _C.call('gn_stats', _r(h), B, C, T, groups, eps, ws, mean, rstd)
    return mean, rstd


def enc_out(h, s, mean, rstd, gamma, beta, P, want_y=True, want_ypl=True, want_upl=True):
    B, C, T = h.shape
    y = cl_empty(B, C, T, h.device) if want_y else None
    ypl = Planes.empty(P, B, C, T, h.device) if want_ypl else None
    upl = Planes.empty(P, B, C, T, h.device) if want_upl else None
    st = getattr(mean, '_vnb_stats', None)
    if st is not None:
        _C.call('enc_out_s', _r(h), _r(s), st[0], st[1], gamma, beta, B, C, T, P, _r(y), ypl.data if ypl else None,
                upl.data if upl else None, mean, rstd)
        del mean._vnb_stats
    else:
        _C.call('enc_out', _r(h), _r(s), mean, rstd, gamma, beta, B, C, T, P, _r(y), ypl.data if ypl else None,
                upl.data if upl else None)
    return y, ypl, upl


def dec_gate(a, mean, rstd, gamma, beta, P):
    B, C2, T = a.shape
    C = C2 // 2
    gpl = Planes.empty(P, B, C, T, a.device)
    st = getattr(mean, '_vnb_stats', None)
    if st is not None:
        _C.call('dec_gate_s', _r(a), st[0], st[1], gamma, beta, B, C, T, P, gpl.data, mean, rstd)
        del mean._vnb_stats
    else:
        _C.call('dec_gate', _r(a), mean, rstd, gamma, beta, B, C, T, P, gpl.data)
    return gpl


def dec_out(r, x, S, P, skip_sum=None, first=True, want_x=True, want_xpl=True):
    """r (B,C+S,T), x (B,C,T) -> x_out, x_planes, skip_sum (allocated when None)."""
    B, C, T = x.shape
    x_out = cl_empty(B, C, T, x.device) if want_x else None
    xpl = Planes.empty(P, B, C, T, x.device) if want_xpl else None
    if skip_sum is None and S > 0:
        skip_sum = cl_empty(B, S, T, x.device)
        first = True
    if skip_sum is not None:
        assert is_cl(skip_sum), 'skip accumulator must be channels-last'
    _C.call('dec_out', _r(r), _r(x), B, C, S, T, P, _r(x_out), xpl.data if xpl else None, _r(skip_sum), int(first))
    return x_out, xpl, skip_sum


# --------------------------------------------------------------------------------------------
# backward elementwise / reductions
# --------------------------------------------------------------------------------------------
def _group_sum_scratch(B, groups, like):
    """zeroed fp64 accumulators + ticket slot for the fused group sums (carved from the per-step arena)"""
    n = B * groups * 2 + 1
    buf, off = _ARENA['buf'], _ARENA['off']
    if buf is not None and off >= 0 and buf.device == like.device and off + n <= buf.numel():
        _ARENA['off'] = off + n
        return buf[off:off + n]
    return torch.zeros((n,), dtype=torch.float64, device=like.device)


def gn_bwd_rows(dn, h, mean, rstd, groups, gamma=None):
    """-> R (B,C,3); with ``gamma`` -> (R, S (B,groups,2) | None): the group sums the apply kernel needs, fused into
    this launch when the shape allows (else None: take S from gn_bwd_finalize)."""
    B, C, T = h.shape
    R = _f32((B, C, 3), h)
    if gamma is not None and (C // groups) % 32 == 0:
        S = _f32((B, groups, 2), h)
        _C.call('gn_bwd_rows_s', _r(dn), _r(h), mean, rstd, gamma, B, C, T, groups, R, _group_sum_scratch(B, groups, h), S)
        return R, S
    _C.call('gn_bwd_rows', _r(dn), _r(h), mean, rstd, B, C, T, groups, R)
    return (R, None) if gamma is not None else R


def gate_bwd_rows(dg, a, mean, rstd, gamma, beta, P=3, want_S=False):
    """``P``: operand mode of the step (1 = bf16 planes: the gate may use the approximate functions).
    ``want_S`` -> (R, S (B,2,2) | None) as gn_bwd_rows."""
    B, C2, T = a.shape
    R = _f32((B, C2, 3), a)
    if want_S and (C2 // 2) % 32 == 0:
        S = _f32((B, 2, 2), a)
        _C.call('gate_bwd_rows_s', _r(dg), _r(a), mean, rstd, gamma, beta, B, C2 // 2, T, P, R, _group_sum_scratch(B, 2, a), S)
        return R, S
    _C.call('gate_bwd_rows_p', _r(dg), _r(a), mean, rstd, gamma, beta, B, C2 // 2, T, P, R)
    return (R, None) if want_S else R


def gn_bwd_finalize(R, gamma, rstd, T, groups, want_sums=False, out=None):
    """-> dgamma (C), dbeta (C), S (B,groups,2), dpre_bsum (B,C)|None, dpre_csum (C)|None.
    ``out``: dict of destination tensors (gradient slots) for 'dgamma', 'dbeta', 'csum' and the optional
    second copies 'dbeta2', 'csum2'; values written there are returned as None."""
    B, C, _ = R.shape
    out = out or {}
    dgamma = out.get('dgamma') if out.get('dgamma') is not None else _f32((C,), R)
    dbeta = out.get('dbeta') if out.get('dbeta') is not None else _f32((C,), R)
    S = _f32((B, groups, 2), R)
    bsum = out.get('bsum') if out.get('bsum') is not None else (_f32((B, C), R) if want_sums else None)
    csum = out.get('csum') if out.get('csum') is not None else (_f32((C,), R) if want_sums else None)
    _C.call('gn_bwd_finalize', R, gamma, rstd, B, C, T, groups, dgamma, dbeta, S, bsum, csum, out.get('dbeta2'),
            out.get('csum2'))
    return (None if out.get('dgamma') is not None else dgamma, None if out.get('dbeta') is not None else dbeta, S,
            bsum, None if out.get('csum') is not None else csum)


def gn_bwd_apply(dn, h, mean, rstd, gamma, S, groups, P, want_dn_planes=False):
    B, C, T = h.shape
    dpre = Planes.empty(P, B, C, T, h.device)
    dnp = Planes.empty(P, B, C, T, h.device) if want_dn_planes else None
    _C.call('gn_bwd_apply', _r(dn), _r(h), mean, rstd, gamma, S, B, C, T, groups, P, dpre.data,
            dnp.data if dnp else None)
    return dpre, dnp


def gate_bwd_apply(dg, a, mean, rstd, gamma, beta, S, P):
    B, C2, T = a.shape
    da = Planes.empty(P, B, C2, T, a.device)
    _C.call('gate_bwd_apply', _r(dg), _r(a), mean, rstd, gamma, beta, S, B, C2 // 2, T, P, da.data)
    return da


def conv_dgrad_act(xpl, w, taps, dil, Cout, mask_src, act, scale=1.0, residual=None):
    """conv_fwd whose epilogue also applies an activation backward: scale * conv(x) * act'(mask_src * scale) + residual
    (= conv_fwd + act_bwd_add in one kernel)."""
    out = cl_empty(xpl.B, Cout, xpl.T, xpl.data.device)
    _C.call('conv_fwd_mask', xpl.data, xpl.P, xpl.B, xpl.C, xpl.T, w, taps, dil, Cout, _r(residual), _r(out), _r(mask_src),
            act, scale)
    return out


def act_bwd_add(du, c, add, act, scale=1.0):
    B, C, T = du.shape
    out = cl_empty(B, C, T, du.device)
    _C.call('act_bwd_add', _r(du), _r(c), _r(add), du.numel(), act, scale, out.transpose(1, 2))
    return out


def channel_sum(x, out=None):
    B, C, T = x.shape
    if out is None:
        out = _f32((C,), x)
    _C.call('channel_sum', _r(x), B, C, T, out)
    return out


# --------------------------------------------------------------------------------------------
# speaker conditioning
# --------------------------------------------------------------------------------------------
def embed_fwd(table, idx):
    n, dim = idx.numel(), table.shape[1]
    out = _f32(tuple(idx.shape) + (dim,), table)
    _C.call('embed_fwd', table, idx.contiguous(), n, dim, out)
    return out


def embed_bwd(dout, idx, num_rows, out=None):
    """scatter-add into a zeroed table (``out``: an already zeroed gradient slot)"""
    dim = dout.shape[-1]
    dtable = out if out is not None else torch.zeros((num_rows, dim), dtype=torch.float32, device=dout.device)
    _C.call('embed_bwd', dout.contiguous(), idx.contiguous(), idx.numel(), dim, dtable)
    return dtable


def linear_fwd(y, W, bias, bias2=None):
    B, I = y.shape
    O = W.shape[0]
    p = _f32((B, O), y)
    _C.call('linear_fwd', y, W, bias, bias2, B, I, O, p)
    return p


def linear_bwd(dp, y, W, dy=None, dW_out=None):
    """-> dW (O,I), dy (B,I); accumulates into ``dy`` when given; dW written to ``dW_out`` when given."""
    B, I = y.shape
    O = W.shape[0]
    dW = dW_out if dW_out is not None else _f32((O, I), y)
    acc = dy is not None
    if dy is None:
        dy = _f32((B, I), y)
    _C.call('linear_bwd', dp, y, W, B, I, O, dW, dy, int(acc))
    return dW, dy


_PTR_TABLES = {}


def _ptr_table(rows, device):
    """(L, 4) int64 device table of raw addresses (None -> 0), cached per address tuple."""
    key = tuple(0 if t is None else t.data_ptr() for row in rows for t in row)
    tab = _PTR_TABLES.get(key)
    if tab is None or tab.device != torch.device(device):
        if len(_PTR_TABLES) > 64:
            _PTR_TABLES.clear()
        tab = torch.tensor(key, dtype=torch.int64).view(len(rows), 4).to(device)
        _PTR_TABLES[key] = tab
    return tab


def linear_fwd_batch(y, Ws, b1s, b2s):
    """p (L,B,O): p[l] = y @ Ws[l].T + b1s[l] + b2s[l] for all L layers in one launch (biases may be None)."""
    B, I = y.shape
    L, O = len(Ws), Ws[0].shape[0]
    tab = _ptr_table([(Ws[l], b1s[l], b2s[l], None) for l in range(L)], y.device)
    p = _f32((L, B, O), y)
    _C.call('linear_fwd_batch', y, tab, L, B, I, O, p)
    return p


def linear_bwd_batch(dp, y, Ws, dWs, want_dy=True):
    """dWs[l] (O,I) <- dp[l].T @ y (written; entries may be None);  -> dy (B,I) = sum_l dp[l] @ Ws[l] (or None)."""
    B, I = y.shape
    L, O = len(Ws), Ws[0].shape[0]
    tab = _ptr_table([(Ws[l], None, None, dWs[l]) for l in range(L)], y.device)
    dy = torch.zeros((B, I), dtype=torch.float32, device=y.device) if want_dy else None
    _C.call('linear_bwd_batch', dp, y, tab, L, B, I, O, dy)
    return dy


# --------------------------------------------------------------------------------------------
# vector quantiser
# --------------------------------------------------------------------------------------------
def vq_prep_codebook(E, normalize, inplace_renorm=True):
    K, D = E.shape
    e_hat = torch.empty_like(E)
    # the pitch padding of the planes (D not a multiple of 8) must be finite: zero-fill only then
    e_pl = (torch.zeros if pitch(D) != D else torch.empty)((3, K, pitch(D)), dtype=torch.bfloat16, device=E.device)
    e_sq = _f32((K,), E)
    _C.call('vq_prep_codebook', E, K, D, int(normalize), int(inplace_renorm), e_hat, e_pl, e_sq)
    return e_hat, e_pl, e_sq


def vq_prep_z(z, normalize):
    B, D, T = z.shape
    zn = cl_empty(B, D, T, z.device)
    zpl = Planes.empty(3, B, D, T, z.device)
    z_sq, z_norm = _f32((B * T,), z), _f32((B * T,), z)
    _C.call('vq_prep_z', _r(z), B, D, T, int(normalize), zn.transpose(1, 2), zpl.data, z_sq, z_norm)
    return zn, zpl, z_sq, z_norm


def vq_lookup(zn, zpl, z_sq, e_hat, e_pl, e_sq):
    B, D, T = zn.shape
    K = e_hat.shape[0]
    idx = torch.empty((B * T,), dtype=torch.int64, device=zn.device)
    _C.call('vq_lookup', _r(zn), zpl.data if zpl is not None else None, z_sq, e_hat, e_pl, e_sq, B, D, T, K, idx)
    return idx


def vq_fused(z, normalize, e_hat, e_pl, e_sq, want_q=True, want_stats=True):
    """Frame preparation + lookup + gather in ONE launch (csrc/tc_vq_fused.cuh).
    -> (zn, z_norm, idx, q | None, sums | None, counts | None), or None when the shape is not taken by the fused kernel
    (the caller then runs vq_prep_z + vq_lookup + vq_gather).  zn is z itself when not normalising."""
    B, D, T = z.shape
    K = e_hat.shape[0]
    if D not in (16, 32, 64, 128) or K < 8 or not z.is_cuda:
        return None
    zc = as_cl(z)
    zn = cl_empty(B, D, T, z.device) if normalize else zc
    z_norm = _f32((B * T,), z) if normalize else None
    idx = torch.empty((B * T,), dtype=torch.int64, device=z.device)
    q = cl_empty(B, D, T, z.device) if want_q else None
    sums = counts = None
    if want_stats:
        buf = torch.zeros((K + 2,), dtype=torch.int32, device=z.device)       # one zero-fill for both accumulators
        sums, counts = buf[:2].view(torch.float32), buf[2:]
    r = _C.call('vq_fused', _r(zc), B, D, T, int(normalize), e_hat, e_pl, e_sq, K, _r(zn) if normalize else None, z_norm,
                idx, _r(q), sums, counts)
    if r == 1:
        return None
    return zn, z_norm, idx, q, sums, counts


def vq_gather(idx, e_hat, zn, z, want_q=True, want_stats=True, shape=None):
    """-> q (B,D,T), sums (2,) [sum (q-zn)^2, sum (zn-z)^2], counts (K,) int32.
    ``zn`` None with ``shape`` = (B, T): rows only (want_stats must be False)."""
    K, D = e_hat.shape
    if zn is None:
        assert shape is not None and not want_stats and z is None
        B, T = shape
    else:
        B, D, T = zn.shape
    dev = e_hat.device
    q = cl_empty(B, D, T, dev) if want_q else None
    sums = torch.zeros((2,), dtype=torch.float32, device=dev) if want_stats else None
    counts = torch.zeros((K,), dtype=torch.int32, device=dev) if want_stats else None
    _C.call('vq_gather', idx, e_hat, _r(zn), _r(z), B, D, T, K, _r(q), sums, counts)
    return q, sums, counts


def vq_finalize(sums, counts, N, scale, normalize, eps):
    scal = _f32((3,), sums)
    _C.call('vq_finalize', sums, counts, counts.numel(), N, scale, int(normalize), eps, scal)
    return scal


def vq_bwd(g_out, g_lq, g_le, idx, z, zn, z_norm, e_hat, E, scale, normalize, want_dE=True):
    B, D, T = zn.shape
    K = e_hat.shape[0]
    dz = cl_empty(B, D, T, zn.device)
    dEh = torch.zeros_like(e_hat) if want_dE else None
    _C.call('vq_bwd', _r(g_out), g_lq, g_le, idx, _r(z), _r(zn), z_norm, e_hat, E, B, D, T, K, scale, int(normalize),
            dz.transpose(1, 2), dEh)
    return dz, dEh


def vq_bwd_codebook(dE_hat, E, normalize, out=None):
    K, D = E.shape
    dE = out if out is not None else torch.empty_like(E)
    _C.call('vq_bwd_codebook', dE_hat, E, K, D, int(normalize), dE)
    return dE


def vq_ema_scatter(idx, z, K):
    """-> zsum (K,D) per-code frame sums, counts (K,) int32"""
    B, D, T = z.shape
    zsum = torch.zeros((K, D), dtype=torch.float32, device=z.device)
    counts = torch.zeros((K,), dtype=torch.int32, device=z.device)
    _C.call('vq_ema_scatter', idx, _r(z), B, D, T, K, zsum, counts)
    return zsum, counts


def vq_ema_update(zsum, counts, n_frames, mu, threshold, rand_rows, emb_sum, emb_elem, embeddings):
    """In-place EMA update of the three buffers; -> stats (4,) [entropy, used_curr, usage, diff_emb]"""
    K, D = embeddings.shape
    stats = _f32((4,), zsum)
    _C.call('vq_ema_update', zsum, counts, K, D, n_frames, mu, threshold, rand_rows, emb_sum, emb_elem,
            embeddings, stats)
    return stats


def vq_frame_rows(z, perm, K):
    B, D, T = z.shape
    out = _f32((K, D), z)
    _C.call('vq_frame_rows', _r(z), perm, B, D, T, K, out)
    return out


def jitter_fwd(x, src):
    B, D, T = x.shape
    out = cl_empty(B, D, T, x.device)
    _C.call('jitter_fwd', _r(x), src, B, D, T, out.transpose(1, 2))
    return out


def jitter_sources(T, p, state):
    """Device-side draws of the jitter source table; ``state`` int64[2] = (seed, call counter)."""
    src = torch.empty((T,), dtype=torch.int64, device=state.device)
    _C.call('jitter_sources', T, float(p), state, src)
    return src


def jitter_bwd(dout, src):
    B, D, T = dout.shape
    din = cl_empty(B, D, T, dout.device)
    _C.call('jitter_bwd', _r(dout), src, B, D, T, din.transpose(1, 2))
    return din


# --------------------------------------------------------------------------------------------
# packed variable-length conversion
# --------------------------------------------------------------------------------------------
def mask_rows(lengths, x=None, *planes):
    """Zero the frames t >= lengths[b] of activation ``x`` (logical (B,C,T)) and of up to three Planes, in place."""
    planes = [p for p in planes if p is not None]
    assert len(planes) <= 3 and (x is not None or planes)
    if x is not None:
        B, C, T = x.shape
        assert is_cl(x), 'mask_rows works on channels-last activations'
    else:
        B, C, T = planes[0].B, planes[0].C, planes[0].T
    P = planes[0].P if planes else 1
    assert all((p.B, p.C, p.T, p.P) == (B, C, T, P) for p in planes)
    d = [p.data for p in planes] + [None] * (3 - len(planes))
    _C.call('mask_rows', None if x is None else x.transpose(1, 2), d[0], d[1], d[2], B, C, T, P, lengths)


def csid(idx, lengths=None):
    """idx (B,T) int64 -> (out (B,T) int64 with consecutive repeats collapsed and -1 fill, out_len (B,) int32)."""
    B, T = idx.shape
    out = torch.empty((B, T), dtype=torch.int64, device=idx.device)
    out_len = torch.empty((B,), dtype=torch.int32, device=idx.device)
    _C.call('csid', idx.contiguous(), B, T, lengths, out, out_len)
    return out, out_len


# --------------------------------------------------------------------------------------------
# losses / optimiser
# --------------------------------------------------------------------------------------------
def logloss_fwd(xhat, x, scale=1.0):
    # elementwise + full reduction: any common layout works, channels-last avoids a copy of xhat
    acc = torch.zeros((1,), dtype=torch.float32, device=x.device)
    _C.call('logloss_fwd', _r(xhat), _r(x), x.numel(), scale, acc)
    return acc


def logloss_bwd(xhat, x, g, scale):
    B, C, T = xhat.shape
    dx = cl_empty(B, C, T, xhat.device)
    _C.call('logloss_bwd', _r(xhat), _r(x), g, x.numel(), scale, dx.transpose(1, 2))
    return dx


def sqnorm(g, acc):
    _C.call('sqnorm', g, g.numel(), acc)
    return acc


def adam_step(p, g, m, v, sqn, max_norm, lr, beta1, beta2, eps, state, step_size, gamma):
    _C.call('adam_step', p, g, m, v, p.numel(), sqn, max_norm, lr, beta1, beta2, eps, state, step_size, gamma)


# --------------------------------------------------------------------------------------------
# hierarchical VQ-VAE2 + style tokens (csrc/hier.cu)
# --------------------------------------------------------------------------------------------
def upsample_cat_fwd(srcs, T):
    """Channel concatenation of nearest-neighbour upsampled sources (vqvae2.py:104-114, 130-143):
    srcs = [(B, C_i, T_i) channels-last, T_i <= T] -> (B, sum C_i, T) channels-last."""
    B = srcs[0].shape[0]
    Ct = sum(s.shape[1] for s in srcs)
    out = cl_empty(B, Ct, T, srcs[0].device)
    off = 0
    for s in srcs:
        _, C, Ts = s.shape
        _C.call('upsample_cat_fwd', _r(s), B, C, Ts, _r(out), Ct, off, T)
        off += C
    return out


def upsample_cat_bwd(dout, shapes):
    """-> [d src_i (B, C_i, T_i)] for ``shapes`` = [(C_i, T_i)]"""
    B, Ct, T = dout.shape
    outs, off = [], 0
    for C, Ts in shapes:
        d = cl_empty(B, C, Ts, dout.device)
        _C.call('upsample_cat_bwd', _r(dout), B, Ct, off, T, _r(d), C, Ts)
        outs.append(d)
        off += C
    return outs


def time_mean_fwd(x):
    B, C, T = x.shape
    out = _f32((B, C), x)
    _C.call('time_mean_fwd', _r(x), B, C, T, out)
    return out


def time_mean_bwd(dy, T):
    B, C = dy.shape
    dx = cl_empty(B, C, T, dy.device)
    _C.call('time_mean_bwd', dy, B, C, T, _r(dx))
    return dx


def gst_attn_fwd(q, K, V, H):
    B, F = q.shape
    N = K.shape[0]
    attn, ctx = _f32((B, H, N), q), _f32((B, F), q)
    _C.call('gst_attn_fwd', q, K, V, B, N, H, F // H, attn, ctx)
    return attn, ctx


def gst_attn_bwd(dctx, q, K, V, attn):
    B, H, N = attn.shape
    F = q.shape[1]
    ds, dq, dK, dV = _f32((B, H, N), q), _f32((B, F), q), _f32((N, F), q), _f32((N, F), q)
    _C.call('gst_attn_bwd', dctx, q, K, V, attn, B, N, H, F // H, ds, dq, dK, dV)
    return dq, dK, dV
