"""torch-CPU restatement of every primitive in vae_npvc_b200/prims.py (the C-ABI contract).

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Two uses:
  * tests monkeypatch this module in place of ``vae_npvc_b200.prims`` to check the host-side
    composition (forward sequencing and every hand-written backward in functional.py) against
    the oracle on CPU, before any GPU time is spent;
  * on the GPU box each CUDA primitive is compared with the function of the same name here.
The product never imports it.
"""
import math

import torch
import torch.nn.functional as F

ACT_NONE, ACT_LRELU, ACT_RELU = 0, 1, 2


def pitch(C):
    return (C + 7) // 8 * 8


# layout helpers of vae_npvc_b200.prims: on the CPU contract side layout is immaterial (every
# function here works on logical (B,C,T) tensors), so they are (near-)identities
def as_cl(x):
    return x


def to_nct(x):
    return x.contiguous()


def is_cl(x):
    return True


def cl_empty(B, C, T, device=None, dtype=torch.float32):
    return torch.empty((B, T, C), dtype=dtype).transpose(1, 2)


def cl_zeros(B, C, T, device=None, dtype=torch.float32):
    return torch.zeros((B, T, C), dtype=dtype).transpose(1, 2)


def _split(x, P):
    """fp32 -> P bf16 planes (hi[/mid/lo])."""
    hi = x.to(torch.bfloat16)
    if P == 1:
        return hi.unsqueeze(0)
    r = x - hi.float()
    mid = r.to(torch.bfloat16)
    if P == 2:
        return torch.stack([hi, mid])
    lo = (r - mid.float()).to(torch.bfloat16)
    return torch.stack([hi, mid, lo])


def _join(data):
    return data.float().sum(0) if data.shape[0] > 1 else data[0].float()


class Planes:
    """Same memory layout as the CUDA side: data (P, B, T, Cp) bf16, channels-last."""
    __slots__ = ('data', 'P', 'B', 'C', 'T', 'Cp')

    def __init__(self, data, P, B, C, T, Cp):
        self.data, self.P, self.B, self.C, self.T, self.Cp = data, P, B, C, T, Cp

    @staticmethod
    def of(x, P):
        B, C, T = x.shape
        Cp = pitch(C)
        data = torch.zeros((P, B, T, Cp), dtype=torch.bfloat16)
        data[..., :C] = _split(x.transpose(1, 2), P)
        return Planes(data, P, B, C, T, Cp)

    def value(self):
        return _join(self.data)[..., :self.C].transpose(1, 2)


class ConvWeights:
    __slots__ = ('wf', 'wb', 'w32', 'taps', 'Cout', 'Cin', 'transposed')

    def __init__(self, wf, wb, w32, taps, Cout, Cin, transposed):
        self.wf, self.wb, self.w32, self.taps, self.Cout, self.Cin, self.transposed = wf, wb, w32, taps, Cout, Cin, transposed


def _act(x, act):
    if act == ACT_LRELU:
        return torch.where(x >= 0, x, 0.2 * x)
    if act == ACT_RELU:
        return torch.relu(x)
    return x


def _act_grad(x, act):
    if act == ACT_LRELU:
        return torch.where(x >= 0, torch.ones_like(x), torch.full_like(x, 0.2))
    if act == ACT_RELU:
        return (x > 0).to(x.dtype)
    return torch.ones_like(x)


# ---- weight norm -----------------------------------------------------------------------------
def wn_prepare(v, g, transposed, P, need_wf=True, need_wb=True, need_w32=False):
    v = v.detach()
    d0, d1, k = v.shape
    w = v if g is None else v * (g.detach().view(d0, 1, 1) / v.pow(2).sum(dim=(1, 2), keepdim=True).sqrt())
    # conv weight in (co, ci, tap) form
    wc = w.transpose(0, 1).flip(2) if transposed else w
    co, ci = wc.shape[0], wc.shape[1]
    wf = wb = None
    if need_wf:
        wf = torch.zeros((P, k, co, pitch(ci)), dtype=torch.bfloat16)
        wf[..., :ci] = _split(wc.permute(2, 0, 1).contiguous(), P)             # [tap][co][ci]
    if need_wb:
        wb = torch.zeros((P, k, ci, pitch(co)), dtype=torch.bfloat16)
        wb[..., :co] = _split(wc.flip(2).permute(2, 1, 0).contiguous(), P)     # [k-1-tap][ci][co]
    return ConvWeights(wf, wb, w.clone() if need_w32 else None, k, co, ci, bool(transposed))


def wn_backward(dw, v, g):
    v, g = v.detach(), g.detach()
    d0 = v.shape[0]
    ss = v.pow(2).sum(dim=(1, 2), keepdim=True)
    norm = ss.sqrt()
    dot = (dw * v).sum(dim=(1, 2), keepdim=True)
    gi = g.view(d0, 1, 1)
    return gi / norm * dw - gi * dot / (norm * ss) * v, (dot / norm).view_as(g)


# ---- GEMM-shaped -----------------------------------------------------------------------------
def conv_fwd(x_pl, w_op, taps, dil, Cout, bias=None, bias_bstride=0, residual=None, out=None, accumulate=False,
             gn_groups=0, eps=1e-5, lengths=None, defer_stats=False):
    r = _conv_fwd(x_pl, w_op, taps, dil, Cout, bias, bias_bstride, residual, out, accumulate)
    if not gn_groups:
        return r
    if lengths is None:
        mean, rstd = gn_stats(r, gn_groups, eps)
    else:
        mean, rstd = gn_stats_masked(r, gn_groups, lengths, eps)
    return r, mean, rstd


def gn_stats_masked(h, groups, lengths, eps=1e-5):
    """GroupNorm statistics over the first lengths[b] frames of each sample (vnb_gn_sums_unpad +
    vnb_gn_finalize_masked): what nn.GroupNorm (layers.py:154,201) computes on the lone utterance."""
    B, C, T = h.shape
    mean, rstd = torch.zeros(B, groups), torch.zeros(B, groups)
    for b in range(B):
        n = int(lengths[b])
        if n <= 0:
            continue
        hb = h[b, :, :n].reshape(groups, -1).double()
        m = hb.mean(dim=1)
        var = (hb - m.unsqueeze(1)).pow(2).mean(dim=1)
        mean[b], rstd[b] = m.float(), torch.rsqrt(var + eps).float()
    return mean, rstd


def _conv_fwd(x_pl, w_op, taps, dil, Cout, bias=None, bias_bstride=0, residual=None, out=None, accumulate=False):
    x = x_pl.value()
    Cin = x_pl.C
    w = _join(w_op)[:, :Cout, :Cin].permute(1, 2, 0).contiguous()               # (co, ci, tap)
    r = F.conv1d(x, w, None, padding=(taps - 1) // 2 * dil, dilation=dil)
    if bias is not None:
        b = bias.detach()
        r = r + (b.view(-1, Cout, 1) if bias_bstride else b.view(1, Cout, 1))
    if residual is not None:
        r = r + residual
    if out is not None:
        if accumulate:
            out += r
        else:
            out.copy_(r)
        return out
    return r


def wn_prepare_batch(flat_p, desc, row2layer, P, wbuf, w32buf, tile2layer=None, max_k=None, row_scale=None):
    present = set(row2layer.tolist())              # a segment call passes the rows of its layers only
    for li, L in enumerate(desc.tolist()):
        if li not in present:
            continue
        v_off, g_off, d0, d1, k, tr, wf_off, wb_off, w32_off = L[:9]
        v = flat_p[v_off:v_off + d0 * d1 * k].view(d0, d1, k)
        g = flat_p[g_off:g_off + d0].view(d0, 1, 1) if g_off >= 0 else None
        W = wn_prepare(v, g, tr, P, wf_off >= 0, wb_off >= 0, w32_off >= 0)
        if wf_off >= 0:
            wbuf[wf_off:wf_off + W.wf.numel()].copy_(W.wf.reshape(-1))
        if wb_off >= 0:
            wbuf[wb_off:wb_off + W.wb.numel()].copy_(W.wb.reshape(-1))
        if w32_off >= 0:
            w32buf[w32_off:w32_off + W.w32.numel()].copy_(W.w32.reshape(-1))


def wn_backward_batch(flat_p, flat_g, desc, row2layer):
    present = set(row2layer.tolist())
    for li, L in enumerate(desc.tolist()):
        if li not in present:
            continue
        v_off, g_off, d0, d1, k = L[:5]
        if g_off < 0:
            continue
        n = d0 * d1 * k
        dv, dg = wn_backward(flat_g[v_off:v_off + n].view(d0, d1, k).clone(), flat_p[v_off:v_off + n].view(d0, d1, k),
                             flat_p[g_off:g_off + d0].view(d0, 1, 1))
        flat_g[v_off:v_off + n].copy_(dv.reshape(-1))
        flat_g[g_off:g_off + d0].copy_(dg.reshape(-1))


def conv_wgrad(dy_pl, u_pl, taps, dil, transposed, out=None):
    dw = _conv_wgrad(dy_pl, u_pl, taps, dil, transposed)
    if out is not None:
        out.copy_(dw)
        return out
    return dw


def _conv_wgrad(dy_pl, u_pl, taps, dil, transposed):
    dy, u = dy_pl.value(), u_pl.value()
    pad = (taps - 1) // 2 * dil
    up = F.pad(u, (pad, pad))
    T = dy.shape[2]
    dw = torch.stack([torch.einsum('bot,bit->oi', dy, up[:, :, j * dil:j * dil + T]) for j in range(taps)], dim=2)
    return dw.transpose(0, 1).flip(2).contiguous() if transposed else dw


# ---- planes ----------------------------------------------------------------------------------
def make_planes(x, P, scale=1.0, act=ACT_NONE, want_id=True, want_act=False, csum=None):
    v = x.detach() * scale
    if csum is not None:
        csum.add_(v.sum(dim=(0, 2)))
    return (Planes.of(v, P) if want_id else None, Planes.of(_act(v, act), P) if want_act else None)


def concat_planes(xa, xb, P, csum=None, Ca=None):
    if xa is None:
        xa = torch.zeros((xb.shape[0], Ca, xb.shape[2]), dtype=xb.dtype)
    cat = torch.cat([xa, xb], dim=1)
    if csum is not None:
        csum.add_(cat.detach().sum(dim=(0, 2)))
    return Planes.of(cat, P)


# ---- forward elementwise ---------------------------------------------------------------------
def gn_stats(h, groups, eps=1e-5):
    B = h.shape[0]
    hg = h.reshape(B, groups, -1)
    mean = hg.mean(dim=2)
    var = (hg - mean.unsqueeze(2)).pow(2).mean(dim=2)
    return mean, torch.rsqrt(var + eps)


def _xhat(h, mean, rstd, groups):
    B, C, T = h.shape
    return ((h.reshape(B, groups, -1) - mean.view(B, groups, 1)) * rstd.view(B, groups, 1)).reshape(B, C, T)


def enc_out(h, s, mean, rstd, gamma, beta, P, want_y=True, want_ypl=True, want_upl=True):
    y = _xhat(h, mean, rstd, 1) * gamma.detach().view(1, -1, 1) + beta.detach().view(1, -1, 1)
    if s is not None:
        y = y + s
    return (y if want_y else None, Planes.of(y, P) if want_ypl else None,
            Planes.of(_act(y, ACT_LRELU), P) if want_upl else None)


def dec_gate(a, mean, rstd, gamma, beta, P):
    C = a.shape[1] // 2
    hn = _xhat(a, mean, rstd, 2) * gamma.detach().view(1, -1, 1) + beta.detach().view(1, -1, 1)
    return Planes.of(torch.tanh(hn[:, :C]) * torch.sigmoid(hn[:, C:]), P)


def dec_out(r, x, S, P, skip_sum=None, first=True, want_x=True, want_xpl=True):
    C = x.shape[1]
    xo = r[:, :C] + x
    if S > 0:
        if skip_sum is None:
            skip_sum = r[:, C:].clone()
        elif first:
            skip_sum.copy_(r[:, C:])
        else:
            skip_sum += r[:, C:]
    return (xo if want_x else None, Planes.of(xo, P) if want_xpl else None, skip_sum)


# ---- packed variable-length conversion -------------------------------------------------------
def mask_rows(lengths, x=None, *planes):
    """vnb_mask_rows: zero the frames t >= lengths[b], in place."""
    for b, n in enumerate(lengths.tolist()):
        if x is not None:
            x[b, :, n:] = 0
        for p in planes:
            if p is not None:
                p.data[:, b, n:, :] = 0


def csid(idx, lengths=None):
    """vnb_csid: per-utterance torch.unique_consecutive (bin/extract_bnf.py:60), -1 filled."""
    B, T = idx.shape
    out = torch.full((B, T), -1, dtype=torch.int64)
    out_len = torch.zeros((B,), dtype=torch.int32)
    for b in range(B):
        n = T if lengths is None else int(lengths[b])
        u = torch.unique_consecutive(idx[b, :n])
        out[b, :u.numel()] = u
        out_len[b] = u.numel()
    return out, out_len


# ---- backward elementwise --------------------------------------------------------------------
def _group_sums(R, gamma, groups):
    B, C, _ = R.shape
    gm = gamma.detach().view(1, C)
    cg = C // groups
    return torch.stack([(gm * R[:, :, 0]).view(B, groups, cg).sum(2), (gm * R[:, :, 1]).view(B, groups, cg).sum(2)], dim=2)


def gn_bwd_rows(dn, h, mean, rstd, groups, gamma=None):
    if gamma is not None:
        R = gn_bwd_rows(dn, h, mean, rstd, groups)
        return R, _group_sums(R, gamma, groups)
    return _gn_bwd_rows(dn, h, mean, rstd, groups)


def _gn_bwd_rows(dn, h, mean, rstd, groups):
    xh = _xhat(h, mean, rstd, groups)
    return torch.stack([dn.sum(2), (dn * xh).sum(2), xh.sum(2)], dim=2)


def _gate_dn(dg, a, mean, rstd, gamma, beta):
    C = a.shape[1] // 2
    xh = _xhat(a, mean, rstd, 2)
    hn = xh * gamma.detach().view(1, -1, 1) + beta.detach().view(1, -1, 1)
    th, sg = torch.tanh(hn[:, :C]), torch.sigmoid(hn[:, C:])
    return torch.cat([dg * sg * (1 - th * th), dg * th * sg * (1 - sg)], dim=1), xh


def gate_bwd_rows(dg, a, mean, rstd, gamma, beta, P=3, want_S=False):
    dn, xh = _gate_dn(dg, a, mean, rstd, gamma, beta)
    R = torch.stack([dn.sum(2), (dn * xh).sum(2), xh.sum(2)], dim=2)
    return (R, _group_sums(R, gamma, 2)) if want_S else R


def gn_bwd_finalize(R, gamma, rstd, T, groups, want_sums=False, out=None):
    out = out or {}
    dgamma, dbeta, S, bsum, csum = _gn_bwd_finalize(R, gamma, rstd, T, groups, want_sums or out.get('csum') is not None
                                                    or out.get('csum2') is not None)
    res = {'dgamma': dgamma, 'dbeta': dbeta, 'csum': csum, 'dbeta2': dbeta, 'csum2': csum, 'bsum': bsum}
    for k, dst in out.items():
        if dst is not None:
            dst.copy_(res[k])
    if out.get('bsum') is not None:
        bsum = out['bsum']
    return (None if out.get('dgamma') is not None else dgamma, None if out.get('dbeta') is not None else dbeta, S,
            bsum if want_sums else None, None if out.get('csum') is not None else (csum if want_sums else None))


def _gn_bwd_finalize(R, gamma, rstd, T, groups, want_sums=False):
    B, C, _ = R.shape
    gm = gamma.detach().view(1, C)
    dgamma, dbeta = R[:, :, 1].sum(0), R[:, :, 0].sum(0)
    cg = C // groups
    S = torch.stack([(gm * R[:, :, 0]).view(B, groups, cg).sum(2), (gm * R[:, :, 1]).view(B, groups, cg).sum(2)], dim=2)
    bsum = csum = None
    if want_sums:
        n = cg * T
        S0 = S[:, :, 0].repeat_interleave(cg, dim=1)
        S1 = S[:, :, 1].repeat_interleave(cg, dim=1)
        rs = rstd.repeat_interleave(cg, dim=1)
        bsum = rs * (gm * R[:, :, 0] - T * S0 / n - R[:, :, 2] * S1 / n)
        csum = bsum.sum(0)
    return dgamma, dbeta, S, bsum, csum


def _gn_dpre(dn, xh, rstd, gamma, S, groups):
    B, C, T = dn.shape
    cg = C // groups
    n = cg * T
    S0 = S[:, :, 0].repeat_interleave(cg, dim=1).view(B, C, 1)
    S1 = S[:, :, 1].repeat_interleave(cg, dim=1).view(B, C, 1)
    rs = rstd.repeat_interleave(cg, dim=1).view(B, C, 1)
    return rs * (gamma.detach().view(1, C, 1) * dn - S0 / n - xh * S1 / n)


def gn_bwd_apply(dn, h, mean, rstd, gamma, S, groups, P, want_dn_planes=False):
    dpre = _gn_dpre(dn, _xhat(h, mean, rstd, groups), rstd, gamma, S, groups)
    return Planes.of(dpre, P), (Planes.of(dn, P) if want_dn_planes else None)


def gate_bwd_apply(dg, a, mean, rstd, gamma, beta, S, P):
    dn, xh = _gate_dn(dg, a, mean, rstd, gamma, beta)
    return Planes.of(_gn_dpre(dn, xh, rstd, gamma, S, 2), P)


def conv_dgrad_act(xpl, w, taps, dil, Cout, mask_src, act, scale=1.0, residual=None):
    """conv_fwd + act_bwd_add in one call (vnb_conv_fwd_mask)."""
    return act_bwd_add(conv_fwd(xpl, w, taps, dil, Cout), mask_src, residual, act, scale)


def act_bwd_add(du, c, add, act, scale=1.0):
    out = scale * du * _act_grad(c * scale, act)
    return out + add if add is not None else out


def channel_sum(x, out=None):
    r = x.sum(dim=(0, 2))
    if out is not None:
        out.copy_(r)
        return out
    return r


# ---- speaker conditioning --------------------------------------------------------------------
def embed_fwd(table, idx):
    return table.detach()[idx]


def embed_bwd(dout, idx, num_rows, out=None):
    dim = dout.shape[-1]
    t = out if out is not None else torch.zeros(num_rows, dim)
    return t.index_add_(0, idx.reshape(-1), dout.reshape(-1, dim))


def linear_fwd(y, W, bias, bias2=None):
    p = y @ W.t()
    if bias is not None:
        p = p + bias.detach()
    if bias2 is not None:
        p = p + bias2.detach()
    return p


def linear_bwd(dp, y, W, dy=None, dW_out=None):
    dW = dp.t() @ y
    if dW_out is not None:
        dW_out.copy_(dW)
        dW = dW_out
    d = dp @ W
    if dy is not None:
        dy += d
        return dW, dy
    return dW, d


# ---- vector quantiser ------------------------------------------------------------------------
def linear_fwd_batch(y, Ws, b1s, b2s):
    return torch.stack([linear_fwd(y, Ws[l], b1s[l], b2s[l]) for l in range(len(Ws))])


def linear_bwd_batch(dp, y, Ws, dWs, want_dy=True):
    dy = torch.zeros_like(y) if want_dy else None
    for l in range(len(Ws)):
        dW, d = linear_bwd(dp[l], y, Ws[l])
        if dWs[l] is not None:
            dWs[l].copy_(dW.view_as(dWs[l]))
        if want_dy:
            dy += d
    return dy


def vq_prep_codebook(E, normalize, inplace_renorm=True):
    K, D = E.shape
    if normalize:
        if inplace_renorm:
            E.mul_(1.0 / E.norm(dim=1, keepdim=True))
        e_hat = E / E.norm(dim=1, keepdim=True)
    else:
        e_hat = E.clone()
    e_pl = torch.zeros((3, K, pitch(D)), dtype=torch.bfloat16)
    e_pl[..., :D] = _split(e_hat, 3)
    return e_hat, e_pl, e_hat.pow(2).sum(1)


def _frames(z):
    B, D, T = z.shape
    return z.transpose(1, 2).reshape(B * T, D)


def vq_prep_z(z, normalize):
    z = z.detach()
    zf = _frames(z)
    nrm = zf.norm(dim=1)
    zn = z / z.norm(dim=1, keepdim=True) if normalize else z.clone()
    return zn, Planes.of(zn, 3), _frames(zn).pow(2).sum(1), nrm


def vq_lookup(zn, zpl, z_sq, e_hat, e_pl, e_sq):
    d = (z_sq.view(-1, 1) + e_sq.view(1, -1)) - 2 * (_frames(zn) @ e_hat.t())
    return torch.argmin(d, dim=1)


def vq_fused(z, normalize, e_hat, e_pl, e_sq, want_q=True, want_stats=True):
    """Contract of vnb_vq_fused = vq_prep_z + vq_lookup + vq_gather."""
    zn, zpl, z_sq, z_norm = vq_prep_z(z, normalize)
    idx = vq_lookup(zn, zpl, z_sq, e_hat, e_pl, e_sq)
    q, sums, counts = vq_gather(idx, e_hat, zn, z if normalize else None, want_q, want_stats)
    return zn, (z_norm if normalize else None), idx, q, sums, counts


def vq_gather(idx, e_hat, zn, z, want_q=True, want_stats=True, shape=None):
    if zn is None:
        (B, T), D = shape, e_hat.shape[1]
    else:
        B, D, T = zn.shape
    q = e_hat[idx].view(B, T, D).transpose(1, 2).contiguous()
    sums = counts = None
    if want_stats:
        s1 = (zn - z).pow(2).sum() if z is not None else torch.zeros(())
        sums = torch.stack([(q - zn).pow(2).sum(), s1])
        counts = torch.bincount(idx, minlength=e_hat.shape[0]).to(torch.int32)
    return (q if want_q else None), sums, counts


def vq_finalize(sums, counts, N, scale, normalize, eps):
    p = counts.float() / N
    return torch.stack([sums[0] * scale, (sums[0] + (sums[1] if normalize else 0.0)) * scale,
                        torch.exp(-(p * torch.log(p + eps)).sum())])


def vq_bwd(g_out, g_lq, g_le, idx, z, zn, z_norm, e_hat, E, scale, normalize, want_dE=True):
    B, D, T = zn.shape
    q = e_hat[idx].view(B, T, D).transpose(1, 2)
    glq = (float(g_lq) if g_lq is not None else 0.0) * 2 * scale
    gle = (float(g_le) if g_le is not None else 0.0) * 2 * scale
    go = g_out if g_out is not None else torch.zeros_like(zn)
    dEh = None
    if want_dE:
        dEh = torch.zeros_like(e_hat).index_add_(0, idx, _frames(glq * (q - zn)))
    if not normalize:
        return go + gle * (zn - q), dEh
    dzn = go + gle * ((zn - q) + (zn - z))
    dot = (zn * dzn).sum(1, keepdim=True)
    dz = (dzn - zn * dot) / z_norm.view(B, 1, T) - gle * (zn - z)
    return dz, dEh


def vq_bwd_codebook(dE_hat, E, normalize, out=None):
    if not normalize:
        r = dE_hat.clone()
    else:
        ss = E.pow(2).sum(1, keepdim=True)
        dot = (E * dE_hat).sum(1, keepdim=True)
        r = (dE_hat - E * dot / ss) / ss.sqrt()
    if out is not None:
        out.copy_(r)
        return out
    return r


def vq_ema_scatter(idx, z, K):
    zf = _frames(z.detach())
    return torch.zeros(K, zf.shape[1]).index_add_(0, idx, zf), torch.bincount(idx, minlength=K).to(torch.int32)


def vq_ema_update(zsum, counts, n_frames, mu, threshold, rand_rows, emb_sum, emb_elem, embeddings):
    K, D = embeddings.shape
    c = counts.float()
    emb_sum.mul_(mu).add_((1 - mu) * zsum)
    emb_elem.mul_(mu).add_((1 - mu) * c)
    alive = (emb_elem >= threshold).view(K, 1)
    new = torch.where(alive, emb_sum / emb_elem.view(K, 1), rand_rows)
    diff = (new - embeddings).norm() / math.sqrt(K * D)
    embeddings.copy_(new)
    p = c / n_frames
    return torch.stack([torch.exp(-(p * torch.log(p + 1e-8)).sum()), (c >= threshold).sum().float(),
                        alive.sum().float(), diff])


def vq_frame_rows(z, perm, K):
    zf = _frames(z.detach())
    return zf[perm[:K] % zf.shape[0]].clone()


def jitter_fwd(x, src):
    return x[:, :, src]


def jitter_sources(T, p, state):
    """Contract of vnb_jitter_sources: same DISTRIBUTION as Jitter's host loop (layers_vq.py:360-378, inverted flag
    included); the stream itself is implementation-defined.  state[1] counts calls."""
    gen = torch.Generator().manual_seed(int(state[0]) * 1000003 + int(state[1]))
    u = torch.rand(T, generator=gen)
    sign = torch.randint(0, 2, (T,), generator=gen) * 2 - 1
    t = torch.arange(T)
    s = torch.where(u >= p, t + sign, t)
    if T > 1:
        s[0] = 1 if u[0] >= p else 0
        s[T - 1] = T - 2 if u[T - 1] >= p else T - 1
    else:
        s[0] = 0
    state[1] += 1
    return s.to(torch.int64)


def jitter_bwd(dout, src):
    keep = (src == torch.arange(src.numel())).view(1, 1, -1)
    return torch.where(keep, dout, torch.zeros_like(dout))


# ---- losses / optimiser ----------------------------------------------------------------------
LOG_2PI = math.log(2.0 * math.pi)


def logloss_fwd(xhat, x, scale=1.0):
    return (scale * (0.5 * (LOG_2PI + (xhat - x).pow(2))).sum()).view(1)


def logloss_bwd(xhat, x, g, scale):
    return (float(g) if g is not None else 1.0) * scale * (xhat - x)


def sqnorm(g, acc):
    acc += g.pow(2).sum()
    return acc


def adam_step(p, g, m, v, sqn, max_norm, lr, beta1, beta2, eps, state, step_size, gamma):
    t = float(state[0]) + 1
    s = float(state[1])
    lr_t = lr * (gamma ** math.floor(s / step_size) if step_size > 0 else 1.0)
    coef = min(1.0, max_norm / (math.sqrt(float(sqn[0])) + 1e-6)) if (max_norm > 0 and sqn is not None) else 1.0
    gi = g * coef
    m.mul_(beta1).add_((1 - beta1) * gi)
    v.mul_(beta2).add_((1 - beta2) * gi * gi)
    p.sub_((lr_t / (1 - beta1 ** t)) * m / (v.sqrt() / math.sqrt(1 - beta2 ** t) + eps))
    state[0] += 1
    state[1] += 1


# ---- hierarchical VQ-VAE2 + style tokens (csrc/hier.cu; vqvae2.py:96,104-114,130-143; layers_gst.py:128-147) --------
def _up_index(T, Ts):
    return torch.clamp(torch.arange(T) // (T // Ts), max=Ts - 1)


def upsample_cat_fwd(srcs, T):
    return torch.cat([s[:, :, _up_index(T, s.shape[2])] for s in srcs], dim=1)


def upsample_cat_bwd(dout, shapes):
    B, Ct, T = dout.shape
    outs, off = [], 0
    for C, Ts in shapes:
        d = torch.zeros(B, C, Ts, dtype=dout.dtype)
        d.index_add_(2, _up_index(T, Ts), dout[:, off:off + C])
        outs.append(d)
        off += C
    return outs


def time_mean_fwd(x):
    return x.mean(dim=2)


def time_mean_bwd(dy, T):
    return (dy / T).unsqueeze(2).expand(-1, -1, T).contiguous()


def gst_attn_fwd(q, K, V, H):
    B, Fd = q.shape
    N, dk = K.shape[0], Fd // H
    s = torch.einsum('bhd,nhd->bhn', q.view(B, H, dk), K.view(N, H, dk)) / math.sqrt(dk)
    attn = torch.softmax(s, dim=-1)
    return attn, torch.einsum('bhn,nhd->bhd', attn, V.view(N, H, dk)).reshape(B, Fd)


def gst_attn_bwd(dctx, q, K, V, attn):
    B, H, N = attn.shape
    Fd = q.shape[1]
    dk = Fd // H
    dc, qh, Kh, Vh = dctx.view(B, H, dk), q.view(B, H, dk), K.view(N, H, dk), V.view(N, H, dk)
    da = torch.einsum('bhd,nhd->bhn', dc, Vh)
    ds = attn * (da - (attn * da).sum(-1, keepdim=True))
    dq = torch.einsum('bhn,nhd->bhd', ds, Kh).reshape(B, Fd) / math.sqrt(dk)
    dK = torch.einsum('bhn,bhd->nhd', ds, qh).reshape(N, Fd) / math.sqrt(dk)
    dV = torch.einsum('bhn,bhd->nhd', attn, dc).reshape(N, Fd)
    return dq, dK, dV
