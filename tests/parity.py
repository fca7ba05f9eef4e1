"""Per-tensor parity of one GEMM operand mode of the CUDA path against the float64 oracle, at the VCC2020 shape.

TEST INFRASTRUCTURE (imports oracle/).  Used by tests/test_gpu_parity_modes.py, which asserts the stated tolerances
and writes the report (committed as profiles/r02_parity_modes.json).

What is compared, on identical synthetic weights and batch (vae_npvc_b200/synth.py, seed 777):
  * training forward: xhat, every loss scalar;  backward: every parameter gradient (norm-wise relative error per
    tensor, and over all tensors concatenated);
  * conversion: Model.encode ids and Model.infer features in eval mode.
The yardstick is the oracle evaluated in FLOAT64.  The reference's own arithmetic is float32, so the report also
carries the float32 oracle against the float64 one ("ref32"): that is the distance between two legitimate
evaluations of the reference (summation order, LeakyReLU pre-activations within round-off of 0 flipping their slope -
oracle/make_golden.py) and the floor below which a comparison says nothing.
"""
import numpy as np
import torch

from oracle import synth
from oracle import vqvae_oracle as O

CASES = {
    'vcc2020_vqnorm': dict(embed_norm=True),
    'vcc2020_vqplain': dict(embed_norm=False),
    'vcc2020_ema': dict(use_ema=True),
}
B, T = 16, 128


def rel(a, b):
    a, b = torch.as_tensor(a).detach().double().cpu(), torch.as_tensor(b).detach().double().cpu()
    return float((a - b).norm() / (b.norm() + 1e-300))


def max_rel(a, b):
    """max |a-b| / max |b|"""
    a, b = torch.as_tensor(a).detach().double().cpu(), torch.as_tensor(b).detach().double().cpu()
    return float((a - b).abs().max() / (b.abs().max() + 1e-300))


def ema_preset(arch, sd, x):
    """Initialised EMA buffers with every code alive (no random restart is drawn): rows of the encoder output."""
    p = O.params_from_state_dict(sd, requires_grad=False)
    z = O.encoder_forward(p, arch['encoder'], x)
    zf = z.transpose(1, 2).reshape(-1, z.shape[1])
    K = arch['z_num']
    rows = zf[torch.arange(K) % zf.shape[0]].clone() + 0.01 * torch.arange(K).view(K, 1) / K
    sd = dict(sd)
    sd['quantizer.emb_init'] = torch.tensor(True)
    sd['quantizer.embeddings'] = rows
    sd['quantizer.emb_sum'] = rows * 3.0
    sd['quantizer.emb_elem'] = torch.full((K,), 3.0)
    return sd


def case_inputs(name):
    arch = synth.arch_config(**CASES[name])
    sd = synth.synthetic_state_dict(arch, 777)
    x, y = synth.synthetic_batch(arch, B, T, 777)
    if arch.get('use_ema'):
        sd = ema_preset(arch, sd, x)
    return arch, sd, x, y


def oracle_eval(arch, sd, x, y, dtype):
    """-> dict(xhat, losses, grads, idx, infer, z, gap) of the oracle in ``dtype``."""
    params = O.params_from_state_dict(sd, dtype=dtype)
    xd = x.to(dtype)
    xhat, loss, losses, idx = O.model_forward(params, arch, xd, y, True)
    loss.backward()
    grads = {n: params[n].grad.detach().clone() for n in O.trainable_names(params) if params[n].grad is not None}
    pe = O.params_from_state_dict(sd, dtype=dtype, requires_grad=False)          # fresh copy: eval-mode conversion
    with torch.no_grad():
        ids = O.model_encode(pe, arch, xd)
        inf = O.model_infer(pe, arch, xd, y)
        # top-2 distance gap of every frame (in the lookup's own space) - where ids must agree bit-exactly
        z = O.encoder_forward(pe, arch['encoder'], xd)
        zf = z.transpose(1, 2).reshape(-1, z.shape[1])
        emb = pe['quantizer.embeddings']
        if not arch.get('use_ema') and arch.get('embed_norm', True):
            zf = zf / zf.norm(dim=1, keepdim=True)
            emb = emb / emb.norm(dim=1, keepdim=True)
        d = O.vq_distances(zf, emb)
        top2 = d.topk(2, dim=1, largest=False).values
        gap = (top2[:, 1] - top2[:, 0])
    return dict(xhat=xhat.detach(), losses={k: float(v) for k, v in losses.items()}, grads=grads, idx=idx.view(-1),
                ids=ids.view(-1), infer=inf, gap=gap)


def cuda_eval(arch, sd, x, y, mode):
    """The drop-in Model on the GPU in operand mode ``mode`` ('bf16' | 'bf16x2' | 'fp32')."""
    import vae_npvc_b200.functional as Fn
    from vae_npvc_b200.model.vqvae import Model
    old = Fn.get_precision()
    Fn.set_precision(mode)
    try:
        m = Model(arch)
        m.load_state_dict({k: v.clone() for k, v in sd.items()})
        m = m.cuda().train()
        xhat, loss, losses = m([x.cuda(), y.cuda()])
        loss.backward()
        grads = {n: p.grad.detach().float().cpu() for n, p in m.named_parameters() if p.grad is not None}
        out = dict(xhat=xhat.detach().float().cpu(), losses={k: float(v) for k, v in losses.items()}, grads=grads)
        m2 = Model(arch)
        m2.load_state_dict({k: v.clone() for k, v in sd.items()})
        m2 = m2.cuda().eval()
        with torch.no_grad():
            out['ids'] = m2.encode(x.cuda()).view(-1).cpu()
            out['infer'] = m2.infer((x.cuda(), y.cuda())).float().cpu()
        return out
    finally:
        Fn.set_precision(old)


def compare(got, want):
    """-> JSON-able report of ``got`` (a cuda_eval / oracle_eval dict) against the float64 oracle ``want``."""
    rep = {}
    rep['losses_rel'] = {k: abs(got['losses'][k] - v) / max(abs(v), 1e-30) for k, v in want['losses'].items()
                         if k in got['losses']}
    rep['xhat_rel'], rep['xhat_max_rel'] = rel(got['xhat'], want['xhat']), max_rel(got['xhat'], want['xhat'])
    per = {n: rel(got['grads'][n], g) for n, g in want['grads'].items() if n in got['grads']}
    missing = [n for n in want['grads'] if n not in got['grads'] and float(want['grads'][n].abs().max()) > 0]
    vals = sorted(per.values())
    num = sum(float((got['grads'][n].double().cpu() - g.double()).pow(2).sum()) for n, g in want['grads'].items()
              if n in got['grads'])
    den = sum(float(g.double().pow(2).sum()) for g in want['grads'].values())
    rep['grad_tensors'] = len(per)
    rep['grad_missing'] = missing
    rep['grad_rel_max'] = vals[-1]
    rep['grad_rel_median'] = vals[len(vals) // 2]
    rep['grad_rel_p90'] = vals[int(0.9 * (len(vals) - 1))]
    rep['grad_rel_global'] = float(np.sqrt(num / den))
    rep['grad_rel_worst5'] = sorted(per.items(), key=lambda kv: -kv[1])[:5]
    rep['grad_rel_over_1e-3'] = sum(1 for v in vals if v > 1e-3)
    ids_ne = (got['ids'].cpu() != want['ids'].cpu())
    rep['ids_mismatch_frac'] = float(ids_ne.float().mean())
    clear = want['gap'].cpu() > 1e-5
    rep['ids_mismatch_where_gap_gt_1e-5'] = int((ids_ne & clear).sum())
    rep['frames_gap_gt_1e-5_frac'] = float(clear.float().mean())
    same = ~ids_ne
    rep['infer_rel'] = rel(got['infer'], want['infer'])
    # converted features restricted to utterances whose every id agrees (a flipped id changes a frame wholesale:
    # that is the index criterion above, not feature precision)
    Bn = got['infer'].shape[0]
    ok_utt = same.view(Bn, -1).all(dim=1)
    rep['infer_utts_all_ids_equal'] = int(ok_utt.sum())
    if bool(ok_utt.any()):
        rep['infer_rel_ids_equal'] = rel(got['infer'][ok_utt], want['infer'][ok_utt])
        rep['infer_max_rel_ids_equal'] = max_rel(got['infer'][ok_utt], want['infer'][ok_utt])
    return rep


def perturbed_sensitivity(arch, sd, x, y, w64, eps, seed=1):
    """Conditioning of the gradients: per-tensor relative change of the EXACT (float64) gradients when every parameter
    is perturbed by a relative ``eps`` (Gaussian).  The model has 21 LeakyReLU / 2 ReLU layers over 2048 x 512
    activations: a pre-activation that crosses 0 flips a slope, and one flip moves an upstream weight gradient by
    ~1e-3 norm-wise.  An implementation whose forward values carry a relative error eps cannot be closer to the
    float64 gradients than this, whatever its backward arithmetic."""
    g = torch.Generator().manual_seed(seed)
    sd2 = {k: (v * (1 + eps * torch.randn(v.shape, generator=g)) if v.is_floating_point() and v.dim() > 0 else v)
           for k, v in sd.items()}
    w2 = oracle_eval(arch, sd2, x, y, torch.float64)
    return {n: rel(w2['grads'][n], gr) for n, gr in w64['grads'].items()}, compare(w2, w64)
