"""Pin oracle/vqvae2_oracle.py against the reference's own vqvae2.Model and write tests/golden/vqvae2_small.npz.

TEST INFRASTRUCTURE.  Runs ONLY in the build container (needs /root/reference):

    python -m oracle.make_golden_vqvae2

Loads oracle.synth weights into the imported reference ``vae_npvc.model.vqvae2.Model`` (through nn.Module.load_state_dict:
the reference's own override reads an undefined attribute, vqvae2.py:160), runs forward + backward, runs the oracle on
the same inputs, asserts they agree (ids exactly, tensors to fp32 round-off) and stores the REFERENCE's outputs.
"""
import os
import sys
import warnings

import numpy as np
import torch

warnings.filterwarnings('ignore')
sys.path.insert(0, '/root/reference')
from vae_npvc.model import vqvae2 as ref_vqvae2           # noqa: E402

from oracle import synth                                  # noqa: E402
from oracle import vqvae2_oracle as O2                    # noqa: E402
from oracle import vqvae_oracle as O                      # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), 'tests', 'golden')
B, T = 3, 128


def rel(a, b):
    a, b = a.detach().double(), b.detach().double()
    return float((a - b).norm() / (b.norm() + 1e-30))


def main():
    arch = synth.vqvae2_small_config()
    sd = synth.vqvae2_state_dict(arch, 777)
    x, y = synth.vqvae2_batch(arch, B, T, 777)
    m = ref_vqvae2.Model(arch)
    assert list(m.state_dict().keys()) == list(sd.keys()), 'state_dict keys / order differ from the reference'
    for k, v in m.state_dict().items():
        assert tuple(v.shape) == tuple(sd[k].shape), k
    torch.nn.Module.load_state_dict(m, {k: v.clone() for k, v in sd.items()})
    m.train()
    xhat, loss, losses = m((x, y))
    loss.backward()
    ref_grads = {k: p.grad.clone() for k, p in m.named_parameters() if p.grad is not None}

    params = O.params_from_state_dict(sd)
    oxhat, oloss, olosses, ids = O2.model_forward(params, arch, x, y)
    oloss.backward()
    assert rel(oxhat, xhat) < 2e-5, rel(oxhat, xhat)
    for k, v in losses.items():
        assert abs(olosses[k] - v) <= 2e-5 * max(1.0, abs(v)), (k, olosses[k], v)
    worst = 0.0
    gscale = max(float(g.norm()) for g in ref_grads.values())
    for k, g in ref_grads.items():
        assert params[k].grad is not None, k
        if float(g.norm()) < 1e-6 * gscale:
            # mathematically zero (the key bias shifts every token's score by the same amount, softmax cancels it):
            # rounding noise on both sides
            assert float(params[k].grad.norm()) < 1e-6 * gscale, k
            continue
        worst = max(worst, rel(params[k].grad, g))
    assert worst < 2e-4, worst
    assert set(k for k, p in params.items() if p.grad is not None) == set(ref_grads)
    # float64 run of the oracle: how far the reference's own fp32 arithmetic is from exact
    p64 = O.params_from_state_dict(sd, dtype=torch.float64)
    xh64, l64, _, ids64 = O2.model_forward(p64, arch, x.double(), y)
    print('oracle vs reference: xhat %.2e, worst grad %.2e; reference fp32 vs oracle fp64: xhat %.2e'
          % (rel(oxhat, xhat), worst, rel(xhat, xh64)))
    out = {'x': x.numpy(), 'y': y.numpy(), 'xhat': xhat.detach().numpy(), 'loss': np.float64(float(loss))}
    for k, v in losses.items():
        out['loss/' + k] = np.float64(v)
    for k, g in ref_grads.items():
        out['grad/' + k] = g.numpy()
    for i, idx in enumerate(ids):
        out['ids/%d' % i] = idx.numpy()
    os.makedirs(OUT, exist_ok=True)
    np.savez_compressed(os.path.join(OUT, 'vqvae2_small.npz'), **out)
    print('wrote tests/golden/vqvae2_small.npz: %d gradient tensors, losses %s' % (len(ref_grads), sorted(losses)))


if __name__ == '__main__':
    main()
