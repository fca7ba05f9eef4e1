"""-m gpu: the batched weight-norm passes over the trainer's flat buffers (vnb_wn_prepare_batch,
vnb_wn_backward_batch) in both generations (option VNB_OPT_WN_V2: full-tile fast path of the tile kernel, warp-per-row
backward) on the real WeightBank of the small and of the VCC2020-shape model (55 layers: k = 1 and 3, Conv1d and
ConvTranspose1d, 80 / 128 / 512 / 640 / 1024-wide matrices, i.e. full and ragged tiles, register and two-pass rows):
  * operand planes: second generation bit-identical to the first (same arithmetic, different data movement), both
    equal to the contract restatement (oracle/prims_ref.py) on the small model
  * (dv, dg): within fp32 round-off of the first generation and of the contract (summation order differs)."""
import pytest
import torch

from conftest import rel_err
from oracle import prims_ref as R
from oracle import synth

pytestmark = pytest.mark.gpu


@pytest.fixture()
def wn_opt():
    from vae_npvc_b200 import _C
    _C.require_cuda()
    old = _C.get_option(_C.OPT_WN_V2)
    yield lambda v: _C.set_option(_C.OPT_WN_V2, v)
    _C.set_option(_C.OPT_WN_V2, old)


def _trainer(arch, precision):
    from vae_npvc_b200.trainer.basic import Trainer
    cfg = dict(arch, model_type='vae_npvc_b200.model.vqvae', precision=precision, cuda_graph=False)
    tr = Trainer(cfg)
    tr.model.load_state_dict(synth.synthetic_state_dict(arch, 777))
    return tr


@pytest.mark.parametrize('shape', ['small', 'vcc2020'])
@pytest.mark.parametrize('precision', ['bf16', 'fp32'])
def test_prepare_and_backward_generations_agree(wn_opt, shape, precision):
    import vae_npvc_b200.functional as Fn
    old = Fn.get_precision()
    try:
        arch = synth.small_config() if shape == 'small' else synth.arch_config()
        tr = _trainer(arch, precision)
        bank = tr.bank
        g = torch.Generator(device='cuda').manual_seed(3)
        dw = torch.randn(tr._flat_g.shape, device='cuda', generator=g)
        planes, w32, grads = {}, {}, {}
        for v2 in (0, 1):
            wn_opt(v2)
            bank.wbuf.zero_()
            bank.w32buf.zero_()
            bank.prepare()
            planes[v2], w32[v2] = bank.wbuf.clone(), bank.w32buf.clone()
            tr._flat_g.copy_(dw)
            bank.finish_backward()
            grads[v2] = tr._flat_g.clone()
        assert torch.equal(planes[0], planes[1]) and torch.equal(w32[0], w32[1])
        assert rel_err(grads[1], grads[0]) < 1e-6
        # per segment (the trainer's side-stream schedule) == whole
        wn_opt(1)
        bank.wbuf.zero_()
        for sidx in reversed(range(len(bank.segments))):          # encoder | decoder | conditioning convs
            bank.prepare(sidx)
        assert torch.equal(bank.wbuf, planes[1])
        tr._flat_g.copy_(dw)
        for sidx in reversed(range(len(bank.segments))):
            bank.finish_backward(sidx)
        assert torch.equal(tr._flat_g, grads[1])
        if shape == 'small':
            # the contract restatement on the same flat buffers
            P = bank.P
            flat_p, flat_g = tr._flat_p.cpu(), dw.cpu().clone()
            wbuf, w32buf = torch.zeros_like(planes[1], device='cpu'), torch.zeros_like(w32[1], device='cpu')
            R.wn_prepare_batch(flat_p, bank.desc.cpu(), bank.row2layer.cpu(), P, wbuf, w32buf)
            assert rel_err(planes[1].float(), wbuf.float()) < 1e-3          # bf16 planes: a last-bit scale difference
            assert rel_err(w32[1], w32buf) < 1e-6
            R.wn_backward_batch(flat_p, flat_g, bank.desc.cpu(), bank.row2layer.cpu())
            assert rel_err(grads[1], flat_g) < 1e-5
    finally:
        Fn.set_precision(old)
