"""-m gpu: the trainer's side-stream schedule of the parameter-sized weight-norm passes (bank.py segments:
decoder operand planes materialised beside the encoder's forward, decoder weight-norm backward beside the encoder's
backward) takes the same steps as the serial schedule - eager and captured into the CUDA graph, plain and EMA
quantiser (whose quantised output carries no gradient, so the early decoder pass never triggers), against the
oracle after three steps."""
import pytest
import torch

from conftest import rel_err
from oracle import synth
from oracle import vqvae_oracle as O

pytestmark = pytest.mark.gpu


def _run(arch, overlap, graph, steps=4):
    from vae_npvc_b200.trainer.basic import Trainer
    cfg = dict(arch, model_type='vae_npvc_b200.model.vqvae', precision='fp32', cuda_graph=graph, graph_warmup=1,
               learning_rate=1e-3, max_grad_norm=10, bank_overlap=overlap)
    tr = Trainer(cfg)
    sd = synth.synthetic_state_dict(arch, 777)
    if arch.get('use_ema'):
        K, D = arch['z_num'], arch['z_dim']
        g = torch.Generator().manual_seed(5)
        rows = torch.randn(K, D, generator=g) * 0.3
        sd.update({'quantizer.embeddings': rows, 'quantizer.emb_sum': rows * 3.0,
                   'quantizer.emb_elem': torch.full((K,), 3.0), 'quantizer.emb_init': torch.tensor(True)})
    tr.model.load_state_dict(sd)
    log = []
    for s in range(steps):
        x, y = synth.synthetic_batch(arch, 4, 24, 100 + s)
        _, d = tr.train_step([x, y], s)
        log.append(d['Total'])
    assert tr.last_step_on_graph == graph
    return log, {n: p.detach().clone() for n, p in tr.model.named_parameters()}, sd


@pytest.mark.parametrize('graph', [False, True])
@pytest.mark.parametrize('use_ema', [False, True])
def test_overlapped_schedule_equals_serial(graph, use_ema):
    import vae_npvc_b200.functional as Fn
    old = Fn.get_precision()
    try:
        arch = synth.small_config(C=32, S=16, use_ema=use_ema)
        la, pa, _ = _run(arch, True, graph)
        lb, pb, _ = _run(arch, False, graph)
        for a, b in zip(la, lb):
            assert abs(a - b) <= 1e-6 * max(1.0, abs(b))
        assert max(rel_err(pa[n], pb[n]) for n in pa) < 1e-6
    finally:
        Fn.set_precision(old)


def test_overlapped_schedule_against_oracle():
    import vae_npvc_b200.functional as Fn
    old = Fn.get_precision()
    try:
        arch = synth.small_config(C=32, S=16)
        log, params, sd = _run(arch, True, True, steps=3)
        po = O.params_from_state_dict(sd)
        opt = O.AdamState(O.trainable_names(po), lr=1e-3, step_size=0)
        for s in range(3):
            x, y = synth.synthetic_batch(arch, 4, 24, 100 + s)
            ol, _, _ = O.train_step(po, arch, opt, x, y, max_grad_norm=10)
            assert abs(log[s] - ol['Total']) <= 1e-4 * max(1.0, abs(ol['Total']))
        worst = max(rel_err(params[n], po[n]) for n in params)
        assert worst < 2e-4, worst
    finally:
        Fn.set_precision(old)
