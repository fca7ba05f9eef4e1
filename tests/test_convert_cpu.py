"""Packed variable-length conversion (vae_npvc_b200/convert.py) on CPU.

The REAL ConvertEngine (packing plan, frozen-weight forward sequencing, masking calls, un-packing) runs with the
CUDA primitives replaced by their torch-CPU contract restatement (oracle/prims_ref.py) and is compared, utterance
by utterance, with the oracle's batch-1 Model.infer / encode / decode - i.e. with what the reference's
decoder/basic.py:58-75 and bin/extract_bnf.py:47-67 loops produce.  The `-m gpu` twin (test_gpu_convert.py) then
checks the kernels under it.
"""
import numpy as np
import pytest
import torch

from conftest import rel_err
from oracle import prims_ref, synth
from oracle import vqvae_oracle as O
from vae_npvc_b200.convert import ConvertEngine, plan_batches

TOL = 2e-5
LENGTHS = [20, 13, 7, 20, 1, 0, 16, 2]


def _setup(use_ema=False, embed_norm=True, **kw):
    from vae_npvc_b200.model.vqvae import Model
    arch = synth.small_config(use_ema=use_ema, embed_norm=embed_norm, **kw)
    sd = synth.synthetic_state_dict(arch, 777)
    if use_ema:
        K, D = arch['z_num'], arch['z_dim']
        g = torch.Generator().manual_seed(5)
        sd['quantizer.embeddings'] = torch.randn(K, D, generator=g) * 0.3
        sd['quantizer.emb_init'] = torch.tensor(True)
    m = Model(arch)
    m.load_state_dict({k: v.clone() for k, v in sd.items()})
    m.eval()
    params = O.params_from_state_dict({k: v.clone() for k, v in m.state_dict().items()}, requires_grad=False)
    X = arch['encoder']['in_channels'][0]
    g = torch.Generator().manual_seed(11)
    feats = [torch.randn(n, X, generator=g).numpy() for n in LENGTHS]
    targets = [int(i % arch['y_num']) for i in range(len(LENGTHS))]
    return arch, m, params, feats, targets


def _oracle_infer(params, arch, f, t):
    x = torch.from_numpy(f).t().unsqueeze(0)                   # decoder/basic.py:62-64
    y = torch.tensor([[t]])
    return O.model_infer(params, arch, x, y)[0].t().numpy()    # :73


def _oracle_ids(params, arch, f):
    return O.model_encode(params, arch, torch.from_numpy(f).t().unsqueeze(0)).view(-1).numpy()


def test_plan_batches_covers_every_utterance_once_within_limits():
    rng = np.random.RandomState(0)
    lengths = [int(v) for v in rng.randint(0, 900, size=300)]
    plan = plan_batches(lengths, max_batch=16, max_batch_frames=8192, max_pad_frac=0.2)
    seen = sorted(i for b in plan for i in b)
    assert seen == [i for i, n in enumerate(lengths) if n > 0]
    for b in plan:
        tmax = lengths[b[0]]
        assert all(lengths[i] <= tmax for i in b)                       # longest first
        assert len(b) <= 16
        if len(b) > 1:
            padded = len(b) * tmax
            assert padded <= 8192
            assert padded - sum(lengths[i] for i in b) <= 0.2 * padded
    assert plan_batches([], 4, 100) == [] and plan_batches([0, 0], 4, 100) == []
    assert plan_batches([5, 5, 5], max_batch=2, max_batch_frames=100) == [[0, 1], [2]]
    # a single over-long utterance still gets its own batch
    assert plan_batches([1000], max_batch=4, max_batch_frames=100) == [[0]]


@pytest.mark.parametrize('mode', ['vqnorm', 'vqplain', 'ema', 'two_layer_dilated'])
def test_packed_convert_equals_per_utterance_reference_loop(mode):
    kw = {'vqnorm': dict(), 'vqplain': dict(embed_norm=False), 'ema': dict(use_ema=True),
          'two_layer_dilated': dict(stack_layers=2, dilation=True)}[mode]
    try:
        arch, m, params, feats, targets = _setup(**kw)
    except TypeError:
        pytest.skip('synth.small_config does not take %s' % sorted(kw))
    eng = ConvertEngine(m, precision='fp32', max_batch=4, max_pad_frac=0.9, prims=prims_ref)
    out = eng.convert(feats, targets)
    ids = eng.encode(feats, 'id')
    assert len(out) == len(feats)
    for f, t, o, i in zip(feats, targets, out, ids):
        assert o.shape == f.shape and o.dtype == np.float32
        if f.shape[0] == 0:
            assert i.shape == (0,)
            continue
        assert np.array_equal(i, _oracle_ids(params, arch, f))
        assert rel_err(o, _oracle_infer(params, arch, f, t)) < TOL


def test_packing_does_not_change_results():
    arch, m, params, feats, targets = _setup()
    packed = ConvertEngine(m, precision='fp32', max_batch=8, max_pad_frac=0.95, prims=prims_ref).convert(feats, targets)
    single = ConvertEngine(m, precision='fp32', max_batch=1, prims=prims_ref).convert(feats, targets)
    for a, b in zip(packed, single):
        assert a.shape == b.shape
        if a.size:
            assert rel_err(a, b) < 1e-6


def test_token_kinds_and_decode():
    arch, m, params, feats, targets = _setup()
    eng = ConvertEngine(m, precision='fp32', max_batch=4, max_pad_frac=0.9, prims=prims_ref)
    ids = eng.encode(feats, 'id')
    cs = eng.encode(feats, 'csid')
    tok = eng.encode(feats, 'token')
    for f, i, c, k in zip(feats, ids, cs, tok):
        want = torch.unique_consecutive(torch.from_numpy(i)).numpy()       # extract_bnf.py:60
        assert np.array_equal(c, want)
        assert k.shape == (1, f.shape[0]) and np.array_equal(k.reshape(-1), i)
    dec = eng.decode(ids, targets)
    for f, t, i, d in zip(feats, targets, ids, dec):
        if f.shape[0] == 0:
            assert d.shape == (0, f.shape[1])
            continue
        want = O.model_decode(params, arch, torch.from_numpy(i).view(1, -1), torch.tensor([[t]]))[0].t().numpy()
        assert rel_err(d, want) < TOL


def test_csid_contract():
    idx = torch.tensor([[3, 3, 3, 1, 1, 2, 2, 2], [5, 5, 5, 5, 5, 5, 5, 5], [1, 2, 3, 4, 5, 6, 7, 8]])
    out, n = prims_ref.csid(idx, torch.tensor([8, 3, 0], dtype=torch.int32))
    assert n.tolist() == [3, 1, 0]
    assert out[0].tolist() == [3, 1, 2, -1, -1, -1, -1, -1]
    assert out[1].tolist() == [5] + [-1] * 7 and out[2].tolist() == [-1] * 8


def test_bad_feature_shape_raises():
    arch, m, params, feats, targets = _setup()
    eng = ConvertEngine(m, precision='fp32', prims=prims_ref)
    with pytest.raises(ValueError):
        eng.convert([np.zeros((4, feats[0].shape[1] + 1), dtype=np.float32)], [0])


def test_plan_shards_balances_frames_and_merge_scp_keeps_trial_order(tmp_path):
    from vae_npvc_b200.convert import plan_shards
    from vae_npvc_b200.decoder.basic import merge_scp
    rng = np.random.RandomState(1)
    lengths = [int(v) for v in rng.randint(50, 900, size=101)]
    for world in (1, 2, 8):
        shards = plan_shards(lengths, world)
        assert len(shards) == world and sorted(i for s in shards for i in s) == list(range(101))
        loads = [sum(lengths[i] for i in s) for s in shards]
        assert max(loads) - min(loads) <= max(lengths)                  # greedy longest-first bound
        assert all(s == sorted(s) for s in shards)
    assert plan_shards([], 2) == [[], []]
    keys = ['u%d' % i for i in range(6)]
    (tmp_path / 'a.scp').write_text(''.join('%s A:%d\n' % (k, i) for i, k in enumerate(keys) if i % 2 == 0))
    (tmp_path / 'b.scp').write_text(''.join('%s B:%d\n' % (k, i) for i, k in enumerate(keys) if i % 2 == 1))
    merge_scp([str(tmp_path / 'a.scp'), str(tmp_path / 'b.scp')], str(tmp_path / 'm.scp'), keys)
    got = (tmp_path / 'm.scp').read_text().splitlines()
    assert [l.split()[0] for l in got] == keys and got[1] == 'u1 B:1' and got[4] == 'u4 A:4'
