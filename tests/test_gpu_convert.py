"""-m gpu: packed variable-length conversion / token extraction on the B200.

  * the new primitives (vnb_mask_rows, vnb_gn_sums_unpad + vnb_gn_finalize_masked behind conv_fwd(lengths=...),
    vnb_csid) against the contract restatement of the same name in oracle/prims_ref.py - csid bit-exact;
  * the ConvertEngine against the oracle's batch-1 Model.infer / encode / decode per utterance (what the
    reference's decoder/basic.py:58-75 and bin/extract_bnf.py:47-67 loops produce): ids bit-exact, features
    <= 2e-5 in precision 'fp32'; 'bf16' (bf16 operands through 21 chained GEMMs) only sanity-bounded at 5e-2;
  * size-independent property at the VCC2020 shape: packing utterances together does not change any of them.
"""
import numpy as np
import pytest
import torch

from conftest import rel_err
from oracle import prims_ref as R
from oracle import synth
from oracle import vqvae_oracle as O

pytestmark = pytest.mark.gpu
TOL = 2e-5
LENGTHS = [20, 13, 7, 20, 5, 0, 16, 3, 33, 31]


@pytest.fixture(scope='module')
def Pm():
    from vae_npvc_b200 import _C, prims
    _C.require_cuda()
    return prims


@pytest.fixture(autouse=True)
def _restore_precision():
    import vae_npvc_b200.functional as Fn
    old = Fn.get_precision()
    yield
    Fn.set_precision(old)


def g(*shape, seed=0):
    return torch.randn(*shape, generator=torch.Generator().manual_seed(seed))


def planes_val(pl):
    d = pl.data.float().cpu()
    return (d.sum(0) if pl.P == 3 else d[0])[..., :pl.C].transpose(1, 2)


@pytest.mark.parametrize('P', [1, 3])
@pytest.mark.parametrize('B,C,T', [(3, 16, 20), (2, 13, 9), (4, 512, 37), (1, 80, 1)])
def test_mask_rows(Pm, P, B, C, T):
    lens = torch.tensor([(7 * b + 3) % (T + 1) for b in range(B)], dtype=torch.int32)
    lens[0] = T                                                       # one full-length sample: nothing touched
    x = g(B, C, T, seed=1)
    xc = Pm.as_cl(x.cuda())
    pa, pb = Pm.make_planes(xc, P, 1.0, Pm.ACT_LRELU, True, True)
    before_a, before_b = pa.data.clone(), pb.data.clone()
    Pm.mask_rows(lens.cuda(), xc, pa, pb)
    want = x.clone()
    for b in range(B):
        want[b, :, int(lens[b]):] = 0
    assert torch.equal(xc.cpu(), want)
    for b in range(B):                        # planes (P, B, T, Cp): valid frames untouched, the rest exactly zero
        n = int(lens[b])
        for pl, before in ((pa, before_a), (pb, before_b)):
            assert torch.equal(pl.data[:, b, :n], before[:, b, :n])
            assert not pl.data[:, b, n:].float().abs().any()
    wa = R.make_planes(want, P, 1.0, R.ACT_LRELU, True, False)[0]
    assert rel_err(planes_val(pa), planes_val(wa)) < 1e-6
    # planes only
    pc = Pm.make_planes(Pm.as_cl(x.cuda()), P)[0]
    Pm.mask_rows(lens.cuda(), None, pc)
    assert torch.equal(pc.data, pa.data)


@pytest.mark.parametrize('P', [1, 3])
@pytest.mark.parametrize('B,Cin,Cout,T,k,groups', [(3, 16, 16, 20, 3, 1), (4, 24, 32, 37, 3, 2), (2, 512, 1024, 50, 3, 2),
                                                  (5, 80, 64, 9, 1, 1)])
def test_masked_groupnorm_statistics(Pm, P, B, Cin, Cout, T, k, groups):
    lens = torch.tensor([T if b == 0 else (0 if b == 1 else (5 * b + 2) % T + 1) for b in range(B)], dtype=torch.int32)
    x = g(B, Cin, T, seed=2)
    for b in range(B):
        x[b, :, int(lens[b]):] = 0
    w = R.wn_prepare(g(Cout, Cin, k, seed=3) * 0.2, None, False, P)
    bias = g(Cout, seed=4)
    rx = R.Planes.of(x, P)
    ro, rmean, rrstd = R.conv_fwd(rx, w.wf, k, 1, Cout, bias, 0, gn_groups=groups, lengths=lens)
    cx = Pm.Planes(rx.data.cuda(), rx.P, rx.B, rx.C, rx.T, rx.Cp)
    co, cmean, crstd = Pm.conv_fwd(cx, w.wf.cuda(), k, 1, Cout, bias.cuda(), 0, gn_groups=groups, lengths=lens.cuda())
    assert rel_err(co, ro) < 1e-5
    assert torch.isfinite(cmean).all() and torch.isfinite(crstd).all()
    assert (cmean.cpu() - rmean).abs().max() < 1e-5
    assert rel_err(crstd, rrstd) < 1e-5
    # an empty slot keeps mean = rstd = 0
    assert float(cmean[1].abs().max()) == 0.0 and float(crstd[1].abs().max()) == 0.0


@pytest.mark.parametrize('B,T,K', [(1, 1, 4), (3, 8, 3), (5, 700, 2), (4, 1000, 512), (2, 256, 1)])
def test_csid_bit_exact(Pm, B, T, K):
    gen = torch.Generator().manual_seed(B * 1000 + T)
    idx = torch.randint(0, K, (B, T), generator=gen)
    lens = torch.tensor([T if b == 0 else int(torch.randint(0, T + 1, (1,), generator=gen)) for b in range(B)],
                        dtype=torch.int32)
    for L in (lens, None):
        ro, rn = R.csid(idx, L)
        co, cn = Pm.csid(idx.cuda(), None if L is None else L.cuda())
        assert torch.equal(cn.cpu(), rn) and torch.equal(co.cpu(), ro)


def _setup(arch, seed=11, lengths=LENGTHS, use_ema=False):
    from vae_npvc_b200.model.vqvae import Model
    sd = synth.synthetic_state_dict(arch, 777)
    if use_ema:
        gen = torch.Generator().manual_seed(5)
        sd['quantizer.embeddings'] = torch.randn(arch['z_num'], arch['z_dim'], generator=gen) * 0.3
        sd['quantizer.emb_init'] = torch.tensor(True)
    m = Model(arch)
    m.load_state_dict({k: v.clone() for k, v in sd.items()})
    m = m.cuda().eval()
    params = O.params_from_state_dict({k: v.detach().cpu().clone() for k, v in m.state_dict().items()},
                                      requires_grad=False)
    X = arch['encoder']['in_channels'][0]
    gen = torch.Generator().manual_seed(seed)
    feats = [torch.randn(n, X, generator=gen).numpy() for n in lengths]
    targets = [int(i % arch['y_num']) for i in range(len(lengths))]
    return m, params, feats, targets


def _oracle_infer(params, arch, f, t):
    x = torch.from_numpy(f).t().unsqueeze(0)
    return O.model_infer(params, arch, x, torch.tensor([[t]]))[0].t().numpy()


def _oracle_ids(params, arch, f):
    return O.model_encode(params, arch, torch.from_numpy(f).t().unsqueeze(0)).view(-1).numpy()


@pytest.mark.parametrize('mode', ['vqnorm', 'vqplain', 'ema', 'two_layer_dilated'])
def test_engine_fp32_matches_per_utterance_oracle(Pm, mode):
    from vae_npvc_b200.convert import ConvertEngine
    kw = {'vqnorm': dict(), 'vqplain': dict(embed_norm=False), 'ema': dict(use_ema=True),
          'two_layer_dilated': dict(stack_layers=2, dilation=True)}[mode]
    arch = synth.small_config(**kw)
    m, params, feats, targets = _setup(arch, use_ema=kw.get('use_ema', False))
    eng = ConvertEngine(m, precision='fp32', max_batch=4, max_pad_frac=0.9)
    out = eng.convert(feats, targets)
    ids = eng.encode(feats, 'id')
    cs = eng.encode(feats, 'csid')
    dec = eng.decode(ids, targets)
    assert eng.launches > 0
    for f, t, o, i, c, d in zip(feats, targets, out, ids, cs, dec):
        assert o.shape == f.shape and o.dtype == np.float32
        if f.shape[0] == 0:
            assert i.shape == (0,) and c.shape == (0,)
            continue
        assert np.array_equal(i, _oracle_ids(params, arch, f))
        assert np.array_equal(c, torch.unique_consecutive(torch.from_numpy(i)).numpy())
        want = _oracle_infer(params, arch, f, t)
        assert rel_err(o, want) < TOL
        assert rel_err(d, want) < TOL                    # decode(encode(x)) == infer(x)


def test_engine_tiny_utterances(Pm):
    """1-4 frame utterances (a k=3 conv then sees padding on both sides of every frame)."""
    from vae_npvc_b200.convert import ConvertEngine
    arch = synth.small_config()
    m, params, feats, targets = _setup(arch, seed=13, lengths=[1, 2, 1, 4, 3])
    for mb in (1, 8):
        out = ConvertEngine(m, precision='fp32', max_batch=mb, max_pad_frac=0.9).convert(feats, targets)
        for f, t, o in zip(feats, targets, out):
            assert rel_err(o, _oracle_infer(params, arch, f, t)) < TOL


def test_engine_equals_model_infer_batch1(Pm):
    """The drop-in Model.infer at batch 1 (the reference's calling pattern) and the packed engine agree."""
    import vae_npvc_b200.functional as Fn
    from vae_npvc_b200.convert import ConvertEngine
    arch = synth.small_config()
    m, params, feats, targets = _setup(arch)
    old = Fn.get_precision()
    Fn.set_precision('fp32')
    try:
        out = ConvertEngine(m, precision='fp32', max_batch=8, max_pad_frac=0.95).convert(feats, targets)
        with torch.no_grad():
            for f, t, o in zip(feats, targets, out):
                if f.shape[0] == 0:
                    continue
                x = torch.from_numpy(f).float().cuda().t().unsqueeze(0)           # decoder/basic.py:62-64
                ref = m.infer((x, torch.tensor([[t]]).cuda()))[0].t().cpu().numpy()
                assert rel_err(o, ref) < TOL
    finally:
        Fn.set_precision(old)


def test_engine_vcc2020_shape_packing_invariance(Pm):
    """VCC2020-shape model (C512, 10+10 blocks, codebook 512x64), utterances of 60-400 frames.
    fp32-faithful planes: packed results equal the same engine run one utterance at a time - the padding is
    invisible; only the GroupNorm sums are accumulated in a different order (fp64 atomics, then the padded share
    taken out again).  bf16 planes: the same comparison is only a sanity bound, because a last-bit difference in a
    GroupNorm mean flips bf16 roundings of the next operand plane and ten blocks amplify that into a few per cent
    of re-assigned codebook ids (measured: 97 % of ids agree, features within 1e-2); the decoder alone is compared
    with the fp32 oracle on the oracle's own ids."""
    from vae_npvc_b200.convert import ConvertEngine
    arch = synth.arch_config()
    rng = np.random.RandomState(3)
    lengths = [int(v) for v in rng.randint(60, 400, size=24)] + [400, 400]
    m, params, feats, targets = _setup(arch, seed=21, lengths=lengths)
    for precision in ('fp32', 'bf16'):
        packed = ConvertEngine(m, precision=precision, max_batch=16, max_pad_frac=0.5)
        single = ConvertEngine(m, precision=precision, max_batch=1)
        a, b = packed.convert(feats, targets), single.convert(feats, targets)
        ia, ib = packed.encode(feats, 'id'), single.encode(feats, 'id')
        assert packed.launches < single.launches
        agree = np.mean([(i == j).mean() for i, j in zip(ia, ib)])
        worst = max(rel_err(x, y) for x, y in zip(a, b))
        assert all(x.shape == y.shape and np.isfinite(x).all() for x, y in zip(a, b))
        if precision == 'fp32':
            assert agree > 0.999 and worst < 1e-2            # a single near-tie flip moves one frame's features
            assert np.median([rel_err(x, y) for x, y in zip(a, b)]) < 1e-5
        else:
            assert agree > 0.9 and worst < 5e-2
    # bf16 operands against the fp32 oracle: the decoder on the oracle's own ids (no near-tie flips involved);
    # a sanity bound - 21 chained bf16-operand GEMMs - the parity bar proper is the fp32 mode above
    for k in (0, 5):
        oid = _oracle_ids(params, arch, feats[k])
        want = O.model_decode(params, arch, torch.from_numpy(oid).view(1, -1), torch.tensor([[targets[k]]]))[0].t().numpy()
        got = packed.decode([oid], [targets[k]])[0]
        assert rel_err(got, want) < 5e-2
        got32 = ConvertEngine(m, precision='fp32').decode([oid], [targets[k]])[0]
        assert rel_err(got32, want) < 1e-4


def test_decoder_mirror_end_to_end(Pm, tmp_path):
    """vae_npvc_b200.decoder.basic.Decoder over real files, as vae_npvc/bin/decode.py drives the reference's:
    checkpoint -> trials + feats.scp (+ spk2spk_id) -> feats.ark / feats.scp, against the per-utterance oracle."""
    from vae_npvc_b200 import kaldi_ark as K
    from vae_npvc_b200.decoder.basic import Decoder
    arch = synth.small_config()
    sd = synth.synthetic_state_dict(arch, 777)
    torch.save({'model': sd, 'optimizer': {}, 'iteration': 1234}, str(tmp_path / 'iter.1234'))
    rng = np.random.RandomState(9)
    feats = {'utt%d' % i: rng.randn(n, 8).astype(np.float32) for i, n in enumerate([25, 11, 25, 18, 40])}
    src = tmp_path / 'src'
    src.mkdir()
    with K.ArkWriter(str(src / 'feats.ark'), str(src / 'feats.scp')) as w:
        for k, m in feats.items():
            w[k] = m
    spk_names = ['TEF1', 'TEM2', 'TFF1', 'TGM1', 'TMF1']
    (src / 'spk2spk_id').write_text(''.join('%s %d\n' % (s, i) for i, s in enumerate(spk_names)))
    trials = [('utt%d' % i, spk_names[(2 * i + 1) % 5]) for i in range(5)]
    (src / 'trials').write_text(''.join('%s %s\n' % t for t in trials))
    cfg = dict(arch, model_type='vae_npvc_b200.model.vqvae', precision='fp32', decode_max_pad_frac=0.9)
    dec = Decoder(cfg)
    assert dec.load_checkpoint(str(tmp_path / 'iter.1234')) == 1234
    assert dec.decode(str(src), str(tmp_path / 'out'), compress=False) == 5
    out = K.read_scp(str(tmp_path / 'out' / 'feats.scp'))
    assert [k for k, _ in out] == [u for u, _ in trials]
    params = O.params_from_state_dict(sd, requires_grad=False)
    for (u, spec), (_, spk) in zip(out, trials):
        got = K.load_mat(spec)
        assert rel_err(got, _oracle_infer(params, arch, feats[u], spk_names.index(spk))) < TOL
    # decode_step: the reference's per-utterance call shape (decoder/basic.py:62-73)
    x = torch.from_numpy(feats['utt1']).float().cuda().t().unsqueeze(0)
    y = torch.tensor([3]).long().cuda().view(1, -1)
    got = dec.decode_step(x, y)[0].t().cpu().numpy()
    assert rel_err(got, _oracle_infer(params, arch, feats['utt1'], 3)) < TOL
    # compressed output (the reference's default, compression_method=1) stays within the codec's error
    dec.decode(str(src), str(tmp_path / 'outc'), compress=True)
    for u, spec in K.read_scp(str(tmp_path / 'outc' / 'feats.scp')):
        ref = K.load_mat(dict(out)[u])
        assert np.abs(K.load_mat(spec) - ref).max() <= (ref.max() - ref.min()) / 64.0 + 1e-6


def test_device_feeder_feeds_trainer(Pm):
    """batches staged through pinned buffers on a copy stream arrive intact and in order,
    and a Trainer fed by the DeviceFeeder takes the same steps as one fed host batches."""
    from vae_npvc_b200.dataset.utt2mel_spk import DeviceFeeder, SyntheticDataset
    from vae_npvc_b200.trainer.basic import Trainer
    arch = synth.small_config()
    cfg = dict(arch, model_type='vae_npvc_b200.model.vqvae', precision='fp32', cuda_graph=True, graph_warmup=1,
               learning_rate=1e-3, max_grad_norm=10, crop_length=24, synthetic_utterances=20)
    ds = SyntheticDataset(None, cfg)
    loader = torch.utils.data.DataLoader(ds, batch_size=4, shuffle=False, drop_last=True)
    host = list(loader)
    feeder = DeviceFeeder(loader)
    assert len(feeder) == 5
    n = 0
    for (hx, hy), (dx, dy) in zip(host, feeder):
        assert dx.is_cuda and dy.is_cuda and torch.equal(dx.cpu(), hx) and torch.equal(dy.cpu(), hy)
        n += 1
    assert n == 5 and feeder.h2d_bytes == 5 * (4 * 8 * 24 * 4 + 4 * 8)
    runs = []
    for fed in (False, True):
        tr = Trainer(cfg)
        tr.model.load_state_dict(synth.synthetic_state_dict(arch, 777))
        log = []
        for batch in (DeviceFeeder(loader) if fed else loader):
            it, detail = tr.train_step(batch, 0)
            log.append(detail['Total'])
        assert tr.last_step_on_graph
        runs.append(log)
    assert len(runs[0]) == 5
    for a, b in zip(*runs):
        assert abs(a - b) <= 1e-5 * max(1.0, abs(a))
