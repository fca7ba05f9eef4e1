"""Host logic either side of the hot path, CPU only: the Kaldi table reader / writer,
the Dataset mirror (vae_npvc/dataset/utt2mel_spk.py:42-74) and the text form of extract_bnf
(vae_npvc/bin/extract_bnf.py:47-67)."""
import random
import struct

import numpy as np
import pytest
import torch

from vae_npvc_b200 import kaldi_ark as K


def _mats(seed=0):
    rng = np.random.RandomState(seed)
    return {'utt_a': rng.randn(100, 80).astype(np.float32), 'utt_b': rng.randn(3, 5).astype(np.float32),
            'utt_c': (rng.randn(9, 4) * 5 + 2).astype(np.float32), 'utt_d': np.full((12, 3), 1.5, dtype=np.float32)}


def test_uncompressed_round_trip_is_bit_exact(tmp_path):
    mats = _mats()
    with K.ArkWriter(str(tmp_path / 'f.ark'), str(tmp_path / 'f.scp')) as w:
        for k, m in mats.items():
            w[k] = m
        w.write('ids', np.arange(7, dtype=np.int64))
        w.write('vec', np.arange(5, dtype=np.float32))
    scp = K.read_scp(str(tmp_path / 'f.scp'))
    assert [k for k, _ in scp] == list(mats) + ['ids', 'vec']            # file order kept
    d = dict(scp)
    for k, m in mats.items():
        r = K.load_mat(d[k])
        assert r.dtype == np.float32 and np.array_equal(r, m)
    assert np.array_equal(K.load_mat(d['ids']), np.arange(7)) and K.load_mat(d['ids']).dtype == np.int32
    assert np.array_equal(K.load_mat(d['vec']), np.arange(5, dtype=np.float32))
    ark = list(K.read_ark(str(tmp_path / 'f.ark')))
    assert [k for k, _ in ark] == list(mats) + ['ids', 'vec']
    assert all(np.array_equal(v, mats[k]) for k, v in ark if k in mats)
    # Kaldi ranges are inclusive: [2:5] = 4 rows (dataset/utt2mel_spk.py:63 passes feat_end-1)
    assert np.array_equal(K.load_mat(d['utt_a'] + '[2:5]'), mats['utt_a'][2:6])
    assert np.array_equal(K.load_mat(d['utt_a'] + '[2:5,10:19]'), mats['utt_a'][2:6, 10:20])
    assert np.array_equal(K.load_mat(d['utt_a'] + '[:,0:0]'), mats['utt_a'][:, :1])


def test_known_answer_bytes(tmp_path):
    """Byte layout of the float-matrix and int-vector objects, written out by hand from kaldi-matrix.cc /
    io-funcs-inl.h: '\\0B' 'FM ' \\4 rows \\4 cols data."""
    m = np.array([[1.0, 2.0], [3.0, 4.0], [5.0, 6.0]], dtype=np.float32)
    want = b'\x00BFM \x04' + struct.pack('<i', 3) + b'\x04' + struct.pack('<i', 2) + m.tobytes()
    assert K.object_bytes(m) == want
    assert K.object_bytes(np.array([7, 8])) == b'\x00B\x04' + struct.pack('<i', 2) + struct.pack('<ii', 7, 8)
    p = tmp_path / 'one.ark'
    p.write_bytes(b'key1 ' + want + b'k2 ' + b'\x00BDM \x04' + struct.pack('<i', 1) + b'\x04' + struct.pack('<i', 2)
                  + np.array([[0.5, -2.0]], dtype='<f8').tobytes())
    got = dict(K.read_ark(str(p)))
    assert np.array_equal(got['key1'], m) and got['k2'].dtype == np.float64 and got['k2'].tolist() == [[0.5, -2.0]]
    assert np.array_equal(K.load_mat('%s:5' % p), m)                     # offset points at the '\0B'


@pytest.mark.parametrize('method', [1, 2, 3])
def test_compressed_round_trip_error_bounds(tmp_path, method):
    mats = _mats(1)
    with K.ArkWriter(str(tmp_path / 'c.ark'), str(tmp_path / 'c.scp'), compression_method=method) as w:
        for k, m in mats.items():
            w[k] = m
    d = dict(K.read_scp(str(tmp_path / 'c.scp')))
    for k, m in mats.items():
        r = K.load_mat(d[k])
        assert r.shape == m.shape and r.dtype == np.float32
        rng = float(m.max() - m.min()) or 1.0
        one_byte = method == 2 or (method == 1 and m.shape[0] > 8)
        # one byte per element inside per-column quartile segments: well under range/64; two bytes: range/65535
        assert np.abs(r - m).max() <= (rng / 64.0 if one_byte else rng / 30000.0) + 1e-6
    raw = (tmp_path / 'c.ark').read_bytes()
    assert (b'CM ' in raw) == (method in (1, 2)) and (b'CM2 ' in raw) == (method in (1, 3))


def _kaldi_dir(tmp_path, lengths, X=6):
    rng = np.random.RandomState(5)
    feats = {'u%d' % i: rng.randn(n, X).astype(np.float32) for i, n in enumerate(lengths)}
    with K.ArkWriter(str(tmp_path / 'feats.ark'), str(tmp_path / 'feats.scp')) as w:
        for k, m in feats.items():
            w[k] = m
    (tmp_path / 'utt2num_frames').write_text(''.join('%s %d\n' % (k, m.shape[0]) for k, m in feats.items()))
    (tmp_path / 'utt2spk_id').write_text(''.join('%s %d\n' % (k, i % 3) for i, k in enumerate(feats)))
    return feats


def test_dataset_mirror_crops_pads_and_matches_reference_draws(tmp_path):
    from vae_npvc_b200.dataset.utt2mel_spk import Dataset
    feats = _kaldi_dir(tmp_path, [50, 16, 9])
    cfg = {'crop_length': 16}
    ds = Dataset(str(tmp_path), cfg)
    assert len(ds) == 3
    random.seed(3)
    x, y = ds[0]
    random.seed(3)
    start = random.randint(0, 50 - 16)                                   # the reference's draw (utt2mel_spk.py:58)
    assert tuple(x.shape) == (6, 16) and x.dtype == torch.float32 and y.tolist() == [0] and y.dtype == torch.int64
    assert np.array_equal(x.numpy(), feats['u0'][start:start + 16].T)
    x, y = ds[1]                                                         # exactly crop_length: whole utterance
    assert np.array_equal(x.numpy(), feats['u1'].T) and y.tolist() == [1]
    x, _ = ds[2]                                                         # short: zero-padded on the right
    assert tuple(x.shape) == (6, 16) and np.array_equal(x.numpy()[:, :9], feats['u2'].T) and not x[:, 9:].any()
    dv = Dataset(str(tmp_path), {'crop_length': 16, 'valid_crop_length': 20}, valid=True)
    x, _ = dv[0]
    assert np.array_equal(x.numpy(), feats['u0'][:20].T)                 # validation crops from frame 0
    # default collate -> the [x (B,X,T), y (B,1)] list Trainer.train_step takes (train.py:69-76)
    xb, yb = next(iter(torch.utils.data.DataLoader(ds, batch_size=3)))
    assert tuple(xb.shape) == (3, 6, 16) and tuple(yb.shape) == (3, 1)


def test_synthetic_dataset_shape_and_determinism():
    from oracle import synth
    from vae_npvc_b200.dataset.utt2mel_spk import SyntheticDataset
    cfg = dict(synth.arch_config(), crop_length=128, synthetic_utterances=10)
    ds = SyntheticDataset(None, cfg)
    x, y = ds[3]
    assert tuple(x.shape) == (80, 128) and y.tolist() == [3] and len(ds) == 10
    assert torch.equal(ds[3][0], x) and not torch.equal(ds[4][0], x)
    assert not torch.equal(SyntheticDataset(None, cfg, valid=True)[3][0], x)


def test_extract_bnf_text_and_ark_outputs(tmp_path):
    from oracle import prims_ref, synth
    from vae_npvc_b200.convert import ConvertEngine
    from vae_npvc_b200.decoder.basic import extract_bnf
    from vae_npvc_b200.model.vqvae import Model
    arch = synth.small_config()
    m = Model(arch)
    m.load_state_dict(synth.synthetic_state_dict(arch, 777))
    m.eval()
    eng = ConvertEngine(m, precision='fp32', prims=prims_ref, max_pad_frac=0.9)
    feats = _kaldi_dir(tmp_path, [12, 7, 12], X=8)
    utts = list(K.read_ark(str(tmp_path / 'feats.ark')))
    res = extract_bnf(eng, utts, 'csid', str(tmp_path / 'bnf.txt'), output_txt=True)
    lines = (tmp_path / 'bnf.txt').read_text().splitlines()
    assert [l.split()[0] for l in lines] == list(feats)                  # reader order, as the reference's loop
    for l in lines:
        u, body = l.split()
        assert body == ''.join('<%d>' % v for v in res[u])               # extract_bnf.py:64-66
    ids = extract_bnf(eng, utts, 'id', 'ark,scp:%s,%s' % (tmp_path / 'id.ark', tmp_path / 'id.scp'), output_txt=False)
    back = dict(K.read_ark(str(tmp_path / 'id.ark')))
    assert all(np.array_equal(back[u], ids[u]) for u in feats)
    for u in feats:
        assert np.array_equal(res[u], torch.unique_consecutive(torch.from_numpy(ids[u])).numpy())


def test_datasets_shard_by_rank_without_launcher_changes():
    """The reference's train.py builds one seeded shuffling DataLoader on every process (bin/train.py:44-46,68-76): the
    datasets map the SAME index to DIFFERENT items per rank, so ranks read disjoint utterances."""
    from vae_npvc_b200.dataset.utt2mel_spk import SyntheticDataset, rank_world
    cfg = {'encoder': {'in_channels': [8]}, 'crop_length': 16, 'y_num': 5, 'synthetic_utterances': 64}
    full = SyntheticDataset(None, cfg)
    parts = [SyntheticDataset(None, dict(cfg, ddp_rank=r, ddp_world=4)) for r in range(4)]
    assert len(full) == 64 and all(len(p) == 16 for p in parts) and rank_world({}) == (0, 1)
    seen = set()
    for r, p in enumerate(parts):
        for i in range(len(p)):
            x, y = p[i]
            fx, fy = full[i * 4 + r]
            assert torch.equal(x, fx) and torch.equal(y, fy)
            seen.add(i * 4 + r)
    assert seen == set(range(64))
