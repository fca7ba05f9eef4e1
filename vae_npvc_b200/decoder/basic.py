"""Drop-in mirror of vae_npvc/decoder/basic.py: ``Decoder`` (select with ``decoder_type:
vae_npvc_b200.decoder.basic``, vae_npvc/bin/decode.py:34).

Same constructor keys (``model_type``, ``use_gpu``), ``decode_step(feat, spk)``, ``decode(decode_dir,
output_dir, compress)`` over the same files (``trials``, ``feats.scp``, optional ``spk2spk_id``; output
``feats.ark`` + ``feats.scp``), ``get_model_info`` and ``load_checkpoint`` as the reference.

B200-native differences behind that API (vae_npvc_b200/convert.py): ``decode`` does not loop ``model.infer`` at
batch 1 - the trial list is read, packed by length and converted in batches on frozen operand planes, with the
same per-utterance results; there is no try/except CPU retry (decoder/basic.py:30-35) because there is no CPU
path.  ``extract_bnf`` is the library form of vae_npvc/bin/extract_bnf.py:47-67 on the same engine.
"""
from importlib import import_module
from pathlib import Path

import numpy as np
import torch

from .. import kaldi_ark
from ..convert import ConvertEngine, plan_shards


def _dist():
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        return dist
    return None


def merge_scp(parts, out_path, order):
    """Concatenate per-rank scp files into one, lines in ``order`` (the trial order)."""
    lines = {}
    for p in parts:
        with open(p) as f:
            for line in f:
                if line.strip():
                    lines[line.split(None, 1)[0]] = line if line.endswith('\n') else line + '\n'
    with open(out_path, 'w') as w:
        for k in order:
            w.write(lines[k])


class Decoder(object):
    def __init__(self, config):
        model_type = config.get('model_type', 'vae_npvc_b200.model.vqvae:Model').split(':')
        if not config.get('use_gpu', True):
            raise RuntimeError('vae_npvc_b200 has no CPU path (use_gpu: false is not supported)')
        from .. import _C
        _C.require_cuda()
        self.device = torch.device('cuda')
        module = import_module(model_type[0], package=None)
        model_name = 'Model' if len(model_type) < 2 else model_type[1]
        model = getattr(module, model_name)(config)
        self.model = model.to(self.device)
        self.model.eval()
        self._engine_cfg = dict(precision=config.get('precision'), max_batch=int(config.get('decode_max_batch', 64)),
                                max_batch_frames=int(config.get('decode_max_batch_frames', 1 << 16)),
                                max_pad_frac=float(config.get('decode_max_pad_frac', 0.25)))
        self._engine = None

    @property
    def engine(self):
        if self._engine is None:
            self._engine = ConvertEngine(self.model, **self._engine_cfg)
        return self._engine

    def decode_step(self, feat, spk):
        """feat (1, X, T), spk (1, n) tensors -> (1, X, T) (decoder/basic.py:27-36), through the drop-in Model."""
        with torch.no_grad():
            return self.model.infer((feat, spk))

    def decode(self, decode_dir, output_dir, compress=True, log=None):
        decode_dir = Path(decode_dir)
        for name in ('trials', 'feats.scp'):
            if not (decode_dir / name).is_file():
                raise FileNotFoundError('No such file {}'.format(decode_dir / name))
        trials = [line.strip().split(None, 1) for line in open(str(decode_dir / 'trials')) if line.strip()]
        feats_scp = dict(kaldi_ark.read_scp(str(decode_dir / 'feats.scp')))
        spk2spk_id = None
        if (decode_dir / 'spk2spk_id').exists():
            spk2spk_id = dict(line.strip().split(None, 1) for line in open(str(decode_dir / 'spk2spk_id')))
        # one process per GPU: the trial list is sharded by frames (utt2num_frames when present, else round-robin);
        # each rank writes feats.<rank>.ark / .scp and rank 0 merges the scp files in trial order - no collective
        # touches the data path
        dist = _dist()
        rank, world = (dist.get_rank(), dist.get_world_size()) if dist else (0, 1)
        mine = list(range(len(trials)))
        if world > 1:
            nf = decode_dir / 'utt2num_frames'
            if nf.exists():
                frames = dict(line.strip().split(None, 1) for line in open(str(nf)) if line.strip())
                lengths = [int(frames.get(u, 1)) for u, _ in trials]
            else:
                lengths = [1] * len(trials)
            mine = plan_shards(lengths, world)[rank]
        feats, targets = [], []
        for i in mine:
            utt, target = trials[i]
            feats.append(np.asarray(kaldi_ark.load_mat(feats_scp[utt]), dtype=np.float32))
            targets.append([int(spk2spk_id[t]) if spk2spk_id else int(t) for t in target.split()])
        out = self.engine.convert(feats, targets)
        Path(output_dir).mkdir(parents=True, exist_ok=True)
        stem = '{}/feats'.format(output_dir) if world == 1 else '{}/feats.{}'.format(output_dir, rank)
        with kaldi_ark.ArkWriter(stem + '.ark', stem + '.scp', compression_method=1 if compress else None) as wf:
            for i, mat in zip(mine, out):                                       # written in trial order
                utt, target = trials[i]
                if log is not None:
                    log('Decode {}: {} to {}'.format(i, utt, target))
                wf[utt] = mat
        if world > 1:
            dist.barrier()
            if rank == 0:
                merge_scp(['{}/feats.{}.scp'.format(output_dir, r) for r in range(world)],
                          '{}/feats.scp'.format(output_dir), [u for u, _ in trials])
            dist.barrier()
        return len(out)

    def get_model_info(self):
        return self.model

    def load_checkpoint(self, checkpoint_file):
        checkpoint_data = torch.load(checkpoint_file, map_location='cpu')
        self.model.load_state_dict(checkpoint_data['model'])
        self.model.to(self.device)
        self._engine = None                       # operand planes are re-materialised from the new weights
        return checkpoint_data['iteration']


def extract_bnf(model_or_engine, utts, bnf_kind='id', wspecifier=None, output_txt=True):
    """vae_npvc/bin/extract_bnf.py:47-67 on the packed engine.  ``utts``: iterable of (utt, (T, X) array) - e.g.
    ``kaldi_ark.read_ark(path)``.  bnf_kind 'id' | 'csid' | 'token'.  Text output (ids / csid only) is the
    reference's ``utt <id><id>...`` lines; otherwise a Kaldi archive of int32 vectors.  Returns {utt: array}."""
    eng = model_or_engine if isinstance(model_or_engine, ConvertEngine) else ConvertEngine(model_or_engine)
    utts = list(utts)
    out = eng.encode([np.asarray(f, dtype=np.float32) for _, f in utts], bnf_kind)
    res = {u: o for (u, _), o in zip(utts, out)}
    if wspecifier is not None:
        if output_txt and bnf_kind in ('id', 'csid'):
            with open(wspecifier, 'w') as w:
                for u, _ in utts:
                    w.write('{} {}\n'.format(u, ''.join('<{}>'.format(int(b)) for b in res[u].reshape(-1))))
        else:
            with kaldi_ark.ArkWriter.from_wspecifier(wspecifier) as w:
                for u, _ in utts:
                    w[u] = res[u].reshape(-1)
    return res
