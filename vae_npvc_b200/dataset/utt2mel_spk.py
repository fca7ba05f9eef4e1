"""Drop-in mirror of vae_npvc/dataset/utt2mel_spk.py (``Dataset``) plus the device feeder of SURVEY.md
section 8 f3.

``Dataset(data_dir, config, valid=False)`` reads the same three Kaldi files (``feats.scp``, ``utt2num_frames``,
``utt2spk_id``), crops ``crop_length`` frames at a random start (start 0 for validation) by a ranged scp read,
zero-pads short utterances and returns ``(feat (X, T) float32, spk (1,) int64)`` - the reference's item
(utt2mel_spk.py:42-74).  kaldiio is replaced by vae_npvc_b200/kaldi_ark.py.

``SyntheticDataset`` is the synthetic mel generator of the recipe's shape that BASELINE.json's north_star puts
in the dataset's place for benchmarking (CMVN-normalised features are ~N(0,1): run.sh:106,115-120).

``DeviceFeeder`` wraps any loader of ``[x, y]`` batches (e.g. torch's DataLoader over either dataset): it stages
each batch in one of two pinned host buffers and copies it to the device on its own stream, one batch ahead of the
trainer, so the H2D transfer of step k+1 overlaps step k (the reference does a synchronous ``x.cuda()`` at the top
of every step, trainer/basic.py:59).  The trainer's ``t.to(device)`` on an already-resident tensor is a no-op.
"""
import os
import random
from pathlib import Path

import numpy as np
import torch

from .. import kaldi_ark


def rank_world(config=None):
    """(rank, world) of this process under torchrun (RANK / WORLD_SIZE), overridable by config keys ``ddp_rank`` /
    ``ddp_world``.  The reference's train.py builds ONE seeded shuffling DataLoader (bin/train.py:44-46,68-76): with
    equal seeds every rank draws the SAME index sequence, so the datasets below map index i to item i * world + rank
    and report len // world - the ranks then read disjoint utterances without any change to the launcher."""
    config = config or {}
    world = int(config.get('ddp_world', os.environ.get('WORLD_SIZE', '1')))
    rank = int(config.get('ddp_rank', os.environ.get('RANK', '0')))
    return (rank, world) if world > 1 else (0, 1)


def load_dict_data(data_file):
    lines = [line.strip().split(None, 1) for line in open(str(data_file)) if line.strip()]
    return {x[0]: x[1] for x in lines}


def load_list_data(data_file):
    return [line.strip().split() for line in open(str(data_file)) if line.strip()]


class Dataset(torch.utils.data.Dataset):
    def __init__(self, data_dir, config, valid=False):
        crop_length = config.get('crop_length', 256)
        self.crop_length = config.get('valid_crop_length' if valid else 'train_crop_length', crop_length)
        self.valid = valid
        data_dir = Path(data_dir)
        self.feats_scp = load_dict_data(data_dir / 'feats.scp')
        self.utt2num_frames = load_dict_data(data_dir / 'utt2num_frames')
        self.utt2spks = load_list_data(data_dir / 'utt2spk_id')
        self.rank, self.world = rank_world(config)
        self.num_data = len(self.utt2spks) // self.world

    def __getitem__(self, index):
        utt, spk = self.utt2spks[index * self.world + self.rank]
        feat_length = int(self.utt2num_frames[utt])
        if feat_length <= self.crop_length:
            feat_start, feat_end = 0, feat_length
        else:
            feat_start = random.randint(0, feat_length - self.crop_length) if not self.valid else 0
            feat_end = feat_start + self.crop_length
        mat = kaldi_ark.load_mat('{}[{}:{}]'.format(self.feats_scp[utt], feat_start, feat_end - 1))
        feat = torch.zeros((mat.shape[1], max(self.crop_length, mat.shape[0])), dtype=torch.float32)
        feat[:, :mat.shape[0]] = torch.from_numpy(np.ascontiguousarray(mat.T)).float()
        return feat, torch.tensor([int(spk)]).long()

    def __len__(self):
        return self.num_data


class SyntheticDataset(torch.utils.data.Dataset):
    """Seeded N(0,1) features of the recipe's shape: item i is always the same (X, crop_length) matrix."""

    def __init__(self, data_dir, config, valid=False, num_data=4096):
        self.x_dim = config['encoder']['in_channels'][0]
        self.crop_length = config.get('valid_crop_length' if valid else 'train_crop_length',
                                      config.get('crop_length', 256))
        self.y_num = config.get('y_num', 10)
        self.rank, self.world = rank_world(config)
        self.num_data = int(config.get('synthetic_utterances', num_data)) // self.world
        self.seed = int(config.get('seed', 777)) + (1 << 20 if valid else 0)

    def __getitem__(self, index):
        index = index * self.world + self.rank
        g = torch.Generator().manual_seed(self.seed + index)
        feat = torch.randn(self.x_dim, self.crop_length, generator=g)
        return feat, torch.tensor([index % self.y_num]).long()

    def __len__(self):
        return self.num_data


class DeviceFeeder(object):
    """Iterate device-resident ``[x, y]`` batches one transfer ahead of the consumer.

    Two pinned staging slots per tensor; slot s is reused for batch k+2 only after the copy of batch k out of it
    has finished (its event).  The consumer's stream waits for the copy event before the batch is handed out, and
    the batch's memory is kept from being recycled under the copy by ``record_stream``."""

    def __init__(self, loader, device='cuda'):
        self.loader = loader
        self.device = torch.device(device)
        if self.device.type != 'cuda':
            raise RuntimeError('DeviceFeeder stages batches for a CUDA device (vae_npvc_b200 has no CPU path)')
        self.stream = torch.cuda.Stream(device=self.device)
        self._slots = [None, None]
        self._events = [None, None]
        self.h2d_bytes = 0

    def __len__(self):
        return len(self.loader)

    def _stage(self, batch, s):
        if self._events[s] is not None:
            self._events[s].synchronize()                      # the previous copy out of this slot is done
        slot = self._slots[s]
        if slot is None or any(a.shape != b.shape or a.dtype != b.dtype for a, b in zip(slot, batch)):
            slot = [torch.empty(t.shape, dtype=t.dtype, pin_memory=True) for t in batch]
            self._slots[s] = slot
        for dst, src in zip(slot, batch):
            dst.copy_(src)
        with torch.cuda.stream(self.stream):
            dev = [t.to(self.device, non_blocking=True) for t in slot]
            ev = torch.cuda.Event()
            ev.record(self.stream)
        self._events[s] = ev
        self.h2d_bytes += sum(t.numel() * t.element_size() for t in slot)
        return dev, ev

    def __iter__(self):
        it = iter(self.loader)
        k = 0
        try:
            nxt = self._stage(next(it), 0)
        except StopIteration:
            return
        while nxt is not None:
            dev, ev = nxt
            k += 1
            try:
                nxt = self._stage(next(it), k & 1)             # queue the next transfer before handing this one out
            except StopIteration:
                nxt = None
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(ev)
            for t in dev:
                t.record_stream(cur)
            yield dev
