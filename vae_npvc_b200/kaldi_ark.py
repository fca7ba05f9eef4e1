"""Kaldi table IO for the data formats either side of the hot path.

The reference reads features with ``kaldiio.load_mat`` on ``feats.scp`` entries, optionally with a frame range
(vae_npvc/dataset/utt2mel_spk.py:42-74, vae_npvc/decoder/basic.py:61) and writes converted features with
``kaldiio.WriteHelper('ark,scp:...', compression_method=1)`` (decoder/basic.py:55-57, 75; bin/extract_bnf.py:45).
kaldiio is not a dependency here; this module restates the published Kaldi binary formats (kaldi-matrix.cc,
compressed-matrix.cc, kaldi-holder-inl.h) in numpy:

  read  : ``FM `` / ``DM `` float / double matrices, ``FV `` / ``DV `` vectors, compressed ``CM `` (speech-feature,
          per-column quartile headers, one byte per element), ``CM2`` (two bytes) and ``CM3`` (one byte) matrices,
          int32 vectors; scp entries ``path:offset`` with an optional ``[r0:r1]`` or ``[r0:r1,c0:c1]`` range
          (Kaldi ranges are INCLUSIVE)
  write : ``ark,scp`` pairs of float matrices - uncompressed, or compression_method 1 (Kaldi's automatic choice:
          ``CM `` for more than 8 rows, else ``CM2``), 2 (speech feature) or 3 (two byte) - and int32 vectors.

Rows of a Kaldi feature matrix are frames, which is exactly this library's channels-last activation layout.
"""
import os
import re
import struct

import numpy as np

_RANGE = re.compile(r'^(.*?)(\[[0-9:,]*\])?$')


# ---- reading ---------------------------------------------------------------------------------------------------
def _read_token(f):
    tok = b''
    while True:
        c = f.read(1)
        if not c:
            raise EOFError('unexpected end of Kaldi stream')
        if c == b' ':
            return tok
        tok += c


def _read_i32(f):
    if f.read(1) != b'\x04':
        raise ValueError('expected a 4-byte integer marker')
    return struct.unpack('<i', f.read(4))[0]


def _uncompress_cm(f):
    """``CM `` body (compressed-matrix.cc, format kOneByteWithColHeaders)."""
    vmin, vrange, rows, cols = struct.unpack('<ffii', f.read(16))
    hdr = np.frombuffer(f.read(cols * 8), dtype='<u2').reshape(cols, 4).astype(np.float32)
    p = vmin + vrange * (1.0 / 65535.0) * hdr                                     # (cols, 4) percentiles 0/25/75/100
    b = np.frombuffer(f.read(rows * cols), dtype=np.uint8).reshape(cols, rows).astype(np.float32)
    p0, p25, p75, p100 = (p[:, i:i + 1] for i in range(4))
    out = np.where(b <= 64, p0 + (p25 - p0) * b * (1.0 / 64.0),
                   np.where(b <= 192, p25 + (p75 - p25) * (b - 64.0) * (1.0 / 128.0),
                            p75 + (p100 - p75) * (b - 192.0) * (1.0 / 63.0)))
    return np.ascontiguousarray(out.T.astype(np.float32))


def read_object(f):
    """One Kaldi object at the current position of binary stream ``f`` (after the key): a float32 / float64 /
    int32 numpy array."""
    if f.read(2) != b'\x00B':
        raise ValueError('only binary Kaldi archives are supported')
    head = f.read(1)
    if head == b'\x04':                                   # vector<int32>: \4 <n> then n x int32
        n = struct.unpack('<i', f.read(4))[0]
        return np.frombuffer(f.read(4 * n), dtype='<i4').copy()
    tok = head + _read_token(f)
    if tok in (b'FM', b'DM'):
        rows, cols = _read_i32(f), _read_i32(f)
        dt = '<f4' if tok == b'FM' else '<f8'
        return np.frombuffer(f.read(rows * cols * int(dt[2])), dtype=dt).reshape(rows, cols).copy()
    if tok in (b'FV', b'DV'):
        n = _read_i32(f)
        dt = '<f4' if tok == b'FV' else '<f8'
        return np.frombuffer(f.read(n * int(dt[2])), dtype=dt).copy()
    if tok == b'CM':
        return _uncompress_cm(f)
    if tok in (b'CM2', b'CM3'):
        vmin, vrange, rows, cols = struct.unpack('<ffii', f.read(16))
        if tok == b'CM2':
            d = np.frombuffer(f.read(rows * cols * 2), dtype='<u2').astype(np.float32) * (1.0 / 65535.0)
        else:
            d = np.frombuffer(f.read(rows * cols), dtype=np.uint8).astype(np.float32) * (1.0 / 255.0)
        return (vmin + vrange * d).reshape(rows, cols).astype(np.float32)
    raise ValueError('unsupported Kaldi object type %r' % tok)


def _apply_range(mat, rng):
    if not rng:
        return mat
    parts = rng[1:-1].split(',')
    sl = []
    for p in parts:
        if p in ('', ':'):
            sl.append(slice(None))
        else:
            a, b = p.split(':')
            sl.append(slice(int(a), int(b) + 1))                          # Kaldi ranges include the end
    return mat[tuple(sl)]


def load_mat(spec):
    """``path:offset[range]`` (an scp value) or a plain ``path`` holding one object -> numpy array
    (the reference's kaldiio.load_mat call sites)."""
    m = _RANGE.match(spec.strip())
    loc, rng = m.group(1), m.group(2)
    path, off = loc, 0
    if ':' in loc and loc.rsplit(':', 1)[1].isdigit():
        path, off = loc.rsplit(':', 1)
        off = int(off)
    with open(path, 'rb') as f:
        f.seek(off)
        return _apply_range(read_object(f), rng)


def read_scp(path):
    """-> list of (key, spec) in file order."""
    out = []
    with open(path) as f:
        for line in f:
            line = line.strip()
            if line:
                k, v = line.split(None, 1)
                out.append((k, v))
    return out


def read_ark(path):
    """Iterate (key, array) over a binary archive (kaldiio.ReadHelper('ark:...'))."""
    with open(path, 'rb') as f:
        while True:
            c = f.read(1)
            if not c:
                return
            key = c
            while True:
                c = f.read(1)
                if c in (b' ', b''):
                    break
                key += c
            yield key.decode(), read_object(f)


# ---- writing ---------------------------------------------------------------------------------------------------
def _u16(vmin, vrange, v):
    return np.clip(np.floor((v - vmin) / vrange * 65535.0 + 0.499), 0, 65535).astype(np.int64)


def _compress_cm(mat):
    """``CM `` body for a (rows > 0, cols) float matrix (compressed-matrix.cc ComputeColHeader / FloatToChar)."""
    rows, cols = mat.shape
    vmin, vmax = float(mat.min()), float(mat.max())
    if vmax == vmin:
        vmax = vmin + (abs(vmin) + 1.0)
    vrange = vmax - vmin
    srt = np.sort(mat, axis=0)
    if rows >= 5:
        q = rows // 4
        pct = np.stack([srt[0], srt[q], srt[3 * q], srt[-1]], axis=1)
    else:                                                   # fewer than 5 rows: the sorted values themselves
        pct = np.stack([srt[min(i, rows - 1)] for i in range(4)], axis=1)
    h = _u16(vmin, vrange, pct)
    h[:, 0] = np.minimum(h[:, 0], 65532)                    # strictly increasing 16-bit quartiles
    h[:, 1] = np.minimum(np.maximum(h[:, 1], h[:, 0] + 1), 65533)
    h[:, 2] = np.minimum(np.maximum(h[:, 2], h[:, 1] + 1), 65534)
    h[:, 3] = np.maximum(h[:, 3], h[:, 2] + 1)
    p = (vmin + vrange * (1.0 / 65535.0) * h.astype(np.float32)).astype(np.float32)
    p0, p25, p75, p100 = (p[:, i:i + 1] for i in range(4))
    v = mat.T.astype(np.float32)                                                  # (cols, rows)
    lo = np.clip(np.floor((v - p0) / (p25 - p0) * 64.0 + 0.5), 0, 64)
    mid = np.clip(64 + np.floor((v - p25) / (p75 - p25) * 128.0 + 0.5), 64, 192)
    hi = np.clip(192 + np.floor((v - p75) / (p100 - p75) * 63.0 + 0.5), 192, 255)
    b = np.where(v < p25, lo, np.where(v < p75, mid, hi)).astype(np.uint8)
    return (b'CM ' + struct.pack('<ffii', vmin, vrange, rows, cols) + h.astype('<u2').tobytes() + b.tobytes())


def _compress_cm2(mat):
    rows, cols = mat.shape
    vmin, vmax = float(mat.min()), float(mat.max())
    if vmax == vmin:
        vmax = vmin + (abs(vmin) + 1.0)
    d = _u16(vmin, vmax - vmin, mat.astype(np.float32))
    return b'CM2 ' + struct.pack('<ffii', vmin, vmax - vmin, rows, cols) + d.astype('<u2').tobytes()


def object_bytes(arr, compression_method=None):
    arr = np.asarray(arr)
    if arr.dtype.kind in 'iu':
        a = np.ascontiguousarray(arr.reshape(-1).astype('<i4'))
        return b'\x00B\x04' + struct.pack('<i', a.size) + a.tobytes()
    if arr.ndim == 1:
        a = np.ascontiguousarray(arr.astype('<f4'))
        return b'\x00BFV \x04' + struct.pack('<i', a.size) + a.tobytes()
    if arr.ndim != 2:
        raise ValueError('Kaldi tables hold vectors and matrices')
    rows, cols = arr.shape
    if compression_method in (None, 0) or arr.size == 0:
        a = np.ascontiguousarray(arr.astype('<f4'))
        return b'\x00BFM \x04' + struct.pack('<i', rows) + b'\x04' + struct.pack('<i', cols) + a.tobytes()
    if compression_method == 1:
        body = _compress_cm(arr) if rows > 8 else _compress_cm2(arr)
    elif compression_method == 2:
        body = _compress_cm(arr)
    elif compression_method == 3:
        body = _compress_cm2(arr)
    else:
        raise ValueError('compression_method %r is not supported (0/None, 1, 2, 3)' % (compression_method,))
    return b'\x00B' + body


class ArkWriter(object):
    """``kaldiio.WriteHelper('ark,scp:A,S', compression_method=...)`` stand-in: ``w[key] = array`` or
    ``w.write(key, array)``; context manager."""

    def __init__(self, ark_path, scp_path=None, compression_method=None):
        self.ark_path = ark_path
        self.ark = open(ark_path, 'wb')
        self.scp = open(scp_path, 'w') if scp_path else None
        self.compression_method = compression_method

    @classmethod
    def from_wspecifier(cls, wspecifier, compression_method=None):
        """'ark,scp:ARK,SCP' | 'ark:ARK' (the two forms the reference uses)."""
        kinds, paths = wspecifier.split(':', 1)
        kinds, paths = kinds.split(','), paths.split(',')
        ark = paths[kinds.index('ark')]
        scp = paths[kinds.index('scp')] if 'scp' in kinds else None
        return cls(ark, scp, compression_method)

    def write(self, key, arr):
        self.ark.write(key.encode() + b' ')
        off = self.ark.tell()
        self.ark.write(object_bytes(arr, self.compression_method))
        if self.scp:
            self.scp.write('%s %s:%d\n' % (key, os.path.abspath(self.ark_path), off))

    __setitem__ = write

    def close(self):
        self.ark.close()
        if self.scp:
            self.scp.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
        return False
