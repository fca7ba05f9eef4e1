"""WeightBank: all conv weights of a model materialised by ONE kernel per optimiser step.

The reference re-derives ``w = g * v / ||v||`` in a forward pre-hook of each of its 55 conv modules
(old-style nn.utils.weight_norm, vae_npvc/model/vqvae.py:203-208, 329-334) and autograd walks 55
separate weight-norm backward graphs.  Here the trainer keeps parameters and gradients in two flat fp32
buffers; the bank holds a device-side table of where every conv's (v, g) sits in them and

  * ``prepare()``          two launches (row scales, then transposing tiles): every layer's bf16 GEMM operand planes (forward operand ``wf``,
                           data-gradient operand ``wb``, fp32 effective weight for the 1x1 speaker-
                           conditioning convs) straight from the flat parameter buffer
  * ``finish_backward()``  one launch: the weight-gradient GEMMs have written raw dW into the v slots of
                           the flat gradient buffer; this turns them into (dv, dg) in place

While a bank is installed (functional.set_bank) the autograd Functions take their operands from it and
write parameter gradients directly into the gradient slots instead of returning them to autograd.
"""
import torch
import torch.nn as nn

from . import prims as Pm


def _align(n, a=64):
    return (n + a - 1) // a * a


class _Entry(object):
    __slots__ = ('W', 'v', 'g')

    def __init__(self, W, v, g):
        self.W, self.v, self.g = W, v, g


class WeightBank(object):
    def __init__(self, model, flat_p, flat_g, P, prims=None):
        self.Pm = prims or Pm
        self.P = P
        self.flat_p, self.flat_g = flat_p, flat_g
        dev = flat_p.device
        base, end = flat_p.data_ptr(), flat_p.data_ptr() + flat_p.numel() * 4
        rows, tiles, desc, entries = [], [], [], []
        w_off = w32_off = row0 = tile0 = 0
        self.max_k = 1
        seg_of_layer = []                 # segment index per layer (see ``segments`` below)
        # Table order = segment order: 0 the leading 'encoder.*' convs, 1 every other conv except 2, the decoders'
        # conditioning 1x1 convs.  The conditioning convs get their weight gradients from ONE batched kernel after the
        # whole decoder has been back-propagated (functional.CondAllFn), i.e. later than the moment the trainer starts
        # the decoder segment's weight-norm backward on a side stream - so they must not be part of that segment.
        convs = [(name, m) for name, m in model.named_modules() if isinstance(m, (nn.Conv1d, nn.ConvTranspose1d))]

        def _seg(name):
            if '.conv_cond' in name:
                return 2
            return 0 if name.startswith('encoder.') else 1
        convs.sort(key=lambda nm: _seg(nm[0]))            # stable: module order inside a segment
        for name, m in convs:
            if hasattr(m, 'weight_g'):
                v, g = m.weight_v, m.weight_g
            else:
                v, g = m.weight, None
            if not (base <= v.data_ptr() < end) or (g is not None and not (base <= g.data_ptr() < end)):
                continue                                  # not a flat-buffer parameter (frozen / foreign)
            if getattr(m, '_vnb_no_bank', False):
                continue                                  # strided conv (model/vqvae2.py): its operand is a re-laid view
            transposed = isinstance(m, nn.ConvTranspose1d)
            d0, d1, k = v.shape
            co, ci = (d1, d0) if transposed else (d0, d1)
            n_wf, n_wb = P * k * co * Pm.pitch(ci), P * k * ci * Pm.pitch(co)
            wf_off, wb_off = w_off, w_off + _align(n_wf)
            w_off = wb_off + _align(n_wb)
            want32 = name.endswith('conv_cond')
            e32 = w32_off if want32 else -1
            if want32:
                w32_off += _align(v.numel())
            seg_of_layer.append(_seg(name))
            desc.append([(v.data_ptr() - base) // 4, (g.data_ptr() - base) // 4 if g is not None else -1, d0, d1, k,
                         int(transposed), wf_off, wb_off, e32, row0, tile0, 0])
            rows += [len(desc) - 1] * d0
            row0 += d0
            n_tiles = ((d0 + 63) // 64) * ((d1 + 63) // 64)
            tiles += [len(desc) - 1] * n_tiles
            tile0 += n_tiles
            self.max_k = max(self.max_k, k)
            entries.append((v, g, transposed, k, co, ci, wf_off, n_wf, wb_off, n_wb, e32))
        self.n_layers = len(desc)
        # zero-filled once: the pitch padding of the operand planes must stay finite
        self.wbuf = torch.zeros(max(w_off, 1), dtype=torch.bfloat16, device=dev)
        self.w32buf = torch.zeros(max(w32_off, 1), dtype=torch.float32, device=dev)
        self.desc = torch.tensor(desc if desc else [[0] * 12], dtype=torch.int64, device=dev)
        self.row2layer = torch.tensor(rows if rows else [0], dtype=torch.int32, device=dev)
        self.tile2layer = torch.tensor(tiles if tiles else [0], dtype=torch.int32, device=dev)
        self.row_scale = torch.empty(max(row0, 1), dtype=torch.float32, device=dev)
        # Segments: contiguous layer ranges (module order) that can be materialised / back-propagated by separate
        # launches, so that the decoder's share of the two parameter-sized passes runs on a side stream beside the
        # encoder's forward / backward (trainer/basic.py).  A segment launch gets views of the row / tile tables that
        # start at its first row / tile and a copy of the descriptor table with row0 / tile0 shifted accordingly.
        self.segments = []
        self.segment_kinds = sorted(set(seg_of_layer))    # which of {0 encoder, 1 other, 2 conditioning} exist, in table order
        for sidx in self.segment_kinds:
            layers = [i for i, s_ in enumerate(seg_of_layer) if s_ == sidx]
            r0, t0 = desc[layers[0]][9], desc[layers[0]][10]
            last = layers[-1]
            r1 = desc[last][9] + desc[last][2]
            t1 = desc[last][10] + ((desc[last][2] + 63) // 64) * ((desc[last][3] + 63) // 64)
            d = self.desc.clone()
            d[:, 9] -= r0
            d[:, 10] -= t0
            self.segments.append(dict(layers=(layers[0], last + 1), rows=(r0, r1), tiles=(t0, t1), desc=d))
        self._pending = {}                # segment index -> CUDA event its prepare() was recorded with
        self.seg_of_ptr = {}
        self.by_ptr = {}
        for li, (v, g, transposed, k, co, ci, wf_off, n_wf, wb_off, n_wb, e32) in enumerate(entries):
            self.seg_of_ptr[v.data_ptr()] = self.segment_kinds.index(seg_of_layer[li])
            wf = self.wbuf[wf_off:wf_off + n_wf].view(P, k, co, Pm.pitch(ci))
            wb = self.wbuf[wb_off:wb_off + n_wb].view(P, k, ci, Pm.pitch(co))
            w32 = self.w32buf[e32:e32 + v.numel()].view_as(v) if e32 >= 0 else None
            self.by_ptr[v.data_ptr()] = _Entry(self.Pm.ConvWeights(wf, wb, w32, k, co, ci, transposed), v, g)
        self.fresh = False

    def has(self, v):
        return v is not None and v.data_ptr() in self.by_ptr

    def weights(self, v):
        if self._pending:
            self.wait_ready(self.seg_of_ptr[v.data_ptr()])
        return self.by_ptr[v.data_ptr()].W

    def wait_ready(self, seg):
        """Make the current stream wait for a segment whose prepare() was issued on another stream."""
        ev = self._pending.pop(seg, None)
        if ev is not None:
            torch.cuda.current_stream().wait_event(ev)

    def prepare(self, seg=None, event=None):
        """Materialise the operand planes of every layer (``seg`` None) or of one segment.  ``event``: a CUDA event
        the caller records after this call on the stream it was issued on; the first use of one of the segment's
        weights on another stream waits for it (weights())."""
        if self.n_layers:
            for i, S in enumerate(self.segments):
                if seg is not None and i != seg:
                    continue
                (r0, r1), (t0, t1) = S['rows'], S['tiles']
                self.Pm.wn_prepare_batch(self.flat_p, S['desc'], self.row2layer[r0:r1], self.P, self.wbuf, self.w32buf,
                                         self.tile2layer[t0:t1], self.max_k, self.row_scale[r0:r1])
                if event is not None:
                    self._pending[i] = event
        if seg is None:
            self.fresh = True

    def finish_backward(self, seg=None):
        if self.n_layers:
            for i, S in enumerate(self.segments):
                if seg is not None and i != seg:
                    continue
                r0, r1 = S['rows']
                self.Pm.wn_backward_batch(self.flat_p, self.flat_g, S['desc'], self.row2layer[r0:r1])
