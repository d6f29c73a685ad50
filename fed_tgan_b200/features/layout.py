"""Flat description of an encoded row, shared by every kernel.

``output_info`` is the reference's list of ``(width, 'tanh'|'softmax')`` spans
(``Server/dtds/features/transformers.py:346-356``).  ``TableLayout`` turns it into integer
tables ("segment table" of SURVEY K4/K5/K9/K10): span starts/widths in the data row, the
offset of every *softmax* span in the conditional vector (``Cond.interval``,
``ctgan.py:123-145``), and which spans are tanh scalars.
"""

from __future__ import annotations

import numpy as np
import torch

TANH, SOFTMAX = 0, 1


class TableLayout:
    def __init__(self, output_info):
        self.output_info = [(int(w), str(k)) for w, k in output_info]
        starts, widths, kinds = [], [], []
        soft_start, soft_width, soft_cstart, tanh_start = [], [], [], []
        st = 0
        cst = 0
        for w, kind in self.output_info:
            starts.append(st)
            widths.append(w)
            if kind == "tanh":
                kinds.append(TANH)
                for j in range(w):
                    tanh_start.append(st + j)
            elif kind == "softmax":
                kinds.append(SOFTMAX)
                soft_start.append(st)
                soft_width.append(w)
                soft_cstart.append(cst)
                cst += w
            else:
                raise ValueError(f"unknown span kind {kind!r}")
            st += w
        self.data_dim = st
        self.n_opt = cst
        self.n_col = len(soft_start)
        self.n_tanh = len(tanh_start)
        self.n_spans = len(starts)
        self.max_width = max(widths) if widths else 0
        i32 = np.int32
        self.span_start = np.asarray(starts, dtype=i32)
        self.span_width = np.asarray(widths, dtype=i32)
        self.span_kind = np.asarray(kinds, dtype=i32)
        self.soft_start = np.asarray(soft_start, dtype=i32)
        self.soft_width = np.asarray(soft_width, dtype=i32)
        self.soft_cstart = np.asarray(soft_cstart, dtype=i32)
        self.tanh_start = np.asarray(tanh_start, dtype=i32)
        # per data column: owning span id and (for softmax spans) softmax-span index
        col_span = np.zeros(self.data_dim, dtype=i32)
        col_soft = np.full(self.data_dim, -1, dtype=i32)
        j = 0
        for s, (a, w, k) in enumerate(zip(starts, widths, kinds)):
            col_span[a:a + w] = s
            if k == SOFTMAX:
                col_soft[a:a + w] = j
                j += 1
        self.col_span = col_span
        self.col_soft = col_soft
        self._dev = {}

    def device_tables(self, device):
        """int32 tensors on ``device`` (cached)."""
        key = str(device)
        if key not in self._dev:
            names = ("span_start", "span_width", "span_kind", "soft_start", "soft_width",
                     "soft_cstart", "tanh_start", "col_span", "col_soft")
            self._dev[key] = {n: torch.from_numpy(getattr(self, n).copy()).to(device) for n in names}
        return self._dev[key]

    def __getstate__(self):
        d = dict(self.__dict__)
        d["_dev"] = {}
        return d

    def __eq__(self, other):
        return isinstance(other, TableLayout) and self.output_info == other.output_info

    def __repr__(self):
        return (f"TableLayout(data_dim={self.data_dim}, n_col={self.n_col}, n_opt={self.n_opt}, "
                f"n_tanh={self.n_tanh})")
