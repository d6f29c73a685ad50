"""Device-resident encoded training shard + sampler indices.

Replaces three host-side pieces of the reference that run Python loops every step
(SURVEY H1/H2/H3):

* the dense encoded matrix + ``Sampler`` row-id lists (``ctgan.py:197-228``)  ->  compact
  table ``vals [R, n_tanh] f32`` / ``codes [R, n_col] i32`` plus a CSR
  ``row_ptr [n_opt+1]`` / ``row_ids [R*n_col]`` of row ids per (span, option);
* ``Cond`` (``ctgan.py:102-172``)  ->  cumulative log-frequency table ``cum_p`` (training
  draws) and cumulative empirical table ``cum_emp`` (``sample_zero``).

Only per-option *counts* ever need to leave a client (for the federator's sampling
distribution) -- not rows (SURVEY 7.1 "Server-side Cond without centralising data").
"""

from __future__ import annotations

import numpy as np
import torch

from ..features.layout import TableLayout


def cumulative_tables(counts: np.ndarray, layout: TableLayout):
    """``(cum_p, cum_emp)`` ``[n_col, max_w]`` from per-option counts ``[n_opt]``."""
    n_col = layout.n_col
    max_w = int(layout.soft_width.max()) if n_col else 0
    cum_p = np.full((n_col, max_w), 2.0, dtype=np.float32)
    cum_e = np.full((n_col, max_w), 2.0, dtype=np.float32)
    for j in range(n_col):
        w, c0 = int(layout.soft_width[j]), int(layout.soft_cstart[j])
        cnt = counts[c0:c0 + w].astype(np.float64)
        lf = np.log(cnt + 1.0)
        p = lf / lf.sum() if lf.sum() > 0 else np.full(w, 1.0 / w)
        e = cnt / cnt.sum() if cnt.sum() > 0 else np.full(w, 1.0 / w)
        cum_p[j, :w] = np.cumsum(p)
        cum_e[j, :w] = np.cumsum(e)
        cum_p[j, w - 1] = 1.0
        cum_e[j, w - 1] = 1.0
    return cum_p, cum_e


class DeviceTable:
    def __init__(self, vals, codes, layout: TableLayout, device="cpu", pinned_host=False):
        vals = np.ascontiguousarray(vals, dtype=np.float32)
        codes = np.ascontiguousarray(codes, dtype=np.int32)
        self.layout = layout
        self.n_rows = int(codes.shape[0])
        self.device = torch.device(device)
        R, n_col, n_opt = self.n_rows, layout.n_col, layout.n_opt

        # CSR of row ids per global option
        g = codes.astype(np.int64) + layout.soft_cstart.astype(np.int64)[None, :]   # [R, n_col]
        flat = g.reshape(-1)
        counts = np.bincount(flat, minlength=n_opt).astype(np.int64)
        order = np.argsort(flat, kind="stable")
        row_ids = (order // max(n_col, 1)).astype(np.int32)
        row_ptr = np.zeros(n_opt + 1, dtype=np.int32)
        np.cumsum(counts, out=row_ptr[1:])
        self.counts = counts
        cum_p, cum_e = cumulative_tables(counts, layout)

        self.host_vals = torch.from_numpy(vals)
        self.host_codes = torch.from_numpy(codes)
        if pinned_host and torch.cuda.is_available():
            self.host_vals = self.host_vals.pin_memory()
            self.host_codes = self.host_codes.pin_memory()
        self.vals = self.host_vals.to(self.device, non_blocking=True)
        self.codes = self.host_codes.to(self.device, non_blocking=True)
        self.row_ptr = torch.from_numpy(row_ptr).to(self.device)
        self.row_ids = torch.from_numpy(row_ids).to(self.device)
        self.cum_p = torch.from_numpy(cum_p).to(self.device)
        self.cum_emp = torch.from_numpy(cum_e).to(self.device)
        self.tables = layout.device_tables(self.device)

    def restage(self, stream=None):
        """Re-copy the compact shard host(pinned) -> device (the per-epoch H2D of the
        end-to-end / out-of-core path)."""
        self.vals.copy_(self.host_vals, non_blocking=True)
        self.codes.copy_(self.host_codes, non_blocking=True)
        return self.host_vals.numel() * 4 + self.host_codes.numel() * 4

    @classmethod
    def from_dense(cls, dense, layout: TableLayout, device="cpu"):
        """Build from a dense one-hot encoded matrix (the reference's representation)."""
        dense = np.asarray(dense)
        vals = dense[:, layout.tanh_start] if layout.n_tanh else np.zeros((len(dense), 0))
        codes = np.stack(
            [dense[:, s:s + w].argmax(1) for s, w in zip(layout.soft_start, layout.soft_width)], axis=1
        ) if layout.n_col else np.zeros((len(dense), 0))
        return cls(vals, codes, layout, device)

    def dense(self):
        """Dense ``[R, data_dim]`` float32 matrix (tests / debugging)."""
        from ..ops import reference

        out = torch.zeros(self.n_rows, self.layout.data_dim)
        rid = torch.arange(self.n_rows, dtype=torch.int32)
        tables = self.layout.device_tables("cpu")
        return reference.expand_rows(self.vals.cpu(), self.codes.cpu(), rid, tables, out)
