"""Flat fp32 parameter arenas.

All parameters and buffers of one network live in ONE contiguous fp32 allocation
(``data``), with gradients and the two Adam moments in parallel allocations of the
trainable prefix.  Why:

* the federated weighted average and fused Adam (K8) are single launches over
  the arena instead of a Python loop over 16 + 6 ``state_dict`` keys
  (``Server/dtds/distributed.py:86-106``);
* ``data`` can be placed inside a peer-mapped symmetric allocation so other GPUs read it
  over NVLink;
* weights are stored with their K dimension padded to a multiple of 4 floats so every
  row is 16-byte aligned (TMA global-stride rule).  Padding elements are zero and stay
  zero (their gradient is identically zero).

The arena converts to and from the reference ``state_dict`` layout
(keys/shapes of ``ctgan.Generator`` / ``ctgan.Discriminator``) through an
index map per slot, so checkpoints are interchangeable with the reference.
"""

from __future__ import annotations

from collections import OrderedDict
from dataclasses import dataclass

import torch

from ..config import CTGANConfig
from ..features.layout import TableLayout
from . import ctgan

ALIGN = 32  # floats (128 B): slot base alignment


def round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


@dataclass
class Slot:
    name: str
    shape: tuple            # logical (reference) shape
    offset: int             # float offset into arena.data
    rows: int               # physical 2-D view: rows x ld
    cols: int               # physical columns carrying data+inner padding
    ld: int
    trainable: bool
    integer: bool = False   # exposed as int64 in the state_dict (num_batches_tracked)
    inner: tuple | None = None   # (pack, width, ld_width) for the packed first D layer

    @property
    def numel_physical(self):
        return self.rows * self.ld


class ParamArena:
    def __init__(self, device="cpu"):
        self.device = torch.device(device)
        self.slots: "OrderedDict[str, Slot]" = OrderedDict()
        self._order: list[str] = []          # state_dict key order
        self._pending = []
        self.data = None
        self.n_trainable = 0

    # ------------------------------------------------------------------ build
    def add(self, name, shape, *, ld=None, trainable=True, integer=False, inner=None):
        self._pending.append((name, tuple(shape), ld, trainable, integer, inner))
        self._order.append(name)

    def finalize(self, storage: torch.Tensor | None = None):
        """Lay out trainables first, then buffers; allocate (or adopt ``storage``)."""
        off = 0
        for want_trainable in (True, False):
            for name, shape, ld, trainable, integer, inner in self._pending:
                if trainable != want_trainable:
                    continue
                if len(shape) == 2:
                    rows = shape[0]
                    if inner is not None:
                        pack, width, ldw = inner
                        cols = pack * ldw
                    else:
                        cols = shape[1]
                    ld_ = ld if ld is not None else round_up(cols, 4)
                elif len(shape) == 1:
                    rows, cols = 1, shape[0]
                    ld_ = round_up(cols, 4)
                else:
                    rows, cols, ld_ = 1, 1, 4
                self.slots[name] = Slot(name, shape, off, rows, cols, ld_, trainable, integer, inner)
                off = round_up(off + rows * ld_, ALIGN)
            if want_trainable:
                self.n_trainable = off
        self.numel = off
        if storage is not None:
            assert storage.numel() >= self.numel and storage.dtype == torch.float32
            self.data = storage[: self.numel]
            self.data.zero_()
        else:
            self.data = torch.zeros(self.numel, dtype=torch.float32, device=self.device)
        self.device = self.data.device
        self.grad = torch.zeros(self.n_trainable, dtype=torch.float32, device=self.device)
        self.exp_avg = torch.zeros_like(self.grad)
        self.exp_avg_sq = torch.zeros_like(self.grad)
        self._index_cache = {}
        return self

    # ------------------------------------------------------------------ views
    def _view_of(self, base: torch.Tensor, name: str):
        s = self.slots[name]
        # as_strided offsets are absolute in the underlying storage: ``base`` may itself be a
        # slice of a larger block (D sits behind G in the symmetric allocation)
        off = base.storage_offset() + s.offset
        if len(s.shape) == 2:
            return torch.as_strided(base, (s.rows, s.cols), (s.ld, 1), off)
        if len(s.shape) == 1:
            return torch.as_strided(base, (s.cols,), (1,), off)
        return torch.as_strided(base, (1,), (1,), off)

    def view(self, name):
        """Physical view for kernels: 2-D ``[rows, cols]`` with row stride ``ld``."""
        return self._view_of(self.data, name)

    def grad_view(self, name):
        return self._view_of(self.grad, name)

    def int_slot_offsets(self):
        return [s.offset for s in self.slots.values() if s.integer]

    # ------------------------------------------------------- reference layout
    def _index(self, name) -> torch.Tensor:
        if name in self._index_cache:
            return self._index_cache[name]
        s = self.slots[name]
        if len(s.shape) == 2:
            r = torch.arange(s.shape[0]).view(-1, 1) * s.ld
            if s.inner is not None:
                pack, width, ldw = s.inner
                c = (torch.arange(pack).view(-1, 1) * ldw + torch.arange(width).view(1, -1)).reshape(1, -1)
            else:
                c = torch.arange(s.shape[1]).view(1, -1)
            idx = (s.offset + r + c).reshape(-1)
        elif len(s.shape) == 1:
            idx = s.offset + torch.arange(s.shape[0])
        else:
            idx = torch.tensor([s.offset])
        idx = idx.to(self.device)
        self._index_cache[name] = idx
        return idx

    def logical(self, name, base=None) -> torch.Tensor:
        s = self.slots[name]
        base = self.data if base is None else base
        t = base[self._index(name)].view(s.shape)
        if s.integer:
            t = t.to(torch.int64)
        return t

    def state_dict(self) -> "OrderedDict[str, torch.Tensor]":
        return OrderedDict((k, self.logical(k).clone()) for k in self._order)

    def load_state_dict(self, sd, strict=True):
        missing = [k for k in self._order if k not in sd]
        extra = [k for k in sd if k not in self.slots]
        if strict and (missing or extra):
            raise KeyError(f"state_dict mismatch: missing={missing} unexpected={extra}")
        with torch.no_grad():
            for k in self._order:
                if k not in sd:
                    continue
                s = self.slots[k]
                v = torch.as_tensor(sd[k]).to(device=self.device, dtype=torch.float32)
                if tuple(v.shape) != tuple(s.shape):
                    raise ValueError(f"{k}: shape {tuple(v.shape)} != {s.shape}")
                self.data[self._index(k)] = v.reshape(-1)

    def grad_state_dict(self):
        return OrderedDict((k, self.logical(k, self.grad).clone()) for k in self._order if self.slots[k].trainable)

    def optimizer_state(self, step: int):
        return {"step": int(step), "exp_avg": self.exp_avg.clone(), "exp_avg_sq": self.exp_avg_sq.clone()}

    def load_optimizer_state(self, st):
        self.exp_avg.copy_(st["exp_avg"].to(self.device))
        self.exp_avg_sq.copy_(st["exp_avg_sq"].to(self.device))
        return int(st["step"])


# ------------------------------------------------------------------- networks
class GeneratorArena(ParamArena):
    """Arena with the ``ctgan.Generator`` parameter set."""

    def __init__(self, cfg: CTGANConfig, layout: TableLayout, device="cpu", storage=None, init=True, seed=None):
        super().__init__(device)
        self.cfg, self.layout = cfg, layout
        self.d0 = cfg.embedding_dim + layout.n_opt
        self.hidden = list(cfg.gen_dim)
        for h in self.hidden:
            if h % 4:
                raise ValueError("gen_dim entries must be multiples of 4")
        self.sum_h = sum(self.hidden)
        width = self.d0
        for l, h in enumerate(self.hidden):
            self.add(f"seq.{l}.fc.weight", (h, width))
            self.add(f"seq.{l}.fc.bias", (h,))
            self.add(f"seq.{l}.bn.weight", (h,))
            self.add(f"seq.{l}.bn.bias", (h,))
            self.add(f"seq.{l}.bn.running_mean", (h,), trainable=False)
            self.add(f"seq.{l}.bn.running_var", (h,), trainable=False)
            self.add(f"seq.{l}.bn.num_batches_tracked", (), trainable=False, integer=True)
            width += h
        L = len(self.hidden)
        self.add(f"seq.{L}.weight", (layout.data_dim, width))
        self.add(f"seq.{L}.bias", (layout.data_dim,))
        self.finalize(storage)
        if init:
            self.load_state_dict(self.reference_module(seed=seed).state_dict())

    def reference_module(self, seed=None) -> ctgan.Generator:
        if seed is not None:
            torch.manual_seed(seed)
        return ctgan.Generator(self.d0, self.hidden, self.layout.data_dim)

    def to_module(self) -> ctgan.Generator:
        m = self.reference_module()
        m.load_state_dict({k: v.cpu() for k, v in self.state_dict().items()})
        return m


class DiscriminatorArena(ParamArena):
    """Arena with the ``ctgan.Discriminator`` parameter set (first layer stored packed)."""

    def __init__(self, cfg: CTGANConfig, layout: TableLayout, device="cpu", storage=None, init=True, seed=None):
        super().__init__(device)
        self.cfg, self.layout = cfg, layout
        self.din = layout.data_dim + layout.n_opt
        self.ld_din = round_up(self.din, 4)
        self.hidden = list(cfg.dis_dim)
        for h in self.hidden:
            if h % 4:
                raise ValueError("dis_dim entries must be multiples of 4")
        width = self.din * cfg.pack
        idx = 0
        for l, h in enumerate(self.hidden):
            inner = (cfg.pack, self.din, self.ld_din) if l == 0 else None
            self.add(f"seq.{idx}.weight", (h, width), inner=inner)
            self.add(f"seq.{idx}.bias", (h,))
            width = h
            idx += 3
        self.add(f"seq.{idx}.weight", (1, width))
        self.add(f"seq.{idx}.bias", (1,))
        self.layer_keys = [f"seq.{3 * l}" for l in range(len(self.hidden) + 1)]
        self.finalize(storage)
        if init:
            self.load_state_dict(self.reference_module(seed=seed).state_dict())

    def reference_module(self, seed=None) -> ctgan.Discriminator:
        if seed is not None:
            torch.manual_seed(seed)
        return ctgan.Discriminator(self.din, self.hidden, pack=self.cfg.pack,
                                   leaky_slope=self.cfg.leaky_slope, dropout_p=self.cfg.dropout_p)

    def to_module(self) -> ctgan.Discriminator:
        m = self.reference_module()
        m.load_state_dict({k: v.cpu() for k, v in self.state_dict().items()})
        return m


def arena_numel(cls, cfg, layout) -> int:
    """Number of floats an arena of type ``cls`` needs (for symmetric allocation sizing)."""
    return cls(cfg, layout, device="cpu", init=False).numel
