"""Synthetic-row generation (the federator's per-epoch ``sample``).

Reference: free function ``sample`` (``Server/dtds/distributed.py:160-208``) -- 81 CPU
batches of 500: ``z``, ``Cond.sample_zero``, eval-mode ``G`` (BN uses the *averaged*
running stats), ``apply_activate``, then ``transformer.inverse_transform``.  Here it is one
op program over large row chunks on the device (eval-mode BN is row independent, so chunk
size is free): empirical cond draw -> GEMM chain -> fused activate -> ``vgm_decode``
(SURVEY K16/K13).  Output is the numeric table ``[n, n_columns]`` (label ids for
categoricals); string decoding / CSV formatting stay on the host.
"""

from __future__ import annotations

import torch

from .. import ops
from ..config import CTGANConfig
from ..features.layout import TableLayout
from ..models.arena import GeneratorArena, round_up

S_GEN_Z, S_GEN_COL, S_GEN_GUMBEL = 32, 33, 35


class RowGenerator:
    def __init__(self, cfg: CTGANConfig, layout: TableLayout, decode_tables: dict, cum_emp, device,
                 chunk: int | None = None, seed: int = 1234):
        self.cfg, self.layout = cfg, layout
        self.device = torch.device(device)
        dev = self.device
        self.chunk = chunk or (8192 if dev.type == "cuda" else 2000)
        self.tables = layout.device_tables(dev)
        self.cum_emp = torch.as_tensor(cum_emp, dtype=torch.float32).to(dev)
        self.dec = {k: torch.as_tensor(v).to(dev) for k, v in decode_tables.items()}
        self.n_columns = int(self.dec["col_is_cont"].numel())
        self.rng = ops.RngState(seed, dev)
        E, n_opt, dd = cfg.embedding_dim, layout.n_opt, layout.data_dim
        self.gh = list(cfg.gen_dim)
        self.sum_h = sum(self.gh)
        self.d0 = E + n_opt
        L = len(self.gh)
        self.off = [0] * (L + 1)
        acc = 0
        for l in range(L, 0, -1):
            self.off[l] = acc
            acc += self.gh[l - 1]
        self.off[0] = acc
        C = self.chunk
        f32 = dict(dtype=torch.float32, device=dev)
        self.h = torch.zeros(C, round_up(self.sum_h + self.d0, 4), **f32)
        self.a = [torch.zeros(C, hl, **f32) for hl in self.gh]
        self.fake = torch.zeros(C, round_up(dd, 4), **f32)
        self.act = torch.zeros(C, round_up(dd, 4), **f32)
        self.col = torch.zeros(C, dtype=torch.int32, device=dev)
        self.opt = torch.zeros(C, dtype=torch.int32, device=dev)

    def _chunk(self, G: GeneratorArena, rows: int, out):
        cfg, lay = self.cfg, self.layout
        E, h0, dd = cfg.embedding_dim, self.off[0], lay.data_dim
        h = self.h[:rows]
        ops.fill_normal(h[:, h0:h0 + E], rng=self.rng, stream=S_GEN_Z)
        ops.sample_condvec(self.cum_emp, self.tables["soft_width"], self.col[:rows], self.opt[:rows],
                           rng=self.rng, stream=S_GEN_COL)
        ops.write_condvec(self.col[:rows], self.opt[:rows], self.tables["soft_cstart"],
                          h[:, h0 + E:h0 + E + lay.n_opt])
        L = len(self.gh)
        for l in range(1, L + 1):
            x = h[:, self.off[l - 1]: h0 + self.d0]
            ops.linear(x, G.view(f"seq.{l - 1}.fc.weight"), self.a[l - 1][:rows], bias=G.view(f"seq.{l - 1}.fc.bias"))
            ops.bn_relu_fwd(self.a[l - 1][:rows], G.view(f"seq.{l - 1}.bn.weight"), G.view(f"seq.{l - 1}.bn.bias"),
                            G.view(f"seq.{l - 1}.bn.running_mean"), G.view(f"seq.{l - 1}.bn.running_var"),
                            h[:, self.off[l]: self.off[l - 1]], None, None, training=False, eps=cfg.bn_eps)
        ops.linear(h[:, : h0 + self.d0], G.view(f"seq.{L}.weight"), self.fake[:rows, :dd], bias=G.view(f"seq.{L}.bias"))
        ops.activate_fwd(self.fake[:rows], self.tables, self.act[:rows], tau=cfg.tau, rng=self.rng, stream=S_GEN_GUMBEL)
        ops.vgm_decode(self.act[:rows], self.tables, self.dec["mu"], self.dec["sd"], self.dec["cat_ids"],
                       self.dec["cat_off"], self.dec["col_is_cont"], out)

    def generate(self, G: GeneratorArena, n: int, out: torch.Tensor | None = None) -> torch.Tensor:
        """Numeric synthetic table ``[n, n_columns]`` on the device."""
        if out is None:
            out = torch.empty(n, self.n_columns, dtype=torch.float32, device=self.device)
        done = 0
        while done < n:
            rows = min(self.chunk, n - done)
            self.rng.tick()
            self._chunk(G, rows, out[done:done + rows])
            done += rows
        return out
