"""The CTGAN local training step as an explicit op program (no autograd).

Reference semantics: ``MDGANClient.train_model`` (``Server/dtds/distributed.py:328-417``),
one discriminator update followed by one generator update per iteration:

  D step   z, condvec, perm, real rows -> G(train-mode BN) -> apply_activate ->
           D(fake), D(real) with independent dropout -> WGAN loss + slerp gradient penalty
           (``ctgan.py:240-258``) -> Adam(D)
  G step   z, condvec -> G -> apply_activate -> D(fake) -> -mean + cond_loss
           (``ctgan.py:174-194``) -> Adam(G, weight_decay=l2scale)

B200 design decisions (vs. the reference's autograd graph of ~300 tiny launches):

* every buffer is preallocated once; ``torch.cat`` disappears because producers write
  straight into column windows of wider buffers (``h = [r_L | ... | r_1 | z | c1]``,
  ``X = [fake_cat ; real_cat ; interpolates]``);
* fake, real and interpolated batches go through the discriminator as ONE 3*B/pack-row
  batch;
* the backward pass is closed form.  In particular the gradient-penalty double backward is
  a short chain of GEMMs that re-uses the dropout/leaky masks of the interpolates'
  forward pass (derivation in DESIGN.md), and the WGAN and penalty weight-gradients are
  produced by the *same* GEMM by stacking their operands along K;
* all randomness is counter-based and keyed by a device-resident step counter, so the
  whole iteration is CUDA-graph capturable.

The same program runs on CPU through ``ops.reference`` (that is what the tests compare
against autograd) and on sm_100a through ``ops.native``.
"""

from __future__ import annotations

import torch

from .. import ops
from ..config import CTGANConfig
from ..features.layout import TableLayout
from ..models.arena import DiscriminatorArena, GeneratorArena, round_up
from .table import DeviceTable

# random-stream ids (one per consumer inside an iteration)
S_Z_D, S_COL_D, S_OPT_D, S_PERM, S_ROW, S_GUMBEL_D, S_ALPHA, S_DROP_D = range(8)
S_Z_G, S_COL_G, S_OPT_G, S_GUMBEL_G, S_DROP_G = range(16, 21)


class _Branch:
    """``with engine.branch(k):`` runs the enclosed launches on side stream ``k`` after forking
    from the current stream; ``engine.join(k)`` makes the current stream wait for them.  In a
    CUDA-graph capture this turns independent kernels into parallel graph branches."""

    def __init__(self, eng, k):
        self.eng, self.k, self.ctx = eng, k, None

    def __enter__(self):
        e = self.eng
        if not e.concurrent:
            return self
        side = e.side_streams[self.k]
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream(e.device))
        side.wait_event(ev)
        self.ctx = torch.cuda.stream(side)
        self.ctx.__enter__()
        return self

    def __exit__(self, *exc):
        if self.ctx is not None:
            self.ctx.__exit__(*exc)
        return False


class TrainEngine:
    def __init__(self, cfg: CTGANConfig, layout: TableLayout, table: DeviceTable | None, device=None,
                 g_arena: GeneratorArena | None = None, d_arena: DiscriminatorArena | None = None,
                 seed: int | None = None):
        self.cfg, self.layout, self.table = cfg, layout, table
        self.device = torch.device(device if device is not None else (table.device if table else "cpu"))
        dev = self.device
        seed = cfg.seed if seed is None else seed
        if layout.n_col == 0:
            raise ValueError("table has no one-hot span: conditional sampling is undefined")
        self.G = g_arena if g_arena is not None else GeneratorArena(cfg, layout, dev, seed=seed)
        self.D = d_arena if d_arena is not None else DiscriminatorArena(cfg, layout, dev, seed=seed + 1)
        self.rng = ops.RngState(seed, dev)
        self.tables = layout.device_tables(dev)
        # independent kernels of an iteration run on side streams (parallel graph branches)
        self.concurrent = dev.type == "cuda"
        self.side_streams = [torch.cuda.Stream(device=dev, priority=-1) for _ in range(3)] if self.concurrent else []

        B, P = cfg.batch_size, cfg.pack
        Bp = B // P
        E, dd, n_opt, n_col = cfg.embedding_dim, layout.data_dim, layout.n_opt, layout.n_col
        self.B, self.Bp, self.E = B, Bp, E
        f32 = dict(dtype=torch.float32, device=dev)
        i32 = dict(dtype=torch.int32, device=dev)

        # ---- generator side
        self.gh = list(cfg.gen_dim)
        self.sum_h = sum(self.gh)
        self.d0 = E + n_opt
        L = len(self.gh)
        # off[l] = first column of r_l's window; off[0] = start of h0; layer l (1-based) reads
        # columns [off[l-1]:] and writes [off[l]:off[l-1]]
        self.off = [0] * (L + 1)
        acc = 0
        for l in range(L, 0, -1):
            self.off[l] = acc
            acc += self.gh[l - 1]
        self.off[0] = acc
        self.ld_h = round_up(self.sum_h + self.d0, 4)
        self.ld_dd = round_up(dd, 4)
        self.h = torch.zeros(B, self.ld_h, **f32)
        self.a = [torch.zeros(B, hl, **f32) for hl in self.gh]
        self.save_mean = [torch.zeros(hl, **f32) for hl in self.gh]
        self.save_invstd = [torch.zeros(hl, **f32) for hl in self.gh]
        self.fake = torch.zeros(B, self.ld_dd, **f32)
        self.dfake = torch.zeros(B, self.ld_dd, **f32)
        self.dh = torch.zeros(B, max(self.sum_h, 4), **f32)
        self.da = [torch.zeros(B, hl, **f32) for hl in self.gh]
        # ---- conditional vector / sampler state
        self.col = torch.zeros(B, **i32)
        self.opt = torch.zeros(B, **i32)
        self.perm = torch.zeros(B, **i32)
        self.rid = torch.zeros(B, **i32)
        self.m1 = torch.zeros(B, n_col, **f32)
        # ---- discriminator side
        self.din = dd + n_opt
        self.ld_din = self.D.ld_din
        self.Din = P * self.ld_din
        self.dhid = list(cfg.dis_dim)
        self.X = torch.zeros(3 * B, self.ld_din, **f32)
        self.Xp = self.X.view(3 * Bp, self.Din)
        self.Lact = [torch.zeros(3 * Bp, hl, **f32) for hl in self.dhid]
        self.Mask = [torch.zeros(3 * Bp, hl, **f32) for hl in self.dhid]
        self.Z = [torch.zeros(3 * Bp, hl, **f32) for hl in self.dhid]
        self.y = torch.zeros(3 * Bp, **f32)
        self.dX = torch.zeros(B, self.ld_din, **f32)
        self.dXp = self.dX.view(Bp, self.Din)
        rs = torch.empty(3 * Bp, **f32)
        rs[:Bp] = 1.0 / Bp
        rs[Bp:2 * Bp] = -1.0 / Bp
        rs[2 * Bp:] = 1.0
        self.rowscale_d = rs
        rl = rs.clone()
        rl[2 * Bp:] = 0.0           # interpolated rows do not enter the WGAN loss value
        self.rowscale_loss = rl
        self.rowscale_g = torch.full((Bp,), -1.0 / Bp, **f32)
        self.alpha = torch.zeros(B, **f32)
        # ---- scalars (device resident; read back only on request)
        self.loss_d = torch.zeros(1, **f32)
        self.pen = torch.zeros(1, **f32)
        self.loss_g_adv = torch.zeros(1, **f32)
        self.ce = torch.zeros(1, **f32)
        self.iterations = 0
        self.dmid_counter = torch.zeros(1, dtype=torch.int32, device=dev)     # last-CTA counter of the fused D middle

    # ------------------------------------------------------------------ helpers
    def branch(self, k):
        return _Branch(self, k)

    def join(self, k):
        if self.concurrent:
            ev = torch.cuda.Event()
            ev.record(self.side_streams[k])
            torch.cuda.current_stream(self.device).wait_event(ev)

    def _gw(self, l):
        return self.G.view(f"seq.{l}.fc.weight")

    def _g(self, l, what):
        return self.G.view(f"seq.{l}.{what}")

    def _dw(self, l):
        return self.D.view(f"{self.D.layer_keys[l]}.weight")

    def _db(self, l):
        return self.D.view(f"{self.D.layer_keys[l]}.bias")

    def _rand(self, inject, key, shape):
        if inject is not None and key in inject:
            return inject[key]
        return None

    # ------------------------------------------------------------- sampling prep
    def _prep(self, inject, prefix, with_real):
        B, E = self.B, self.E
        t = self.table
        h0 = self.off[0]
        rng = self.rng
        if self._fused_prep(inject):
            return self._prep_fused(with_real)
        R = lambda k: self._rand(inject, k, None)  # noqa: E731
        ops.fill_normal(self.h[:, h0:h0 + E], z=R(f"z_{prefix}"), rng=rng, stream=S_Z_D if with_real else S_Z_G)
        ops.sample_condvec(t.cum_p, self.tables["soft_width"], self.col, self.opt,
                           u_col=R(f"u_col_{prefix}"), u_opt=R(f"u_opt_{prefix}"), rng=rng,
                           stream=S_COL_D if with_real else S_COL_G)
        c1 = self.h[:, h0 + E:h0 + E + self.layout.n_opt]
        ops.write_condvec(self.col, self.opt, self.tables["soft_cstart"], c1, None if with_real else self.m1)
        if with_real:
            ops.argsort_perm(self.perm, u=R("u_perm"), rng=rng, stream=S_PERM)
            ops.sample_rows(t.row_ptr, t.row_ids, self.tables["soft_cstart"], self.col, self.opt, self.rid,
                            perm=self.perm, u=R("u_row"), rng=rng, stream=S_ROW)
            real = self.X[B:2 * B]
            ops.expand_rows(t.vals, t.codes, self.rid, self.tables, real)
            dd = self.layout.data_dim
            ops.write_condvec(self.col, self.opt, self.tables["soft_cstart"],
                              real[:, dd:dd + self.layout.n_opt], None, self.perm)
        return c1

    def _fused_prep(self, inject):
        return (inject is None and self.device.type == "cuda" and ops.native_available()
                and not ops._force_reference())

    def _prep_fused(self, with_real):
        from ..ops import native

        B, E, h0 = self.B, self.E, self.off[0]
        z = self.h[:, h0:h0 + E]
        c1 = self.h[:, h0 + E:h0 + E + self.layout.n_opt]
        if with_real:
            native.prep_d_batch(self.table, self.tables, z, c1, self.X[B:2 * B], self.col, self.opt, self.perm,
                                self.rid, self.rng, S_Z_D, S_COL_D, S_PERM, S_ROW, tick=True)
        else:
            native.prep_g_batch(self.table, self.tables, z, c1, self.m1, self.col, self.opt, self.rng, S_Z_G, S_COL_G)
        return c1

    # --------------------------------------------------------------- generator
    def g_forward(self, training=True, h=None, a=None, fake=None):
        h = self.h if h is None else h
        a = self.a if a is None else a
        fake = self.fake if fake is None else fake
        L = len(self.gh)
        for l in range(1, L + 1):
            x = h[:, self.off[l - 1]: self.off[0] + self.d0]
            if training:
                # Linear + BatchNorm(batch stats) + ReLU + concat-store: one launch (GEMM epilogue)
                ops.linear_bn_relu(
                    x, self._gw(l - 1), a[l - 1], h[:, self.off[l]: self.off[l - 1]],
                    bias=self._g(l - 1, "fc.bias"), gamma=self._g(l - 1, "bn.weight"), beta=self._g(l - 1, "bn.bias"),
                    running_mean=self._g(l - 1, "bn.running_mean"), running_var=self._g(l - 1, "bn.running_var"),
                    save_mean=self.save_mean[l - 1], save_invstd=self.save_invstd[l - 1], eps=self.cfg.bn_eps,
                    momentum=self.cfg.bn_momentum, num_batches_tracked=self._g(l - 1, "bn.num_batches_tracked"))
                continue
            ops.linear(x, self._gw(l - 1), a[l - 1], bias=self._g(l - 1, "fc.bias"))
            ops.bn_relu_fwd(
                a[l - 1], self._g(l - 1, "bn.weight"), self._g(l - 1, "bn.bias"),
                self._g(l - 1, "bn.running_mean"), self._g(l - 1, "bn.running_var"),
                h[:, self.off[l]: self.off[l - 1]],
                self.save_mean[l - 1] if training else None, self.save_invstd[l - 1] if training else None,
                training=training, eps=self.cfg.bn_eps, momentum=self.cfg.bn_momentum,
                num_batches_tracked=self._g(l - 1, "bn.num_batches_tracked") if training else None,
            )
        ops.linear(h[:, : self.off[0] + self.d0], self.G.view(f"seq.{L}.weight"), fake[:, : self.layout.data_dim],
                   bias=self.G.view(f"seq.{L}.bias"))
        return fake

    def g_backward(self):
        """Given ``dfake`` (grad wrt pre-activation output), fill ``G.grad``."""
        L = len(self.gh)
        dd = self.layout.data_dim
        dfake = self.dfake[:, :dd]
        hfull = self.h[:, : self.off[0] + self.d0]
        with self.branch(0):        # weight/bias gradient of the output layer: off the critical path
            ops.linear(dfake, hfull, self.G.grad_view(f"seq.{L}.weight"), a_t=True, b_t=True)
            ops.colsum(dfake, self.G.grad_view(f"seq.{L}.bias"))
        dh = self.dh[:, : self.sum_h]
        ops.linear(dfake, self.G.view(f"seq.{L}.weight")[:, : self.sum_h], dh, b_t=True)
        for l in range(L, 0, -1):
            lo, hi = self.off[l], self.off[l - 1]
            ops.bn_relu_bwd(
                dh[:, lo:hi], self.a[l - 1], self.save_mean[l - 1], self.save_invstd[l - 1],
                self._g(l - 1, "bn.weight"), self._g(l - 1, "bn.bias"), self.da[l - 1],
                self.G.grad_view(f"seq.{l - 1}.bn.weight"), self.G.grad_view(f"seq.{l - 1}.bn.bias"),
                dbias_fc=self.G.grad_view(f"seq.{l - 1}.fc.bias"),
            )
            x = self.h[:, hi: self.off[0] + self.d0]
            if l > 1:
                with self.branch(1):    # wgrad in parallel with the dgrad that feeds the next layer
                    ops.linear(self.da[l - 1], x, self.G.grad_view(f"seq.{l - 1}.fc.weight"), a_t=True, b_t=True)
                width = self.sum_h - hi      # columns of W_l that multiply r_{l-1}..r_1
                ops.linear(self.da[l - 1], self._gw(l - 1)[:, :width], dh[:, hi: self.sum_h], b_t=True,
                           accumulate=True)
            else:
                ops.linear(self.da[l - 1], x, self.G.grad_view(f"seq.{l - 1}.fc.weight"), a_t=True, b_t=True)
        self.join(0)
        self.join(1)

    # ------------------------------------------------------------ discriminator
    def _fused_dmid(self):
        """Layer 1 + head + both backward signals as ONE row-local fp32 kernel (``ftg_d_mid``): standard two-hidden-
        layer discriminator on the device with hidden widths <= 512.  Opt-in (``FEDTGAN_FUSED_DMID=1``): it removes two
        launches per D pass but is not faster than them yet."""
        if len(self.dhid) != 2 or self.device.type != "cuda" or not ops.native_available() or ops._force_reference():
            return False
        from ..ops import native

        return native.d_mid_enabled() and native.d_mid_supported(self.dhid[0], self.dhid[1])

    def d_forward(self, rows, inject, key, stream, rowscale=None, loss_out=None, rs_sig=None):
        """Forward rows ``[0:rows)`` of the packed batch; fills Lact/Mask/y, the fused scalar
        ``loss_out = sum rowscale * y`` and the head's backward signal ``Z_{L-1}`` (with the fused middle kernel
        also ``Z_0``: then ``d_signals`` has nothing left to do)."""
        cfg = self.cfg
        x = self.Xp[:rows]
        nl = len(self.dhid)
        fused = self._fused_dmid()
        self._signals_done = False
        for l, hl in enumerate(self.dhid):
            u = None
            if inject is not None and key in inject:
                u = inject[key][l]
            if fused and l == 1:
                ops.d_mid(x, self._dw(1), self._db(1), self.Lact[1][:rows], self.Mask[1][:rows], self.Mask[0][:rows],
                          self._dw(nl).view(-1), self._db(nl), self.y[:rows], rowscale, loss_out, rs_sig,
                          self.Z[1][:rows], self.Z[0][:rows], u=u, rng=self.rng, stream=stream + 1,
                          slope=cfg.leaky_slope, p=cfg.dropout_p, counter=self.dmid_counter)
                self._signals_done = True
                return
            ops.linear(x, self._dw(l), self.Lact[l][:rows], bias=self._db(l), act="leaky_dropout",
                       mask_out=self.Mask[l][:rows], u=u, rng=self.rng, stream=stream + l,
                       slope=cfg.leaky_slope, p=cfg.dropout_p)
            x = self.Lact[l][:rows]
        ops.d_head(x, self._dw(nl), self._db(nl), self.y[:rows], rowscale, loss_out, rs_sig,
                   self.Mask[nl - 1][:rows], self.Z[nl - 1][:rows])

    def d_signals(self, rows, rowscale):
        """Backward signals ``Z_l = dLoss/d(pre-activation_l)`` for ``loss = sum rowscale*y``."""
        if getattr(self, "_signals_done", False):
            return
        nl = len(self.dhid)                 # Z[nl-1] was produced by d_head inside d_forward
        for l in range(nl - 1, 0, -1):
            ops.linear(self.Z[l][:rows], self._dw(l), self.Z[l - 1][:rows], b_t=True, mask_in=self.Mask[l - 1][:rows])

    # ------------------------------------------------------------------- steps
    def d_step(self, inject=None, prefetch_g=False):
        cfg, B, Bp = self.cfg, self.B, self.Bp
        dd, n_opt, din = self.layout.data_dim, self.layout.n_opt, self.din
        c1 = self._prep(inject, "d", with_real=True)
        self.g_forward(training=True)
        ops.activate_fwd(self.fake, self.tables, self.X[:B], tau=cfg.tau, u=self._rand(inject, "u_gumbel_d", None),
                         rng=self.rng, stream=S_GUMBEL_D, cond=c1, slerp_real=self.X[B:2 * B, :din],
                         slerp_out=self.X[2 * B:, :din], alpha=self._rand(inject, "alpha", None), slerp_stream=S_ALPHA)
        if prefetch_g:
            # The G step's batch preparation and generator forward only depend on buffers this D
            # step is done with (h, a, fake) and on G's weights, which the D step never changes:
            # run them on a side stream underneath the discriminator's forward/backward.
            with self.branch(2):
                self._g_c1 = self._prep(inject, "g", with_real=False)
                self.g_forward(training=True)
        self.d_forward(3 * Bp, inject, "u_drop_d", S_DROP_D, self.rowscale_loss, self.loss_d, self.rowscale_d)
        # backward signals for all rows (interp rows carry the d y / d x seed)
        self.d_signals(3 * Bp, self.rowscale_d)
        nl = len(self.dhid)
        I = slice(2 * Bp, 3 * Bp)
        # input-gradient of the interpolates, penalty, and its gradient seed (in place)
        ops.linear(self.Z[0][I], self._dw(0), self.Xp[I], b_t=True)
        ops.gp_penalty(self.Xp[I], self.pen, lambda_=cfg.gp_lambda)
        # tangent pass: F_l = (In_l @ V_l^T) * Mask_l, stored over the interp rows of Lact
        src = self.Xp[I]
        for l in range(nl):
            ops.linear(src, self._dw(l), self.Lact[l][I], mask_in=self.Mask[l][I])
            src = self.Lact[l][I]
        # weight gradients: WGAN rows and penalty rows stacked along K; the per-layer GEMMs and the
        # bias / head reductions are independent of each other -> parallel branches
        with self.branch(0):
            inp = self.Xp
            for l in range(nl):
                if l > 0:
                    ops.linear(self.Z[l], inp, self.D.grad_view(f"{self.D.layer_keys[l]}.weight"), a_t=True, b_t=True)
                inp = self.Lact[l]
        with self.branch(1):
            for l in range(nl):
                ops.colsum(self.Z[l][:2 * Bp], self.D.grad_view(f"{self.D.layer_keys[l]}.bias"))
            ops.colsum(self.Lact[nl - 1], self.D.grad_view(f"{self.D.layer_keys[nl]}.weight").view(-1),
                       rowscale=self.rowscale_d)
        ops.linear(self.Z[0], self.Xp, self.D.grad_view(f"{self.D.layer_keys[0]}.weight"), a_t=True, b_t=True)
        self.join(0)
        self.join(1)
        # d loss_d / d b_head = sum(rowscale over fake+real) = 0 exactly: its grad slot stays 0
        ops.adam_step(self.D.data[: self.D.n_trainable], self.D.grad, self.D.exp_avg, self.D.exp_avg_sq,
                      self.rng.state[1:2], lr=cfg.lr, beta1=cfg.betas[0], beta2=cfg.betas[1], eps=cfg.adam_eps)

    def g_step(self, inject=None, prefetched=False):
        cfg, B, Bp = self.cfg, self.B, self.Bp
        dd = self.layout.data_dim
        if prefetched:
            self.join(2)
            c1 = self._g_c1
        else:
            c1 = self._prep(inject, "g", with_real=False)
            self.g_forward(training=True)
        ops.activate_fwd(self.fake, self.tables, self.X[:B], tau=cfg.tau, u=self._rand(inject, "u_gumbel_g", None),
                         rng=self.rng, stream=S_GUMBEL_G, cond=c1)
        self.d_forward(Bp, inject, "u_drop_g", S_DROP_G, self.rowscale_g, self.loss_g_adv, self.rowscale_g)
        self.d_signals(Bp, self.rowscale_g)
        ops.linear(self.Z[0][:Bp], self._dw(0), self.dXp, b_t=True)
        ops.gen_output_bwd(self.dX[:, :dd], self.X[:B, :dd], self.fake, c1, self.m1, self.tables, self.dfake, self.ce,
                           tau=cfg.tau)
        self.g_backward()
        ops.adam_step(self.G.data[: self.G.n_trainable], self.G.grad, self.G.exp_avg, self.G.exp_avg_sq,
                      self.rng.state[1:2], lr=cfg.lr, beta1=cfg.betas[0], beta2=cfg.betas[1], eps=cfg.adam_eps,
                      weight_decay=cfg.l2scale)

    def iteration(self, inject=None):
        if not self._fused_prep(inject):
            self.rng.tick()          # (the fused D-step prep kernel advances the step counter itself)
        overlap = self.concurrent
        self.d_step(inject, prefetch_g=overlap)
        self.g_step(inject, prefetched=overlap)
        self.iterations += 1

    def train(self, steps: int):
        for _ in range(int(steps)):
            self.iteration()

    # ----------------------------------------------------------------- logging
    def losses(self):
        """Host copy of the last iteration's losses (synchronises)."""
        ld, pen, adv, ce = (float(t.item()) for t in (self.loss_d, self.pen, self.loss_g_adv, self.ce))
        return {"loss_d": ld, "pen": pen, "loss_g": adv + ce, "cond_loss": ce}
