"""Shared test helpers: tiny synthetic tables and an autograd oracle of one CTGAN iteration."""

import numpy as np
import torch
from torch.nn import functional as F

from fed_tgan_b200.config import CTGANConfig
from fed_tgan_b200.engine import DeviceTable, TrainEngine
from fed_tgan_b200.features.layout import TableLayout
from fed_tgan_b200.models import ctgan


def small_setup(device="cpu", batch=40, pack=10, rows=300, seed=0, gen_dim=(32, 32), dis_dim=(24, 16), emb=12,
                block=False):
    rng = np.random.default_rng(seed)
    output_info = [(1, "tanh"), (3, "softmax"), (5, "softmax"), (1, "tanh"), (2, "softmax"), (4, "softmax")]
    layout = TableLayout(output_info)
    vals = rng.uniform(-0.9, 0.9, size=(rows, layout.n_tanh)).astype(np.float32)
    codes = np.stack([rng.integers(0, w, size=rows) for w in layout.soft_width], axis=1).astype(np.int32)
    cfg = CTGANConfig(embedding_dim=emb, gen_dim=gen_dim, dis_dim=dis_dim, batch_size=batch, pack=pack, seed=seed)
    table = DeviceTable(vals, codes, layout, device)
    if not block:
        eng = TrainEngine(cfg, layout, table, device)
        return cfg, layout, table, eng
    # both arenas inside one allocation, behind a non-zero offset (the symmetric-memory layout)
    from fed_tgan_b200.models.arena import DiscriminatorArena, GeneratorArena

    probe = TrainEngine(cfg, layout, table, device)
    ng, nd = probe.G.numel, probe.D.numel
    buf = torch.full((64 + ng + nd + 64,), float("nan"), device=device)
    G = GeneratorArena(cfg, layout, device, storage=buf[64: 64 + ng], seed=seed)
    D = DiscriminatorArena(cfg, layout, device, storage=buf[64 + ng: 64 + ng + nd], seed=seed + 1)
    eng = TrainEngine(cfg, layout, table, device, G, D, seed=seed)
    eng.G.data.copy_(probe.G.data)
    eng.D.data.copy_(probe.D.data)
    eng._block = buf
    return cfg, layout, table, eng


def make_inject(eng, seed=1):
    g = torch.Generator().manual_seed(seed)
    B, Bp, E = eng.B, eng.Bp, eng.E
    dd = eng.layout.data_dim
    dev = eng.device
    U = lambda *s: torch.rand(*s, generator=g).to(dev)  # noqa: E731
    N = lambda *s: torch.randn(*s, generator=g).to(dev)  # noqa: E731
    return {
        "z_d": N(B, E), "u_col_d": U(B), "u_opt_d": U(B), "u_perm": U(B), "u_row": U(B),
        "u_gumbel_d": U(B, dd), "alpha": U(B),
        "u_drop_d": [U(3 * Bp, h) for h in eng.dhid],
        "z_g": N(B, E), "u_col_g": U(B), "u_opt_g": U(B), "u_gumbel_g": U(B, dd),
        "u_drop_g": [U(Bp, h) for h in eng.dhid],
    }


def d_forward_masked(D, x, us, p=0.5):
    """ctgan.Discriminator forward with explicit dropout uniforms (keep = u >= p)."""
    h = x.view(-1, D.packdim)
    mods = list(D.seq)
    li = 0
    i = 0
    while i < len(mods) - 1:
        h = mods[i](h)
        h = mods[i + 1](h)
        h = h * (us[li] >= p).to(h.dtype) / (1 - p)
        li += 1
        i += 3
    return mods[-1](h)


def gumbel_activate(fake, output_info, u, tau):
    noise = -torch.log(-torch.log(u.clamp(1e-10, 1 - 1e-7)))
    parts, st = [], 0
    for w, k in output_info:
        if k == "tanh":
            parts.append(torch.tanh(fake[:, st:st + w]))
        else:
            parts.append(torch.softmax((fake[:, st:st + w] + noise[:, st:st + w]) / tau, dim=1))
        st += w
    return torch.cat(parts, 1)


def oracle_iteration(eng, inj, G, D, optG, optD):
    """One reference-semantics iteration (distributed.py:334-413) driven by injected randoms.
    Returns dict of losses; G/D/opt are updated in place."""
    cfg, lay, t = eng.cfg, eng.layout, eng.table
    B, Bp = eng.B, eng.Bp
    dev = eng.device
    tabs = lay.device_tables(dev)
    cst = tabs["soft_cstart"].long()
    sw = tabs["soft_width"].long()

    def condvec(u_col, u_opt):
        col = torch.clamp((u_col * lay.n_col).long(), max=lay.n_col - 1)
        opt = (t.cum_p[col] > u_opt.view(-1, 1)).float().argmax(1)
        opt = torch.minimum(opt, sw[col] - 1)
        c = torch.zeros(B, lay.n_opt, device=dev)
        c[torch.arange(B), cst[col] + opt] = 1
        m = torch.zeros(B, lay.n_col, device=dev)
        m[torch.arange(B), col] = 1
        return c, m, col, opt

    dense = t.dense().to(dev)
    # ---- D step
    c1, m1, col, opt = condvec(inj["u_col_d"], inj["u_opt_d"])
    fakez = torch.cat([inj["z_d"], c1], 1)
    perm = torch.argsort(inj["u_perm"])
    g = cst[col[perm]] + opt[perm]
    lo = t.row_ptr.long()[g]
    cnt = t.row_ptr.long()[g + 1] - lo
    k = torch.clamp((inj["u_row"] * cnt).long(), max=(cnt - 1).clamp_min(0))
    rid = t.row_ids.long()[lo + k]
    real = dense[rid]
    c2 = c1[perm]
    fake = G(fakez)
    fakeact = gumbel_activate(fake, lay.output_info, inj["u_gumbel_d"], cfg.tau)
    fake_cat = torch.cat([fakeact, c1], 1)
    real_cat = torch.cat([real, c2], 1)
    us = inj["u_drop_d"]
    y_fake = d_forward_masked(D, fake_cat, [u[:Bp] for u in us])
    y_real = d_forward_masked(D, real_cat, [u[Bp:2 * Bp] for u in us])
    loss_d = -(y_real.mean() - y_fake.mean())
    interp = ctgan.slerp(inj["alpha"].view(-1, 1), real_cat, fake_cat)
    disc = d_forward_masked(D, interp, [u[2 * Bp:] for u in us])
    grads = torch.autograd.grad(disc, interp, torch.ones_like(disc), create_graph=True, retain_graph=True)[0]
    pen = ((grads.view(-1, cfg.pack * real_cat.size(1)).norm(2, dim=1) - 1) ** 2).mean() * cfg.gp_lambda
    optD.zero_grad()
    (loss_d + pen).backward()
    d_grads = {k: v.grad.clone() for k, v in D.named_parameters()}
    optD.step()
    # ---- G step
    c1, m1, col, opt = condvec(inj["u_col_g"], inj["u_opt_g"])
    fakez = torch.cat([inj["z_g"], c1], 1)
    fake = G(fakez)
    fakeact = gumbel_activate(fake, lay.output_info, inj["u_gumbel_g"], cfg.tau)
    y_fake = d_forward_masked(D, torch.cat([fakeact, c1], 1), inj["u_drop_g"])
    ce = ctgan.cond_loss(fake, lay.output_info, c1, m1)
    loss_g = -y_fake.mean() + ce
    optG.zero_grad()
    loss_g.backward()
    g_grads = {k: v.grad.clone() for k, v in G.named_parameters()}
    optG.step()
    return {"loss_d": loss_d.item(), "pen": pen.item(), "loss_g": loss_g.item(), "cond_loss": ce.item(),
            "d_grads": d_grads, "g_grads": g_grads}
