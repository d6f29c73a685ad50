"""The closed-form train step must equal autograd on the reference-layout modules."""

import torch

from helpers import make_inject, oracle_iteration, small_setup


def _close(a, b, tol=2e-4):
    a, b = a.double(), b.double()
    return (a - b).abs().max().item() <= tol * (1 + b.abs().max().item())


def test_iteration_matches_autograd():
    torch.manual_seed(0)
    cfg, lay, table, eng = small_setup()
    G, D = eng.G.to_module(), eng.D.to_module()
    G.train(); D.train()
    optG = torch.optim.Adam(G.parameters(), lr=cfg.lr, betas=cfg.betas, weight_decay=cfg.l2scale)
    optD = torch.optim.Adam(D.parameters(), lr=cfg.lr, betas=cfg.betas)
    for it in range(3):
        inj = make_inject(eng, seed=10 + it)
        ref = oracle_iteration(eng, inj, G, D, optG, optD)
        eng.iteration(inj)
        got = eng.losses()
        for k in ("loss_d", "pen", "loss_g", "cond_loss"):
            assert abs(got[k] - ref[k]) <= 2e-4 * (1 + abs(ref[k])), (it, k, got[k], ref[k])
        gd = eng.D.grad_state_dict()
        for k, v in ref["d_grads"].items():
            assert _close(gd[k], v), (it, "D grad", k, (gd[k] - v).abs().max())
        gg = eng.G.grad_state_dict()
        for k, v in ref["g_grads"].items():
            assert _close(gg[k], v), (it, "G grad", k, (gg[k] - v).abs().max())
        for k, v in G.state_dict().items():
            assert _close(eng.G.state_dict()[k].float(), v.float(), 1e-4), (it, "G param", k)
        for k, v in D.state_dict().items():
            assert _close(eng.D.state_dict()[k].float(), v.float(), 1e-4), (it, "D param", k)


def test_arenas_inside_one_block_match_autograd():
    """G and D laid out back to back inside one allocation at a non-zero offset (how they sit in the
    peer-mapped block): every kernel-facing view must address its own arena, nothing outside."""
    cfg, lay, table, eng = small_setup(block=True)
    G, D = eng.G.to_module(), eng.D.to_module()
    G.train(); D.train()
    optG = torch.optim.Adam(G.parameters(), lr=cfg.lr, betas=cfg.betas, weight_decay=cfg.l2scale)
    optD = torch.optim.Adam(D.parameters(), lr=cfg.lr, betas=cfg.betas)
    assert eng.D.view("seq.0.weight").data_ptr() >= eng.D.data.data_ptr()
    for it in range(2):
        inj = make_inject(eng, seed=20 + it)
        ref = oracle_iteration(eng, inj, G, D, optG, optD)
        eng.iteration(inj)
        for k, v in ref["d_grads"].items():
            assert _close(eng.D.grad_state_dict()[k], v), (it, "D grad", k)
        for k, v in D.state_dict().items():
            assert _close(eng.D.state_dict()[k].float(), v.float(), 1e-4), (it, "D param", k)
        for k, v in G.state_dict().items():
            assert _close(eng.G.state_dict()[k].float(), v.float(), 1e-4), (it, "G param", k)
    guard = eng._block
    assert torch.isnan(guard[:64]).all() and torch.isnan(guard[-64:]).all()      # nothing spilled outside


def test_three_layer_nets():
    cfg, lay, table, eng = small_setup(gen_dim=(16, 24, 8), dis_dim=(16, 12, 8), batch=20, pack=5)
    G, D = eng.G.to_module(), eng.D.to_module()
    optG = torch.optim.Adam(G.parameters(), lr=cfg.lr, betas=cfg.betas, weight_decay=cfg.l2scale)
    optD = torch.optim.Adam(D.parameters(), lr=cfg.lr, betas=cfg.betas)
    inj = make_inject(eng, seed=3)
    ref = oracle_iteration(eng, inj, G, D, optG, optD)
    eng.iteration(inj)
    gd = eng.D.grad_state_dict()
    for k, v in ref["d_grads"].items():
        assert _close(gd[k], v), ("D grad", k)
    gg = eng.G.grad_state_dict()
    for k, v in ref["g_grads"].items():
        assert _close(gg[k], v), ("G grad", k)


def test_free_running_iteration_is_finite():
    cfg, lay, table, eng = small_setup()
    eng.train(5)
    out = eng.losses()
    assert all(abs(v) < 1e6 for v in out.values())
    assert int(eng.G.state_dict()["seq.0.bn.num_batches_tracked"]) == 10
