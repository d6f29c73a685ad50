"""sm_100a kernels vs the fp32 PyTorch oracle of the same op (``ops.reference``)."""

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from fed_tgan_b200.ops import reference as R  # noqa: E402


@pytest.fixture(scope="module")
def N():
    from fed_tgan_b200.ops import native

    return native


def dev():
    return torch.device("cuda:0")


def rnd(*shape, seed=0, scale=1.0):
    g = torch.Generator().manual_seed(seed)
    return (torch.randn(*shape, generator=g) * scale).to(dev())


def uni(*shape, seed=0):
    g = torch.Generator().manual_seed(seed)
    return torch.rand(*shape, generator=g).to(dev())


def padded(rows, cols, ld, seed=0):
    """[rows, cols] view with row stride ld (multiple of 4)."""
    buf = rnd(rows, ld, seed=seed)
    return buf[:, :cols]


def close(a, b, rtol, atol):
    err = (a.double() - b.double()).abs().max().item()
    ref = b.double().abs().max().item()
    assert err <= atol + rtol * ref, (err, ref)


GEMM_SHAPES = [
    # M, N, K, a_t, b_t
    (500, 256, 368, False, False), (50, 256, 5020, False, False), (150, 256, 256, False, True),
    (256, 5020, 150, True, True), (50, 5020, 256, False, True), (500, 262, 880, False, False),
    (262, 880, 500, True, True), (500, 512, 262, False, True), (37, 19, 45, False, False),
    (129, 65, 33, True, False), (64, 64, 32, False, False),
]


def _gemm_case(N, M, Nn, K, a_t, b_t, prec, **epi):
    r4 = lambda x: (x + 3) // 4 * 4  # noqa: E731
    a = padded(K, M, r4(M) + 4, 1) if a_t else padded(M, K, r4(K) + 4, 1)
    b = padded(K, Nn, r4(Nn), 2) if b_t else padded(Nn, K, r4(K), 2)
    out_n = padded(M, Nn, r4(Nn) + 8, 3).clone()
    out_r = out_n.clone()
    kw_n, kw_r = dict(epi), dict(epi)
    if epi.get("mask_out") is not None:
        kw_n["mask_out"] = torch.zeros(M, Nn, device=dev())
        kw_r["mask_out"] = torch.zeros(M, Nn, device=dev())
    N.linear(a, b, out_n, a_t=a_t, b_t=b_t, precision=prec, **kw_n)
    R.linear(a, b, out_r, a_t=a_t, b_t=b_t, **kw_r)
    torch.cuda.synchronize()
    # fp32x3 (3xTF32 split on the tensor cores) is held to the SAME tolerance as the exact-fp32 kernel
    tol = dict(rtol=2e-5, atol=1e-4) if prec in ("fp32", "fp32x3") else dict(rtol=4e-3, atol=4e-3 * (K ** 0.5) / 8)
    close(out_n, out_r, **tol)
    return kw_n, kw_r, tol


@pytest.mark.parametrize("shape", GEMM_SHAPES, ids=lambda s: "x".join(map(str, s)))
def test_gemm_fp32(N, shape):
    _gemm_case(N, *shape, "fp32")


@pytest.mark.parametrize("shape", GEMM_SHAPES, ids=lambda s: "x".join(map(str, s)))
def test_gemm_tf32_tcgen05(N, shape):
    _gemm_case(N, *shape, "tf32")


@pytest.mark.parametrize("shape", GEMM_SHAPES, ids=lambda s: "x".join(map(str, s)))
def test_gemm_fp32x3_tcgen05(N, shape):
    """3xTF32 hi/lo split on tcgen05 reproduces fp32 products (reference: torch fp32 matmul)."""
    before = N.simt_fallback_count()
    _gemm_case(N, *shape, "fp32x3")
    assert N.simt_fallback_count() == before, "fp32x3 must run on the tensor cores, not on the SIMT fallback"


@pytest.mark.parametrize("shape", GEMM_SHAPES[:6], ids=lambda s: "x".join(map(str, s)))
def test_gemm_fp32x3_full_split_variant(N, shape):
    """Splitter variant 1 (hi rewritten as rna_tf32 in place + lo copy) must meet the same fp32 tolerance as the
    default variant 2 (hi = the tensor core's own truncation, only lo written)."""
    N.set_x3_mode(1)
    try:
        _gemm_case(N, *shape, "fp32x3")
    finally:
        N.set_x3_mode(2)


def test_gemm_fp32x3_beats_tf32_error(N):
    """On one shape: the error of fp32x3 vs an fp64 product is >= 100x smaller than single-pass tf32's."""
    M, Nn, K = 500, 256, 880
    a, b = padded(M, K, K, 1), padded(Nn, K, K, 2)
    exact = a.double() @ b.double().t()
    errs = {}
    for prec in ("tf32", "fp32x3"):
        out = torch.empty(M, Nn, device=dev())
        N.linear(a, b, out, precision=prec)
        errs[prec] = (out.double() - exact).abs().max().item()
    fp32_err = ((a @ b.t()).double() - exact).abs().max().item()
    assert errs["fp32x3"] * 100 < errs["tf32"], errs
    assert errs["fp32x3"] < 8 * fp32_err + 1e-6, (errs, fp32_err)


@pytest.mark.parametrize("prec", ["fp32", "tf32", "fp32x3"])
def test_gemm_epilogues(N, prec):
    M, Nn, K = 150, 256, 640
    bias = rnd(Nn, seed=5)
    u = uni(M, Nn, seed=6)
    kw_n, kw_r, tol = _gemm_case(N, M, Nn, K, False, False, prec, bias=bias, act="leaky_dropout", u=u, mask_out=True)
    # masks may differ only where the pre-activation is within rounding of 0
    diff = (kw_n["mask_out"] != kw_r["mask_out"]).float().mean().item()
    assert diff < (5e-3 if prec == "tf32" else 1e-5)
    mask = (uni(M, Nn, seed=7) > 0.3).float() * 2
    _gemm_case(N, M, Nn, K, False, True, prec, mask_in=mask)
    _gemm_case(N, M, Nn, K, False, False, prec, accumulate=True, alpha=0.5)
    _gemm_case(N, 50, 256, 5020, False, False, prec, bias=bias, mask_in=mask[:50])


@pytest.mark.parametrize("prec", ["fp32x3", "tf32"])
@pytest.mark.parametrize("shape", [(500, 256, 250), (500, 256, 506), (300, 64, 100), (130, 70, 64), (512, 250, 762),
                                   (40, 32, 52)], ids=lambda s: "x".join(map(str, s)))
def test_linear_bn_relu_fused_epilogue(N, shape, prec):
    """Linear + BatchNorm1d(train) + ReLU + concat-store as ONE tcgen05 launch (cluster over the M-tiles,
    DSMEM statistics) vs the fp32 torch composition of the same ops (reference Residual, ctgan.py:40-44)."""
    B, H, K = shape
    r4 = lambda x: (x + 3) // 4 * 4  # noqa: E731
    assert N._gemm_tc_bn_plan(B, K) > 0
    wide = rnd(B, r4(K) + r4(H) + 8, seed=1)                  # [ out window | x ] in one buffer, like `h`
    w = padded(H, K, r4(K), 2).mul_(0.2)      # in place: keeps the padded leading dimension
    bias, gamma, beta = rnd(H, seed=3), rnd(H, seed=4).abs() + 0.5, rnd(H, seed=5) * 0.3
    res = {}
    for name, fn in (("native", N.linear_bn_relu), ("ref", R.linear_bn_relu)):
        buf = wide.clone()
        xx = buf[:, r4(H) + 4: r4(H) + 4 + K]
        out = buf[:, :H]
        a = torch.full((B, r4(H)), 7.0, device=dev())[:, :H]
        rm, rv = rnd(H, seed=6) * 0.1, rnd(H, seed=7).abs() + 0.5
        sm, si = torch.zeros(H, device=dev()), torch.zeros(H, device=dev())
        nbt = torch.tensor([3.0], device=dev())
        n0 = N.launch_count()
        fn(xx, w, a, out, bias=bias, gamma=gamma, beta=beta, running_mean=rm, running_var=rv, save_mean=sm,
           save_invstd=si, eps=1e-5, momentum=0.1, num_batches_tracked=nbt, **({"precision": prec} if name == "native" else {}))
        torch.cuda.synchronize()
        if name == "native":
            assert N.launch_count() - n0 == 1, "fused path must be a single launch"
        res[name] = (a.clone(), out.clone(), rm, rv, sm, si, nbt, buf)
    tight = prec == "fp32x3"
    rt, at = (2e-5, 2e-4) if tight else (4e-3, 2e-2)
    for i, what in enumerate(("pre-activation", "relu(bn)", "running_mean", "running_var", "save_mean", "save_invstd")):
        got, want = res["native"][i], res["ref"][i]
        err = (got.double() - want.double()).abs().max().item()
        assert err <= at + rt * want.abs().max().item(), (what, err)
    assert res["native"][6].item() == 4.0
    # nothing outside the output window was touched
    assert torch.equal(res["native"][7][:, H:], wide[:, H:])


@pytest.mark.parametrize("prec", ["tf32", "fp32x3"])
def test_cluster_splitk_push_equals_pull(N, prec):
    """The two in-cluster reduction protocols (st.async push / remote-load pull) sum in the same rank order:
    results are bitwise identical, for every epilogue and for the fused BatchNorm statistics."""
    shapes = [(500, 256, 506, False, False), (150, 256, 2520, False, False), (50, 2520, 256, False, True),
              (256, 640, 150, True, True)]
    for M, Nn, K, a_t, b_t in shapes:
        r4 = lambda x: (x + 3) // 4 * 4  # noqa: E731
        a = padded(K, M, r4(M), 1) if a_t else padded(M, K, r4(K), 1)
        b = padded(K, Nn, r4(Nn), 2) if b_t else padded(Nn, K, r4(K), 2)
        bias = rnd(Nn, seed=3)
        outs = []
        for push in (True, False):
            N.set_cluster_push(push)
            try:
                o = torch.zeros(M, Nn, device=dev())
                N.linear(a, b, o, a_t=a_t, b_t=b_t, bias=bias, precision=prec)
                torch.cuda.synchronize()
                outs.append(o)
            finally:
                N.set_cluster_push(False)
        assert torch.equal(outs[0], outs[1]), (M, Nn, K)
        ref = torch.zeros(M, Nn, device=dev())
        R.linear(a, b, ref, a_t=a_t, b_t=b_t, bias=bias)
        close(outs[0], ref, **(dict(rtol=2e-5, atol=1e-4) if prec == "fp32x3" else dict(rtol=4e-3, atol=4e-3 * (K ** 0.5) / 8)))


def test_gemm_splitk_explicit(N):
    M, Nn, K = 150, 256, 5020
    a, b = padded(M, K, K + 4, 1), padded(Nn, K, K, 2)
    ref = torch.empty(M, Nn, device=dev())
    R.linear(a, b, ref)
    for sk in (1, 2, 7, 16):
        out = torch.zeros(M, Nn, device=dev())
        N.linear(a, b, out, precision="tf32", split_k=sk)
        close(out, ref, rtol=4e-3, atol=0.05)
    # determinism of the fixed-order split-K reduction
    o1, o2 = torch.zeros(M, Nn, device=dev()), torch.zeros(M, Nn, device=dev())
    N.linear(a, b, o1, precision="tf32", split_k=16)
    N.linear(a, b, o2, precision="tf32", split_k=16)
    assert torch.equal(o1, o2)


def test_gemm_dropout_philox_rate(N):
    M, Nn, K = 150, 256, 64
    a, b = padded(M, K, K, 1), padded(Nn, K, K, 2)
    rng = R.RngState(3, dev())
    rng.tick()
    out, mask = torch.empty(M, Nn, device=dev()), torch.empty(M, Nn, device=dev())
    N.linear(a, b, out, act="leaky_dropout", mask_out=mask, rng=rng, stream=9, precision="fp32")
    keep = (mask != 0).float().mean().item()
    assert abs(keep - 0.5) < 0.02
    assert all(min(abs(v - t) for t in (0.0, 0.4, 2.0)) < 1e-6 for v in torch.unique(mask).tolist())


def test_small_dense_ops(N):
    a = padded(150, 256, 260, 1)
    w, bias = rnd(256, seed=2), rnd(1, seed=3)
    o1, o2 = torch.empty(150, device=dev()), torch.empty(150, device=dev())
    N.rowdot(a, w, bias, o1); R.rowdot(a, w, bias, o2)
    close(o1, o2, 1e-5, 1e-4)
    rs = rnd(150, seed=4)
    c1, c2 = torch.empty(256, device=dev()), torch.empty(256, device=dev())
    N.colsum(a, c1, rowscale=rs); R.colsum(a, c2, rowscale=rs)
    close(c1, c2, 1e-5, 1e-4)
    N.colsum(a, c1, accumulate=True); R.colsum(a, c2, accumulate=True)
    close(c1, c2, 1e-5, 1e-4)
    m = uni(150, 256, seed=5)
    s1, s2 = torch.empty(150, 256, device=dev()), torch.empty(150, 256, device=dev())
    N.scale_outer_mask(rs, w, m, s1); R.scale_outer_mask(rs, w, m, s2)
    close(s1, s2, 1e-6, 1e-6)


def test_bn_relu(N):
    B, H = 500, 256
    a = padded(B, H, H, 1) * 2 + 0.3
    gamma, beta = rnd(H, seed=2) * 0.5 + 1, rnd(H, seed=3) * 0.2
    st = [dict(rm=rnd(H, seed=4), rv=uni(H, seed=5) + 0.5, nbt=torch.zeros(1, device=dev())) for _ in range(2)]
    outs, saves = [], []
    for fn, s in zip((N.bn_relu_fwd, R.bn_relu_fwd), st):
        buf = torch.zeros(B, H + 64, device=dev())
        sm, si = torch.empty(H, device=dev()), torch.empty(H, device=dev())
        fn(a, gamma, beta, s["rm"], s["rv"], buf[:, 32:32 + H], sm, si, training=True, num_batches_tracked=s["nbt"])
        outs.append(buf); saves.append((sm, si))
    close(outs[0], outs[1], 1e-5, 1e-5)
    close(saves[0][0], saves[1][0], 1e-5, 1e-6); close(saves[0][1], saves[1][1], 1e-4, 1e-5)
    close(st[0]["rm"], st[1]["rm"], 1e-5, 1e-6); close(st[0]["rv"], st[1]["rv"], 1e-5, 1e-6)
    assert st[0]["nbt"].item() == 1.0
    dout = rnd(B, H, seed=7)
    res = []
    for fn in (N.bn_relu_bwd, R.bn_relu_bwd):
        da, dg, db = torch.empty(B, H, device=dev()), torch.empty(H, device=dev()), torch.empty(H, device=dev())
        fn(dout, a, saves[1][0], saves[1][1], gamma, beta, da, dg, db)
        res.append((da, dg, db))
    for x, y in zip(*res):
        close(x, y, 2e-5, 2e-5)
    # eval mode, many rows
    a2 = rnd(3000, H, seed=9)
    e1, e2 = torch.empty(3000, H, device=dev()), torch.empty(3000, H, device=dev())
    N.bn_relu_fwd(a2, gamma, beta, st[1]["rm"], st[1]["rv"], e1, None, None, training=False)
    R.bn_relu_fwd(a2, gamma, beta, st[1]["rm"], st[1]["rv"], e2, None, None, training=False)
    close(e1, e2, 1e-5, 1e-5)


def _layout():
    from fed_tgan_b200.features.layout import TableLayout

    info = []
    rng = np.random.default_rng(0)
    for i in range(30):
        if i % 3 == 0:
            info += [(1, "tanh"), (int(rng.integers(1, 9)), "softmax")]
        else:
            info += [(int(rng.integers(1, 40)), "softmax")]
    return TableLayout(info)


def test_activate_and_cond_loss(N):
    lay = _layout()
    T = lay.device_tables(dev())
    B, dd, n_opt = 500, lay.data_dim, lay.n_opt
    fake = padded(B, dd, (dd + 3) // 4 * 4, 1)
    u = uni(B, dd, seed=2)
    cond = torch.zeros(B, n_opt, device=dev())
    col = torch.randint(0, lay.n_col, (B,), device=dev())
    opt = (torch.rand(B, device=dev()) * T["soft_width"][col]).long()
    cond[torch.arange(B), T["soft_cstart"][col].long() + opt] = 1
    m = torch.zeros(B, lay.n_col, device=dev()); m[torch.arange(B), col] = 1
    o1, o2 = torch.zeros(B, dd + n_opt + 3, device=dev()), torch.zeros(B, dd + n_opt + 3, device=dev())
    N.activate_fwd(fake, T, o1, tau=0.2, u=u, cond=cond)
    R.activate_fwd(fake, T, o2, tau=0.2, u=u, cond=cond)
    close(o1, o2, 1e-4, 2e-5)
    dout = rnd(B, dd, seed=3)
    d1, d2 = torch.zeros(B, dd + 1, device=dev()), torch.zeros(B, dd + 1, device=dev())
    N.activate_bwd(dout, o2, T, d1, tau=0.2); R.activate_bwd(dout, o2, T, d2, tau=0.2)
    close(d1, d2, 1e-4, 1e-5)
    l1, l2 = torch.zeros(1, device=dev()), torch.zeros(1, device=dev())
    N.cond_loss_fwd_bwd(fake, cond, m, T, d1, l1, accumulate=True)
    R.cond_loss_fwd_bwd(fake, cond, m, T, d2, l2, accumulate=True)
    close(l1, l2, 1e-4, 1e-5); close(d1, d2, 1e-4, 1e-5)
    # philox gumbel path: rows sum to one on softmax spans
    rng = R.RngState(1, dev()); rng.tick()
    N.activate_fwd(fake, T, o1, tau=0.2, rng=rng, stream=5)
    st, w = int(lay.soft_start[3]), int(lay.soft_width[3])
    assert torch.allclose(o1[:, st:st + w].sum(1), torch.ones(B, device=dev()), atol=1e-4)


def test_slerp_gp_adam(N):
    B, din = 500, 417
    low, high = padded(B, din, 420, 1), padded(B, din, 420, 2)
    alpha = uni(B, seed=3)
    o1, o2 = torch.zeros(B, 420, device=dev()), torch.zeros(B, 420, device=dev())
    N.slerp(low, high, o1[:, :din], alpha=alpha); R.slerp(low, high, o2[:, :din], alpha=alpha)
    close(o1, o2, 1e-4, 1e-4)
    g1 = rnd(50, 4200, seed=4) * 0.02; g2 = g1.clone()
    p1, p2 = torch.zeros(1, device=dev()), torch.zeros(1, device=dev())
    N.gp_penalty(g1, p1, lambda_=10.0); R.gp_penalty(g2, p2, lambda_=10.0)
    close(p1, p2, 1e-4, 1e-5); close(g1, g2, 1e-4, 1e-7)
    n = 10007 * 4
    rng = R.RngState(0, dev())
    P = [rnd(n, seed=5).clone() for _ in range(2)]
    G = rnd(n, seed=6) * 0.1
    Mo = [torch.zeros(n, device=dev()) for _ in range(2)]
    V = [torch.zeros(n, device=dev()) for _ in range(2)]
    for t in range(3):
        rng.tick()
        N.adam_step(P[0], G, Mo[0], V[0], rng.state[1:2], lr=2e-4, beta1=0.5, beta2=0.9, eps=1e-8, weight_decay=1e-6)
        R.adam_step(P[1], G, Mo[1], V[1], rng.state[1:2], lr=2e-4, beta1=0.5, beta2=0.9, eps=1e-8, weight_decay=1e-6)
    close(P[0], P[1], 1e-6, 1e-6)


def test_sampling_ops(N):
    from fed_tgan_b200.engine import DeviceTable

    lay = _layout()
    rng = np.random.default_rng(1)
    rows = 4000
    vals = rng.uniform(-1, 1, (rows, lay.n_tanh)).astype(np.float32)
    codes = np.stack([rng.integers(0, w, rows) for w in lay.soft_width], 1).astype(np.int32)
    t = DeviceTable(vals, codes, lay, dev())
    T = t.tables
    B = 500
    ucol, uopt, uperm, urow = uni(B, seed=1), uni(B, seed=2), uni(B, seed=3), uni(B, seed=4)
    i32 = lambda: torch.zeros(B, dtype=torch.int32, device=dev())  # noqa: E731
    c1, o1, c2, o2 = i32(), i32(), i32(), i32()
    N.sample_condvec(t.cum_p, T["soft_width"], c1, o1, u_col=ucol, u_opt=uopt)
    R.sample_condvec(t.cum_p, T["soft_width"], c2, o2, u_col=ucol, u_opt=uopt)
    assert torch.equal(c1, c2) and torch.equal(o1, o2)
    p1, p2 = i32(), i32()
    N.argsort_perm(p1, u=uperm); R.argsort_perm(p2, u=uperm)
    assert torch.equal(p1, p2)
    r1, r2 = i32(), i32()
    N.sample_rows(t.row_ptr, t.row_ids, T["soft_cstart"], c1, o1, r1, perm=p1, u=urow)
    R.sample_rows(t.row_ptr, t.row_ids, T["soft_cstart"], c1, o1, r2, perm=p1, u=urow)
    assert torch.equal(r1, r2)
    x1, x2 = torch.ones(B, lay.data_dim + lay.n_opt + 2, device=dev()), torch.ones(B, lay.data_dim + lay.n_opt + 2, device=dev())
    N.expand_rows(t.vals, t.codes, r1, T, x1); R.expand_rows(t.vals, t.codes, r1, T, x2)
    dd = lay.data_dim
    N.write_condvec(c1, o1, T["soft_cstart"], x1[:, dd:dd + lay.n_opt], None, p1)
    R.write_condvec(c1, o1, T["soft_cstart"], x2[:, dd:dd + lay.n_opt], None, p1)
    assert torch.equal(x1, x2)
    m1, m2 = torch.ones(B, lay.n_col, device=dev()), torch.ones(B, lay.n_col, device=dev())
    N.write_condvec(c1, o1, T["soft_cstart"], x1[:, dd:dd + lay.n_opt], m1)
    R.write_condvec(c1, o1, T["soft_cstart"], x2[:, dd:dd + lay.n_opt], m2)
    assert torch.equal(m1, m2) and torch.equal(x1, x2)
    # philox paths: distributional sanity
    st = R.RngState(7, dev()); st.tick()
    z = torch.zeros(500, 128, device=dev())
    N.fill_normal(z, rng=st, stream=0)
    assert abs(z.mean().item()) < 0.02 and abs(z.std().item() - 1) < 0.02
    N.argsort_perm(p1, rng=st, stream=3)
    assert sorted(p1.tolist()) == list(range(B))
    N.sample_condvec(t.cum_p, T["soft_width"], c1, o1, rng=st, stream=1)
    assert int(c1.max()) < lay.n_col and bool((o1 < T["soft_width"][c1.long()]).all())
    st.tick()
    z2 = torch.zeros(500, 128, device=dev())
    N.fill_normal(z2, rng=st, stream=0)
    assert not torch.equal(z, z2)


def test_vgm_codec(N):
    from fed_tgan_b200.features.transformers import BGM_CTGAN_Transformer

    rng = np.random.default_rng(0)
    n = 3000
    data = np.stack([np.concatenate([rng.normal(0, 1, n // 2), rng.normal(8, 1, n // 2)]),
                     rng.integers(0, 5, n).astype(float), rng.exponential(2, n)], 1)
    t = BGM_CTGAN_Transformer(backend="vb", seed=0)
    t.fit(data, (1,))
    dense = torch.tensor(t.transform(data), dtype=torch.float32, device=dev())
    dec = {k: torch.as_tensor(v).to(dev()) for k, v in t.decode_tables().items()}
    T = t.layout.device_tables(dev())
    o1 = torch.empty(n, 3, device=dev()); o2 = torch.empty(n, 3, device=dev())
    N.vgm_decode(dense, T, dec["mu"], dec["sd"], dec["cat_ids"], dec["cat_off"], dec["col_is_cont"], o1)
    R.vgm_decode(dense, T, dec["mu"], dec["sd"], dec["cat_ids"], dec["cat_off"], dec["col_is_cont"], o2)
    close(o1, o2, 1e-5, 1e-5)
    gm, mask = t.model[0], t.components[0]
    x = torch.tensor(data[:, 0], dtype=torch.float32, device=dev())
    lc = torch.tensor(gm.logit_const(), dtype=torch.float32, device=dev())
    mu = torch.tensor(gm.means, dtype=torch.float32, device=dev())
    var = torch.tensor(gm.covs, dtype=torch.float32, device=dev())
    act = torch.tensor(mask, device=dev())
    u = uni(n, seed=3)
    v1, v2 = torch.empty(n, device=dev()), torch.empty(n, device=dev())
    k1, k2 = torch.empty(n, dtype=torch.int32, device=dev()), torch.empty(n, dtype=torch.int32, device=dev())
    N.vgm_encode(x, lc, mu, var, act, u, v1, k1); R.vgm_encode(x, lc, mu, var, act, u, v2, k2)
    assert (k1 != k2).float().mean().item() < 2e-3
    same = k1 == k2
    close(v1[same], v2[same], 1e-4, 1e-4)


def test_fedavg_local_and_stats(N):
    world, n = 4, 4096 * 3 + 64
    arenas = [rnd(n, seed=i) for i in range(world)]
    for a in arenas:
        a[100] = 7.0 + torch.rand(1).item()
    w = [0.1, 0.2, 0.3, 0.4]
    ref = [a.clone() for a in arenas]
    R.fedavg(ref, w, [100])
    N.fedavg_local(arenas, w, [100])
    torch.cuda.synchronize()
    for a, r in zip(arenas, ref):
        close(a, r, 1e-6, 1e-6)
    assert float(arenas[0][100]) == float(int(arenas[0][100]))
    # stats -> weights vs numpy restatement
    from fed_tgan_b200.parallel import weights as W

    rng = np.random.default_rng(0)
    sizes = [3, 7, 2]
    hist = [np.concatenate([rng.integers(1, 50, k) for k in sizes]).astype(np.float32) for _ in range(world)]
    cat_off = np.concatenate([[0], np.cumsum(sizes)]).astype(np.int32)
    rows = np.array([1000, 3000, 500, 1500], dtype=np.float32)
    samples = [[np.sort(rng.normal(i * 0.5 + c, 1 + 0.1 * c, int(rows[i] / 10))).astype(np.float32) for i in range(world)]
               for c in range(2)]
    pool = np.stack([np.sort(np.concatenate(s)) for s in samples])
    flat = np.concatenate([x for s in samples for x in s])
    off = np.concatenate([[0], np.cumsum([len(x) for s in samples for x in s])]).astype(np.int32)
    t = lambda x: torch.as_tensor(x).to(dev())  # noqa: E731
    jsd, wd, wt = N.stats_to_weights(t(np.stack(hist)), t(cat_off), t(pool), t(flat), t(off), t(rows))
    cv = [[h[cat_off[c]:cat_off[c + 1]] for h in hist] for c in range(3)]
    jr, wr = W.categorical_similarity(cv), W.continuous_similarity(samples)
    assert np.allclose(jsd.cpu().numpy(), jr, atol=2e-4)
    assert np.allclose(wd.cpu().numpy(), wr, atol=2e-3)
    assert np.allclose(wt.cpu().numpy(), W.final_weights(jr, wr, rows), atol=1e-4)


def test_fused_prep_and_heads(N):
    """Fused prep kernels: structural invariants (randomness is Philox, so no bitwise oracle)."""
    from fed_tgan_b200.engine import DeviceTable

    lay = _layout()
    rng = np.random.default_rng(2)
    rows, B, E = 5000, 500, 128
    vals = rng.uniform(-1, 1, (rows, lay.n_tanh)).astype(np.float32)
    codes = np.stack([rng.integers(0, w, rows) for w in lay.soft_width], 1).astype(np.int32)
    t = DeviceTable(vals, codes, lay, dev())
    T = t.tables
    st = R.RngState(11, dev()); st.tick()
    dd, n_opt = lay.data_dim, lay.n_opt
    h = torch.full((B, E + n_opt + 4), 7.0, device=dev())
    X = torch.full((B, dd + n_opt + 3), 7.0, device=dev())
    i32 = lambda: torch.zeros(B, dtype=torch.int32, device=dev())  # noqa: E731
    col, opt, perm, rid = i32(), i32(), i32(), i32()
    N.prep_d_batch(t, T, h[:, :E], h[:, E:E + n_opt], X, col, opt, perm, rid, st, 0, 1, 3, 4)
    torch.cuda.synchronize()
    z, c1 = h[:, :E], h[:, E:E + n_opt]
    assert abs(z.mean().item()) < 0.03 and abs(z.std().item() - 1) < 0.03
    assert sorted(perm.tolist()) == list(range(B))
    g = T["soft_cstart"][col.long()].long() + opt.long()
    assert torch.equal(c1.argmax(1), g) and torch.allclose(c1.sum(1), torch.ones(B, device=dev()))
    # real rows: exactly the table rows `rid`, whose code in the conditioned span equals the permuted option
    dense = t.dense().to(dev())
    assert torch.equal(X[:, :dd], dense[rid.long()])
    gp = g[perm.long()]
    assert torch.equal(X[:, dd:dd + n_opt].argmax(1), gp)
    assert bool((X[torch.arange(B), T["soft_start"][col[perm.long()].long()].long() + opt[perm.long()].long()] == 1).all())
    assert bool((h[:, E + n_opt:] == 7).all()) and bool((X[:, dd + n_opt:] == 7).all())     # padding untouched
    m1 = torch.zeros(B, lay.n_col, device=dev())
    N.prep_g_batch(t, T, h[:, :E], h[:, E:E + n_opt], m1, col, opt, st, 16, 17)
    assert torch.equal(m1.argmax(1), col.long()) and torch.allclose(m1.sum(1), torch.ones(B, device=dev()))
    # log-frequency marginal of the option draw for column 0
    sel = col == 1
    w = int(lay.soft_width[1])
    freq = torch.bincount(opt[sel].long(), minlength=w).float() / max(int(sel.sum()), 1)
    # fused rowdot + loss, bn backward bias
    a = padded(150, 256, 256, 1); wv, bias = rnd(256, seed=2), rnd(1, seed=3); rs = rnd(150, seed=4)
    y1, y2 = torch.empty(150, device=dev()), torch.empty(150, device=dev())
    l1, l2 = torch.zeros(1, device=dev()), torch.zeros(1, device=dev())
    N.rowdot(a, wv, bias, y1, rowscale=rs, loss_out=l1); R.rowdot(a, wv, bias, y2, rowscale=rs, loss_out=l2)
    close(y1, y2, 1e-5, 1e-4); close(l1, l2, 1e-5, 1e-3)
    Bn, H = 500, 256
    x = padded(Bn, H, H, 5); gamma, beta = rnd(H, seed=6) + 1, rnd(H, seed=7) * 0.1
    sm, si = x.mean(0), torch.rsqrt(x.var(0, unbiased=False) + 1e-5)
    dout = rnd(Bn, H, seed=8)
    outs = []
    for fn in (N.bn_relu_bwd, R.bn_relu_bwd):
        da, dg, db, dbf = (torch.empty(Bn, H, device=dev()), torch.empty(H, device=dev()), torch.empty(H, device=dev()),
                           torch.empty(H, device=dev()))
        fn(dout, x, sm, si, gamma, beta, da, dg, db, dbias_fc=dbf)
        outs.append((da, dg, db, dbf))
    for k in range(3):
        close(outs[0][k], outs[1][k], 2e-5, 2e-5)
    assert outs[0][3].abs().max().item() < 1e-3          # analytically zero


def test_fused_activate_slerp_and_output_bwd(N):
    lay = _layout()
    T = lay.device_tables(dev())
    B, dd, n_opt = 500, lay.data_dim, lay.n_opt
    din = dd + n_opt
    ld = (din + 3) // 4 * 4
    fake = padded(B, dd, (dd + 3) // 4 * 4, 1)
    u = uni(B, dd, seed=2)
    alpha = uni(B, seed=5)
    cond = torch.zeros(B, n_opt, device=dev())
    col = torch.randint(0, lay.n_col, (B,), device=dev())
    opt = (torch.rand(B, device=dev()) * T["soft_width"][col]).long()
    cond[torch.arange(B), T["soft_cstart"][col].long() + opt] = 1
    m = torch.zeros(B, lay.n_col, device=dev()); m[torch.arange(B), col] = 1
    X1 = rnd(3 * B, ld, seed=7).abs() + 0.1
    X2 = X1.clone()
    for fn, X in ((N.activate_fwd, X1), (R.activate_fwd, X2)):
        fn(fake, T, X[:B], tau=0.2, u=u, cond=cond, slerp_real=X[B:2 * B, :din], slerp_out=X[2 * B:, :din], alpha=alpha)
    close(X1[:B, :din], X2[:B, :din], 1e-4, 2e-5)
    close(X1[2 * B:, :din], X2[2 * B:, :din], 2e-4, 1e-4)
    assert torch.equal(X1[B:2 * B], X2[B:2 * B])
    dout = rnd(B, dd, seed=3)
    d1, d2 = torch.zeros(B, dd + 1, device=dev()), torch.zeros(B, dd + 1, device=dev())
    l1, l2 = torch.zeros(1, device=dev()), torch.zeros(1, device=dev())
    N.gen_output_bwd(dout, X2[:B, :dd], fake, cond, m, T, d1, l1, tau=0.2)
    R.gen_output_bwd(dout, X2[:B, :dd], fake, cond, m, T, d2, l2, tau=0.2)
    close(d1, d2, 1e-4, 1e-5); close(l1, l2, 1e-4, 1e-5)


@pytest.mark.parametrize("prec", ["fp32", "tf32", "fp32x3"])
def test_gemm_never_writes_outside_its_output_view(N, prec):
    """compute-sanitizer is closed on this pool, so out-of-bounds stores are hunted with canaries:
    the output is a window inside a sentinel-filled buffer, for ragged shapes, all operand majors,
    every epilogue specialisation and both split-K implementations."""
    SENT = 12345.0
    cases = [(150, 256, 640, False, False), (50, 250, 256, False, True), (130, 70, 150, True, True),
             (500, 128, 764, False, False), (37, 19, 44, True, False)]
    for M, Nn, K, a_t, b_t in cases:
        r4 = lambda x: (x + 3) // 4 * 4  # noqa: E731
        a = padded(K, M, r4(M), 1) if a_t else padded(M, K, r4(K), 1)
        b = padded(K, Nn, r4(Nn), 2) if b_t else padded(Nn, K, r4(K), 2)
        bias = rnd(Nn, seed=3)
        mask = uni(M, r4(Nn), seed=4)[:, :Nn]
        for epi in (dict(), dict(bias=bias), dict(mask_in=mask), dict(accumulate=True),
                    dict(bias=bias, act="leaky_dropout", u=uni(M, r4(Nn), seed=5)[:, :Nn])):
            for split in (None, 3):
                if prec == "fp32" and split is not None:
                    continue
                big = torch.full((M + 4, r4(Nn) + 8), SENT, device=dev())
                out = big[2:M + 2, 4:4 + Nn]
                out.zero_()
                kw = dict(epi)
                if "act" in kw:
                    kw["mask_out"] = torch.zeros(M, r4(Nn), device=dev())[:, :Nn]
                N.linear(a, b, out, a_t=a_t, b_t=b_t, precision=prec, split_k=split, **kw)
                torch.cuda.synchronize()
                guard = big.clone()
                guard[2:M + 2, 4:4 + Nn] = SENT
                assert bool((guard == SENT).all()), (M, Nn, K, a_t, b_t, list(epi), split)
                assert bool(torch.isfinite(out).all())


def test_stats_kernel_near_identical_histograms(N):
    """Binary columns whose client distributions differ in the 4th digit (Covertype-like): the
    per-column normalised JSD must still match the fp64 host restatement."""
    from fed_tgan_b200.parallel import weights as W

    rng = np.random.default_rng(3)
    world, n_cat = 2, 12
    rows = np.array([250000.0, 331012.0], dtype=np.float32)
    hist = []
    for i in range(world):
        parts = []
        for c in range(n_cat):
            p = 0.03 + 0.9 * (c / n_cat) + 1e-4 * rng.normal()
            k = float(np.round(rows[i] * p))
            parts.append(np.array([rows[i] - k, k], dtype=np.float32))
        hist.append(np.concatenate(parts))
    cat_off = (np.arange(n_cat + 1) * 2).astype(np.int32)
    t = lambda x: torch.as_tensor(x).to(dev())  # noqa: E731
    jsd, wd, wt = N.stats_to_weights(t(np.stack(hist)), t(cat_off), None, None, None, t(rows))
    cv = [[h[cat_off[c]:cat_off[c + 1]].astype(np.float64) for h in hist] for c in range(n_cat)]
    jr = W.categorical_similarity(cv)
    assert np.allclose(jsd.cpu().numpy(), jr, atol=2e-3), (jsd.cpu().numpy(), jr)
    assert np.allclose(wt.cpu().numpy(), W.final_weights(jr, np.zeros((world, 0)), rows), atol=1e-3)


def test_gemm_persistent_variant_is_bitwise_identical(N):
    """Opt-in persistent kernel (one CTA per SM, double-buffered TMEM accumulator): same MMAs in the same
    order per tile, so multi-wave GEMMs must reproduce the default kernel bit for bit."""
    torch.manual_seed(0)
    for (M, Nn, K, a_t, b_t) in [(1624, 3700, 500, True, True), (8192, 256, 252, False, False),
                                 (52, 31480, 1024, False, True)]:
        a = torch.randn((K, M) if a_t else (M, K), device=dev())
        b = torch.randn((K, Nn) if b_t else (Nn, K), device=dev())
        bias = torch.randn(Nn, device=dev())
        outs = []
        try:
            for mode in (False, True):
                N.set_gemm_persistent(mode)
                out = torch.full((M, Nn), float("nan"), device=dev())
                N.linear(a, b, out, a_t=a_t, b_t=b_t, bias=bias, precision="tf32")
                outs.append(out)
        finally:
            N.set_gemm_persistent(False)
        torch.cuda.synchronize()
        assert torch.equal(outs[0], outs[1]), (M, Nn, K)
        ref = (a.t() if a_t else a) @ (b if b_t else b.t()) + bias
        assert ((outs[1] - ref).abs().max() / ref.abs().max()).item() < 5e-3


# ------------------------------------------------------------------ federated statistics on the device (K14 / K15)
def test_sort_rows_in_kernel_bitonic(N):
    """In-kernel bitonic network (no library sort): ragged sizes, duplicates, more than one shared-memory tile."""
    for C_, n in ((3, 1000), (2, 4096), (2, 5000), (1, 70001)):
        g = torch.Generator().manual_seed(n)
        x = torch.randn(C_, n, generator=g).to(dev())
        x[:, : n // 7] = x[:, n // 7: 2 * (n // 7)]                    # duplicates
        out = N.sort_rows(x.contiguous())
        assert out.shape == (C_, N.sort_pad(n))
        assert torch.equal(out[:, :n], torch.sort(x, dim=1).values)
        assert bool(torch.isinf(out[:, n:]).all())


def test_gmm_sample_sort_moments_and_order(N):
    """Device mixture sampler: sorted output, mixture moments and component proportions match
    (reference: client_gmm.sample(n), Server/dtds/distributed.py:732-734)."""
    from fed_tgan_b200.features.gmm import GMMParams

    K = 10
    w = np.array([0.5, 0.3, 0.2] + [0.0] * 7)
    mu = np.array([0.0, 6.0, -4.0] + [0.0] * 7)
    var = np.array([1.0, 0.25, 2.0] + [1.0] * 7)
    g = GMMParams(mu, var, w, np.ones(K), np.ones(K), np.ones(K), np.ones(K))
    arr = torch.as_tensor(np.stack([g.to_array(), g.to_array()])).to(dev())
    n = 60000
    seg = N.gmm_sample_sort(arr, n, seed=11, client_id=3)
    s = seg[:, :n]
    assert bool((s[:, 1:] >= s[:, :-1]).all()) and bool(torch.isfinite(s).all())
    m_true = float((w * mu).sum())
    v_true = float((w * (var + mu ** 2)).sum() - m_true ** 2)
    for c in range(2):
        assert abs(float(s[c].mean()) - m_true) < 0.05
        assert abs(float(s[c].var()) - v_true) < 0.03 * v_true
        assert abs(float((s[c] > 3.5).float().mean()) - 0.3) < 0.01       # the mu = 6 component
    assert not torch.equal(s[0], s[1])                                     # columns draw independent streams
    again = N.gmm_sample_sort(arr, n, seed=11, client_id=3)
    assert torch.equal(seg, again)                                         # counter-based: reproducible


def test_fed_stats_to_weights_vs_scipy(N):
    """Fused statistics kernel on the worked-example structure of SURVEY 3.2: JSD vs scipy
    jensenshannon, WD vs scipy wasserstein_distance ON THE SAME device-drawn samples, weights vs the numpy
    restatement; pooled quantile seeds and pooled sample are exact."""
    from scipy.spatial.distance import jensenshannon
    from scipy.stats import wasserstein_distance

    from fed_tgan_b200.features.gmm import GMMParams
    from fed_tgan_b200.parallel import weights as W
    from fed_tgan_b200.parallel.devstats import local_statistics, sample_counts

    K = 10
    one = np.ones(K)

    def gm(mus, sds, ws):
        pad = K - len(mus)
        return GMMParams(np.array(list(mus) + [0.0] * pad), np.array([s * s for s in sds] + [1.0] * pad),
                         np.array(list(ws) + [0.0] * pad), one, one, one, one)

    rows = [1000, 3000, 2500]
    models = [[None, gm([0.0], [1.0], [1.0]), gm([1.0, 5.0], [0.5, 1.0], [0.6, 0.4])],
              [None, gm([2.0], [1.0], [1.0]), gm([1.2, 5.0], [0.5, 1.0], [0.5, 0.5])],
              [None, gm([1.0], [2.0], [1.0]), gm([0.8, 4.0], [0.6, 1.5], [0.7, 0.3])]]
    count_vectors = [[np.array([90.0, 10.0]), np.array([10.0, 90.0]), np.array([50.0, 50.0])],
                     [np.array([5.0, 5.0, 990.0]), np.array([1000.0, 1000.0, 1000.0]), np.array([0.0, 100.0, 2400.0])]]
    st = local_statistics(models, [1, 2], count_vectors, rows, dev(), seed=5, K=K, want_pool=True)
    jsd, wd, w = st.host()
    n_i = sample_counts(rows)
    # --- JSD
    for c, col in enumerate(count_vectors):
        pooled = np.sum(col, axis=0)
        raw = np.array([jensenshannon(pooled, v) for v in col])
        assert np.allclose(jsd[:, c], raw / raw.sum(), atol=2e-5), (c, jsd[:, c], raw / raw.sum())
    # --- WD on the very samples the kernel used (segments are kept alive by the result object)
    segs = [t for t in st._keep if torch.is_tensor(t) and t.dim() == 2]
    assert len(segs) == 3
    for c in range(2):
        per = [segs[i][c, : n_i[i]].double().cpu().numpy() for i in range(3)]
        pool = np.concatenate(per)
        raw = np.array([wasserstein_distance(pool, p) for p in per])
        assert np.allclose(wd[:, c], raw / raw.sum(), rtol=2e-3, atol=2e-4), (c, wd[:, c], raw / raw.sum())
        # pooled sample = concatenation in client order; quantile seeds = pooled order statistics
        assert np.array_equal(st.pool[c].cpu().numpy(), pool.astype(np.float32))
        srt = np.sort(pool.astype(np.float32))
        q = np.round((np.arange(K) + 0.5) / K * (len(pool) - 1)).astype(int)
        assert np.array_equal(st.centers[c].cpu().numpy(), srt[q])
    # --- final weights vs the numpy restatement of calculate_final_weights_for_aggregation
    host = W.final_weights(jsd, wd, rows)
    assert np.allclose(w, host, atol=1e-5) and abs(w.sum() - 1) < 1e-5
    # --- the aggregation kernel accepts the device weights as they are
    reps = [torch.randn(4096, device=dev()) for _ in range(3)]
    want = sum(float(w[i]) * reps[i] for i in range(3))
    N.fedavg_local(reps, st.weights, [])
    torch.cuda.synchronize()
    assert torch.allclose(reps[0], want, atol=1e-5) and torch.equal(reps[0], reps[1])


@pytest.mark.parametrize("shape", [(150, 256, 256), (50, 256, 256), (7, 64, 32), (33, 512, 100)], ids=lambda s: "x".join(map(str, s)))
def test_d_mid_fused_vs_composition(N, shape):
    """Fused discriminator middle (layer 1 + head + both backward signals, one row-local fp32 kernel) vs the
    composition it replaces: linear(leaky_dropout) -> d_head -> linear(b_t, mask_in)  (reference: Discriminator,
    ctgan.py:22-30, and its autograd backward)."""
    M, H0, H1 = shape
    assert N.d_mid_supported(H0, H1)
    a0 = padded(M, H0, H0 + 4, 1)
    w1 = padded(H1, H0, H0 + 8, 2).mul_(0.1)
    b1, wh, bh = rnd(H1, seed=3) * 0.1, rnd(H1, seed=4), rnd(1, seed=5)
    mask0 = (uni(M, H0, seed=6) > 0.5).float() * 2.0
    u = uni(M, H1, seed=7)
    rs_loss, rs_sig = rnd(M, seed=8), rnd(M, seed=9)
    out = {}
    for name, fn in (("native", N.d_mid), ("ref", R.d_mid)):
        a1, m1 = torch.full((M, H1), 9.0, device=dev()), torch.full((M, H1), 9.0, device=dev())
        z1, z0 = torch.full((M, H1), 9.0, device=dev()), torch.full((M, H0 + 4), 9.0, device=dev())[:, :H0]
        y, loss = torch.zeros(M, device=dev()), torch.zeros(1, device=dev())
        n0 = N.launch_count()
        fn(a0, w1, b1, a1, m1, mask0, wh, bh, y, rs_loss, loss, rs_sig, z1, z0, u=u, slope=0.2, p=0.5)
        torch.cuda.synchronize()
        if name == "native":
            assert N.launch_count() - n0 == 1
        out[name] = (a1, m1, y, loss, z1, z0.clone())
    for i, what in enumerate(("a1", "mask1", "y", "loss", "z1", "z0")):
        g, r = out["native"][i], out["ref"][i]
        if what == "mask1":
            assert (g != r).float().mean().item() < 1e-4, what          # only where the pre-activation rounds across 0
            continue
        err = (g.double() - r.double()).abs().max().item()
        assert err <= 2e-5 * (1.0 + r.abs().max().item()) * (4.0 if what in ("loss", "z0") else 1.0), (what, err)
    # Philox dropout: same keep rate and value set as the GEMM epilogue, and a second launch re-arms the loss counter
    rng = R.RngState(3, dev())
    rng.tick()
    a1, m1 = torch.empty(M, H1, device=dev()), torch.empty(M, H1, device=dev())
    z1, z0 = torch.empty(M, H1, device=dev()), torch.empty(M, H0, device=dev())
    y, loss = torch.zeros(M, device=dev()), torch.zeros(1, device=dev())
    for _ in range(2):
        N.d_mid(a0, w1, b1, a1, m1, mask0, wh, bh, y, rs_loss, loss, rs_sig, z1, z0, rng=rng, stream=9)
    torch.cuda.synchronize()
    assert abs((m1 != 0).float().mean().item() - 0.5) < 0.06
    assert abs(loss.item() - float((rs_loss * y).sum())) <= 1e-4 * (1 + abs(loss.item()))
    ref_mask = torch.empty(M, H1, device=dev())
    N.linear(a0, w1, torch.empty(M, H1, device=dev()), bias=b1, act="leaky_dropout", mask_out=ref_mask, rng=rng, stream=9,
             precision="fp32")
    assert (ref_mask != m1).float().mean().item() < 1e-4        # same Philox draw per element as the GEMM epilogue
