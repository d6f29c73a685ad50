"""Plain fp32 PyTorch restatement of every op in the kernel set.

Each function here is (a) the numerics oracle that ``tests/`` compare the sm_100a kernel of
the same name against and (b) the CPU execution path (gloo / RPC tests, config 1).  All
functions write into caller-provided output tensors (possibly strided views into larger
activation buffers) exactly like the kernels do, so the engine code is identical on both
paths.

Randomness: ops that consume random numbers accept explicit uniforms (``u=...``) so tests
can drive kernel and oracle with the same numbers; when ``u`` is ``None`` they draw from
``rng`` (an ``RngState``).
"""

from __future__ import annotations

import math

import torch
from torch.nn import functional as F

TINY = 1e-10


# ------------------------------------------------------------------ random state
class RngState:
    """Seed + step counter.  On CUDA the counter lives in device memory so that CUDA-graph
    replays advance it and the Philox kernels key every draw by (seed, step, site).  The
    torch restatement draws from a ``torch.Generator`` that is re-seeded from (seed, step) at
    every tick, so on both paths the random stream is a pure function of the two checkpointed
    integers (bit-exact resume)."""

    def __init__(self, seed: int, device):
        self.seed = int(seed)
        self.device = torch.device(device)
        # [seed, step] as int64 in device memory (read by kernels)
        self.state = torch.tensor([self.seed, 0], dtype=torch.int64, device=self.device)
        self.gen = torch.Generator(device="cpu")
        self.resync()

    def resync(self):
        """Re-derive the host generator from ``state`` (after a tick or a checkpoint load)."""
        step = int(self.state[1])
        self.gen.manual_seed((self.seed * 0x9E3779B1 + step * 0x85EBCA77 + 0x27D4EB2F) & 0x7FFFFFFFFFFFFFFF)

    def uniform(self, *shape):
        return torch.rand(*shape, generator=self.gen).to(self.device)

    def normal(self, *shape):
        return torch.randn(*shape, generator=self.gen).to(self.device)

    def tick(self):
        if self.state.is_cuda:
            import os
            if os.environ.get("FEDTGAN_FORCE_REFERENCE", "0") != "1":
                from . import native

                native.tick(self.state)
                return
        self.state[1] += 1
        self.resync()


# ------------------------------------------------------------------------ GEMM
def linear(a, b, out, *, a_t=False, b_t=False, bias=None, act=None, mask_in=None, mask_out=None,
           u=None, rng=None, stream=0, accumulate=False, alpha=1.0, slope=0.2, p=0.5):
    """``out = epilogue(alpha * A @ B^T)`` with logical ``A [M,K]``, ``B [N,K]``.

    ``a_t``: ``a`` is stored ``[K,M]``;  ``b_t``: ``b`` is stored ``[K,N]``.
    Epilogue order: ``+bias[n]`` -> activation -> ``*mask_in`` -> (``+=`` | ``=``) ``out``.
    ``act``: ``None`` | ``"leaky"`` | ``"leaky_dropout"`` (writes the multiplier
    ``slope_or_1 * keep/(1-p)`` to ``mask_out``; ``keep = u >= p``).
    """
    A = a.t() if a_t else a
    B = b if b_t else b.t()
    v = alpha * (A @ B)
    if bias is not None:
        v = v + bias.view(1, -1)
    if act is not None:
        mult = torch.where(v > 0, torch.ones_like(v), torch.full_like(v, slope))
        if act == "leaky_dropout":
            if u is None:
                u = rng.uniform(*v.shape)
            mult = mult * (u >= p).to(v.dtype) / (1.0 - p)
        elif act != "leaky":
            raise ValueError(act)
        if mask_out is not None:
            mask_out.copy_(mult)
        v = v * mult
    if mask_in is not None:
        v = v * mask_in
    if accumulate:
        out.add_(v)
    else:
        out.copy_(v)
    return out


def rowdot(a, w, bias, out, rowscale=None, loss_out=None):
    """``out[m] = a[m,:] . w + bias`` (the discriminator's 1-wide head); optionally also
    ``loss_out = sum_m rowscale[m] * out[m]``."""
    out.copy_(a @ w.reshape(-1) + (bias.reshape(()) if bias is not None else 0.0))
    if loss_out is not None:
        loss_out.copy_((out * rowscale).sum().reshape(1))
    return out


def d_head(a, w, bias, y, rs_loss, loss_out, rs_sig, mask, z):
    """Discriminator head: ``y = a.w + b``, ``loss = sum rs_loss*y`` and the first backward signal
    ``z = rs_sig (x) w (.) mask`` (one fused launch on the device)."""
    rowdot(a, w, bias, y, rowscale=rs_loss, loss_out=loss_out)
    scale_outer_mask(rs_sig, w.reshape(-1), mask, z)
    return y


def colsum(a, out, rowscale=None, accumulate=False):
    """``out[n] = sum_m rowscale[m] * a[m,n]``."""
    v = a.sum(0) if rowscale is None else (a * rowscale.view(-1, 1)).sum(0)
    if accumulate:
        out.add_(v)
    else:
        out.copy_(v)
    return out


def scale_outer_mask(rowscale, v, mask, out):
    """``out[m,n] = rowscale[m] * v[n] * mask[m,n]``."""
    out.copy_(rowscale.view(-1, 1) * v.view(1, -1) * mask)
    return out


# ------------------------------------------------------------------- batch norm
def bn_relu_fwd(a, gamma, beta, running_mean, running_var, out, save_mean, save_invstd, *,
                training=True, eps=1e-5, momentum=0.1, num_batches_tracked=None):
    """``out = relu(batchnorm(a))``; training mode updates running stats (unbiased var) and
    bumps ``num_batches_tracked`` (a 1-element float tensor)."""
    if training:
        if num_batches_tracked is not None:
            num_batches_tracked.add_(1.0)
        n = a.shape[0]
        mean = a.mean(0)
        var = a.var(0, unbiased=False)
        running_mean.mul_(1 - momentum).add_(momentum * mean)
        running_var.mul_(1 - momentum).add_(momentum * var * (n / max(n - 1, 1)))
    else:
        mean, var = running_mean, running_var
    invstd = torch.rsqrt(var + eps)
    if save_mean is not None:
        save_mean.copy_(mean)
        save_invstd.copy_(invstd)
    out.copy_(torch.relu((a - mean) * invstd * gamma + beta))
    return out


def d_mid(a0, w1, b1, a1, mask1, mask0, w_head, b_head, y, rs_loss, loss_out, rs_sig, z1, z0, *, u=None, rng=None,
          stream=0, slope=0.2, p=0.5, counter=None):
    """Layer 1 of the two-hidden-layer discriminator, its head and the backward signals of both (the composition the
    fused ``ftg_d_mid`` kernel replaces): ``linear(leaky_dropout)`` -> ``d_head`` -> ``linear(b_t, mask_in)``."""
    linear(a0, w1, a1, bias=b1, act="leaky_dropout", mask_out=mask1, u=u, rng=rng, stream=stream, slope=slope, p=p)
    d_head(a1, w_head, b_head, y, rs_loss, loss_out, rs_sig, mask1, z1)
    linear(z1, w1, z0, b_t=True, mask_in=mask0)
    return y


def linear_bn_relu(x, w, a_out, out, *, bias, gamma, beta, running_mean, running_var, save_mean=None, save_invstd=None,
                   eps=1e-5, momentum=0.1, num_batches_tracked=None, precision=None):
    """``a_out = x @ w.T + bias``; ``out = relu(batchnorm_train(a_out))`` (Residual block, ``ctgan.py:40-44``)."""
    linear(x, w, a_out, bias=bias)
    return bn_relu_fwd(a_out, gamma, beta, running_mean, running_var, out, save_mean, save_invstd, training=True,
                       eps=eps, momentum=momentum, num_batches_tracked=num_batches_tracked)


def bn_relu_bwd(dout, a, save_mean, save_invstd, gamma, beta, da, dgamma, dbeta, dbias_fc=None):
    """Backward of ``relu(bn_train(a))``.  ``dgamma``/``dbeta`` are overwritten; ``dbias_fc``
    (optional) receives the column sum of ``da`` = gradient of the preceding Linear's bias."""
    n = a.shape[0]
    xhat = (a - save_mean) * save_invstd
    dy = dout * ((xhat * gamma + beta) > 0).to(a.dtype)
    db = dy.sum(0)
    dg = (dy * xhat).sum(0)
    da.copy_(gamma * save_invstd / n * (n * dy - db - xhat * dg))
    dgamma.copy_(dg)
    dbeta.copy_(db)
    if dbias_fc is not None:
        dbias_fc.copy_(da.sum(0))
    return da


# ----------------------------------------------------------------- apply_activate
def _gumbel_from_u(u):
    u = u.clamp(TINY, 1.0 - 1e-7)
    return -torch.log(-torch.log(u))


def activate_fwd(fake, tables, out, *, tau=0.2, u=None, rng=None, stream=0, cond=None, hard_noise=True,
                 slerp_real=None, slerp_out=None, alpha=None, slerp_stream=0):
    """``out[:, :data_dim] = apply_activate(fake)``; optionally ``out[:, data_dim:] = cond``.

    ``tables`` = ``TableLayout.device_tables``.  ``u`` are the uniforms behind the gumbel
    noise (same shape as the data part).  ``hard_noise=False`` disables the noise (test)."""
    dd = tables["col_span"].numel()
    x = fake[:, :dd]
    if hard_noise:
        if u is None:
            u = rng.uniform(*x.shape)
        noise = _gumbel_from_u(u)
    else:
        noise = torch.zeros_like(x)
    starts = tables["span_start"].tolist()
    widths = tables["span_width"].tolist()
    kinds = tables["span_kind"].tolist()
    for st, w, k in zip(starts, widths, kinds):
        if k == 0:
            out[:, st:st + w] = torch.tanh(x[:, st:st + w])
        else:
            out[:, st:st + w] = torch.softmax((x[:, st:st + w] + noise[:, st:st + w]) / tau, dim=1)
    if cond is not None:
        out[:, dd:dd + cond.shape[1]] = cond
    if slerp_real is not None:      # fused on the device: interpolates = slerp(alpha, real_cat, fake_cat)
        din = slerp_real.shape[1]
        slerp(slerp_real, out[:, :din], slerp_out, alpha=alpha, rng=rng, stream=slerp_stream)
    return out


def gen_output_bwd(dout, act, fake, cond, mask, tables, dfake, loss_out, *, tau=0.2):
    """``activate_bwd`` followed by the ``cond_loss`` value and gradient (one launch on the device)."""
    activate_bwd(dout, act, tables, dfake, tau=tau)
    cond_loss_fwd_bwd(fake, cond, mask, tables, dfake, loss_out, accumulate=True)
    return dfake


def activate_bwd(dout, act, tables, dfake, *, tau=0.2, accumulate=False):
    """Gradient of ``apply_activate`` wrt its pre-activation input, given its output ``act``."""
    dd = tables["col_span"].numel()
    g = torch.empty(dout.shape[0], dd, dtype=dout.dtype, device=dout.device)
    starts = tables["span_start"].tolist()
    widths = tables["span_width"].tolist()
    kinds = tables["span_kind"].tolist()
    for st, w, k in zip(starts, widths, kinds):
        y = act[:, st:st + w]
        d = dout[:, st:st + w]
        if k == 0:
            g[:, st:st + w] = d * (1 - y * y)
        else:
            g[:, st:st + w] = y * (d - (d * y).sum(1, keepdim=True)) / tau
    if accumulate:
        dfake[:, :dd].add_(g)
    else:
        dfake[:, :dd].copy_(g)
    return dfake


def cond_loss_fwd_bwd(fake, cond, mask, tables, dfake, loss_out, *, accumulate=True):
    """Masked CE of ``ctgan.cond_loss`` and its gradient wrt ``fake`` (added to ``dfake``)."""
    B = fake.shape[0]
    s_start = tables["soft_start"].tolist()
    s_width = tables["soft_width"].tolist()
    s_cst = tables["soft_cstart"].tolist()
    total = fake.new_zeros(())
    if not accumulate:
        dfake[:, : tables["col_span"].numel()].zero_()
    for j, (st, w, cst) in enumerate(zip(s_start, s_width, s_cst)):
        logits = fake[:, st:st + w]
        target = torch.argmax(cond[:, cst:cst + w], dim=1)
        lse = torch.logsumexp(logits, dim=1)
        ce = lse - logits.gather(1, target.view(-1, 1)).squeeze(1)
        mj = mask[:, j]
        total = total + (ce * mj).sum()
        g = torch.softmax(logits, dim=1)
        g = g - F.one_hot(target, w).to(g.dtype)
        dfake[:, st:st + w].add_(g * (mj / B).view(-1, 1))
    loss_out.copy_(total / B)
    return loss_out


# ------------------------------------------------------------- slerp / penalty
def slerp(low, high, out, *, alpha=None, rng=None, stream=0):
    """Row-wise spherical interpolation ``alpha`` of the way from ``low`` to ``high``."""
    if alpha is None:
        alpha = rng.uniform(low.shape[0])
    ln = low / low.norm(dim=1, keepdim=True)
    hn = high / high.norm(dim=1, keepdim=True)
    omega = torch.acos((ln * hn).sum(1)).view(-1, 1)
    so = torch.sin(omega)
    a = alpha.view(-1, 1)
    out.copy_((torch.sin((1.0 - a) * omega) / so) * low + (torch.sin(a * omega) / so) * high)
    return out


def gp_penalty(g, pen_out, *, lambda_=10.0):
    """``g`` is ``[Bp, Din]`` (input-gradient per pack).  Writes
    ``pen = lambda * mean((|g|-1)^2)`` and overwrites ``g`` with ``d pen / d g``."""
    bp = g.shape[0]
    nrm = g.norm(2, dim=1)
    pen_out.copy_(((nrm - 1) ** 2).mean() * lambda_)
    coef = (2.0 * lambda_ / bp) * (nrm - 1) / nrm.clamp_min(1e-30)
    g.mul_(coef.view(-1, 1))
    return pen_out


# ------------------------------------------------------------------------ Adam
def adam_step(param, grad, exp_avg, exp_avg_sq, step, *, lr, beta1, beta2, eps, weight_decay=0.0):
    """``torch.optim.Adam`` (non-amsgrad, L2-in-gradient weight decay) over flat tensors.
    ``step`` is the 1-based step count (python int or 0-dim/1-elem tensor)."""
    t = float(step.item()) if torch.is_tensor(step) else float(step)
    g = grad + weight_decay * param if weight_decay != 0.0 else grad
    exp_avg.mul_(beta1).add_(g, alpha=1 - beta1)
    exp_avg_sq.mul_(beta2).addcmul_(g, g, value=1 - beta2)
    bc1 = 1 - beta1 ** t
    bc2 = 1 - beta2 ** t
    denom = (exp_avg_sq.sqrt() / math.sqrt(bc2)).add_(eps)
    param.addcdiv_(exp_avg, denom, value=-lr / bc1)
    return param


# --------------------------------------------------------------------- sampling
def fill_normal(out, *, z=None, rng=None, stream=0):
    out.copy_(rng.normal(*out.shape) if z is None else z)
    return out


def sample_condvec(cum_p, soft_width, col_out, opt_out, *, u_col=None, u_opt=None, rng=None, stream=0):
    """Column ~ U{0..n_col-1}; option ~ the column's cumulative table ``cum_p``
    (``Cond.sample``, ``ctgan.py:147-161``): first index whose cumulative prob exceeds u."""
    n = col_out.shape[0]
    if u_col is None:
        u_col = rng.uniform(n)
    if u_opt is None:
        u_opt = rng.uniform(n)
    n_col = cum_p.shape[0]
    col = torch.clamp((u_col * n_col).long(), max=n_col - 1)
    opt = (cum_p[col] > u_opt.view(-1, 1)).float().argmax(dim=1)
    opt = torch.minimum(opt, soft_width.long()[col] - 1)
    col_out.copy_(col.to(col_out.dtype))
    opt_out.copy_(opt.to(opt_out.dtype))
    return col_out, opt_out


def write_condvec(col, opt, soft_cstart, c_out, m_out=None, perm=None):
    """One-hot conditional vector (and column mask) from ``(col, opt)``; with ``perm`` row i
    uses entry ``perm[i]``."""
    if perm is not None:
        col, opt = col[perm.long()], opt[perm.long()]
    c_out.zero_()
    rows = torch.arange(col.shape[0], device=col.device)
    c_out[rows, soft_cstart.long()[col.long()] + opt.long()] = 1.0
    if m_out is not None:
        m_out.zero_()
        m_out[rows, col.long()] = 1.0
    return c_out


def sample_rows(row_ptr, row_ids, soft_cstart, col, opt, rid_out, *, perm=None, u=None, rng=None, stream=0):
    """Uniformly random training row having option ``opt`` in span ``col``
    (``Sampler.sample``, ``ctgan.py:221-228``)."""
    if perm is not None:
        col, opt = col[perm.long()], opt[perm.long()]
    if u is None:
        u = rng.uniform(col.shape[0])
    g = soft_cstart.long()[col.long()] + opt.long()
    lo = row_ptr.long()[g]
    cnt = row_ptr.long()[g + 1] - lo
    k = torch.clamp((u * cnt).long(), max=(cnt - 1).clamp_min(0))
    rid_out.copy_(row_ids.long()[(lo + k).clamp(max=row_ids.numel() - 1)].to(rid_out.dtype))
    return rid_out


def expand_rows(vals, codes, rid, tables, out):
    """Compact encoded rows -> dense ``[tanh scalar | one-hot ...]`` rows in ``out[:, :data_dim]``."""
    dd = tables["col_span"].numel()
    r = rid.long()
    out[:, :dd] = 0
    ts = tables["tanh_start"].long()
    if ts.numel():
        out[:, ts] = vals[r]
    rows = torch.arange(r.shape[0], device=out.device)
    ss = tables["soft_start"].long()
    cc = codes[r].long()
    for j in range(ss.numel()):
        out[rows, ss[j] + cc[:, j]] = 1.0
    return out


def argsort_perm(perm_out, *, u=None, rng=None, stream=0):
    """Random permutation = argsort of iid uniforms."""
    if u is None:
        u = rng.uniform(perm_out.shape[0])
    perm_out.copy_(torch.argsort(u).to(perm_out.dtype))
    return perm_out


# ------------------------------------------------------------------ VGM codec
def vgm_decode(act, tables, mu, sd, cat_ids, cat_off, col_is_cont, out):
    """``BGM_CTGAN_Transformer.inverse_transform`` on activated rows -> ``[rows, n_columns]``."""
    st = 0
    ci = 0
    ki = 0
    for j in range(col_is_cont.numel()):
        if int(col_is_cont[j]):
            u = act[:, st].clamp(-1, 1)
            # width of the following softmax span
            w = int(tables["span_width"][torch.nonzero(tables["span_start"] == st + 1)[0, 0]])
            pick = act[:, st + 1: st + 1 + w].argmax(1)
            out[:, j] = u * 4.0 * sd[ci][pick] + mu[ci][pick]
            st += 1 + w
            ci += 1
        else:
            lo, hi = int(cat_off[ki]), int(cat_off[ki + 1])
            w = hi - lo
            pick = act[:, st: st + w].argmax(1)
            out[:, j] = cat_ids[lo:hi][pick]
            st += w
            ki += 1
    return out


def vgm_encode(x, logit_const, mu, var, active, u, val_out, code_out, rng=None, stream=0):
    """One continuous column: posterior over active modes, sampled mode, normalised scalar.

    ``logit_const, mu, var`` are ``[K]``; ``active`` a bool ``[K]`` mask; ``u`` uniforms."""
    if u is None:
        u = rng.uniform(x.shape[0])
    lc, m, v = logit_const[active], mu[active], var[active]
    logit = lc.view(1, -1) - 0.5 * (x.view(-1, 1) - m.view(1, -1)) ** 2 / v.view(1, -1)
    p = torch.softmax(logit, dim=1) + 1e-6
    p = p / p.sum(1, keepdim=True)
    sel = (p.cumsum(1) > u.view(-1, 1)).float().argmax(1)
    val_out.copy_(((x - m[sel]) / (4.0 * v[sel].sqrt())).clamp(-0.99, 0.99))
    code_out.copy_(sel.to(code_out.dtype))
    return val_out, code_out


# ------------------------------------------------------------ federated average
def fedavg(arenas, weights, int_offsets=()):
    """``theta <- sum_i w_i theta_i`` written back into every replica, fixed client order.
    Reference semantics: ``average_model`` (``Server/dtds/distributed.py:86-106``) incl. the
    truncation of ``num_batches_tracked`` back to an integer."""
    if len(arenas) == 1:
        return arenas
    acc = arenas[0] * float(weights[0])
    for a, w in zip(arenas[1:], weights[1:]):
        acc += a * float(w)
    for off in int_offsets:
        acc[off] = torch.trunc(acc[off])
    for a in arenas:
        a.copy_(acc)
    return arenas
