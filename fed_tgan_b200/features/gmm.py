"""1-D variational (Dirichlet-process) Gaussian mixtures as fixed-layout arrays.

The reference ships whole pickled ``sklearn.mixture.BayesianGaussianMixture`` objects over
RPC (``Server/dtds/distributed.py:697,765``).  Here a column's mixture is a
``GMMParams`` record of ``[K]`` arrays -- the only statistics the encoder, the sampler and
the posterior need -- so it can live in a device struct / symmetric buffer and be read by
the ``fed_stats_to_weights`` and ``vgm_encode`` kernels.

``fit_vbgmm_1d`` is an in-house batched variational-Bayes fit (same model and priors as
``BayesianGaussianMixture(n_components=10, weight_concentration_prior_type=
'dirichlet_process', weight_concentration_prior=1e-3, n_init=1)`` used at
``Server/dtds/features/transformers.py:334-340`` and ``distributed.py:725-730``): all
continuous columns of a table are fitted in one batched tensor program that runs on the
GPU when given CUDA tensors.  sklearn stays available as the oracle
(``backend="sklearn"``).
"""

from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np
import torch
from scipy.special import digamma

_FIELDS = ("means", "covs", "weights", "dof", "mean_prec", "wc_a", "wc_b")


@dataclass
class GMMParams:
    means: np.ndarray       # [K]
    covs: np.ndarray        # [K] variances (sklearn ``covariances_`` squeezed)
    weights: np.ndarray     # [K] normalised stick-breaking expectation (``weights_``)
    dof: np.ndarray         # [K] Wishart degrees of freedom
    mean_prec: np.ndarray   # [K] mean precision (kappa)
    wc_a: np.ndarray        # [K] Beta(a, b) stick parameters
    wc_b: np.ndarray

    # ---- sklearn-compatible attribute names (read-only views) -----------------
    @property
    def means_(self):
        return self.means.reshape(-1, 1)

    @property
    def covariances_(self):
        return self.covs.reshape(-1, 1, 1)

    @property
    def weights_(self):
        return self.weights

    @property
    def n_components(self):
        return int(self.means.shape[0])

    @property
    def stds(self):
        return np.sqrt(self.covs)

    # ---- construction / wire format ------------------------------------------
    @classmethod
    def from_sklearn(cls, gm) -> "GMMParams":
        if hasattr(gm, "weight_concentration_") and isinstance(gm.weight_concentration_, tuple):
            a, b = gm.weight_concentration_
            dof, mp = gm.degrees_of_freedom_, gm.mean_precision_
        else:  # plain GaussianMixture: no variational posterior, emulate with sharp priors
            k = gm.means_.shape[0]
            a = b = np.ones(k)
            dof = np.full(k, 1e12)
            mp = np.full(k, 1e12)
        return cls(
            means=np.asarray(gm.means_, dtype=np.float64).reshape(-1).copy(),
            covs=np.asarray(gm.covariances_, dtype=np.float64).reshape(-1).copy(),
            weights=np.asarray(gm.weights_, dtype=np.float64).copy(),
            dof=np.asarray(dof, dtype=np.float64).copy(),
            mean_prec=np.asarray(mp, dtype=np.float64).copy(),
            wc_a=np.asarray(a, dtype=np.float64).copy(),
            wc_b=np.asarray(b, dtype=np.float64).copy(),
        )

    @classmethod
    def coerce(cls, obj) -> "GMMParams":
        if obj is None or isinstance(obj, cls):
            return obj
        return cls.from_sklearn(obj)

    def to_array(self) -> np.ndarray:
        """``[7, K]`` float64 block (the fixed wire / device layout)."""
        return np.stack([getattr(self, f) for f in _FIELDS]).astype(np.float64)

    @classmethod
    def from_array(cls, arr) -> "GMMParams":
        arr = np.asarray(arr, dtype=np.float64)
        return cls(*[arr[i].copy() for i in range(len(_FIELDS))])

    # ---- posterior -------------------------------------------------------------
    def logit_const(self) -> np.ndarray:
        """``A_k`` with ``log r_k(x) = A_k - 0.5 (x-mu_k)^2 / cov_k + const``.

        Restates sklearn's ``_estimate_log_weights`` + ``_estimate_log_prob`` for
        ``n_features == 1``; terms identical for all k (``log 2pi``, ``log 2``) are dropped.
        """
        dsum = digamma(self.wc_a + self.wc_b)
        e_log_pi = digamma(self.wc_a) - dsum + np.concatenate(([0.0], np.cumsum(digamma(self.wc_b) - dsum)[:-1]))
        return (
            e_log_pi
            - 0.5 * np.log(self.covs)
            - 0.5 * np.log(self.dof)
            + 0.5 * (digamma(0.5 * self.dof) - 1.0 / self.mean_prec)
        )

    def predict_proba(self, x: np.ndarray) -> np.ndarray:
        x = np.asarray(x, dtype=np.float64).reshape(-1, 1)
        logit = self.logit_const()[None, :] - 0.5 * (x - self.means[None, :]) ** 2 / self.covs[None, :]
        logit -= logit.max(axis=1, keepdims=True)
        p = np.exp(logit)
        return p / p.sum(axis=1, keepdims=True)

    # ---- sampling (sklearn ``BaseMixture.sample`` semantics) ------------------------
    def sample(self, n: int, rng: np.random.Generator | None = None) -> np.ndarray:
        rng = np.random.default_rng() if rng is None else rng
        w = self.weights / self.weights.sum()
        counts = rng.multinomial(int(n), w)
        parts = [rng.normal(m, s, size=c) for m, s, c in zip(self.means, self.stds, counts)]
        return np.concatenate(parts) if parts else np.zeros(0)


# --------------------------------------------------------------------------- fit
def _kmeans_1d(x: torch.Tensor, k: int, iters: int = 50) -> torch.Tensor:
    """Batched 1-D Lloyd iterations. x: [C, n] -> hard labels [C, n]."""
    C, n = x.shape
    q = torch.linspace(0.5 / k, 1 - 0.5 / k, k, device=x.device, dtype=x.dtype)
    centers = torch.quantile(x, q, dim=1).T.contiguous()           # [C, k]
    labels = None
    for _ in range(iters):
        d = (x[:, :, None] - centers[:, None, :]).abs()           # [C, n, k]
        labels = d.argmin(dim=2)
        onehot = torch.nn.functional.one_hot(labels, k).to(x.dtype)
        cnt = onehot.sum(dim=1)
        s = (onehot * x[:, :, None]).sum(dim=1)
        centers = torch.where(cnt > 0, s / cnt.clamp_min(1), centers)
    return labels


def _betaln(a, b):
    return torch.lgamma(a) + torch.lgamma(b) - torch.lgamma(a + b)


@torch.no_grad()
def fit_vbgmm_1d(
    x,
    n_components: int = 10,
    dp_prior: float = 1e-3,
    max_iter: int = 100,
    tol: float = 1e-3,
    reg_covar: float = 1e-6,
    device=None,
    use_kernel: bool = True,
    init_centers=None,
):
    """Fit DP-GMMs to every row of ``x`` (``[C, n]``: C columns, n samples) at once.

    Returns a list of ``C`` ``GMMParams``.  Priors follow sklearn's defaults:
    ``mean_precision_prior=1``, ``mean_prior=mean(x)``, ``degrees_of_freedom_prior=1``,
    ``covariance_prior=var(x, ddof=1)``.
    """
    x = torch.as_tensor(np.asarray(x) if not torch.is_tensor(x) else x)
    if x.dim() == 1:
        x = x[None, :]
    dev = torch.device(device) if device is not None else x.device
    if dev.type == "cuda" and n_components <= 12 and use_kernel:
        fitted = _fit_vbgmm_kernel(x.to(dev), n_components, dp_prior, max_iter, tol, reg_covar, init_centers)
        if fitted is not None:
            return fitted
    x = x.to(device=dev, dtype=torch.float64)
    C, n = x.shape
    K = n_components
    eps10 = 10 * torch.finfo(torch.float64).eps

    mean_prior = x.mean(dim=1)                                      # [C]
    cov_prior = x.var(dim=1, unbiased=True).clamp_min(1e-12)        # [C]
    kappa0, nu0 = 1.0, 1.0

    labels = _kmeans_1d(x, K)
    resp = torch.nn.functional.one_hot(labels, K).to(torch.float64)  # [C, n, K]

    def m_step(resp):
        nk = resp.sum(dim=1) + eps10                                # [C, K]
        xk = (resp * x[:, :, None]).sum(dim=1) / nk
        sk = (resp * (x[:, :, None] - xk[:, None, :]) ** 2).sum(dim=1) / nk + reg_covar
        a = 1.0 + nk
        tail = torch.flip(torch.cumsum(torch.flip(nk, [1]), 1), [1])  # sum_{j>=k} n_j
        b = dp_prior + torch.cat([tail[:, 1:], torch.zeros(C, 1, dtype=nk.dtype, device=nk.device)], 1)
        kappa = kappa0 + nk
        mu = (kappa0 * mean_prior[:, None] + nk * xk) / kappa
        nu = nu0 + nk
        diff = xk - mean_prior[:, None]
        cov = (cov_prior[:, None] + nk * sk + nk * kappa0 / kappa * diff * diff) / nu
        return a, b, kappa, mu, nu, cov

    def e_step(a, b, kappa, mu, nu, cov):
        dsum = torch.digamma(a + b)
        cum = torch.cumsum(torch.digamma(b) - dsum, 1)
        e_log_pi = torch.digamma(a) - dsum + torch.cat([torch.zeros(C, 1, dtype=a.dtype, device=a.device), cum[:, :-1]], 1)
        const = (
            e_log_pi - 0.5 * math.log(2 * math.pi) - 0.5 * torch.log(cov) - 0.5 * torch.log(nu)
            + 0.5 * (math.log(2.0) + torch.digamma(0.5 * nu) - 1.0 / kappa)
        )
        wlp = const[:, None, :] - 0.5 * (x[:, :, None] - mu[:, None, :]) ** 2 / cov[:, None, :]
        lse = torch.logsumexp(wlp, dim=2, keepdim=True)
        return wlp - lse

    def lower_bound(log_resp, a, b, kappa, nu, cov):
        log_det_chol = -0.5 * torch.log(cov) - 0.5 * torch.log(nu)
        log_wishart = -(nu * log_det_chol + nu * 0.5 * math.log(2.0) + torch.lgamma(0.5 * nu)).sum(1)
        log_norm_w = -_betaln(a, b).sum(1)
        r = log_resp.exp()
        return -(r * log_resp).sum(dim=(1, 2)) - log_wishart - log_norm_w - 0.5 * torch.log(kappa).sum(1)

    params = m_step(resp)
    prev = torch.full((C,), -float("inf"), dtype=torch.float64, device=x.device)
    active = torch.ones(C, dtype=torch.bool, device=x.device)
    for _ in range(max_iter):
        log_resp = e_step(*params)
        new = m_step(log_resp.exp())
        # freeze columns that already converged (sklearn stops each fit independently)
        params = tuple(torch.where(active[:, None], nw, old) for nw, old in zip(new, params))
        lb = lower_bound(log_resp, params[0], params[1], params[2], params[4], params[5])
        done = (lb - prev).abs() < tol
        prev = torch.where(active, lb, prev)
        active = active & ~done
        if not bool(active.any()):
            break

    a, b, kappa, mu, nu, cov = (p.cpu().numpy() for p in params)
    out = []
    for c in range(C):
        s = a[c] + b[c]
        w = a[c] / s * np.concatenate(([1.0], np.cumprod((b[c] / s)[:-1])))
        w = w / w.sum()
        out.append(GMMParams(mu[c].copy(), cov[c].copy(), w, nu[c].copy(), kappa[c].copy(), a[c].copy(), b[c].copy()))
    return out


def _fit_vbgmm_kernel(x, K, dp_prior, max_iter, tol, reg_covar, init_centers=None):
    """One CTA per column on the device (``csrc/vbgmm.cu``); ``None`` if the native library is missing.
    ``init_centers`` ([C, K] quantile seeds, e.g. from the fused statistics kernel) avoids sorting here."""
    try:
        from ..ops import native
    except Exception:
        return None
    xf = x.to(torch.float32).contiguous()
    if init_centers is None:
        init_centers = native.quantile_centers(xf, K)
    out, _ = native.vbgmm_fit(xf, init_centers, dp_prior, max_iter, tol, reg_covar)
    arr = out.cpu().numpy()
    return [GMMParams.from_array(arr[c]) for c in range(arr.shape[0])]


def fit_sklearn_1d(col: np.ndarray, n_components=10, dp_prior=1e-3, random_state=None) -> GMMParams:
    from sklearn.mixture import BayesianGaussianMixture

    gm = BayesianGaussianMixture(
        n_components=n_components,
        weight_concentration_prior_type="dirichlet_process",
        weight_concentration_prior=dp_prior,
        n_init=1,
        random_state=random_state,
    )
    gm.fit(np.asarray(col, dtype=np.float64).reshape(-1, 1))
    return GMMParams.from_sklearn(gm)
