"""Similarity-weighted aggregation weights (the heart of Fed-TGAN's initialisation).

Reference: ``MDGANServer.uniform_meta_category`` (JSD part, ``Server/dtds/distributed.py:
592-657``), ``uniform_continuous_gmm`` (WD part, ``:689-760``) and
``calculate_final_weights_for_aggregation`` (``:767-783``):

    SD_i  = sum_cols JSD_norm[i, col] + sum_cols WD_norm[i, col]
    w_i   = (1 - SD_i / sum_j SD_j) * rows_i / sum rows ;   weights = softmax(w)

The distances are restated here (numpy, no scipy on the path; scipy is the test oracle)
and as a fused device kernel (``ops.native.fed_stats_to_weights``).
"""

from __future__ import annotations

import numpy as np


def jensenshannon(p, q) -> float:
    """Jensen-Shannon *distance* (sqrt of the divergence, natural log), inputs are
    un-normalised count vectors (scipy ``distance.jensenshannon`` semantics)."""
    p = np.asarray(p, dtype=np.float64)
    q = np.asarray(q, dtype=np.float64)
    p = p / p.sum()
    q = q / q.sum()
    m = 0.5 * (p + q)

    def kl(a, b):
        mask = a > 0
        return float(np.sum(a[mask] * np.log(a[mask] / b[mask])))

    js = 0.5 * (kl(p, m) + kl(q, m))
    return float(np.sqrt(max(js, 0.0)))


def wasserstein_1d(u, v) -> float:
    """First Wasserstein distance between two 1-D empirical samples (scipy
    ``wasserstein_distance`` semantics: integral of |CDF_u - CDF_v|)."""
    u = np.sort(np.asarray(u, dtype=np.float64))
    v = np.sort(np.asarray(v, dtype=np.float64))
    allv = np.concatenate([u, v])
    allv.sort(kind="mergesort")
    deltas = np.diff(allv)
    cu = np.searchsorted(u, allv[:-1], side="right") / u.size
    cv = np.searchsorted(v, allv[:-1], side="right") / v.size
    return float(np.sum(np.abs(cu - cv) * deltas))


def normalise_columns(dist: np.ndarray, zero_fill: float | None) -> np.ndarray:
    """Divide every column by its sum over clients.  Columns summing to 0: ``zero_fill``
    (categorical rule: 1/N, ``distributed.py:650-657``) or left at 0 (continuous rule,
    ``:754-760``)."""
    dist = np.array(dist, dtype=np.float64)
    tot = dist.sum(axis=0)
    nz = tot != 0
    dist[:, nz] = dist[:, nz] / tot[nz]
    if zero_fill is not None:
        dist[:, ~nz] = zero_fill
    return dist


def categorical_similarity(count_vectors) -> np.ndarray:
    """``count_vectors[col][client]`` = count vector over the global categories.
    Returns normalised JSD ``[n_clients, n_cat_cols]``."""
    n_cols = len(count_vectors)
    n_cli = len(count_vectors[0]) if n_cols else 0
    out = np.zeros((n_cli, n_cols))
    for c, per_client in enumerate(count_vectors):
        glob = np.sum(per_client, axis=0)
        for i, vec in enumerate(per_client):
            out[i, c] = jensenshannon(glob, vec)
    return normalise_columns(out, zero_fill=1.0 / max(n_cli, 1))


def continuous_similarity(samples) -> np.ndarray:
    """``samples[col][client]`` = 1-D sample drawn from that client's column mixture.
    Returns normalised WD ``[n_clients, n_cont_cols]`` (pooled vs. each client)."""
    n_cols = len(samples)
    n_cli = len(samples[0]) if n_cols else 0
    out = np.zeros((n_cli, n_cols))
    for c, per_client in enumerate(samples):
        pool = np.concatenate(per_client)
        for i, s in enumerate(per_client):
            out[i, c] = wasserstein_1d(pool, s) if len(s) else 0.0
    return normalise_columns(out, zero_fill=None)


def softmax(v):
    e = np.exp(np.asarray(v, dtype=np.float64))
    return e / e.sum()


def final_weights(jsd_norm: np.ndarray, wd_norm: np.ndarray, rows) -> np.ndarray:
    rows = np.asarray(rows, dtype=np.float64)
    frac = rows / rows.sum()
    sd = np.asarray(wd_norm).sum(axis=1) + np.asarray(jsd_norm).sum(axis=1)
    total = sd.sum()
    w = (1.0 - (sd / total if total != 0 else 0.0)) * frac
    return softmax(w)
