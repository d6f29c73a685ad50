"""Statistical similarity of synthetic vs. real tables: Avg_JSD / Avg_WD.

Parity: ``Server/similarity_analysis.py:15-117``.  Categorical columns: Jensen-Shannon
distance (base 2) between category frequency vectors over the real table's categories;
continuous columns: Wasserstein distance after min-max scaling fitted on the real column.
The column list and paths are parameters (the reference hard-codes Intrusion's).
"""

from __future__ import annotations

import argparse

import numpy as np
import pandas as pd

from ..data.schemas import INTRUSION_CATEGORICAL
from ..parallel.weights import wasserstein_1d


def _jsd_base2(p, q):
    p = np.asarray(p, dtype=np.float64)
    q = np.asarray(q, dtype=np.float64)
    p, q = p / p.sum(), q / q.sum()
    m = 0.5 * (p + q)

    def kl(a, b):
        k = a > 0
        return np.sum(a[k] * np.log2(a[k] / b[k]))

    return float(np.sqrt(max(0.5 * (kl(p, m) + kl(q, m)), 0.0)))


def column_similarity(real: pd.DataFrame, fake: pd.DataFrame, cat_cols):
    cat_cols = list(cat_cols or [])
    stats = {}
    for col in real.columns:
        if col in cat_cols:
            rp = real[col].value_counts(normalize=True)
            fp = fake[col].value_counts(normalize=True)
            cats = sorted(rp.index.tolist(), key=str)
            stats[col] = _jsd_base2([rp[c] for c in cats], [fp.get(c, 0.0) for c in cats])
        else:
            r = real[col].to_numpy(dtype=np.float64)
            f = fake[col].to_numpy(dtype=np.float64)
            lo, hi = r.min(), r.max()
            scale = (hi - lo) if hi > lo else 1.0
            stats[col] = wasserstein_1d((r - lo) / scale, (f - lo) / scale)
    return stats


def stat_sim_normalize(real_path, fake_path, cat_cols=None):
    """``(Avg_JSD, Avg_WD)`` for one synthetic CSV."""
    real = pd.read_csv(real_path) if isinstance(real_path, str) else real_path
    fake = pd.read_csv(fake_path) if isinstance(fake_path, str) else fake_path
    stats = column_similarity(real, fake, cat_cols)
    cat = [v for k, v in stats.items() if k in (cat_cols or [])]
    num = [v for k, v in stats.items() if k not in (cat_cols or [])]
    return (float(np.mean(cat)) if cat else float("nan"), float(np.mean(num)) if num else float("nan"))


def analyse_run(real_path, result_prefix, n_epoch, cat_cols, timestamp_csv="timestamp_experiment.csv",
                out_csv="Intrusion_statistical_similarity_analysis.csv"):
    rows = []
    for i in range(int(n_epoch)):
        jsd, wd = stat_sim_normalize(real_path, f"{result_prefix}{i}.csv", cat_cols)
        rows.append([i, jsd, wd])
    res = pd.DataFrame(rows, columns=["Epoch_No.", "Avg_JSD", "Avg_WD"])
    try:
        ts = pd.read_csv(timestamp_csv, header=None)
        res["time_stamp"] = ts.iloc[:, 0].cumsum().iloc[: len(res)].to_numpy()
    except FileNotFoundError:
        pass
    res.to_csv(out_csv, index=False)
    return res


def main(argv=None):
    p = argparse.ArgumentParser()
    p.add_argument("-nepoch", help="number of training epoch", required=True)
    p.add_argument("-name", default="Intrusion")
    p.add_argument("-real_path", default=None)
    p.add_argument("-categorical_list", default=None, help="comma separated; default = Intrusion's")
    a = p.parse_args(argv)
    cats = a.categorical_list.split(",") if a.categorical_list else list(INTRUSION_CATEGORICAL)
    real = a.real_path or f"data/raw/{a.name}_train.csv"
    res = analyse_run(real, f"{a.name}_result/{a.name}_synthesis_epoch_", a.nepoch, cats,
                      out_csv=f"{a.name}_statistical_similarity_analysis.csv")
    print(res.to_string(index=False))
    return res
