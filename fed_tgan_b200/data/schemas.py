"""Synthetic tables of the shapes named in BASELINE.json (there is no network: every run
uses synthetic data of the named schema/shape).

* ``intrusion``  KDD-Intrusion schema: 42 columns = 22 continuous + 20 categorical with the
                 category sizes documented in ``Server/models/Intrusion_train.json``
                 (3,63,11,2,3,1,17,2,2,11,2,3,12,7,3,5,1,1,2,23) and the column names /
                 defaults of the reference CLI (``Server/dtds/distributed.py:912-932``)
* ``adult``      ~48 842 rows x 14 mixed columns
* ``covertype``  ~581 012 rows x 54 columns (10 continuous + 44 binary/label)
* ``credit``     ~284 807 rows x 30 continuous + imbalanced label
* ``wide``       1 000 000 rows x 200 mixed columns
"""

from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np
import pandas as pd

INTRUSION_COLUMNS = [
    'duration', 'protocol_type', 'service', 'flag', 'src_bytes', 'dst_bytes', 'land', 'wrong_fragment',
    'urgent', 'hot', 'num_failed_logins', 'logged_in', 'num_compromised', 'root_shell', 'su_attempted',
    'num_root', 'num_file_creations', 'num_shells', 'num_access_files', 'num_outbound_cmds',
    'is_host_login', 'is_guest_login', 'count', 'srv_count', 'serror_rate', 'srv_serror_rate',
    'rerror_rate', 'srv_rerror_rate', 'same_srv_rate', 'diff_srv_rate', 'srv_diff_host_rate',
    'dst_host_count', 'dst_host_srv_count', 'dst_host_same_srv_rate', 'dst_host_diff_srv_rate',
    'dst_host_same_src_port_rate', 'dst_host_srv_diff_host_rate', 'dst_host_serror_rate',
    'dst_host_srv_serror_rate', 'dst_host_rerror_rate', 'dst_host_srv_rerror_rate', 'class',
]
INTRUSION_CATEGORICAL = [
    'protocol_type', 'service', 'flag', 'land', 'wrong_fragment', 'urgent', 'hot', 'num_failed_logins',
    'logged_in', 'num_compromised', 'root_shell', 'su_attempted', 'num_root', 'num_file_creations',
    'num_shells', 'num_access_files', 'num_outbound_cmds', 'is_host_login', 'is_guest_login', 'class',
]
INTRUSION_CAT_SIZES = [3, 63, 11, 2, 3, 1, 17, 2, 2, 11, 2, 3, 12, 7, 3, 5, 1, 1, 2, 23]
INTRUSION_NONNEG = ['dst_bytes', 'src_bytes']


@dataclass
class TableSchema:
    name: str
    rows: int
    columns: list
    categorical: list
    cat_sizes: list
    nonnegative: list = field(default_factory=list)
    target: str = ""
    modes: int = 3           # mixture modes per continuous column of the generator
    label_imbalance: float | None = None

    @property
    def continuous(self):
        cats = set(self.categorical)
        return [c for c in self.columns if c not in cats]


def _schema_generic(name, rows, n_cont, cat_sizes, target_last=True, modes=3, label_imbalance=None):
    cont = [f"x{i}" for i in range(n_cont)]
    cats = [f"c{i}" for i in range(len(cat_sizes))]
    cols = cont + cats
    return TableSchema(name, rows, cols, cats, list(cat_sizes), [], cats[-1] if (cats and target_last) else "",
                       modes, label_imbalance)


def get_schema(name: str, rows: int | None = None) -> TableSchema:
    name = name.lower()
    if name == "intrusion":
        s = TableSchema("Intrusion", 40000, list(INTRUSION_COLUMNS), list(INTRUSION_CATEGORICAL),
                        list(INTRUSION_CAT_SIZES), list(INTRUSION_NONNEG), "class")
    elif name == "adult":
        # 6 continuous + 8 categorical (workclass 9, education 16, marital 7, occupation 15,
        # relationship 6, race 5, sex 2, income 2) = 14 columns
        s = _schema_generic("Adult", 48842, 6, [9, 16, 7, 15, 6, 5, 2, 2])
    elif name == "covertype":
        s = _schema_generic("Covertype", 581012, 10, [2] * 43 + [7])
    elif name == "credit":
        s = _schema_generic("Credit", 284807, 30, [2], modes=5, label_imbalance=0.00173)
    elif name == "wide":
        s = _schema_generic("Wide", 1000000, 100, [2 + (i % 9) for i in range(100)])
    else:
        raise KeyError(name)
    if rows is not None:
        s.rows = int(rows)
    return s


def make_table(schema: TableSchema, seed=0, rows=None, skew: float = 0.0, correlated: float = 0.0) -> pd.DataFrame:
    """Random table of the schema.  ``skew`` in [0,1) tilts category frequencies and mode
    weights (used to build non-IID client shards).  ``correlated`` > 0 redraws the target
    column from a logistic model over the latent codes (mode / category index) of the first
    few feature columns, so that classifiers have something to learn (ML-utility runs); the
    default keeps all columns independent (benchmarks only need the shapes)."""
    rng = np.random.default_rng(seed)
    n = int(rows if rows is not None else schema.rows)
    data = {}
    codes = {}
    cats = dict(zip(schema.categorical, schema.cat_sizes))
    base_rng = np.random.default_rng(12345)     # column-level structure is shared across shards
    for ci, col in enumerate(schema.columns):
        if col in cats:
            k = cats[col]
            p = base_rng.dirichlet(np.ones(k) * 1.5)
            if schema.label_imbalance is not None and col == schema.target and k == 2:
                p = np.array([1 - schema.label_imbalance, schema.label_imbalance])
            if skew > 0:
                tilt = rng.dirichlet(np.ones(k) * 0.8)
                p = (1 - skew) * p + skew * tilt
            ids = rng.choice(k, size=n, p=p / p.sum())
            codes[col] = (ids, k, p / p.sum())
            if correlated > 0 and col == schema.target:
                struct = np.random.default_rng(4242)
                logits = np.tile(np.log(p / p.sum()), (n, 1))
                drivers = [c for c in schema.columns if c != col and c in codes][:4]
                for d in drivers:
                    d_ids, d_k, _ = codes[d]
                    logits += correlated * struct.normal(0, 1.5, size=(d_k, k))[d_ids]
                g = rng.gumbel(size=logits.shape)
                ids = np.argmax(logits + g, axis=1)
            if col == "class" or schema.name != "Intrusion":
                data[col] = np.array([f"{col}_v{i}" for i in range(k)], dtype=object)[ids]
            else:
                data[col] = ids if k > 3 else np.array([f"{col[:3]}{i}" for i in range(k)], dtype=object)[ids]
        else:
            m = schema.modes
            mu = base_rng.normal(0, 4, size=m)
            sd = base_rng.uniform(0.2, 1.2, size=m)
            w = base_rng.dirichlet(np.ones(m) * 2)
            if skew > 0:
                w = (1 - skew) * w + skew * rng.dirichlet(np.ones(m) * 0.8)
            comp = rng.choice(m, size=n, p=w / w.sum())
            codes[col] = (comp, m, w / w.sum())
            x = rng.normal(mu[comp], sd[comp])
            if col in schema.nonnegative:
                x = np.round(np.exp(np.abs(x)))          # heavy-tailed non-negative integers
            data[col] = x
    return pd.DataFrame(data, columns=schema.columns)


def split_table(df: pd.DataFrame, n_clients: int, mode="iid", seed=0):
    """Horizontal partition.  ``iid``: equal random shards.  ``skewed``: sorted by a key
    column then cut into unequal contiguous shards (non-IID in features and size)."""
    rng = np.random.default_rng(seed)
    if mode == "iid":
        perm = rng.permutation(len(df))
        return [df.iloc[idx].reset_index(drop=True) for idx in np.array_split(perm, n_clients)]
    key = df.columns[0]
    order = np.argsort(df[key].to_numpy(), kind="stable")
    fr = np.linspace(1.0, 2.5, n_clients)
    cuts = np.floor(np.cumsum(fr / fr.sum()) * len(df)).astype(int)[:-1]
    return [df.iloc[idx].reset_index(drop=True) for idx in np.split(order, cuts)]
