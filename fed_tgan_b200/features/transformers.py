"""Table encoders.

``BGM_CTGAN_Transformer`` is the federated VGM ("mode-specific normalisation") encoder
(parity: ``Server/dtds/features/transformers.py:310-464``): per continuous column a
10-component DP-GMM, components with weight <= 0.005 masked out, a value is emitted as
``[(x-mu_k)/(4 sigma_k) clipped to +-0.99, one-hot(k)]`` with ``k`` *sampled* from the
posterior; categoricals are one-hot in global-frequency order.  ``refit`` swaps in the
federator's global mixtures / masks / category order so that every client has the same
``output_info``.

Differences from the reference (behaviour-preserving):

* mixtures are ``GMMParams`` arrays, not pickled sklearn objects (sklearn objects are
  accepted and converted);
* ``transform`` is vectorised -- the reference loops over rows x columns with
  ``np.random.choice`` (``:406-410``) and ``list.index`` per cell (``:424``);
* ``transform_compact`` emits the compact form ``(vals[rows,n_tanh] f32,
  codes[rows,n_col] i32)`` that the device sampler expands on the fly, so the one-hot
  matrix never has to exist in HBM;
* the fit backend is selectable: ``"sklearn"`` (oracle) or ``"vb"`` (in-house batched fit
  that runs on the GPU).

The SDGym-heritage encoders the reference carries but never calls
(``DiscretizeTransformer``, ``GeneralTransformer``, ``GMMTransformer``, ``BGMTransformer``,
``TableganTransformer``, ``decode_train_data``; ``transformers.py:82-305,467-699``) are
provided with the same public surface.
"""

from __future__ import annotations

import datetime
import json

import numpy as np
import pandas as pd

from ..data.constants import CATEGORICAL, CONTINUOUS, ORDINAL, EMPTY_TOKEN
from ..data.date import Date
from .gmm import GMMParams, fit_sklearn_1d, fit_vbgmm_1d
from .layout import TableLayout


# ----------------------------------------------------------------------------- meta
def _value_order(column: np.ndarray):
    """Values sorted by descending frequency (pandas ``value_counts`` order)."""
    return pd.Series(column).value_counts().index.tolist()


class Transformer:
    """Base: column meta lists (parity ``transformers.py:13-79``)."""

    @staticmethod
    def get_metadata(data, categorical_columns=tuple(), ordinal_columns=tuple()):
        data = np.asarray(data)
        meta = []
        for j in range(data.shape[1]):
            col = data[:, j]
            if j in categorical_columns:
                order = _value_order(col)
                meta.append({"name": j, "type": CATEGORICAL, "size": len(order), "i2s": order})
            elif j in ordinal_columns:
                order = _value_order(col)
                meta.append({"name": j, "type": ORDINAL, "size": len(order), "i2s": order})
            else:
                meta.append({"name": j, "type": CONTINUOUS, "min": col.min(), "max": col.max()})
        return meta

    @staticmethod
    def get_metadata_refit(data, meta_file, label_encoder, categorical_columns=tuple(), ordinal_columns=tuple()):
        """Meta whose categorical ``i2s`` is the label-id list in *global* frequency order
        (``transformers.py:42-71``)."""
        data = np.asarray(data)
        meta = []
        cursor = 0
        for j in range(data.shape[1]):
            col = data[:, j]
            if j in categorical_columns:
                names = meta_file["columns"][j]["i2s"]
                ids = label_encoder[cursor].transform(names).tolist()
                cursor += 1
                meta.append({"name": j, "type": CATEGORICAL, "size": len(names), "i2s": ids})
            elif j in ordinal_columns:
                order = _value_order(col)
                meta.append({"name": j, "type": ORDINAL, "size": len(order), "i2s": order})
            else:
                meta.append({"name": j, "type": CONTINUOUS, "min": col.min(), "max": col.max()})
        return meta

    def fit(self, data, categorical_columns=tuple(), ordinal_columns=tuple()):
        raise NotImplementedError

    def transform(self, data):
        raise NotImplementedError

    def inverse_transform(self, data):
        raise NotImplementedError


def _onehot_by_lookup(col, i2s):
    """One-hot of ``col`` in the slot order given by ``i2s`` (vectorised ``list.index``)."""
    i2s_arr = np.asarray(i2s)
    order = np.argsort(i2s_arr, kind="stable")
    pos = np.searchsorted(i2s_arr[order], col)
    pos = np.clip(pos, 0, len(i2s_arr) - 1)
    slot = order[pos]
    if not np.array_equal(i2s_arr[slot], np.asarray(col)):
        bad = np.asarray(col)[i2s_arr[slot] != np.asarray(col)][:3]
        raise ValueError(f"value(s) {bad!r} not present in category table")
    return slot


# -------------------------------------------------------------------- federated VGM
class BGM_CTGAN_Transformer(Transformer):
    def __init__(self, n_clusters=10, eps=0.005, backend="sklearn", dp_prior=1e-3, device=None, seed=None):
        self.meta = None
        self.n_clusters = n_clusters
        self.eps = eps
        self.backend = backend
        self.dp_prior = dp_prior
        self.device = device
        self.rng = np.random.default_rng(seed)
        self.model = []
        self.components = []
        self.output_info = []
        self.output_dim = 0

    # ---- fitting ---------------------------------------------------------------
    def _fit_columns(self, data, cont_ids):
        if not cont_ids:
            return {}
        if self.backend == "vb":
            fitted = fit_vbgmm_1d(np.ascontiguousarray(data[:, cont_ids].T), self.n_clusters, self.dp_prior,
                                  device=self.device)
            return dict(zip(cont_ids, fitted))
        return {j: fit_sklearn_1d(data[:, j], self.n_clusters, self.dp_prior) for j in cont_ids}

    def _rebuild_output_info(self):
        self.output_info = []
        for j, info in enumerate(self.meta):
            if info["type"] == CONTINUOUS:
                self.output_info += [(1, "tanh"), (int(np.sum(self.components[j])), "softmax")]
            else:
                self.output_info += [(int(info["size"]), "softmax")]
        self.output_dim = int(sum(w for w, _ in self.output_info))
        self.layout = TableLayout(self.output_info)

    def fit(self, data, categorical_columns=tuple(), ordinal_columns=tuple()):
        data = np.asarray(data, dtype=np.float64)
        self.meta = self.get_metadata(data, categorical_columns, ordinal_columns)
        cont = [j for j, info in enumerate(self.meta) if info["type"] == CONTINUOUS]
        fitted = self._fit_columns(data, cont)
        self.model, self.components = [], []
        for j, info in enumerate(self.meta):
            if info["type"] == CONTINUOUS:
                gm = fitted[j]
                self.model.append(gm)
                self.components.append(np.asarray(gm.weights > self.eps))
            else:
                self.model.append(None)
                self.components.append(None)
        self._rebuild_output_info()

    def refit(self, data, meta_file, label_encoder, categorical_columns=tuple(), ordinal_columns=tuple(),
              model=None, components=None):
        self.meta = self.get_metadata_refit(np.asarray(data), meta_file, label_encoder,
                                            categorical_columns, ordinal_columns)
        self.set_model(model, components)
        self._rebuild_output_info()

    def get_information(self):
        return self.model, self.components, self.meta

    def set_model(self, model, components):
        self.model = [GMMParams.coerce(m) for m in model]
        self.components = [None if c is None else np.asarray(c, dtype=bool) for c in components]

    # ---- encode ----------------------------------------------------------------
    def _encode_continuous(self, col, gm: GMMParams, mask):
        probs = gm.predict_proba(col)[:, mask]
        pp = probs + 1e-6
        pp /= pp.sum(axis=1, keepdims=True)
        u = self.rng.random(len(col))[:, None]
        sel = (pp.cumsum(axis=1) > u).argmax(axis=1)
        mu = gm.means[mask][sel]
        sd = gm.stds[mask][sel]
        val = np.clip((col - mu) / (4.0 * sd), -0.99, 0.99)
        return val, sel

    def transform_compact_device(self, data, device):
        """Same as ``transform_compact`` but the continuous columns are encoded by the
        ``vgm_encode`` kernel (posterior over active modes, Philox mode draw, normalised scalar)
        writing straight into the compact table on ``device``; returns device tensors."""
        import torch

        from .. import ops

        data = np.asarray(data, dtype=np.float64)
        n = len(data)
        lay = self.layout
        dev = torch.device(device)
        vals = torch.zeros(n, max(lay.n_tanh, 1), dtype=torch.float32, device=dev)[:, : lay.n_tanh]
        codes = torch.zeros(n, lay.n_col, dtype=torch.int32, device=dev)
        rng = ops.RngState(int(self.rng.integers(1 << 30)), dev)
        rng.tick()
        t = s = 0
        f32 = lambda a: torch.as_tensor(np.ascontiguousarray(a), dtype=torch.float32).to(dev)  # noqa: E731
        for j, info in enumerate(self.meta):
            if info["type"] == CONTINUOUS:
                gm, mask = self.model[j], self.components[j]
                ops.vgm_encode(f32(data[:, j]), f32(gm.logit_const()), f32(gm.means), f32(gm.covs),
                               torch.as_tensor(np.asarray(mask, dtype=bool)).to(dev), None, vals[:, t], codes[:, s],
                               rng=rng, stream=j)
                t += 1
                s += 1
            else:
                codes[:, s] = torch.as_tensor(_onehot_by_lookup(data[:, j], info["i2s"]).astype(np.int32)).to(dev)
                s += 1
        return vals, codes

    def transform_compact(self, data):
        """``(vals [rows, n_tanh] float32, codes [rows, n_col] int32)``."""
        data = np.asarray(data, dtype=np.float64)
        vals, codes = [], []
        for j, info in enumerate(self.meta):
            col = data[:, j]
            if info["type"] == CONTINUOUS:
                v, sel = self._encode_continuous(col, self.model[j], self.components[j])
                vals.append(v)
                codes.append(sel)
            else:
                codes.append(_onehot_by_lookup(col, info["i2s"]))
        n = len(data)
        vals = np.stack(vals, axis=1).astype(np.float32) if vals else np.zeros((n, 0), np.float32)
        codes = np.stack(codes, axis=1).astype(np.int32) if codes else np.zeros((n, 0), np.int32)
        return vals, codes

    def expand(self, vals, codes):
        """Compact form -> dense ``[rows, output_dim]`` (what the reference's transform returns)."""
        lay = self.layout
        out = np.zeros((len(codes), lay.data_dim), dtype=np.float64)
        if lay.n_tanh:
            out[:, lay.tanh_start] = vals
        rows = np.arange(len(codes))
        for s in range(lay.n_col):
            out[rows, lay.soft_start[s] + codes[:, s]] = 1.0
        return out

    def transform(self, data, ispositive=False, positive_list=()):
        if ispositive and positive_list:
            # the reference's (unused) variant encodes |x-mu| for the listed columns
            data = np.asarray(data, dtype=np.float64)
            vals, codes = self.transform_compact(data)
            t = 0
            for j, info in enumerate(self.meta):
                if info["type"] == CONTINUOUS:
                    if j in positive_list:
                        vals[:, t] = np.abs(vals[:, t])
                    t += 1
            return self.expand(vals, codes)
        return self.expand(*self.transform_compact(data))

    # ---- decode ----------------------------------------------------------------
    def decode_tables(self):
        """Per-column decode constants for the device decode kernel (``vgm_decode``).

        Returns dict with ``mu/sd [n_cont, n_clusters]`` restricted to active modes (packed
        left) and ``cat_ids`` (flat label-id table + offsets) for categorical columns.
        """
        mus, sds = [], []
        cat_flat, cat_off = [], [0]
        for j, info in enumerate(self.meta):
            if info["type"] == CONTINUOUS:
                m = self.components[j]
                mu = np.zeros(self.n_clusters)
                sd = np.ones(self.n_clusters)
                k = int(m.sum())
                mu[:k] = self.model[j].means[m]
                sd[:k] = self.model[j].stds[m]
                mus.append(mu)
                sds.append(sd)
            else:
                cat_flat.extend(float(v) for v in info["i2s"])
                cat_off.append(len(cat_flat))
        return {
            "mu": np.asarray(mus, dtype=np.float32).reshape(-1, self.n_clusters),
            "sd": np.asarray(sds, dtype=np.float32).reshape(-1, self.n_clusters),
            "cat_ids": np.asarray(cat_flat, dtype=np.float32),
            "cat_off": np.asarray(cat_off, dtype=np.int32),
            "col_is_cont": np.asarray([info["type"] == CONTINUOUS for info in self.meta], dtype=np.int32),
        }

    def inverse_transform(self, data, sigmas=None):
        data = np.asarray(data)
        out = np.zeros((len(data), len(self.meta)))
        st = 0
        for j, info in enumerate(self.meta):
            if info["type"] == CONTINUOUS:
                m = self.components[j]
                k = int(m.sum())
                u = data[:, st]
                if sigmas is not None:
                    u = self.rng.normal(u, sigmas[st])
                u = np.clip(u, -1, 1)
                pick = np.argmax(data[:, st + 1: st + 1 + k], axis=1)
                out[:, j] = u * 4.0 * self.model[j].stds[m][pick] + self.model[j].means[m][pick]
                st += 1 + k
            else:
                w = info["size"]
                pick = np.argmax(data[:, st: st + w], axis=1)
                out[:, j] = np.asarray(info["i2s"], dtype=np.float64)[pick]
                st += w
        return out


class BGMTransformer(BGM_CTGAN_Transformer):
    """Non-federated twin used by the stand-alone ``CTGANSynthesizer.fit``
    (``transformers.py:467-586``): identical maths, no ``refit``."""


# --------------------------------------------------------------- SDGym-heritage set
class DiscretizeTransformer(Transformer):
    """Uniform binning of continuous columns to ints (``transformers.py:82-133``)."""

    def __init__(self, n_bins):
        self.n_bins = n_bins
        self.meta = None
        self.column_index = None
        self.edges = None

    def fit(self, data, categorical_columns=tuple(), ordinal_columns=tuple()):
        data = np.asarray(data, dtype=np.float64)
        self.meta = self.get_metadata(data, categorical_columns, ordinal_columns)
        self.column_index = [i for i, info in enumerate(self.meta) if info["type"] == CONTINUOUS]
        self.edges = {j: np.linspace(data[:, j].min(), data[:, j].max(), self.n_bins + 1) for j in self.column_index}

    def transform(self, data):
        data = np.array(data, dtype=np.float64)
        for j in self.column_index:
            e = self.edges[j]
            data[:, j] = np.clip(np.searchsorted(e[1:-1], data[:, j], side="right"), 0, self.n_bins - 1)
        return data.astype("int")

    def inverse_transform(self, data):
        data = np.array(data, dtype="float32")
        for j in self.column_index:
            e = self.edges[j]
            centers = 0.5 * (e[:-1] + e[1:])
            data[:, j] = centers[np.clip(data[:, j].astype(int), 0, self.n_bins - 1)]
        return data


class GeneralTransformer(Transformer):
    """Min-max scalars (+ one-hot categoricals) (``transformers.py:136-215``)."""

    def __init__(self, act="sigmoid"):
        self.act = act
        self.meta = None
        self.output_dim = None

    def fit(self, data, categorical_columns=tuple(), ordinal_columns=tuple()):
        self.meta = self.get_metadata(data, categorical_columns, ordinal_columns)
        self.output_dim = sum(1 if i["type"] in (CONTINUOUS, ORDINAL) else i["size"] for i in self.meta)

    def _squash(self, x):
        return x * 2 - 1 if self.act == "tanh" else x

    def _unsquash(self, x):
        return (x + 1) / 2 if self.act == "tanh" else x

    def transform(self, data):
        data = np.asarray(data, dtype=np.float64)
        parts, self.output_info = [], []
        for j, info in enumerate(self.meta):
            col = data[:, j]
            if info["type"] == CONTINUOUS:
                parts.append(self._squash((col - info["min"]) / (info["max"] - info["min"]))[:, None])
                self.output_info.append((1, self.act))
            elif info["type"] == ORDINAL:
                parts.append(self._squash(col / info["size"])[:, None])
                self.output_info.append((1, self.act))
            else:
                hot = np.zeros((len(data), info["size"]))
                hot[np.arange(len(data)), _onehot_by_lookup(col, info["i2s"])] = 1
                parts.append(hot)
                self.output_info.append((info["size"], "softmax"))
        return np.concatenate(parts, axis=1)

    def inverse_transform(self, data):
        data = np.asarray(data)
        out = np.zeros((len(data), len(self.meta)))
        st = 0
        for j, info in enumerate(self.meta):
            if info["type"] == CONTINUOUS:
                cur = np.clip(self._unsquash(data[:, st]), 0, 1)
                out[:, j] = cur * (info["max"] - info["min"]) + info["min"]
                st += 1
            elif info["type"] == ORDINAL:
                cur = self._unsquash(data[:, st]) * info["size"]
                out[:, j] = np.round(cur).clip(0, info["size"] - 1)
                st += 1
            else:
                pick = np.argmax(data[:, st: st + info["size"]], axis=1)
                out[:, j] = np.asarray(info["i2s"], dtype=np.float64)[pick]
                st += info["size"]
        return out


class GMMTransformer(Transformer):
    """Plain 5-component GMM encoder, ``/(2 sigma)``, soft posteriors (``:218-305``)."""

    def __init__(self, n_clusters=5):
        self.meta = None
        self.n_clusters = n_clusters

    def fit(self, data, categorical_columns=tuple(), ordinal_columns=tuple()):
        from sklearn.mixture import GaussianMixture

        data = np.asarray(data, dtype=np.float64)
        self.meta = self.get_metadata(data, categorical_columns, ordinal_columns)
        self.model, self.output_info, self.output_dim = [], [], 0
        for j, info in enumerate(self.meta):
            if info["type"] == CONTINUOUS:
                gm = GaussianMixture(self.n_clusters).fit(data[:, j].reshape(-1, 1))
                self.model.append(gm)
                self.output_info += [(1, "tanh"), (self.n_clusters, "softmax")]
                self.output_dim += 1 + self.n_clusters
            else:
                self.model.append(None)
                self.output_info += [(info["size"], "softmax")]
                self.output_dim += info["size"]

    def transform(self, data):
        data = np.asarray(data, dtype=np.float64)
        parts = []
        for j, info in enumerate(self.meta):
            col = data[:, j]
            if info["type"] == CONTINUOUS:
                gm = self.model[j]
                mu = gm.means_.reshape(-1)
                sd = np.sqrt(gm.covariances_).reshape(-1)
                probs = gm.predict_proba(col.reshape(-1, 1))
                k = probs.argmax(axis=1)
                feat = np.clip((col - mu[k]) / (2 * sd[k]), -0.99, 0.99)
                parts += [feat[:, None], probs]
            else:
                hot = np.zeros((len(data), info["size"]))
                hot[np.arange(len(data)), _onehot_by_lookup(col, info["i2s"])] = 1
                parts.append(hot)
        return np.concatenate(parts, axis=1)

    def inverse_transform(self, data, sigmas=None):
        data = np.asarray(data)
        out = np.zeros((len(data), len(self.meta)))
        st = 0
        for j, info in enumerate(self.meta):
            if info["type"] == CONTINUOUS:
                u = data[:, st]
                if sigmas is not None:
                    u = np.random.normal(u, sigmas[st])
                u = np.clip(u, -1, 1)
                k = np.argmax(data[:, st + 1: st + 1 + self.n_clusters], axis=1)
                mu = self.model[j].means_.reshape(-1)
                sd = np.sqrt(self.model[j].covariances_).reshape(-1)
                out[:, j] = u * 2 * sd[k] + mu[k]
                st += 1 + self.n_clusters
            else:
                pick = np.argmax(data[:, st: st + info["size"]], axis=1)
                out[:, j] = np.asarray(info["i2s"], dtype=np.float64)[pick]
                st += info["size"]
        return out


class TableganTransformer(Transformer):
    """Min-max to [-1,1] and reshape to ``side x side`` images (``:589-626``)."""

    def __init__(self, side):
        self.height = side

    def fit(self, data, categorical_columns=tuple(), ordinal_columns=tuple()):
        self.meta = self.get_metadata(data, categorical_columns, ordinal_columns)
        lo, hi = [], []
        for info in self.meta:
            if info["type"] == CONTINUOUS:
                lo.append(info["min"] - 1e-3)
                hi.append(info["max"] + 1e-3)
            else:
                lo.append(-1e-3)
                hi.append(info["size"] - 1 + 1e-3)
        self.minn, self.maxx = np.asarray(lo, dtype=np.float64), np.asarray(hi, dtype=np.float64)

    def transform(self, data):
        x = (np.asarray(data, dtype="float32") - self.minn) / (self.maxx - self.minn) * 2 - 1
        area = self.height * self.height
        if area > x.shape[1]:
            x = np.concatenate([x, np.zeros((len(x), area - x.shape[1]))], axis=1)
        return x.reshape(-1, 1, self.height, self.height)

    def inverse_transform(self, data):
        flat = np.asarray(data).reshape(-1, self.height * self.height)
        out = np.zeros((len(flat), len(self.meta)))
        for j, info in enumerate(self.meta):
            out[:, j] = (flat[:, j] + 1) / 2 * (self.maxx[j] - self.minn[j]) + self.minn[j]
            if info["type"] in (CATEGORICAL, ORDINAL):
                out[:, j] = np.round(out[:, j])
        return out


def decode_train_data(data, meta_json, LEs, model_name, pr=None, save=True, output_dir="data/generated/"):
    """Decode an encoded table to user values incl. integer casting (``:629-699``)."""
    import os

    if isinstance(meta_json, dict):
        meta = meta_json
    else:
        with open(meta_json) as fh:
            meta = json.load(fh)
    names = [c["column_name"] for c in meta["columns"]]
    cats = {c["column_name"] for c in meta["columns"] if c["type"] == CATEGORICAL}
    ints = list(meta.get("integer_info", []))
    non_neg = set(meta.get("non_negative_cols", []))
    df = pd.DataFrame(np.asarray(data), columns=names)
    for item in LEs:
        name = item["column_name"]
        df[name] = item["label_encoder"].inverse_transform(df[name].astype(int))
    for name in names:
        if name in non_neg:
            y = np.exp(df[name].astype(np.float64).to_numpy()) - 1
            y = np.where(y < 0, np.ceil(y), y)
            if name in ints:
                y = y.astype(np.int64)
            col = pd.Series(y, dtype=object)
            col[y == -1] = EMPTY_TOKEN
            df[name] = col
    if meta.get("date_info"):
        df = Date.inverse_transform(df, meta["date_info"])
    df = df.replace(EMPTY_TOKEN, " ")
    for name in ints:
        if name not in df:
            continue
        if name in cats:
            df[name] = df[name].map(lambda v: int(float(v)) if v != " " else " ")
        else:
            df[name] = df[name].map(lambda v: int(v) if v != " " else " ")
    if save:
        os.makedirs(output_dir, exist_ok=True)
        stamp = str(datetime.datetime.now().timestamp()).replace(".", "")
        df.to_csv(os.path.join(output_dir, f"{model_name.split('-')[0]}_{stamp}.csv"), index=False)
    return df.head(pr) if pr is not None else df
