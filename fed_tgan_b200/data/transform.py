"""Decode a generated numeric table back to user-facing values.

Parity: ``Server/dtds/data/utils/transform.py:8-69`` (``Transform.inverse``): column names
from the meta JSON, ``LabelEncoder.inverse_transform`` per categorical column,
``exp(x)-1`` on non-negative columns (ceil when negative; ``-1`` -> ``"empty"``), date
re-assembly, ``"empty"`` -> ``" "``.  Returns ``(df, time_stamp, column_names)``.

``decode_arrays`` is the vectorised core used by the fast CSV path: it avoids the
per-cell pandas ``apply`` lambdas of the reference (hot: 40 000 x cols every epoch,
SURVEY H8).
"""

from __future__ import annotations

import datetime
import json

import numpy as np
import pandas as pd

from .constants import EMPTY_TOKEN
from .date import Date


def _load_meta(meta_json):
    if isinstance(meta_json, dict):
        return meta_json
    with open(meta_json) as fh:
        return json.load(fh)


def _expm1_column(x: np.ndarray):
    """Vectorised form of the reference's two ``apply`` passes on a log-space column."""
    y = np.exp(x.astype(np.float64)) - 1.0
    y = np.where(y < 0, np.ceil(y), y)
    return y


class Transform:
    @staticmethod
    def decode_arrays(data, meta_json, LEs, as_codes=False):
        """Return ``(column_names, list_of_column_arrays)`` with decoded values.  With
        ``as_codes`` categorical columns come back as ``(int32 codes, class strings)`` pairs
        (what the native CSV writer consumes) instead of object arrays."""
        meta = _load_meta(meta_json)
        names = [c["column_name"] for c in meta["columns"]]
        data = np.asarray(data)
        le_by_col = {item["column_name"]: item["label_encoder"] for item in LEs}
        non_neg = set(meta.get("non_negative_cols", []))
        cols = []
        for j, name in enumerate(names):
            raw = data[:, j]
            if name in le_by_col and as_codes:
                classes = [" " if str(c) == EMPTY_TOKEN else str(c) for c in le_by_col[name].classes_]
                codes = np.clip(raw.astype(np.int64), 0, len(classes) - 1).astype(np.int32)
                cols.append((codes, classes))
            elif name in le_by_col:
                vals = le_by_col[name].inverse_transform(raw.astype(np.int64))
                if vals.dtype.kind in "OUS":
                    vals = np.where(vals == EMPTY_TOKEN, " ", vals)
                cols.append(vals)
            elif name in non_neg:
                y = _expm1_column(raw)
                if np.any(y == -1):
                    obj = y.astype(object)
                    obj[y == -1] = " "
                    cols.append(obj)
                else:
                    cols.append(y)
            else:
                # generated values carry float32 precision: keep them float32 on the fast path
                cols.append(np.ascontiguousarray(raw) if (as_codes and raw.dtype == np.float32) else raw.astype(np.float64))
        return names, cols

    @staticmethod
    def inverse(data, meta_json, LEs, epochs=None, fname=None):
        meta = _load_meta(meta_json)
        names, cols = Transform.decode_arrays(data, meta, LEs)
        df = pd.DataFrame({n: c for n, c in zip(names, cols)}, columns=names)
        if meta.get("date_info"):
            df = Date.inverse_transform(df, meta["date_info"])
            df = df.replace(EMPTY_TOKEN, " ")
        stamp = str(datetime.datetime.now().timestamp()).replace(".", "")
        return df, stamp, names
