"""Per-client table ingestion: cleaning, file-level meta, label encoding.

Parity target: ``Server/dtds/data/utils/file_generator.py:59-265`` (``FileGenerator``) and
its ``NumpyEncoder`` (``:18-56``).  Behaviour kept:

* integer-valued columns are recorded in ``integer_info`` (``:104-110``);
* blank cells (``" "``/NaN) become the token ``"empty"`` (``:115-116``);
* non-categorical, non-date columns named in ``non_negative_columns`` are moved to log
  space with ``log(x+1)`` (``:123-126``);
* optional date explosion (``:129-133``);
* ``generate_meta_data()`` returns the dict layout of ``:191-231`` (``type`` is
  ``'categorical'`` or the historical ``'continous'``; ``i2s`` = ``{str value: count}``);
* ``generate_data(meta, label_encoders, timestamp)`` label-encodes categoricals with the
  *global* encoders and writes ``<out>/<stem>_CTGANSynthesizer-<ts>/{json,npz,csv}``
  (``:156-188``).

Differences (deliberate): the side-effect files can be switched off
(``write_side_files=False``) and the encoded table is also returned directly, so the hot
init path needs no disk round trip.
"""

from __future__ import annotations

import datetime
import json
import logging
import os
import pickle

import numpy as np
import pandas as pd

from .constants import CATEGORICAL, EMPTY_TOKEN, FILE_META_CONTINUOUS
from .date import Date

logger = logging.getLogger(__name__)


class NumpyEncoder(json.JSONEncoder):
    """JSON encoder that understands numpy scalars/arrays (NumPy 2.x safe: no ``np.float_``)."""

    def default(self, obj):  # noqa: D401
        if isinstance(obj, np.integer):
            return int(obj)
        if isinstance(obj, np.floating):
            return float(obj)
        if isinstance(obj, np.complexfloating):
            return {"real": float(obj.real), "imag": float(obj.imag)}
        if isinstance(obj, np.ndarray):
            return obj.tolist()
        if isinstance(obj, np.bool_):
            return bool(obj)
        if isinstance(obj, np.void):
            return None
        return super().default(obj)


def _timestamp_token() -> str:
    return str(datetime.datetime.now().timestamp()).replace(".", "")


def _is_integer_valued(series: pd.Series) -> bool:
    present = series[~series.isnull()]
    kind = str(present.dtype)
    if "int" in kind:
        return True
    if "float" in kind:
        vals = present.to_numpy()
        with np.errstate(invalid="ignore"):
            return bool(np.array_equal(vals, vals.astype(np.int64)))
    return False


class FileGenerator:
    """Cleans one client's dataframe and produces meta / encoded arrays."""

    def __init__(
        self,
        dataframe: pd.DataFrame,
        file_name: str,
        problem_type: str,
        target_col: str,
        categorical_list: list,
        non_negative_columns: list,
        date_columns: dict | None = None,
        integer_columns: list | None = None,
        synthesizer_used: str = "CTGANSynthesizer",
        output_dir: str = "dtds/saved_models/",
        write_side_files: bool = True,
    ):
        self.file_name = file_name.strip()
        self.problem_type = (problem_type or "").strip()
        self.target_col = (target_col or "").strip()
        self.categorical_list = list(categorical_list)
        self.non_negative_columns = list(non_negative_columns)
        self.date_columns = dict(date_columns or {})
        self.synthesizer_used = synthesizer_used.strip()
        self.output_dir = output_dir.strip()
        self.write_side_files = write_side_files
        self.output_name = f"{self.file_name}_{self.synthesizer_used}-{_timestamp_token()}"
        self.output_path = None

        df = dataframe.copy()
        self.integer_columns = [c for c in df.columns if _is_integer_valued(df[c])]

        df = df.replace(r" ", np.nan).fillna(EMPTY_TOKEN)
        skip = set(self.categorical_list) | set(self.date_columns)
        for col in df.columns:
            if col not in skip and col in self.non_negative_columns:
                df[col] = np.log(df[col].astype(np.float64) + 1.0)
        if self.date_columns:
            self.categorical_list.extend(self.date_columns.keys())
            df = Date.convert(df, self.date_columns, self.categorical_list)
        self.df = df

    # ------------------------------------------------------------------ meta
    def generate_meta_data(self) -> dict:
        columns = []
        for idx, name in enumerate(self.df.columns):
            entry = {"column_name": name}
            if name in self.categorical_list:
                as_str = self.df[name].astype(str)
                counts = as_str.value_counts()
                entry["type"] = CATEGORICAL
                entry["size"] = int(counts.shape[0])
                entry["i2s"] = {str(k): int(v) for k, v in counts.items()}
            else:
                entry["type"] = FILE_META_CONTINUOUS
                entry["min"] = np.min(self.df[name])
                entry["max"] = np.max(self.df[name])
            entry["column no"] = idx
            columns.append(entry)
        meta = {
            "columns": columns,
            "problem_type": self.problem_type,
            "name": self.output_name,
            "date_info": self.date_columns,
            "integer_info": self.integer_columns,
            "non_negative_cols": self.non_negative_columns,
        }
        if self.target_col:
            meta["target"] = self.target_col
        return meta

    # -------------------------------------------------------------- encoding
    def encode(self, label_encoders) -> np.ndarray:
        """Label-encode categoricals with the global encoders; return the float table."""
        cursor = 0
        for name in self.df.columns:
            if name in self.categorical_list:
                self.df[name] = label_encoders[cursor].transform(self.df[name].astype(str))
                cursor += 1
        return self.df.to_numpy(dtype=np.float64)

    def generate_data(self, meta_file, label_encoder, timestamp):
        """Encode + (optionally) write json / npz / csv like the reference; returns ``train``."""
        self.output_name = f"{self.file_name}_{self.synthesizer_used}-{timestamp}"
        train = self.encode(label_encoder)
        self.output_path = os.path.join(self.output_dir, self.output_name) + "/"
        if self.write_side_files:
            os.makedirs(self.output_path, mode=0o760, exist_ok=True)
            stem = os.path.join(self.output_path, self.output_name)
            with open(stem + ".json", "w") as fh:
                json.dump(meta_file, fh, sort_keys=True, indent=4, separators=(",", ": "), cls=NumpyEncoder)
            np.savez(stem + ".npz", train=train, test=train[len(train):])
            self.df.to_csv(stem + ".csv", index=False)
        return train

    # ------------------------------------------------ rarely used utilities
    def write_data_to_disk(self, meta_data, ratio=1.0):
        self.output_path = os.path.join(self.output_dir, self.output_name) + "/"
        os.makedirs(self.output_path, mode=0o760, exist_ok=True)
        with open(os.path.join(self.output_path, self.output_name + ".json"), "w") as fh:
            json.dump(meta_data, fh, sort_keys=True, indent=4, separators=(",", ": "), cls=NumpyEncoder)

    def generate_sample_csv_data(self, generated_sample, label_encoders=None):
        from .transform import Transform

        if label_encoders is None:
            path = os.path.join(self.output_dir, f"model_{self.output_name}", f"label_encoders_{self.output_name}.pickle")
            with open(path, "rb") as fh:
                label_encoders = pickle.load(fh)
        meta_json = os.path.join(self.output_path, self.output_name + ".json")
        return Transform.inverse(generated_sample, meta_json, label_encoders)

    def save_synthesizer_model_and_label_encoders(self, synthesizer, label_encoders):
        folder = os.path.join(self.output_dir, self.output_name)
        os.makedirs(folder, mode=0o760, exist_ok=True)
        with open(os.path.join(folder, self.output_name + ".pickle"), "wb") as fh:
            pickle.dump(synthesizer, fh, -1)
        with open(os.path.join(folder, f"label_encoders_{self.output_name}.pickle"), "wb") as fh:
            pickle.dump(label_encoders, fh, protocol=pickle.HIGHEST_PROTOCOL)
