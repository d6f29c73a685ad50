"""CSV -> meta / encoded table entry points.

Parity: ``Server/dtds/data/load.py:1-91`` (``prepare_data``,
``encode_data_with_meta_labelencoder``, ``load_datapath``) and the client copy's
``load_dataset`` (``Client/.../dtds/data/load.py:93-105``).
"""

from __future__ import annotations

import json
import os
from glob import glob

import numpy as np
import pandas as pd

from .constants import CATEGORICAL, ORDINAL
from .file_generator import FileGenerator

SYNTHESIZER_NAME = "CTGANSynthesizer"


def _stem(path: str) -> str:
    return os.path.basename(path).split(".")[0]


def _columns_by_kind(meta: dict):
    cat, ordi = [], []
    for i, col in enumerate(meta["columns"]):
        if col["type"] == CATEGORICAL:
            cat.append(i)
        elif col["type"] == ORDINAL:
            ordi.append(i)
    return cat, ordi


def _make_generator(filename, selected_variables, categorical_list, nonnegative_list, date_dic,
                    target_column, problem_type, dataframe=None, **kw):
    df = pd.read_csv(filename) if dataframe is None else dataframe
    target = "" if target_column in (None, "none") else target_column
    return FileGenerator(
        dataframe=df[list(selected_variables)],
        file_name=_stem(filename),
        problem_type=problem_type or "",
        target_col=target,
        categorical_list=list(categorical_list),
        non_negative_columns=list(nonnegative_list),
        date_columns=dict(date_dic or {}),
        synthesizer_used=SYNTHESIZER_NAME,
        **kw,
    )


def prepare_data(filename, selected_variables, categorical_list, nonnegative_list, date_dic,
                 target_column, problem_type, dataframe=None):
    """Local file-level meta incl. per-category counts (what ``get_meta`` returns)."""
    return _make_generator(filename, selected_variables, categorical_list, nonnegative_list,
                           date_dic, target_column, problem_type, dataframe).generate_meta_data()


def encode_data_with_meta_labelencoder(filename, meta_file, label_encoder, selected_variables,
                                       categorical_list, nonnegative_list, date_dic, target_column,
                                       problem_type, timestamp, dataframe=None, write_side_files=True):
    """Encode the client's CSV with the *global* label encoders.

    Returns ``(train, categorical_column_indices, ordinal_column_indices)`` like the
    reference, which gets there via an npz round trip (``load.py:72-91``).
    """
    gen = _make_generator(filename, selected_variables, categorical_list, nonnegative_list, date_dic,
                          target_column, problem_type, dataframe, write_side_files=write_side_files)
    train = gen.generate_data(meta_file, label_encoder, timestamp)
    cat, ordi = _columns_by_kind(meta_file)
    return train, cat, ordi


def load_datapath(path, benchmark=False):
    """Read back a ``<dir>/*.npz`` + ``*.json`` pair written by ``generate_data``."""
    data = np.load(glob(f"{path}/*.npz")[0], allow_pickle=True)
    with open(glob(f"{path}/*.json")[0]) as fh:
        meta = json.load(fh)
    cat, ordi = _columns_by_kind(meta)
    if benchmark:
        return data["train"], data["test"], meta, cat, ordi
    return data["train"], cat, ordi


def load_dataset(name, benchmark=False, root="data/processed"):
    """Load a pre-processed dataset directory ``<root>/<name>*`` (client-copy helper)."""
    hits = sorted(glob(os.path.join(root, f"{name}*")))
    if not hits:
        raise FileNotFoundError(f"no processed dataset matching {name!r} under {root}")
    return load_datapath(hits[0], benchmark=benchmark)
