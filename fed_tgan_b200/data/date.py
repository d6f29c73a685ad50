"""Date columns <-> categorical parts.

Behavioural parity with ``Server/dtds/data/utils/date.py:6-200``: a date column described by
a spec ``"<orig>|<parts>"`` (e.g. ``"yymmdd|YYYY-MM-DD"``) or just ``"<parts>"`` is exploded
into categorical columns ``<col>-year/-month/-day/-hour/-minute/-second`` and re-joined on
decode, with impossible days clamped.  Dormant unless ``date_dic`` is non-empty.
"""

from __future__ import annotations

import pandas as pd

from .constants import EMPTY_TOKEN

_PART_SUFFIX = {"YYYY": "-year", "MM": "-month", "DD": "-day", "hh": "-hour", "mm": "-minute", "ss": "-second"}
_PART_STRFTIME = {"YYYY": "%y", "MM": "%m", "DD": "%d", "hh": "%H", "mm": "%M", "ss": "%S"}
_DAYS_IN_MONTH = {1: 31, 2: 28, 3: 31, 4: 30, 5: 31, 6: 30, 7: 31, 8: 31, 9: 30, 10: 31, 11: 30, 12: 31}


def _parse_spec(spec):
    pieces = spec.split("|")
    if len(pieces) == 2:
        return pieces[0], pieces[1]
    return None, pieces[0]


class Date:
    @staticmethod
    def convert(df: pd.DataFrame, date_columns: dict, categorical_list: list):
        """Turn each date column into its categorical parts (in place on ``categorical_list``)."""
        for name, spec in date_columns.items():
            if name in categorical_list:
                categorical_list.remove(name)
            orig_fmt, parts_fmt = _parse_spec(spec)
            col = df[name]
            if orig_fmt == "yymmdd":
                parsed = col.map(lambda v: pd.to_datetime(str(int(v))) if v != EMPTY_TOKEN else EMPTY_TOKEN)
            else:
                parsed = col.astype(str).map(lambda v: pd.to_datetime(v) if v != EMPTY_TOKEN else EMPTY_TOKEN)
            df = df.copy()
            df[name] = parsed
            present = [i for i, v in enumerate(parsed) if not (isinstance(v, str) and v == EMPTY_TOKEN)]
            df = Date.split(df, name, parts_fmt, categorical_list, present)
        return df

    @staticmethod
    def split(df: pd.DataFrame, column_name: str, format_: str, categorical_list: list, non_empty_list: list):
        rows = df.index[non_empty_list]
        stamps = pd.to_datetime(df.loc[rows, column_name])
        for part in format_.split("-"):
            part_key = part if part in _PART_SUFFIX else "ss"
            out = column_name + _PART_SUFFIX[part_key]
            df[out] = EMPTY_TOKEN
            df.loc[rows, out] = stamps.dt.strftime(_PART_STRFTIME[part_key]).values
            categorical_list.append(out)
        return df.drop(columns=[column_name])

    @staticmethod
    def _fix_day(text: str) -> str:
        if EMPTY_TOKEN in text:
            return EMPTY_TOKEN
        bits = text.split("-")
        if len(bits) >= 3:
            month = int(bits[1])
            if int(bits[2]) > _DAYS_IN_MONTH[month]:
                if month == 2:
                    year = int(bits[0])
                    # the reference's (deliberately kept) leap rule: only years divisible by 400
                    bits[2] = "29" if (year % 4 == 0 and year % 100 == 0 and year % 400 == 0) else "28"
                else:
                    bits[2] = "30"
        return "-".join(bits)

    @staticmethod
    def inverse_transform(df: pd.DataFrame, date_columns: dict):
        part_cols = []
        for name, spec in date_columns.items():
            orig_fmt, parts_fmt = _parse_spec(spec)
            part_cols = [name + _PART_SUFFIX[p] for p in parts_fmt.split("-")]
            joined = df[part_cols].astype(str).agg("-".join, axis=1).map(Date._fix_day)
            stamps = joined.map(lambda v: pd.to_datetime(v) if v != EMPTY_TOKEN else EMPTY_TOKEN)
            if orig_fmt == "yymmdd":
                stamps = stamps.map(
                    lambda v: int(v.strftime("%y%m%d")) if not isinstance(v, str) else EMPTY_TOKEN
                )
            df[name] = stamps
            df = df.drop(columns=part_cols)
        return df
