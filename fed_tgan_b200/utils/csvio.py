"""CSV output of the synthetic table.

``DataFrame.to_csv`` of 40 000 x 42 sits inside the reference's timed round.
The fast path formats columns with the native multi-threaded writer
(``csrc/csv_writer.cpp``: shortest round-trip ``std::to_chars``) when the library is
available, else falls back to pandas.  Output is ``pandas.read_csv``-compatible and keeps
the header/column order of ``Transform.inverse``.
"""

from __future__ import annotations

import numpy as np
import pandas as pd


def _native():
    try:
        from ..ops import hostlib

        return hostlib
    except Exception:
        return None


def write_table_csv(path, names, cols, n_threads=None):
    lib = _native()
    if lib is not None and lib.available():
        try:
            lib.write_csv(path, names, cols, n_threads=n_threads)
            return path
        except Exception:
            pass
    plain = [np.asarray(c[1], dtype=object)[c[0]] if isinstance(c, tuple) else c for c in cols]
    pd.DataFrame({n: c for n, c in zip(names, plain)}, columns=list(names)).to_csv(path, index=False)
    return path
