"""Host-side native helpers (CSV writer) from the same in-tree shared object."""

from __future__ import annotations

import ctypes as C
import os

import numpy as np

_lib = None


def _load():
    global _lib
    if _lib is None:
        from . import build

        path = build.LIB
        if not os.path.exists(path):
            path = build.build()
        _lib = C.CDLL(path)
        _lib.ftg_csv_write.argtypes = [C.c_char_p, C.c_int, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int),
                                       C.POINTER(C.c_void_p), C.POINTER(C.POINTER(C.c_char_p)), C.c_int]
        _lib.ftg_csv_write.restype = C.c_int
        _lib.ftg_csv_write_strided.argtypes = [C.c_char_p, C.c_int, C.c_int, C.POINTER(C.c_char_p), C.POINTER(C.c_int),
                                               C.POINTER(C.c_void_p), C.POINTER(C.c_int), C.POINTER(C.c_int),
                                               C.POINTER(C.POINTER(C.c_char_p)), C.c_int]
        _lib.ftg_csv_write_strided.restype = C.c_int
    return _lib


def available() -> bool:
    try:
        _load()
        return True
    except Exception:
        return False


def write_csv(path, names, cols, n_threads=None):
    """``cols[i]`` is a float/int ndarray or a ``(codes int32, classes list[str])`` pair."""
    lib = _load()
    n_cols = len(cols)
    n_rows = len(cols[0][0]) if isinstance(cols[0], tuple) else len(cols[0])
    c_names = (C.c_char_p * n_cols)(*[str(n).encode() for n in names])
    kinds = (C.c_int * n_cols)()
    ptrs = (C.c_void_p * n_cols)()
    tables = (C.POINTER(C.c_char_p) * n_cols)()
    keep = []
    for i, col in enumerate(cols):
        if isinstance(col, tuple):
            codes = np.ascontiguousarray(col[0], dtype=np.int32)
            tab = (C.c_char_p * len(col[1]))(*[str(s).encode() for s in col[1]])
            keep += [codes, tab]
            kinds[i], ptrs[i] = 3, codes.ctypes.data
            tables[i] = C.cast(tab, C.POINTER(C.c_char_p))
        else:
            arr = np.asarray(col)
            if arr.dtype == np.float32:
                arr = np.ascontiguousarray(arr)
                kinds[i] = 1
            elif arr.dtype.kind in "iu":
                arr = np.ascontiguousarray(arr, dtype=np.int64)
                kinds[i] = 2
            elif arr.dtype.kind == "f":
                arr = np.ascontiguousarray(arr, dtype=np.float64)
                kinds[i] = 0
            else:
                raise TypeError(f"column {names[i]}: unsupported dtype {arr.dtype}")
            keep.append(arr)
            ptrs[i] = arr.ctypes.data
    nt = n_threads or min(4, os.cpu_count() or 1)   # several CSVs are in flight at once (writer pool)
    rc = lib.ftg_csv_write(str(path).encode(), n_rows, n_cols, c_names, kinds, ptrs, tables, nt)
    if rc != 0:
        raise OSError(f"ftg_csv_write({path}) failed: {rc}")
    return path


class TablePlan:
    """Everything about a table's CSV that does not change from epoch to epoch (names, kinds, class strings), so
    that writing one epoch is a single native call over the generator's row-major float32 table."""

    def __init__(self, names, specs):
        """``specs[j]``: ``("real",)`` float column, or ``("cat", classes)`` float-coded category column."""
        self.n_cols = len(names)
        self.c_names = (C.c_char_p * self.n_cols)(*[str(n).encode() for n in names])
        self.kinds = (C.c_int * self.n_cols)()
        self.n_classes = (C.c_int * self.n_cols)()
        self.tables = (C.POINTER(C.c_char_p) * self.n_cols)()
        self._keep = []
        for j, sp in enumerate(specs):
            if sp[0] == "cat":
                tab = (C.c_char_p * len(sp[1]))(*[str(x).encode() for x in sp[1]])
                self._keep.append(tab)
                self.kinds[j], self.n_classes[j] = 4, len(sp[1])
                self.tables[j] = C.cast(tab, C.POINTER(C.c_char_p))
            else:
                self.kinds[j], self.n_classes[j] = 1, 0

    def write(self, path, table: np.ndarray, n_threads=4, header=True):
        assert table.dtype == np.float32 and table.ndim == 2 and table.shape[1] == self.n_cols
        assert table.strides[1] == 4, "row-major float32 table expected"
        lib = _load()
        ld = table.strides[0] // 4
        ptrs = (C.c_void_p * self.n_cols)(*[table.ctypes.data + 4 * j for j in range(self.n_cols)])
        strides = (C.c_int * self.n_cols)(*([ld] * self.n_cols))
        rc = lib.ftg_csv_write_strided(str(path).encode(), table.shape[0], self.n_cols, self.c_names, self.kinds, ptrs,
                                       strides, self.n_classes, self.tables, int(n_threads) if header else -int(n_threads))
        if rc != 0:
            raise OSError(f"ftg_csv_write_strided({path}) failed: {rc}")
        return path
