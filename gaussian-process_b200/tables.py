"""Result tables of the GP workflow: the outer merge of per-experiment lattices on their
kinematic columns that GP_fit.py:79,103,160 and main.py:48 write as

    reduce(lambda l, r: pd.merge(l, r, on=dim_labels, how="outer"), dfs).fillna(float("inf"))

A chain of hash joins re-hashes the growing left table once per experiment.  When every table has unique keys (the
lattice of fill_convex_hull is de-duplicated, convex_hull.py:34) the result is simply: union of the keys in
lexicographic order, each experiment's columns scattered into place, +inf elsewhere -- one sort of all keys.
`merge_outer` does that and returns exactly the DataFrame the pandas expression returns (same rows, row order, column
order, dtypes, index); anything it does not cover (duplicate keys, non-float columns, NaN keys, overlapping value column
names) goes through the pandas expression itself.

Row N3: `write_csv(df, path)` writes what `df.to_csv(path, index=False)` writes (GP_fit.py:63,81,104;
calc_prob_surf.py:84), byte for byte, with the float formatting done by libgpk's threaded formatter
(`gpk_csv_format_f64`) instead of pandas' per-value Python objects."""
from functools import reduce

import numpy as np
import pandas as pd


def _pandas_merge(dfs, dim_labels):
    return reduce(lambda left, right: pd.merge(left, right, on=dim_labels, how="outer"), dfs).fillna(float("inf"))


def _eligible(dfs, dim_labels):
    seen = set(dim_labels)
    for df in dfs:
        if not isinstance(df.index, pd.RangeIndex) or df.columns.has_duplicates:
            return False
        if any(c not in df.columns for c in dim_labels):
            return False
        if any(df[c].dtype != np.float64 for c in df.columns):
            return False
        for c in df.columns:
            if c in dim_labels:
                continue
            if c in seen:             # pandas would add _x/_y suffixes
                return False
            seen.add(c)
    return True


def _group_ids(keys):
    """keys (n, nd) float64 -> (gid (n,) int64 = rank of each row's key in lexicographic order of the distinct keys,
    first (n_out,) = lowest row index holding each distinct key).  One radix-style pass per column on the GPU when
    there is one (torch.unique on a few million keys is milliseconds there), else on the host."""
    import torch
    dev = torch.device("cuda") if torch.cuda.is_available() else torch.device("cpu")
    k = torch.from_numpy(np.ascontiguousarray(keys)).to(dev)
    code = torch.zeros(k.shape[0], dtype=torch.int64, device=dev)
    span = 1
    for j in range(k.shape[1]):
        vals, inv = torch.unique(k[:, j], sorted=True, return_inverse=True)      # -0.0 == 0.0: one value
        span *= int(vals.numel())
        if span >= 2 ** 62:
            return None
        code = code * int(vals.numel()) + inv
    _, gid = torch.unique(code, sorted=True, return_inverse=True)
    n_out = int(gid.max().item()) + 1 if gid.numel() else 0
    first = torch.full((n_out,), k.shape[0], dtype=torch.int64, device=dev)
    first.scatter_reduce_(0, gid, torch.arange(k.shape[0], device=dev), reduce="amin")
    return gid.cpu().numpy(), first.cpu().numpy()


def merge_outer(dfs, dim_labels):
    """Outer merge of `dfs` on `dim_labels` with missing values set to +inf (see module docstring)."""
    dfs = list(dfs)
    dim_labels = list(dim_labels)
    if len(dfs) == 0:
        raise TypeError("reduce() of empty iterable with no initial value")
    if len(dfs) == 1:
        return dfs[0].fillna(float("inf"))
    if not _eligible(dfs, dim_labels):
        return _pandas_merge(dfs, dim_labels)
    nd = len(dim_labels)
    keys = np.concatenate([df[dim_labels].to_numpy(dtype=np.float64).reshape(len(df), nd) for df in dfs])
    if len(keys) == 0 or np.isnan(keys).any():
        return _pandas_merge(dfs, dim_labels)
    ids = _group_ids(keys)
    if ids is None:
        return _pandas_merge(dfs, dim_labels)
    gid, first = ids
    n_out = len(first)
    # unique keys inside every table?  (otherwise a join multiplies rows)
    off = 0
    for df in dfs:
        g = gid[off:off + len(df)]
        if len(g) and np.bincount(g, minlength=n_out).max() > 1:
            return _pandas_merge(dfs, dim_labels)
        off += len(df)
    out = {}
    for j, c in enumerate(dim_labels):
        out[c] = keys[first, j]                          # the key as the first table holding it spells it (-0.0 / 0.0)
    off = 0
    for df in dfs:
        g = gid[off:off + len(df)]
        for c in df.columns:
            if c in dim_labels:
                continue
            col = np.full(n_out, np.inf)
            v = df[c].to_numpy(dtype=np.float64)
            col[g] = np.where(np.isnan(v), np.inf, v)
            out[c] = col
        off += len(df)
    return pd.DataFrame(out)


_CSV_BLOCK_ROWS = 1 << 20


def _plain_header(columns):
    """Header line if no column name needs CSV quoting, else None."""
    names = []
    for c in columns:
        if not isinstance(c, str) or c == "" or any(ch in c for ch in ',"\r\n'):
            return None
        names.append(c)
    return ",".join(names) + "\n"


def format_block(block, threads=0):
    """float64 (rows, cols) C-contiguous block -> the bytes of its to_csv(index=False, header=False) text
    (libgpk's threaded formatter, gpk_csv_format_f64)."""
    import ctypes as C
    from . import _lib
    lib = _lib.load()
    block = np.ascontiguousarray(block, dtype=np.float64)
    rows, ncol = block.shape
    cap = lib.gpk_csv_bound(rows, ncol)
    out = np.empty(cap, dtype=np.uint8)
    n = lib.gpk_csv_format_f64(block.ctypes.data_as(C.c_void_p), rows, ncol, ncol,
                               out.ctypes.data_as(C.c_void_p), cap, threads)
    if n < 0:
        raise _lib.GpkError("gpk_csv_format_f64 rejected its arguments")
    return memoryview(out)[:n]


def format_surface_block(coords, y, p, threads=0):
    """Bytes of the long-format surface rows of `coords` (k, ndims), `y`, `p` (k, points): what format_block returns for
    the expanded (k * points, ndims + 2) table, with the coordinate text formatted once per merged row
    (gpk_csv_format_surface)."""
    import ctypes as C
    from . import _lib
    lib = _lib.load()
    coords = np.ascontiguousarray(coords, dtype=np.float64)
    y = np.ascontiguousarray(y, dtype=np.float64)
    p = np.ascontiguousarray(p, dtype=np.float64)
    k, points = y.shape
    ndims = coords.shape[1]
    cap = lib.gpk_csv_surface_bound(k, ndims, points)
    out = np.empty(cap, dtype=np.uint8)
    n = lib.gpk_csv_format_surface(coords.ctypes.data_as(C.c_void_p), y.ctypes.data_as(C.c_void_p),
                                   p.ctypes.data_as(C.c_void_p), k, ndims, points, out.ctypes.data_as(C.c_void_p), cap, threads)
    if n < 0:
        raise _lib.GpkError("gpk_csv_format_surface rejected its arguments")
    return memoryview(out)[:n]


def plain_header(columns):
    """Header line of to_csv for column names that need no quoting, else None."""
    return _plain_header(columns)


def write_csv(df, path, threads=0):
    """df.to_csv(path, index=False) for an all-float64 table; anything else is handed to pandas."""
    header = _plain_header(df.columns)
    if header is None or len(df.columns) == 0 or any(dt != np.float64 for dt in df.dtypes):
        df.to_csv(path, index=False)
        return
    with open(path, "wb") as f:
        f.write(header.encode("utf-8"))
        for r0 in range(0, len(df), _CSV_BLOCK_ROWS):
            f.write(format_block(df.iloc[r0:r0 + _CSV_BLOCK_ROWS].to_numpy(dtype=np.float64), threads))
