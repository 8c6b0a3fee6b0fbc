"""Probability-surface step: host side of calc_prob_surf.py on top of the fused GPU reduction.

The row loop of generate_prob_surf (calc_prob_surf.py:67-81: finite mask, linspace y-axis,
sum_gaussians) is one `gpk_prob_surface` launch over the whole merged table; the frame assembly (:83) stays in
pandas (surface_frame) for callers that want the table; generate_prob_surf itself streams pass by pass into libgpk's
threaded formatter (the frozen text format of `to_csv`, :84) without ever holding the long-format table."""
import os

import numpy as np
import pandas as pd
import yaml

from . import engine, tables

POINTS = 100                      # calc_prob_surf.py:45
DEFAULT_FILENAME = "prob_surf.txt"


def sum_gaussians(temp_y, temp_gaus):
    """Average Gaussian bin probability at the points temp_y for interleaved (mean, std) pairs
    (calc_prob_surf.py:15-30), evaluated by the surface kernel's inner routine."""
    temp_y = np.asarray(temp_y, dtype=np.float64)
    dy = abs(max(temp_y) - min(temp_y)) / len(temp_y)                 # :23
    return engine.sum_gaussians(temp_y, temp_gaus, dy / 2)


def resolve_output_path(options_path):
    """calc_prob_surf.py:50-65."""
    output_file = DEFAULT_FILENAME
    if os.path.exists(options_path):
        with open(options_path, 'r') as f:
            options = yaml.safe_load(f) or {}
        output_path = options.get("out_file_name", "").strip()
        if output_path:
            output_file = os.path.join(output_path, DEFAULT_FILENAME) if os.path.isdir(output_path) else output_path
    return output_file


def surface_frame(df, ndims, points=POINTS):
    """Merged GP-result frame -> long-format frame (dims..., quantity, prob)."""
    values = df.to_numpy(dtype=np.float64)
    coords, payload = values[:, :ndims], values[:, ndims:]
    idx, y, p = engine.prob_surface(payload, points)
    out = np.empty((idx.shape[0] * points, ndims + 2))
    out[:, :ndims] = np.repeat(coords[idx], points, axis=0)
    out[:, ndims] = y.reshape(-1)
    out[:, ndims + 1] = p.reshape(-1)
    return pd.DataFrame(out, columns=df.columns[:ndims].tolist() + ["quantity", "prob"])


STREAM_ROWS = 1 << 15             # merged rows per CSV block of the streaming writer (3.3 M output rows, ~160 MB of float64)


def write_surface_csv(df, ndims, output_file, points=POINTS, rows_per_pass=None, passes_fn=None, row_range=None,
                      header=True):
    """The file generate_prob_surf writes (:80-84), produced pass by pass: each pass of the GPU reduction is formatted as
    its long-format rows (coords repeated `points` times, quantity, prob) by libgpk's threaded CSV formatter (the
    coordinate text once per merged row) and appended, while the next pass runs on the device.  Host memory stays bounded by the pass size, so the
    64^4-row configuration (1.7e9 output rows) never exists as one frame.  Returns the number of output rows.
    `row_range` = (lo, hi) restricts the file to those merged rows and `header=False` drops the header line (the parts
    of write_surface_csv_sharded); `passes_fn` replaces engine.prob_surface_passes (tests)."""
    columns = df.columns[:ndims].tolist() + ["quantity", "prob"]
    header_line = tables.plain_header(columns)
    if header_line is None:                                # column names that need quoting: let pandas write the file
        if row_range is not None:
            raise ValueError("column names that need CSV quoting are not supported by the sharded writer")
        out = surface_frame(df, ndims, points)
        out.to_csv(output_file, index=False)
        return len(out)
    values = df.to_numpy(dtype=np.float64)
    if row_range is not None:
        values = values[row_range[0]:row_range[1]]
    coords, payload = values[:, :ndims], values[:, ndims:]
    passes_fn = passes_fn or engine.prob_surface_passes
    n_rows = 0
    with open(output_file, "wb") as f:
        if header:
            f.write(header_line.encode("utf-8"))
        for idx, y, p in passes_fn(payload, points, rows_per_pass):
            for b0 in range(0, idx.shape[0], STREAM_ROWS):
                sel = idx[b0:b0 + STREAM_ROWS]
                k = sel.shape[0]
                f.write(tables.format_surface_block(coords[sel], y[b0:b0 + k], p[b0:b0 + k]))
                n_rows += k * points
    return n_rows


def write_surface_csv_sharded(df, ndims, output_file, points=POINTS, rows_per_pass=None, passes_fn=None):
    """The same file written by all ranks of the process group (surface rows): rank r reduces and formats the
    contiguous block shard_slice(len(df), r, size) of the merged rows into `<output_file>.part<r>` (rank 0's part carries
    the header), then rank 0 concatenates the parts in rank order: the bytes are those of the single-process writer,
    because every output row depends on its own merged row only.  The per-rank output row counts are exchanged
    (sharding.surface_offsets) and the total is returned on every rank."""
    import shutil
    from . import sharding
    rank, size = sharding.dist_world()
    if size == 1:
        return write_surface_csv(df, ndims, output_file, points, rows_per_pass, passes_fn)
    import torch.distributed as dist
    lo, hi = sharding.shard_slice(len(df), rank, size)
    part = f"{output_file}.part{rank}"
    n_local = write_surface_csv(df, ndims, part, points, rows_per_pass, passes_fn, row_range=(lo, hi), header=(rank == 0))
    _offset, total = sharding.surface_offsets(n_local)
    dist.barrier()
    if rank == 0:
        with open(output_file, "wb") as out:
            for r in range(size):
                with open(f"{output_file}.part{r}", "rb") as src:
                    shutil.copyfileobj(src, out, 1 << 24)
                os.remove(f"{output_file}.part{r}")
    dist.barrier()
    return total


def generate_prob_surf(df, ndims, options_path="options.yaml"):
    """Drop-in for calc_prob_surf.generate_prob_surf (:39-88): same prints, same CSV.  Under an initialised process
    group the merged rows shard across the ranks (every rank calls this with the same table, as main.py does)."""
    from . import sharding
    rank, size = sharding.dist_world()
    if rank == 0:
        print("Calculating Probability")
    output_file = resolve_output_path(options_path)
    if size > 1:
        write_surface_csv_sharded(df, ndims, output_file)
    else:
        write_surface_csv(df, ndims, output_file)          # the bytes of to_csv(index=False), :84
    if rank == 0:
        print(f"Probability results written to {output_file}")
    return
