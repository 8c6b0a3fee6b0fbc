"""Drop-in for the reference's GP_fit module (same names, signatures, prints, files and return values;
GP_fit.py:20-164), SURVEY 8f row N1.  What changes underneath:

* `len_scale_opt` and `GP` are the B200 path (dropin/find_len_scales.py, dropin/GP_func.py);
* under `torchrun` with at least as many experiments as ranks, whole experiments are the sharded unit: rank r fits
  experiments r, r + G, ... with its swarm kept local (no collective on the data path), the per-experiment tables are
  gathered once, and rank 0 alone writes files.  With fewer experiments than ranks every rank walks all experiments and
  the swarm shards inside `len_scale_opt` as before;
* the chains of pandas outer merges (GP_fit.py:79,103,160) are one key sort (`tables.merge_outer`, identical result);
* files are written by `tables.write_csv` (row N3: libgpk's threaded formatter, the same bytes as `to_csv`).

`read_in.read_yaml` and `convex_hull.fill_convex_hull` are the reference's own modules, found further down sys.path."""
import _bootstrap  # noqa: F401
from pathlib import Path

import numpy as np
import pandas as pd

from find_len_scales import len_scale_opt
from convex_hull import fill_convex_hull
from GP_func import GP
from read_in import read_yaml

from gaussian_process_b200 import sharding
from gaussian_process_b200.tables import merge_outer, write_csv


def process_experiment(x_known, y_known, e_known, resolution, dim_labels, filename, exp_idx, total_exps, PSO_progress):
    """GP fit of one experiment -> DataFrame of the lattice with the fitted value and uncertainty (GP_fit.py:20-46)."""
    if x_known.shape[1] == 0:
        print(f"  Skipping experiment: no valid data points")
        return None
    len_scale = len_scale_opt(x_known, y_known, e_known, PSO_progress)
    x_fit = fill_convex_hull(x_known.T, resolution)
    y_fit, e_fit = GP(x_known, y_known, e_known, x_fit.T, len_scale)
    df = pd.DataFrame(x_fit, columns=dim_labels)
    if total_exps == 1:
        quant_col, err_col = f"{filename}", f"{filename}_unc"
    else:
        quant_col, err_col = f"{filename}_exp{exp_idx}", f"{filename}_unc{exp_idx}"
    df[quant_col] = y_fit.flatten()
    df[err_col] = e_fit.flatten()
    return df


def _output_folder(out_path):
    folder = out_path if out_path.is_dir() else out_path.parent
    folder.mkdir(parents=True, exist_ok=True)
    return folder


def write_individual_file(df, filename, out_path, total_exps, exp_idx):
    """One experiment -> <file>[_exp<i>]_GP_results.txt (GP_fit.py:50-66)."""
    folder = _output_folder(out_path)
    if total_exps == 1:
        output_path = folder / f"{filename}_GP_results.txt"
    else:
        output_path = folder / f"{filename}_exp{exp_idx}_GP_results.txt"
    write_csv(df, output_path)
    print(f"Written individual output file: {output_path}")
    return


def write_grouped_file(file_dfs, filename, out_path, dim_labels):
    """All experiments of one input file merged -> <file>_GP_results.txt (GP_fit.py:70-85)."""
    folder = _output_folder(out_path)
    merged_df = merge_outer(file_dfs, dim_labels)
    output_path = folder / f"{filename}_GP_results.txt"
    write_csv(merged_df, output_path)
    print(f"Written grouped output file: {output_path}")
    return


def write_combined_file(experiment_dfs, out_path, dim_labels):
    """Every experiment merged into one file (GP_fit.py:89-108)."""
    if str(out_path).endswith("/"):
        output_folder = out_path
        output_folder.mkdir(parents=True, exist_ok=True)
        combined_output_file = output_folder / "GP_results.txt"
    else:
        combined_output_file = out_path
        combined_output_file.parent.mkdir(parents=True, exist_ok=True)
    merged_df = merge_outer(experiment_dfs, dim_labels)
    write_csv(merged_df, combined_output_file)
    print(f"Combined results written to {combined_output_file}")
    return merged_df


def create_GP():
    """All experiments of options.yaml: fit, write the configured files, return (merged table, number of dimensions)
    (GP_fit.py:112-164)."""
    resolution, PSO_progress, out_file_name, labels, data_list, write_ind, group_exps = read_yaml()

    num_dims = len(resolution)
    dim_labels = labels[:num_dims]
    out_path = Path(out_file_name) if out_file_name else Path("GP_results.txt")
    rank, size = sharding.dist_world()
    writer = rank == 0

    if writer:
        if write_ind:
            print("Writing individual output files")
            if group_exps:
                print("Writing experiments from the same file together")
            else:
                print("Writing experiments from the same file separately")
        else:
            print("Writing combined output file")

    jobs = []
    for file_idx, (file_path, x_known_list, exp_pairs, labels_out) in enumerate(data_list, start=1):
        filename = Path(file_path).stem
        for idx, (x_known, (y_known, e_known)) in enumerate(zip(x_known_list, exp_pairs), start=1):
            jobs.append((file_idx, filename, idx, len(x_known_list), x_known, y_known, e_known))

    def run(job):
        file_idx, filename, idx, total, x_known, y_known, e_known = job
        print(f"Processing experiment {idx}/{total} from file {file_idx}/{len(data_list)}: {filename}")
        return process_experiment(x_known, y_known, e_known, resolution, dim_labels, filename, idx, total, PSO_progress)

    by_experiment = size > 1 and len(jobs) >= size
    if by_experiment:
        # a failing experiment (LinAlgError from GP_func.py:22, a library error) must not leave the other ranks waiting in
        # the gather: the error travels with the results and is raised on every rank, as the reference aborts the run
        mine, failure = {}, None
        with sharding.local_swarm():
            for j in range(rank, len(jobs), size):
                try:
                    mine[j] = run(jobs[j])
                except Exception as ex:              # noqa: BLE001 -- re-raised below, on every rank
                    failure = (j, type(ex).__name__, str(ex), ex if isinstance(ex, np.linalg.LinAlgError) else None)
                    break
        results = {}
        first = None
        for part, fail in sharding.gather_objects((mine, failure)):
            results.update(part)
            if fail is not None and (first is None or fail[0] < first[0]):
                first = fail
        if first is not None:
            j, kind, text, ex = first
            if kind == "LinAlgError":
                raise np.linalg.LinAlgError(text)
            raise RuntimeError(f"experiment job {j} failed with {kind}: {text}")
        dfs = [results[j] for j in range(len(jobs))]
    else:
        dfs = None                    # fitted inside the loop below, so that stdout reads exactly like the reference's

    # the reference's writing order, on one rank
    experiment_dfs = []
    pos = 0
    for file_idx, (file_path, x_known_list, exp_pairs, labels_out) in enumerate(data_list, start=1):
        filename = Path(file_path).stem
        file_dfs = []
        for idx in range(1, len(x_known_list) + 1):
            df = dfs[pos] if dfs is not None else run(jobs[pos])
            pos += 1
            if df is None:
                continue
            experiment_dfs.append(df)
            file_dfs.append(df)
            if writer and write_ind and not group_exps:
                write_individual_file(df, filename, out_path, len(x_known_list), idx)
        if writer and write_ind and group_exps and file_dfs:
            write_grouped_file(file_dfs, filename, out_path, dim_labels)

    if not write_ind and experiment_dfs:
        if writer:
            return write_combined_file(experiment_dfs, out_path, dim_labels), num_dims
        return merge_outer(experiment_dfs, dim_labels), num_dims

    if experiment_dfs:
        return merge_outer(experiment_dfs, dim_labels), num_dims
    print("No experiment data to return")
    return pd.DataFrame(), num_dims
