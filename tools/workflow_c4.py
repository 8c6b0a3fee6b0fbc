"""End-to-end timing of the drop-in workflow on BASELINE config C4: synthetic 2-D file, N=8192 rows, 8 experiments per
file with inf-masked gaps (each row finite in each experiment with probability 0.5), GP fit of every experiment on
its convex-hull lattice, outer merge, probability surface.  Everything above the kernels is the drop-in modules the
reference's main.py would import (GP_fit, find_len_scales, convex_hull, GP_func, calc_prob_surf); read_in is stubbed
with the synthetic file.  Under torchrun the experiments shard across the ranks (GP_fit.create_GP).
    python tools/workflow_c4.py [rows] [experiments] [resolution]"""
import contextlib
import importlib
import io
import json
import os
import sys
import tempfile
import time
import types

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "gaussian-process_b200", "dropin"))
import numpy as np

rows = int(sys.argv[1]) if len(sys.argv) > 1 else 8192
n_exp = int(sys.argv[2]) if len(sys.argv) > 2 else 8
res = float(sys.argv[3]) if len(sys.argv) > 3 else 0.005

rng = np.random.default_rng(4)
x = rng.random((2, rows))
x[:, :4] = np.array([[0, 0, 1, 1], [0, 1, 0, 1]], dtype=np.float64)          # box corners: the hull is the whole box
x_list, pairs = [], []
for k in range(n_exp):
    keep = rng.random(rows) < 0.5
    keep[:4] = True
    xk = np.ascontiguousarray(x[:, keep])
    yk = np.sin(2 * np.pi * xk[0]) + np.sin(2 * np.pi * xk[1]) + 0.3 * k + 0.1 * rng.standard_normal(xk.shape[1])
    ek = rng.uniform(0.05, 0.15, xk.shape[1])
    x_list.append(xk)
    pairs.append((yk, ek))
labels = ["a", "b", "q", "err"]
data_list = [("data/c4.txt", x_list, pairs, labels)]

read_in = types.ModuleType("read_in")
read_in.read_yaml = lambda: ([res, res], False, "results/GP_results.txt", labels, data_list, False, False)
sys.modules["read_in"] = read_in

import torch
import torch.distributed as dist
world = int(os.environ.get("WORLD_SIZE", "1"))
rank = int(os.environ.get("RANK", "0"))
if world > 1:
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0"))))

gp_fit = importlib.import_module("GP_fit")
cps = importlib.import_module("calc_prob_surf")
phase = {}


def timed(name, fn):
    def wrapper(*a, **k):
        t0 = time.perf_counter()
        out = fn(*a, **k)
        torch.cuda.synchronize()
        phase[name] = phase.get(name, 0.0) + time.perf_counter() - t0
        return out
    return wrapper


def seeded_opt(inner):
    def f(xk, yk, ek, progress):
        np.random.seed(0)
        return inner(xk, yk, ek, progress, lhs_random_state=0)
    return f


gp_fit.len_scale_opt = timed("len_scale_opt (KMeans subsample + PSO)", seeded_opt(gp_fit.len_scale_opt))
gp_fit.fill_convex_hull = timed("fill_convex_hull (lattice)", gp_fit.fill_convex_hull)
gp_fit.GP = timed("GP fit + predict", gp_fit.GP)
gp_fit.merge_outer = timed("outer merge", gp_fit.merge_outer)
gp_fit.write_csv = timed("CSV of the merged table", gp_fit.write_csv)

holder = tempfile.TemporaryDirectory() if rank == 0 else None
shared = [holder.name if rank == 0 else None]
if world > 1:
    dist.broadcast_object_list(shared, src=0)            # one working directory for all ranks (sharded CSV parts)
if True:
    tmp = shared[0]
    os.chdir(tmp)
    # warm-up on a small problem (module load, workspaces, pinned buffers)
    small = [("data/w.txt", [x_list[0][:, :300]], [(pairs[0][0][:300], pairs[0][1][:300])], labels)]
    gp_fit.read_yaml = lambda: ([0.05, 0.05], False, "results/w.txt", labels, small, False, False)
    with contextlib.redirect_stdout(io.StringIO()):
        gp_fit.create_GP()
    phase.clear()
    gp_fit.read_yaml = lambda: ([res, res], False, "results/GP_results.txt", labels, data_list, False, False)
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()):
        df, nd = gp_fit.create_GP()
    t_fit = time.perf_counter() - t0
    if rank == 0:
        open("options.yaml", "w").write('out_file_name: "results/prob_surf.txt"\n')
    if world > 1:
        dist.barrier()
    t1 = time.perf_counter()
    with contextlib.redirect_stdout(io.StringIO()):
        cps.generate_prob_surf(df, nd)                   # all ranks: the merged rows shard across them
    t_surf = time.perf_counter() - t1
    surf_bytes = os.path.getsize("results/prob_surf.txt") if rank == 0 else 0
    if world > 1:
        dist.barrier()
    os.chdir(ROOT)
if rank == 0:
    print(json.dumps({
        "workload": f"C4 workflow: {rows} rows, d=2, {n_exp} experiments (~{int(np.mean([v.shape[1] for v in x_list]))} points each), "
                    f"resolution {res} ({len(df)} lattice rows)",
        "n_gpus": world, "create_GP_seconds": t_fit, "generate_prob_surf_seconds": t_surf, "prob_surf_csv_bytes": surf_bytes,
        "phases_rank0_seconds": {k: round(v, 3) for k, v in phase.items()},
        "merged_table_shape": list(df.shape)}))
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
if holder is not None:
    holder.cleanup()
