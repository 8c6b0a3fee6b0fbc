#!/usr/bin/env python
"""Benchmark of the GP-fitting hot path: PSO loss evaluations / second (batched GP fits) at N=4096.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  torchrun --nproc-per-node N bench.py --gpus N ...          (one rank per GPU, NCCL)

Workload (BASELINE.json configs[2], SURVEY 8d "C3"): synthetic 3-D problem with N=4096 training points,
a swarm of 256 particles PER GPU (weak scaling: particles shard with no data-path collective; only the
per-particle losses are all-gathered).  A step = one swarm evaluation = 256 loss evaluations per GPU,
each a full GP fit (covariance build, blocked FP64 Cholesky, triangular inverse, self-prediction,
count reduction) -- no KMeans subsample, nothing cached between steps.

One JSON line is printed by rank 0; see the task contract for the keys.  `--impl reference` times the
CPU implementation of the same path (the NumPy oracle port: the reference itself is NumPy) on the
host cores."""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

WORKLOAD = dict(n_points=4096, dims=3, particles_per_gpu=256, n_sigma=1000)
INT8_PEAK_TOPS = 4497.0
# the same instruction stream with uniformly random signed digits as operands and no memory traffic at all: the board
# reaches its 1000 W cap and holds 3780 TOP/s (tools/umma_power.cu, profiles/r01_umma_power.log)
INT8_SUSTAINED_RANDOM_TOPS = 3780.0
INT8_PEAK_SOURCE = ("int8 tcgen05.mma issue-rate probe on this pool's B200 (tools/umma_probe.cu, 148 CTAs, N=256: 4497 TOP/s, "
                    "profiles/r01_umma_probe.log; nominal 4500).  MEASURED_PEAKS.json has only bf16 (1645 burst / 1379 "
                    "sustained TFLOP/s, cuBLAS); twice that, 3291 / 2757, is the library-GEMM-equivalent int8 figure")
METRIC = "pso_loss_evals_per_sec_n4096"
UNIT = "loss_evals/s"


def algorithmic_flops_per_eval(n):
    """SURVEY 8(d), reduced form: Cholesky N^3/3 + triangular inverse N^3/3."""
    return 2.0 / 3.0 * n ** 3


def make_problem(seed=0):
    from gaussian_process_b200 import synthetic            # the measured arm never imports oracle/
    return synthetic.problem(WORKLOAD["n_points"], WORKLOAD["dims"], seed)


def make_particles(x, p, seed):
    from gaussian_process_b200 import synthetic
    return synthetic.particles(x, p, seed)


# --------------------------------------------------------------------------------------- clocks
class ClockSampler:
    QUERY = "clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.QUERY}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, smax, power, reasons = [], [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            f = [v.strip() for v in r.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])); smax.append(float(f[1])); power.append(float(f[2]))
            except ValueError:
                continue
            for nme, v in zip(names, f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(nme)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}


# --------------------------------------------------------------------------------------- CPU side
def _cpu_eval(args):
    """One reference-algorithm loss evaluation at the bench workload (runs in a worker process)."""
    seed, k = args
    from oracle import gp_oracle as orc
    x, y, e = orc.synthetic_problem(WORKLOAD["n_points"], WORKLOAD["dims"], seed)
    pos, lo, hi = orc.synthetic_particles(x, WORKLOAD["particles_per_gpu"], seed)
    s, ex = orc.sigma_grid()
    t0 = time.perf_counter()
    v = orc.evaluate_loss(pos[k], x, y, e, s, ex, lo, hi)
    return v, time.perf_counter() - t0


class CpuPool:
    """The reference's own parallel structure (find_len_scales.py:73-77: a process pool over particles,
    one BLAS thread each, GP_func.py:3-7), widened to every host core."""

    def __init__(self, workers, blas_threads=1):
        import multiprocessing as mp
        self.workers = workers
        self.blas_threads = blas_threads
        saved = {k: os.environ.get(k) for k in ("OMP_NUM_THREADS", "OPENBLAS_NUM_THREADS", "MKL_NUM_THREADS")}
        for k in saved:
            os.environ[k] = str(blas_threads)          # inherited by the spawned workers before they import numpy
        self.pool = mp.get_context("spawn").Pool(workers)
        self.pool.map(_noop, range(workers))
        for k, v in saved.items():
            if v is None:
                os.environ.pop(k, None)
            else:
                os.environ[k] = v

    def step(self, n_evals, seed=0):
        t0 = time.perf_counter()
        out = self.pool.map(_cpu_eval, [(seed, k) for k in range(n_evals)], chunksize=1)
        return time.perf_counter() - t0, out

    def close(self):
        self.pool.terminate()
        self.pool.join()


def _noop(i):
    import numpy  # noqa: F401
    from oracle import gp_oracle  # noqa: F401
    return i


def cpu_baseline_sample():
    """cpu_baseline of the 'ours' line: one pool pass, one evaluation per host core."""
    cores = os.cpu_count() or 1
    pool = CpuPool(cores, 1)
    try:
        wall, out = pool.step(cores)
    finally:
        pool.close()
    return {"value": cores / wall, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{cores} of the {WORKLOAD['particles_per_gpu']} particles, one per process (1 BLAS thread each, "
                      f"the reference's pool structure), N=4096 d=3; {wall:.1f} s wall, "
                      f"{np.mean([o[1] for o in out]):.1f} s per evaluation"}


def cpu_aux_baselines():
    """SURVEY 8(d) CPU legs for the other quantities of the metric, bounded samples, one process (the reference runs
    these parts single-process): (iii) GP() fit + predict on a 10 000-point slice of the C3 grid, (iv) the surface
    reduction on 2 000 merged rows, and the default-workflow loss evaluation (N=100 after the KMeans subsample)."""
    from oracle import gp_oracle as orc
    out = {}
    N, d = WORKLOAD["n_points"], WORKLOAD["dims"]
    x, y, e = orc.synthetic_problem(N, d, 1)
    xfit = np.random.default_rng(0).random((d, 10000))
    t0 = time.perf_counter()
    orc.gp(x, y, e, xfit, np.full(d, 0.25))
    dt = time.perf_counter() - t0
    out["grid_predict"] = {"value": 10000 / dt, "unit": "grid_pts/s", "cores": os.cpu_count(), "kind": "port",
                           "sample": f"GP() fit + predict of one 10 000-point batch at N={N}, d={d}: {dt:.1f} s (NumPy/LAPACK threads as configured)"}
    t = orc.synthetic_surface_table(2000, 16, seed=3)
    t0 = time.perf_counter()
    orc.prob_surface_rows(t)
    dt = time.perf_counter() - t0
    out["prob_surface"] = {"value": 2000 / dt, "unit": "rows/s", "cores": 1, "kind": "port",
                           "sample": f"2 000 merged rows x 16 experiments, row loop of generate_prob_surf: {dt:.2f} s"}
    x, y, e = orc.synthetic_problem(100, 2, 0)
    pos, lo, hi = orc.synthetic_particles(x, 40, 0)
    sv, ex = orc.sigma_grid()
    t0 = time.perf_counter()
    for k in range(200):
        orc.evaluate_loss(pos[k % 40], x, y, e, sv, ex, lo, hi)
    dt = time.perf_counter() - t0
    out["len_scale_opt_default"] = {"value": 200 / dt, "unit": "loss_evals/s", "cores": 1, "kind": "port",
                                    "sample": f"200 evaluate_loss calls at N=100, d=2 in one process: {dt:.2f} s; the reference maps "
                                              f"them over max(1, cpu_count // 4) pool workers"}
    return out


def run_reference(args):
    """--impl reference: the CPU implementation of the path (NumPy oracle port) on all host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    pool = CpuPool(cores, 1)
    mode = f"process pool, {cores} workers x 1 BLAS thread, {cores} evaluations per step"
    evals_per_step = cores
    budget = 240.0
    times = []
    try:
        for i in range(args.warmup):
            wall, _ = pool.step(evals_per_step)
            if i == 0 and wall * (args.steps + args.warmup) > budget:
                # too slow for the requested step count: one evaluation per step on all cores' BLAS threads
                pool.close()
                pool = CpuPool(1, cores)
                evals_per_step = 1
                mode = f"1 worker x {cores} BLAS threads, 1 evaluation per step (pool steps would exceed {budget:.0f} s)"
        t_all = time.perf_counter()
        for _ in range(args.steps):
            wall, _ = pool.step(evals_per_step)
            times.append(wall)
        total = time.perf_counter() - t_all
    finally:
        pool.close()
    value = evals_per_step * args.steps / total
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": "C3 loss evaluation: N=4096 points, d=3, swarm of 256 particles, S=1000 sigma scales",
                   **WORKLOAD, "cpu_mode": mode},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{evals_per_step} of the 256 particles per step; {mode}"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------- GPU side
def measure_fp64_peak(torch, dev):
    """FP64 dense peak measured live with a plain library DGEMM (cuBLAS via torch.matmul, 8192^3): the
    roofline denominator for the FP64 tensor (DMMA) work; MEASURED_PEAKS.json only has HBM and bf16."""
    n = 8192
    a = torch.randn(n, n, dtype=torch.float64, device=dev)
    b = torch.randn(n, n, dtype=torch.float64, device=dev)
    for _ in range(2):
        a @ b
    torch.cuda.synchronize()
    best = 1e30
    for _ in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); a @ b; e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    del a, b
    torch.cuda.empty_cache()
    return 2.0 * n ** 3 / (best * 1e-3) / 1e12


# DRAM bytes per launch of the digit-GEMM roles, from the ncu launch list of this command (dram__bytes_read.sum +
# dram__bytes_write.sum, mean over the launches of one role; 64-particle groups): profiles/r01_launch_summary_v12.csv
NCU_TRAFFIC_PER_LAUNCH = {"k_oz_invu": 1409.57e6 + 219.37e6, "k_oz_syrk": 1445.11e6 + 107.60e6}


def dram_traffic(kernel, N, P, launches):
    """roofline.traffic for the top kernel (measured by ncu earlier, same command) next to its algorithmic bytes."""
    if kernel not in NCU_TRAFFIC_PER_LAUNCH or N != 4096 or P != 256:
        return {"traffic": None}
    T = N // 128
    pairs = sum(n * (n + 1) // 2 for n in range(T))            # 128 x 128 operand blocks read per matrix per phase
    per_matrix = pairs * 128 * 128 * 7                          # A digits: 7 bytes per element, each block read once
    if kernel == "k_oz_invu":
        per_matrix += sum(range(T)) * 128 * 128 * 7             # B digits: the phase-2 panel (SYRK's B is one of its A row blocks)
    if kernel == "k_oz_syrk":
        per_matrix += sum(range(T)) * 128 * 128 * (8 + 7)       # covariance tile read, digits of the Schur block written
        per_matrix += T * 128 * 128 * 8                         # FP64 copy of the diagonal Schur blocks only
    else:
        per_matrix += sum(range(T)) * 128 * 128 * (8 + 7)       # FP64 output tile + its digits for the next contraction
    return {"traffic": NCU_TRAFFIC_PER_LAUNCH[kernel],
            "traffic_unit": "DRAM bytes per launch (64-particle group)",
            "traffic_source": "ncu launch list of this command, profiles/r01_launch_summary_v12.csv",
            "algorithmic_bytes_per_launch": per_matrix * 64 / (T - 1)}


def aux_measurements(torch, engine):
    """The other two quantities BASELINE.json's metric names, on one GPU: grid points/s of the predictive mean/sd
    (C3: N=4096, d=3, 100^3 grid) and rows/s of the probability surface (16 experiments with gaps), each device-resident
    and end to end through the public call with host arrays."""
    from gaussian_process_b200 import synthetic
    out = {}
    N, d, M = WORKLOAD["n_points"], WORKLOAD["dims"], 100 ** 3
    x, y, e = synthetic.problem(N, d, 1)
    ls = np.full(d, 0.25)
    xfit = np.random.default_rng(0).random((d, M))
    engine.gp(x, y, e, xfit[:, :20000], ls)                                   # warm-up (workspaces, module load)
    t0 = time.perf_counter()
    mu, sd = engine.gp(x, y, e, xfit, ls)
    e2e_s = time.perf_counter() - t0
    fit = engine.GPFit(x, y, e, ls)
    grid = engine.to_dev(xfit)
    fit.predict_device(grid, M)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    fit.predict_device(grid, M)
    e1.record()
    torch.cuda.synchronize()
    out["grid_predict"] = {"workload": f"C3 prediction: N={N}, d={d}, {M} grid points", "value": M / (e0.elapsed_time(e1) * 1e-3),
                           "unit": "grid_pts/s", "e2e": {"value": M / e2e_s, "unit": "grid_pts/s", "includes": "fit + H2D of the grid + predict + D2H of mean and sd",
                                                         "h2d_bytes": int(xfit.nbytes), "d2h_bytes": int(mu.nbytes + sd.nbytes)},
                           "precision": {0: "fp64", 1: "emu", 2: "emu28"}[fit.mode], "sd_checksum": float(sd.sum())}
    del grid, fit
    R, E = 1 << 22, 16
    t = synthetic.surface_table(1 << 18, E, seed=3)
    table = np.tile(t, (R // t.shape[0], 1))
    engine.prob_surface(table[:1 << 16])
    t0 = time.perf_counter()
    idx, yy, pp = engine.prob_surface(table)
    s_e2e = time.perf_counter() - t0
    out["prob_surface"] = {"workload": f"surface reduction: {R} merged rows, {E} experiments, 25% gaps, 100 points per row",
                           "e2e": {"value": R / s_e2e, "unit": "rows/s", "h2d_bytes": int(table.nbytes), "d2h_bytes": int(yy.nbytes + pp.nbytes + idx.nbytes)},
                           "valid_rows": int(len(idx))}
    del idx, yy, pp
    # the same reduction as the workflow uses it: generate_prob_surf streaming into the CSV text of the output file
    import pandas as pd
    import tempfile
    from gaussian_process_b200 import surface
    Rc = 1 << 16
    df = pd.DataFrame(np.hstack([np.random.default_rng(1).random((Rc, 4)), table[:Rc]]),
                      columns=["d0", "d1", "d2", "d3"] + [f"c{i}" for i in range(2 * E)])
    with tempfile.TemporaryDirectory() as tmp:
        path = os.path.join(tmp, "prob_surf.txt")
        t0 = time.perf_counter()
        n_out = surface.write_surface_csv(df, 4, path)
        s_csv = time.perf_counter() - t0
        nbytes = os.path.getsize(path)
    out["prob_surface"]["csv_e2e"] = {"value": Rc / s_csv, "unit": "rows/s", "merged_rows": Rc, "output_rows": int(n_out),
                                      "csv_bytes": int(nbytes), "host_threads": os.cpu_count(),
                                      "includes": "H2D, reduction, D2H, long-format expansion, CSV text (to_csv bytes), file write"}
    # BASELINE config C2: N=1024, d=2, 64 particles (loss), 200 x 200 grid (prediction)
    from gaussian_process_b200 import pso as _pso
    x2, y2, e2 = synthetic.problem(1024, 2, 0)
    pos2, _lo2, _hi2 = synthetic.particles(x2, 64, 0)
    sv2, _ex2 = _pso.sigma_grid()
    ctx2 = engine.LossContext(x2, y2, e2, sv2)
    dpos2 = engine.to_dev(pos2)
    dact2 = torch.ones(64, dtype=torch.uint8, device=dpos2.device)
    for _ in range(3):
        ctx2.counts_device(dpos2, dact2, 64)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        ctx2.counts_device(dpos2, dact2, 64)
    e1.record()
    torch.cuda.synchronize()
    g = np.linspace(0.0, 1.0, 200)
    grid2 = np.ascontiguousarray(np.stack(np.meshgrid(g, g, indexing="ij")).reshape(2, -1))
    engine.gp(x2, y2, e2, grid2, np.array([0.25, 0.25]))
    t0 = time.perf_counter()
    engine.gp(x2, y2, e2, grid2, np.array([0.25, 0.25]))
    t_c2 = time.perf_counter() - t0
    out["config_c2"] = {"workload": "C2: N=1024, d=2, 64 particles; 200 x 200 prediction grid",
                        "loss": {"value": 64 * 20 / (e0.elapsed_time(e1) * 1e-3), "unit": "loss_evals/s", "ms_per_swarm": e0.elapsed_time(e1) / 20,
                                 "precision": {0: "fp64", 1: "emu", 2: "emu28"}[ctx2.mode]},
                        "predict_e2e": {"value": grid2.shape[1] / t_c2, "unit": "grid_pts/s", "includes": "fit + H2D + predict + D2H"}}
    del ctx2
    # the reference's real PSO configuration: KMeans subsample to 100 points, 40 particles, 500 iterations
    import contextlib
    import io
    from gaussian_process_b200 import pso
    x, y, e = synthetic.problem(1024, 2, 0)
    trace = []
    np.random.seed(0)
    with contextlib.redirect_stdout(io.StringIO()):
        pso.len_scale_opt(x, y, e, False, lhs_random_state=0, max_iter=20)           # warm-up (module load, pinned buffers)
        np.random.seed(0)
        t0 = time.perf_counter()
        best = pso.len_scale_opt(x, y, e, False, lhs_random_state=0, trace=trace)
        dt = time.perf_counter() - t0
    n_evals = sum(len(sc) for _p, sc in trace)
    out["len_scale_opt_default"] = {"workload": "len_scale_opt as GP_fit calls it: C2 points (N=1024, d=2), KMeans subsample to 100, 40 particles, 500 iterations",
                                    "seconds": dt, "swarm_evaluations": len(trace), "loss_evaluations": n_evals,
                                    "e2e": {"value": n_evals / dt, "unit": "loss_evals/s",
                                            "includes": "KMeans subsample, LHS, host PSO update, H2D positions, loss kernel, D2H counts, host finish"},
                                    "best_length_scales": [float(v) for v in best]}
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist
    from gaussian_process_b200 import engine, pso, sharding, _lib

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    cpu_base = cpu_aux = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu_base = cpu_baseline_sample()            # before CUDA is initialised in this process
        cpu_aux = None if args.no_aux else cpu_aux_baselines()

    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    json_fd = None
    if world > 1:
        # NCCL writes its version banner to stdout at communicator creation: keep stdout for the one JSON line
        sys.stdout.flush()
        json_fd = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
    _lib.load()
    engine.set_precision(args.precision)
    N, d, P, S = WORKLOAD["n_points"], WORKLOAD["dims"], args.particles, WORKLOAD["n_sigma"]
    x, y, e = make_problem(0)
    # every rank owns a different slice of the global swarm (world * P particles)
    pos_all, lower, upper = make_particles(x, world * P, 0)
    pos = np.ascontiguousarray(pos_all[rank * P:(rank + 1) * P])
    sigma_vals, expected = pso.sigma_grid()

    fp64_peak = measure_fp64_peak(torch, dev) if rank == 0 else None
    ctx = engine.LossContext(x, y, e, sigma_vals)
    dpos = engine.to_dev(pos)
    dact = torch.ones(P, dtype=torch.uint8, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput ("value") ----
    for _ in range(args.warmup):
        counts, info = ctx.counts_device(dpos, dact, P)
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = _lib.launch_count()
    st = torch.cuda.current_stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(args.steps):
        counts, info = ctx.counts_device(dpos, dact, P)
    e1.record(st)
    barrier()
    launches = _lib.launch_count() - launches0
    clocks = sampler.stop() if rank == 0 else None
    dev_ms = e0.elapsed_time(e1)
    assert int(info.abs().sum().item()) == 0, "covariance not positive definite in the bench workload"

    # ---- the same region with the native FP64 (DMMA) contractions, for reference next to the emulated default ----
    native = None
    if args.precision != "fp64" and not args.no_native:
        engine.set_precision("fp64")
        for _ in range(2):
            ctx.counts_device(dpos, dact, P)
        torch.cuda.synchronize()
        n0, n1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n0.record(st)
        for _ in range(2):
            c_nat, _i = ctx.counts_device(dpos, dact, P)
        n1.record(st)
        torch.cuda.synchronize()
        native = {"value": P * 2 / (n0.elapsed_time(n1) * 1e-3), "unit": UNIT, "per": "GPU", "steps": 2,
                  "counts_identical_to_default": bool(torch.equal(c_nat, counts))}
        engine.set_precision(args.precision)
        ctx.counts_device(dpos, dact, P)
        torch.cuda.synchronize()

    # ---- and with 28 instead of 34 digit products (twice the error of the native FP64 kernels on well-conditioned
    #      problems, profiles/r01_small_e_accuracy_v2.log; an option, not the default) ----
    emu28 = None
    if args.precision == "emu" and not args.no_native:
        engine.set_precision("emu28")
        for _ in range(2):
            ctx.counts_device(dpos, dact, P)
        torch.cuda.synchronize()
        n0, n1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n0.record(st)
        for _ in range(2):
            c_28, _i = ctx.counts_device(dpos, dact, P)
        n1.record(st)
        torch.cuda.synchronize()
        emu28 = {"value": P * 2 / (n0.elapsed_time(n1) * 1e-3), "unit": UNIT, "per": "GPU", "steps": 2,
                 "counts_identical_to_default": bool(torch.equal(c_28, counts))}
        engine.set_precision(args.precision)
        ctx.counts_device(dpos, dact, P)
        torch.cuda.synchronize()

    # ---- per-kernel-class device time: one extra, separately instrumented step (single stream, events after
    #      every launch); the timed region above runs uninstrumented with concurrent particle groups ----
    _lib.timing_begin(st.cuda_stream)
    ctx.counts_device(dpos, dact, P)
    torch.cuda.synchronize()
    classes = _lib.timing_end()

    # ---- end to end through the public API with host buffers ("e2e") ----
    def e2e_step():
        local_loss = pso.evaluate_loss_batch(pos, x, y, e, sigma_vals, expected, lower, upper)
        return sharding.all_gather_losses(local_loss, world * P) if world > 1 else local_loss

    for _ in range(min(args.warmup, 2)):
        losses = e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        losses = e2e_step()
    barrier()
    e2e_s = time.perf_counter() - t0
    h2d = (d * N + 2 * N + S + P * d) * 8 + P
    d2h = P * S * 4 + P * 4

    aux = None
    if rank == 0 and world == 1 and not args.no_aux:
        del ctx
        engine.release_workspaces()
        aux = aux_measurements(torch, engine)
        for k, v in (cpu_aux or {}).items():
            aux.setdefault(k, {})["cpu_baseline"] = v

    if world > 1:
        t = torch.tensor([dev_ms, e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms, e2e_s = float(t[0]), float(t[1])
    if rank == 0:
        evals = world * P * args.steps
        value = evals / (dev_ms * 1e-3)
        top = max(classes.items(), key=lambda kv: kv[1][0])
        total_cls_ms = sum(v[0] for v in classes.values())
        # algorithmic FP64 flops of the top kernel class over the timed region (this rank)
        T = (N + 127) // 128
        alg = {"k_syrk_update": N ** 3 / 3.0, "k_inv_update": N ** 3 / 3.0, "k_oz_syrk": N ** 3 / 3.0, "k_oz_invu": N ** 3 / 3.0}
        flops_top = alg.get(top[0], 0.0) * P
        achieved = flops_top / (top[1][0] * 1e-3) / 1e12 if flops_top else None
        emulated = top[0].startswith("k_oz_")
        if emulated:
            # the digit GEMM does `prods` int8 multiply-adds per FP64 multiply-add: report it as int8 tensor work
            prods = 28 if args.precision == "emu28" else 34
            achieved = achieved * prods
            flops_top = flops_top * prods
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": {"workload": "C3 loss evaluation: N=4096 points, d=3, swarm of 256 particles per GPU, S=1000 sigma scales",
                       **WORKLOAD, "particles_per_gpu": P, "algorithm": "Cholesky + triangular inverse (2/3 N^3 flop/eval)", "precision": args.precision,
                       "arithmetic": ("FP64 emulated on the int8 tcgen05 tensor cores: error-free 7 x 8-bit digit split, "
                                      "34 digit products, exact int32 accumulation; same rounding level as FP64 (DESIGN.md 3.1)")
                       if args.precision != "fp64" else "native FP64 (DMMA)",
                       "l2_policy": "working set 34 GB per step >> 126 MB L2 (inputs larger than L2)"},
            "e2e": {"value": evals / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": {"bound": "tensor", "kernel": top[0], "achieved": achieved,
                         "peak": INT8_PEAK_TOPS if emulated else fp64_peak, "unit": "TOP/s (int8)" if emulated else "TFLOP/s",
                         "frac": (achieved / (INT8_PEAK_TOPS if emulated else fp64_peak)) if achieved else None,
                         **dram_traffic(top[0], N, P, top[1][1]),
                         "peak_source": INT8_PEAK_SOURCE if emulated else
                                        "FP64 cuBLAS DGEMM 8192^3 measured in this run (MEASURED_PEAKS.json has no FP64 figure; "
                                        "DMMA issue-rate probe: 37.1 TFLOP/s, profiles/r01_fp64_peak_probe.txt)",
                         "fp64_dgemm_tflops_measured": fp64_peak,
                         "peak_sustained_random_operands": INT8_SUSTAINED_RANDOM_TOPS if emulated else None,
                         "frac_of_sustained": (achieved / INT8_SUSTAINED_RANDOM_TOPS) if (emulated and achieved) else None,
                         "algorithmic_flops_per_launch": flops_top / top[1][1] if flops_top else None,
                         "avg_launch_ms": top[1][0] / top[1][1],
                         "whole_step_tflops": algorithmic_flops_per_eval(N) * P * args.steps / (dev_ms * 1e-3) / 1e12,
                         "kernel_share_of_step": {k: round(v[0] / total_cls_ms, 4) for k, v in sorted(classes.items(), key=lambda kv: -kv[1][0])}},
            "cpu_baseline": cpu_base,
            "native_fp64": native,
            "emu28": emu28,
            "aux": aux,
            "loss_checksum": float(np.sum(losses[:P])),
        }
        if json_fd is not None:
            os.write(json_fd, (json.dumps(line) + "\n").encode())
        else:
            print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--particles", type=int, default=WORKLOAD["particles_per_gpu"], help="particles per GPU")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-aux", action="store_true", help="skip the grid-prediction and surface measurements")
    ap.add_argument("--no-native", action="store_true", help="skip the short native-FP64 (DMMA) comparison run")
    ap.add_argument("--precision", default="emu", choices=["fp64", "emu", "emu28"],
                    help="fp64: native DMMA contractions; emu: FP64 emulated on the int8 tcgen05 tensor cores "
                         "(34 digit products); emu28: 28 digit products")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
