#!/usr/bin/env python
"""Benchmark of the GP-fitting hot path: PSO loss evaluations / second (batched GP fits) at N=4096.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  torchrun --nproc-per-node N bench.py --gpus N ...          (one rank per GPU, NCCL)

Workload (BASELINE.json configs[2], SURVEY 8d "C3"): synthetic 3-D problem with N=4096 training points and ONE swarm of
256 particles, sharded over the ranks (STRONG scaling: 256 / N particles per GPU; the only exchange is the all-gather of
the per-particle losses, inside the end-to-end region).  A step = one swarm evaluation = 256 loss evaluations, each a full
GP fit (covariance build, blocked Cholesky, triangular inverse, self-prediction, count reduction) -- no KMeans subsample
(SURVEY F3), nothing cached between steps.  `aux` carries the other quantities of the metric at every N (C4: experiments
sharded through create_GP; C5: surface rows sharded) and, at N > 1, the weak-scaling number (256 particles per GPU).

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

WORKLOAD = dict(n_points=4096, dims=3, particles_total=256, n_sigma=1000)
# the `config` object both arms print (the driver compares them): the workload only, nothing arm specific
CONFIG = {"workload": "C3 loss evaluation: N=4096 points, d=3, one swarm of 256 particles sharded over the GPUs, S=1000 sigma scales",
          **WORKLOAD, "l2_policy": "working set per step (134 MB per particle) >> 126 MB L2: inputs larger than L2"}
INT8_PEAK_TOPS = 4497.0
# the same instruction stream with uniformly random signed digits as operands and no memory traffic at all: the board
# reaches its 1000 W cap and holds 3780 TOP/s (tools/umma_power.cu, profiles/r01_umma_power.log)
INT8_SUSTAINED_RANDOM_TOPS = 3780.0
INT8_PEAK_SOURCE = ("int8 tcgen05.mma issue-rate probe on this pool's B200 (tools/umma_probe.cu, 148 CTAs, N=256: 4497 TOP/s, "
                    "profiles/r01_umma_probe.log; nominal 4500).  MEASURED_PEAKS.json has only bf16 (cuBLAS); twice its "
                    "sustained figure is reported beside it as peak_from_measured_peaks")
METRIC = "pso_loss_evals_per_sec_n4096"
UNIT = "loss_evals/s"
PRODUCTS = {"emu": 34, "emu28": 28, "emu21": 21, "emu15": 15, "emu28_34": 31, "emu34_28": 31}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f)
    except Exception:
        return {}


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
    pos, lo, hi = orc.synthetic_particles(x, WORKLOAD["particles_total"], seed)
    s, ex = orc.sigma_grid()
    t0 = time.perf_counter()
    v = orc.evaluate_loss(pos[k % len(pos)], x, y, e, s, ex, lo, hi)
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

    def throughput(self, n_evals, seed=0):
        """Steady-state throughput of the saturated pool: n_evals + workers evaluations are submitted and the clock stops when
        n_evals of them have completed.  (A blocking map waits for its slowest task, and with one worker per core one
        evaluation in ~25 takes several times longer than the others -- 78 s against 12 s in the build container,
        7.5 s typical on the GPU boxes -- which understated the reference arm by a third in round-2 runs.)  The
        outstanding evaluations are abandoned: call close() afterwards.  -> (seconds, [(value, seconds of the evaluation)])."""
        tasks = [(seed, k) for k in range(n_evals + self.workers)]
        t0 = time.perf_counter()
        out = []
        for res in self.pool.imap_unordered(_cpu_eval, tasks, chunksize=1):
            out.append(res)
            if len(out) >= n_evals:
                break
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
    n = 2 * cores
    try:
        wall, out = pool.throughput(n)
    finally:
        pool.close()
    return {"value": n / wall, "unit": UNIT, "cores": cores, "kind": "port",
            "sample": f"{n} of the {WORKLOAD['particles_total']} particles through a pool of {cores} processes (1 BLAS thread each, "
                      f"the reference's pool structure) kept saturated, N=4096 d=3; {wall:.1f} s wall, "
                      f"{np.median([o[1] for o in out]):.1f} s per evaluation (median)"}


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
    """--impl reference: the CPU implementation of the path (NumPy oracle port) on all host cores, in the reference's own
    parallel structure on EVERY box: a process pool over the particles, one BLAS thread per process
    (find_len_scales.py:73-77, GP_func.py:3-7), widened to every core.  A step is a bounded sample of the 256-particle swarm
    (`chunk` particles); the K steps are handed to the pool as one stream so that it stays saturated (no barrier between
    steps), and `chunk` is sized so that the timed region is about four full rounds of the pool whatever K is."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    pool = CpuPool(cores, 1)
    steps = max(1, args.steps)
    chunk = max(1, -(-4 * cores // steps))
    mode = (f"process pool, {cores} workers x 1 BLAS thread (the reference's structure); a step = {chunk} of the 256 "
            f"particles, steps streamed through the saturated pool, clock stopped at the last evaluation of the last step")
    try:
        pool.step(max(cores, min(args.warmup, 3) * chunk))          # warm-up: at least one full round
        total, out = pool.throughput(steps * chunk)
    finally:
        pool.close()
    value = steps * chunk / total
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / steps, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": dict(CONFIG),
        "impl_detail": {"cpu_mode": mode, "seconds_per_evaluation_median": float(np.median([o[1] for o in out])),
                        "seconds_per_evaluation_max": float(np.max([o[1] for o in out]))},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port",
                         "sample": f"{steps * chunk} evaluations ({chunk} of the 256 particles per step); {mode}"},
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
# dram__bytes_write.sum, mean over the launches of one role; 128-particle groups): profiles/r02_launch_summary_dram.csv
NCU_TRAFFIC_PROFILE = ("profiles/r02_launch_summary_dram.csv (ncu launch list of the bench command, round-2 code, default two "
                       "groups of 128 particles: mean over the captured launches of the role)")
NCU_TRAFFIC_GROUP = 128
NCU_TRAFFIC_PER_LAUNCH = {"k_oz_invu": 2819.19e6 + 462.36e6, "k_oz_syrk": 2876.18e6 + 225.17e6}


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
    # avg_launch_ms comes from the instrumented step, which runs the whole slice as ONE group (events after every launch need
    # one stream); DRAM bytes scale linearly with the particles of a launch (each particle has its own operands), so the ncu
    # figure of the 128-particle launches is scaled to that launch size: both fields describe the same launch
    scale = P / NCU_TRAFFIC_GROUP
    return {"traffic": NCU_TRAFFIC_PER_LAUNCH[kernel] * scale,
            "traffic_unit": f"DRAM bytes per launch of {P} particles (the instrumented step's launches) = {scale:g} x the "
                            f"{NCU_TRAFFIC_GROUP}-particle launches ncu measured",
            "traffic_measured_in_run": False,
            "traffic_from_profile": NCU_TRAFFIC_PROFILE,
            "algorithmic_bytes_per_launch": per_matrix * P / (T - 1)}


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
                           "precision": engine.MODE_NAMES[fit.mode], "fell_back_to_fp64": bool(fit.fell_back), "probe": float(fit.probe), "sd_checksum": float(sd.sum())}
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
                                 "precision": engine.MODE_NAMES[ctx2.mode], "fallbacks": int(ctx2.last_fallback.sum().item())},
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


def timed_steps(torch, fn, steps, warmup, barrier):
    """Device time (ms) of `steps` calls of fn after `warmup` untimed ones, CUDA events on the current stream, barrier +
    synchronize on both sides."""
    for _ in range(warmup):
        fn()
    barrier()
    st = torch.cuda.current_stream()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(st)
    for _ in range(steps):
        out = fn()
    e1.record(st)
    barrier()
    return e0.elapsed_time(e1), out


def aux_c4_experiments(torch, dist, rank, world):
    """BASELINE config C4 through the drop-in GP_fit.create_GP: 8 experiments of ~4096 points (8192-row file with
    inf-masked gaps), KMeans subsample + PSO + lattice + GP fit / predict per experiment, outer merge; whole experiments shard
    over the ranks (no collective on the data path, one object gather of the result tables)."""
    import contextlib
    import importlib
    import io
    import tempfile
    import types
    from gaussian_process_b200 import synthetic, sharding
    dropin = os.path.join(ROOT, "gaussian-process_b200", "dropin")
    if dropin not in sys.path:
        sys.path.insert(0, dropin)
    x_list, pairs = synthetic.c4_experiments(8192, 8, 4)
    labels = ["a", "b", "q", "err"]
    res = 0.005
    data_list = [("data/c4.txt", x_list, pairs, labels)]
    read_in = types.ModuleType("read_in")                  # the reference's reader, stubbed with the synthetic file
    read_in.read_yaml = lambda: ([res, res], False, "results/GP_results.txt", labels, data_list, False, False)
    sys.modules["read_in"] = read_in
    gp_fit = importlib.import_module("GP_fit")
    phase = {}
    pso_split = {}

    def timed(name, fn):
        def wrapper(*a, **k):
            t0 = time.perf_counter()
            out = fn(*a, **k)
            torch.cuda.synchronize()
            phase[name] = phase.get(name, 0.0) + time.perf_counter() - t0
            return out
        return wrapper

    inner_opt = gp_fit.len_scale_opt

    def seeded_opt(xk, yk, ek, progress):
        np.random.seed(0)
        return inner_opt(xk, yk, ek, progress, lhs_random_state=0, timings=pso_split)

    saved = (gp_fit.len_scale_opt, gp_fit.fill_convex_hull, gp_fit.GP, gp_fit.merge_outer, gp_fit.write_csv, gp_fit.read_yaml)
    gp_fit.len_scale_opt = timed("len_scale_opt", seeded_opt)
    gp_fit.fill_convex_hull = timed("fill_convex_hull", gp_fit.fill_convex_hull)
    gp_fit.GP = timed("GP_fit_predict", gp_fit.GP)
    gp_fit.merge_outer = timed("outer_merge", gp_fit.merge_outer)
    gp_fit.write_csv = timed("csv", gp_fit.write_csv)
    holder = tempfile.TemporaryDirectory() if rank == 0 else None
    tmp = sharding.broadcast_object(holder.name if rank == 0 else None)
    cwd = os.getcwd()
    try:
        os.chdir(tmp)
        small = [("data/w.txt", [x_list[0][:, :300]], [(pairs[0][0][:300], pairs[0][1][:300])], labels)]
        gp_fit.read_yaml = lambda: ([0.05, 0.05], False, "results/w.txt", labels, small, False, False)
        with contextlib.redirect_stdout(io.StringIO()):
            gp_fit.create_GP()                             # warm-up: module load, workspaces, pinned buffers
        phase.clear()
        pso_split.clear()
        gp_fit.read_yaml = read_in.read_yaml
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        with contextlib.redirect_stdout(io.StringIO()):
            df, _nd = gp_fit.create_GP()
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        dt = time.perf_counter() - t0
    finally:
        os.chdir(cwd)
        (gp_fit.len_scale_opt, gp_fit.fill_convex_hull, gp_fit.GP, gp_fit.merge_outer, gp_fit.write_csv, gp_fit.read_yaml) = saved
        sys.modules.pop("read_in", None)
        if world > 1:
            dist.barrier()
        if holder is not None:
            holder.cleanup()
    n_pts = int(np.mean([v.shape[1] for v in x_list]))
    return {"workload": f"C4 through create_GP: 8192-row 2-D file, 8 experiments (~{n_pts} points each), resolution {res} "
                        f"({len(df)} lattice rows), experiments sharded over {world} rank(s)",
            "value": 8 / dt, "unit": "experiments/s", "seconds": dt, "grid_pts_per_s": 8 * len(df) / dt,
            "rank0_phase_seconds": {k: round(v, 4) for k, v in phase.items()},
            "rank0_len_scale_opt_split_seconds": {k: round(v, 4) for k, v in pso_split.items()},
            "merged_table_shape": list(df.shape)}


def aux_c5_surface(torch, dist, rank, world, engine):
    """Surface reduction (C5 shape: E=16, 25 % gaps) on 2^22 merged rows, rows sharded over the ranks through
    sharding.sharded_prob_surface (host table in, host arrays out; the only exchange is surface_offsets: one int64 per rank)."""
    from gaussian_process_b200 import synthetic, sharding
    R, E = 1 << 22, 16
    t = synthetic.surface_table(1 << 18, E, seed=3)
    table = np.tile(t, (R // t.shape[0], 1))
    engine.prob_surface(table[:1 << 16])
    if world > 1:
        dist.barrier()
    t0 = time.perf_counter()
    offset, total, idx, yy, pp = sharding.sharded_prob_surface(engine.prob_surface, table)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    dt = time.perf_counter() - t0
    out = {"workload": f"surface reduction: {R} merged rows, {E} experiments, 25% gaps, 100 points per row, rows sharded over "
                       f"{world} rank(s), host arrays in and out",
           "value": R / dt, "unit": "rows/s", "seconds": dt, "valid_rows_total": int(total), "rank0_rows": int(len(idx)),
           "h2d_bytes_rank0": int(table.nbytes // world), "d2h_bytes_rank0": int(yy.nbytes + pp.nbytes + idx.nbytes)}
    del idx, yy, pp, table
    return out


def run_ours(args):
    import torch
    import torch.distributed as dist
    from gaussian_process_b200 import engine, pso, sharding, _lib

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    cpu_base = cpu_aux = None
    if rank == 0 and not args.no_cpu_baseline:
        cpu_base = cpu_baseline_sample()            # before CUDA is initialised in this process; the other ranks wait
        cpu_aux = None if (args.no_aux or world > 1) else cpu_aux_baselines()

    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    json_fd = None
    if world > 1:
        # NCCL writes its version banner to stdout at communicator creation: keep stdout for the one JSON line
        import datetime
        sys.stdout.flush()
        json_fd = os.dup(1)
        os.dup2(2, 1)
        dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev, timeout=datetime.timedelta(minutes=20))
    if world > 1:
        # torchrun pins OMP_NUM_THREADS=1 per rank: give the host-side copies (pinned staging -> result arrays) their share
        torch.set_num_threads(max(1, (os.cpu_count() or 1) // world))
    _lib.load()
    engine.set_precision(args.precision)
    N, d, S = WORKLOAD["n_points"], WORKLOAD["dims"], WORKLOAD["n_sigma"]
    P_total = args.particles
    x, y, e = make_problem(0)
    pos_all, lower, upper = make_particles(x, P_total, 0)              # ONE swarm; every rank owns a contiguous slice of it
    lo_p, hi_p = sharding.shard_slice(P_total, rank, world)
    P = hi_p - lo_p
    pos = np.ascontiguousarray(pos_all[lo_p:hi_p])
    sigma_vals, expected = pso.sigma_grid()

    fp64_peak = measure_fp64_peak(torch, dev) if rank == 0 else None
    ctx = engine.LossContext(x, y, e, sigma_vals)
    dpos = engine.to_dev(pos)
    dact = torch.ones(P, dtype=torch.uint8, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput ("value"): every rank evaluates its slice of the one swarm ----
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

    # ---- checked mixed precision: the a-posteriori probe of this rank's slice and the number of FP64 fallbacks ----
    _c, _i, probe_vals = ctx.counts(pos, want_probe=True)
    checked = {"check_tol": engine.get_check_tol(), "fallbacks": int(ctx.last_fallback.sum().item()), "particles": P,
               "probe_max": float(probe_vals.max()), "probe_median": float(np.median(probe_vals)),
               "what": "rho = |W^T W K v - v|_inf / |v|_inf per particle; rho > check_tol -> re-evaluated by the native FP64 kernels in the same call"}

    # ---- other arithmetic modes on the same slice (N = 1: the whole swarm), 2 steps each: native FP64 (DMMA), and fewer
    #      digit products.  Each emulated mode is measured UNCHECKED (check_tol = 0: what the arithmetic alone delivers --
    #      throughput, errors against the native FP64 result of the same particles, count differences) and then under the
    #      default check (how many particles the probe sends back to the native kernels) ----
    modes = None
    if world == 1 and not args.no_native:
        tol0 = engine.get_check_tol()
        engine.set_precision("fp64")
        nat_ms, _o = timed_steps(torch, lambda: ctx.counts_device(dpos, dact, P), 2, 1, barrier)
        c64, _i64, r64, s64 = ctx.counts_device(dpos[:16], dact[:16], 16, want_detail=True, want_probe=True)
        p64 = float(ctx.last_probe.max().item())
        c64_all, _ = ctx.counts_device(dpos, dact, P)
        modes = {"fp64": {"value": P * 2 / (nat_ms * 1e-3), "unit": UNIT, "steps": 2, "probe_max_16_particles": p64,
                          "counts_identical_to_default": bool(torch.equal(c64_all, counts))}}
        rmax = float(r64.abs().max())
        for m in ("emu", "emu28", "emu21", "emu15"):
            engine.set_precision(m)
            engine.set_check_tol(0.0)
            ms, _o = timed_steps(torch, lambda: ctx.counts_device(dpos, dact, P), 2, 1, barrier)
            cm, _im = ctx.counts_device(dpos, dact, P)
            cm = cm.clone()
            cd, _id, rd, sd_ = ctx.counts_device(dpos[:16], dact[:16], 16, want_detail=True, want_probe=True)
            pm = float(ctx.last_probe.max().item())
            engine.set_check_tol(tol0)
            ms_chk, _o = timed_steps(torch, lambda: ctx.counts_device(dpos, dact, P), 2, 1, barrier)
            fb_all = int(ctx.last_fallback.sum().item())
            modes[m] = {"digit_products": PRODUCTS[m],
                        "unchecked": {"value": P * 2 / (ms * 1e-3), "unit": UNIT, "steps": 2,
                                      "counts_identical_to_fp64": bool(torch.equal(cm, c64_all)),
                                      "particles_with_any_count_difference_vs_fp64": int((cm != c64_all).any(dim=1).sum().item()),
                                      "max_rel_err_resid_vs_fp64_16_particles": float((rd - r64).abs().max()) / rmax,
                                      "max_rel_err_sd_vs_fp64_16_particles": float((sd_ / s64 - 1).abs().max()),
                                      "probe_max_16_particles": pm},
                        "checked": {"check_tol": tol0, "value": P * 2 / (ms_chk * 1e-3), "unit": UNIT, "steps": 2,
                                    "fallbacks": fb_all, "of_particles": P}}
        engine.set_precision(args.precision)
        engine.set_check_tol(tol0)
        ctx.counts_device(dpos, dact, P)
        torch.cuda.synchronize()

    # ---- per-kernel-class device time: one extra, separately instrumented step (single stream, events after
    #      every launch); the timed region above runs uninstrumented with concurrent particle groups ----
    _lib.timing_begin(st.cuda_stream)
    ctx.counts_device(dpos, dact, P)
    torch.cuda.synchronize()
    classes = _lib.timing_end()

    # ---- end to end through the public API with host buffers ("e2e"): H2D of the inputs, kernels, D2H of the counts, host
    #      finish, and (N > 1) the NCCL all-gather of the losses ----
    def e2e_step():
        local_loss = pso.evaluate_loss_batch(pos, x, y, e, sigma_vals, expected, lower, upper)
        return sharding.all_gather_losses(local_loss, P_total) if world > 1 else local_loss

    for _ in range(min(args.warmup, 2)):
        losses = e2e_step()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        losses = e2e_step()
    barrier()
    e2e_s = time.perf_counter() - t0
    h2d = (d * N + 2 * N + S + P * d) * 8 + P
    d2h = P * S * 4 + P * 4 + P

    # ---- N ranks == 1 rank, bit for bit: rank 0 evaluates the whole swarm alone and compares with the gathered losses ----
    identical = None
    if world > 1:
        if rank == 0:
            alone = pso.evaluate_loss_batch(pos_all, x, y, e, sigma_vals, expected, lower, upper)
            identical = bool(np.array_equal(alone, losses))
        dist.barrier()

    # ---- weak scaling beside it (N > 1): 256 particles on every GPU, no collective in the region ----
    weak = None
    if world > 1 and not args.no_aux:
        pos_w, _lw, _uw = make_particles(x, world * P_total, 0)
        dpos_w = engine.to_dev(np.ascontiguousarray(pos_w[rank * P_total:(rank + 1) * P_total]))
        dact_w = torch.ones(P_total, dtype=torch.uint8, device=dev)
        w_ms, _o = timed_steps(torch, lambda: ctx.counts_device(dpos_w, dact_w, P_total), 3, 1, barrier)
        t = torch.tensor([w_ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        weak = {"value": world * P_total * 3 / (float(t[0]) * 1e-3), "unit": UNIT, "particles_per_gpu": P_total, "steps": 3,
                "scaling": "weak"}
        del dpos_w

    # ---- how the strong-scaling slices behave on ONE GPU (the limiter of the N-GPU line): the same swarm cut to the
    #      per-rank sizes of N = 2, 4, 8, by number of concurrent particle groups ----
    slices = None
    if world == 1 and not args.no_aux:
        lib = _lib.load()
        slices = {}
        for pp in (256, 128, 64, 32):
            if pp > P:
                continue
            row = {}
            for G in (1, 2, 4):
                lib.gpk_set_groups(G, 1)
                ms, _o = timed_steps(torch, lambda: ctx.counts_device(dpos[:pp], dact[:pp], pp), 3, 1, barrier)
                row[f"groups_{G}"] = round(pp * 3 / (ms * 1e-3), 1)
            lib.gpk_set_groups(0, 0)                                   # the built-in rule
            ms, _o = timed_steps(torch, lambda: ctx.counts_device(dpos[:pp], dact[:pp], pp), 3, 1, barrier)
            row["default"] = round(pp * 3 / (ms * 1e-3), 1)
            slices[f"particles_{pp}"] = row
        _lib.timing_begin(st.cuda_stream)
        ctx.counts_device(dpos[:32], dact[:32], 32)
        torch.cuda.synchronize()
        c32 = _lib.timing_end()
        slices["class_ms_at_32_particles"] = {k: round(v[0], 3) for k, v in sorted(c32.items(), key=lambda kv: -kv[1][0])}
        slices["unit"] = "loss_evals/s on one GPU; N-GPU strong-scaling efficiency ~ value(256 / N) / value(256)"

    aux = {}
    if not args.no_aux:
        del ctx
        engine.release_workspaces()
        if rank == 0 and world == 1:
            aux = aux_measurements(torch, engine)
            for k, v in (cpu_aux or {}).items():
                aux.setdefault(k, {})["cpu_baseline"] = v
        aux["c4_experiments"] = aux_c4_experiments(torch, dist, rank, world)
        aux["c5_surface"] = aux_c5_surface(torch, dist, rank, world, engine)
        if weak is not None:
            aux["weak"] = weak
        if slices is not None:
            aux["strong_scaling_slices_1gpu"] = slices

    if world > 1:
        t = torch.tensor([dev_ms, e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms, e2e_s = float(t[0]), float(t[1])
        fb = torch.tensor([checked["fallbacks"]], dtype=torch.int64, device=dev)
        dist.all_reduce(fb)
        checked["fallbacks_all_ranks"] = int(fb[0])
    if rank == 0:
        evals = P_total * args.steps
        value = evals / (dev_ms * 1e-3)
        top = max(classes.items(), key=lambda kv: kv[1][0])
        total_cls_ms = sum(v[0] for v in classes.values())
        # algorithmic FP64 flops of the top kernel class over one step (this rank's slice)
        alg = {"k_syrk_update": N ** 3 / 3.0, "k_inv_update": N ** 3 / 3.0, "k_oz_syrk": N ** 3 / 3.0, "k_oz_invu": N ** 3 / 3.0}
        flops_top = alg.get(top[0], 0.0) * P
        achieved = flops_top / (top[1][0] * 1e-3) / 1e12 if flops_top else None
        emulated = top[0].startswith("k_oz_")
        prods = PRODUCTS.get(args.precision, 34)
        if emulated:
            # the digit GEMM does `prods` int8 multiply-adds per FP64 multiply-add: report it as int8 tensor work
            achieved = achieved * prods
            flops_top = flops_top * prods
        mp_ = measured_peaks()
        peak_mp = 2.0 * mp_["bf16_tflops_sustained"] if "bf16_tflops_sustained" in mp_ else None
        step_tops = algorithmic_flops_per_eval(N) * evals / (dev_ms * 1e-3) / 1e12 / world
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": dev_ms / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic",
            "config": dict(CONFIG) if P_total == WORKLOAD["particles_total"] else {**CONFIG, "particles_total": P_total},
            "impl_detail": {"particles_per_gpu": P, "algorithm": "Cholesky + triangular inverse (2/3 N^3 flop/eval)",
                            "precision": args.precision,
                            "arithmetic": (f"FP64 emulated on the int8 tcgen05 tensor cores: error-free 7 x 8-bit digit split, "
                                           f"{prods} digit products, exact int32 accumulation, a-posteriori residual check with "
                                           f"native-FP64 fallback (DESIGN.md 3.1)")
                            if args.precision != "fp64" else "native FP64 (DMMA)"},
            "e2e": {"value": evals / e2e_s, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "collective": "NCCL all-gather of the per-particle losses (inside the region)" if world > 1 else None},
            "gpu_launches": int(launches),
            "clocks": clocks,
            "checked": checked,
            "multirank_identical_to_1rank": identical,
            "roofline": {"bound": "tensor", "kernel": top[0], "achieved": achieved,
                         "peak": INT8_PEAK_TOPS if emulated else fp64_peak, "unit": "TOP/s (int8)" if emulated else "TFLOP/s",
                         "frac": (achieved / (INT8_PEAK_TOPS if emulated else fp64_peak)) if achieved else None,
                         **dram_traffic(top[0], N, P, top[1][1]),
                         "peak_source": INT8_PEAK_SOURCE if emulated else
                                        "FP64 cuBLAS DGEMM 8192^3 measured in this run (MEASURED_PEAKS.json has no FP64 figure; "
                                        "DMMA issue-rate probe: 37.1 TFLOP/s, profiles/r01_fp64_peak_probe.txt)",
                         "peak_from_measured_peaks": peak_mp,
                         "frac_of_peak_from_measured_peaks": (achieved / peak_mp) if (emulated and achieved and peak_mp) else None,
                         "fp64_dgemm_tflops_measured": fp64_peak,
                         "peak_sustained_random_operands": INT8_SUSTAINED_RANDOM_TOPS if emulated else None,
                         "frac_of_sustained": (achieved / INT8_SUSTAINED_RANDOM_TOPS) if (emulated and achieved) else None,
                         "algorithmic_flops_per_launch": flops_top / top[1][1] if flops_top else None,
                         "avg_launch_ms": top[1][0] / top[1][1],
                         "avg_launch_ms_source": "this run: one instrumented step of this rank's slice, one stream, an event after every launch",
                         "whole_step_tflops_fp64_equivalent_per_gpu": step_tops,
                         "whole_step_int8_tops_per_gpu": step_tops * prods if emulated else None,
                         "whole_step_frac_of_peak": (step_tops * prods / INT8_PEAK_TOPS) if emulated else (step_tops / fp64_peak),
                         "kernel_share_of_step": {k: round(v[0] / total_cls_ms, 4) for k, v in sorted(classes.items(), key=lambda kv: -kv[1][0])}},
            "cpu_baseline": cpu_base,
            "modes": modes,
            "aux": aux or None,
            "loss_checksum": float(np.sum(losses)),
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
    ap.add_argument("--particles", type=int, default=WORKLOAD["particles_total"], help="particles of the one swarm (all GPUs together)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-aux", action="store_true", help="skip the C4 / C5 / prediction / surface / scaling-slice measurements")
    ap.add_argument("--no-native", action="store_true", help="skip the short runs in the other arithmetic modes")
    ap.add_argument("--precision", default="emu", choices=["fp64", "emu", "emu28", "emu21", "emu15", "emu28_34", "emu34_28"],
                    help="fp64: native DMMA contractions; emu: FP64 emulated on the int8 tcgen05 tensor cores "
                         "(34 digit products); emu28 / emu21 / emu15: fewer digit products (checked a posteriori)")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
