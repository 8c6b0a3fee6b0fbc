"""Per-kernel measurements for the (a) covariance, (c) predict and (d) surface rows of SURVEY 8 (the loss path
(b) is bench.py).  Device-resident timing with CUDA events; prints one JSON line per kernel."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np, torch
from gaussian_process_b200 import synthetic
from gaussian_process_b200 import engine, _lib

HBM = 6533.2   # GB/s, MEASURED_PEAKS.json
FP64 = 37.15   # TFLOP/s, tools/fp64_probe.cu (DMMA issue rate)


def timed(fn, reps=3, warm=1):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def cov():
    lib = _lib.load()
    n, d, p = 4096, 3, 64
    x, y, e = synthetic.problem(n, d, 0)
    pos, lo, hi = synthetic.particles(x, p, 0)
    dx, de, dls = engine.to_dev(x), engine.to_dev(e), engine.to_dev(pos)
    out = torch.empty((p, n, n), dtype=torch.float64, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    ms = timed(lambda: _lib.check(lib.gpk_cov_batched(dx.data_ptr(), dx.data_ptr(), de.data_ptr(), dls.data_ptr(), p, n, n, d,
                                                       out.data_ptr(), n, n * n, st), "cov"))
    gb = 8.0 * n * n * p / 1e9
    print(json.dumps({"kernel": "k_cov_full (gpk_cov_batched)", "shape": f"P={p} N={n} d={d}", "ms": ms, "algorithmic_GB": gb,
                      "achieved_GBps": gb / ms * 1e3, "frac_of_hbm_measured": gb / ms * 1e3 / HBM,
                      "exp_per_s": n * n * p / ms * 1e3}), flush=True)


def predict(n, d, m, label):
    x, y, e = synthetic.problem(n, d, 1)
    ls = np.full(d, 0.25)
    t0 = time.perf_counter(); fit = engine.GPFit(x, y, e, ls); torch.cuda.synchronize(); fit_s = time.perf_counter() - t0
    ms_fit = timed(lambda: engine.GPFit(x, y, e, ls), reps=2, warm=0)
    grid = engine.to_dev(np.random.default_rng(0).random((d, m)))
    ms = timed(lambda: fit.predict_device(grid, m), reps=2, warm=1)
    fl = float(n) * n * m
    print(json.dumps({"kernel": "gpk_predict", "precision": engine.get_precision(), "config": label, "shape": f"N={n} d={d} M={m}",
                      "ms": ms, "grid_pts_per_s": m / ms * 1e3, "algorithmic_TFLOPs": fl / ms * 1e3 / 1e12,
                      "frac_of_fp64_dmma_peak": fl / ms * 1e3 / 1e12 / FP64, "out_GBps": 16.0 * m / ms * 1e3 / 1e9,
                      "fit_ms_single_matrix": ms_fit}), flush=True)


def surface(r, e_):
    lib = _lib.load()
    t = synthetic.surface_table(min(r, 1 << 18), e_, seed=3)
    reps_ = -(-r // t.shape[0])
    dt = engine.to_dev(t).repeat(reps_, 1)[:r].contiguous()
    y = torch.empty((r, 100), dtype=torch.float64, device="cuda"); p = torch.empty_like(y)
    valid = torch.empty(r, dtype=torch.int32, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    ms = timed(lambda: _lib.check(lib.gpk_prob_surface(dt.data_ptr(), r, 2 * e_, 100, y.data_ptr(), p.data_ptr(), valid.data_ptr(), st), "surf"))
    nv = int(valid.sum().item())
    byts = r * 2 * e_ * 8 + nv * 1600 + r * 4
    pairs = float((torch.isfinite(dt[:, 0::2]).sum()).item())
    print(json.dumps({"kernel": "k_prob_surface", "shape": f"R={r} E={e_} (25% gaps)", "ms": ms, "rows_per_s": r / ms * 1e3,
                      "valid_rows": nv, "algorithmic_GB": byts / 1e9, "achieved_GBps": byts / ms * 1e3 / 1e9,
                      "frac_of_hbm_measured": byts / ms * 1e3 / 1e9 / HBM, "cdf_evals_per_s": pairs * 200 / ms * 1e3}), flush=True)


if __name__ == "__main__":
    which = sys.argv[1:] or ["cov", "predict", "surface"]
    if "cov" in which: cov()
    if "predict" in which:
        for mode in ("fp64", "emu"):
            engine.set_precision(mode)
            predict(1024, 2, 40000, "C2")
            predict(4096, 3, 1000000, "C3")
        engine.set_precision("auto")
    if "surface" in which:
        surface(1 << 22, 16)
