"""Digit-product counts per phase (include/gpk.h GPK_PREC_*): on the C3 bench swarm (N=4096, d=3), for each emulated mode the
unchecked throughput of the whole 256-particle swarm, the a-posteriori probe over all particles, the errors of mu - y and sd
of 16 particles against the native FP64 result and the particles whose counts differ.
    python tools/mixed_modes.py > profiles/r02_mixed_modes.json"""
import json
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from gaussian_process_b200 import engine, pso, synthetic

N, d, P = 4096, 3, 256
x, y, e = synthetic.problem(N, d, 0)
pos, lo, hi = synthetic.particles(x, P, 0)
s, ex = pso.sigma_grid()
ctx = engine.LossContext(x, y, e, s)
dpos = engine.to_dev(pos)
dact = torch.ones(P, dtype=torch.uint8, device="cuda")


def timed(fn, steps=3):
    fn(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


engine.set_check_tol(0.0)
engine.set_precision("fp64")
c64, _ = ctx.counts_device(dpos, dact, P)
c64 = c64.clone()
_c, _i, r64, s64 = ctx.counts_device(dpos[:16], dact[:16], 16, want_detail=True)
rmax = float(r64.abs().max())
out = {"workload": f"C3 bench swarm: N={N}, d={d}, {P} particles; unchecked (check_tol = 0)"}
for mode in ("emu", "emu28_34", "emu34_28", "emu28"):
    engine.set_precision(mode)
    ms = timed(lambda: ctx.counts_device(dpos, dact, P))
    cm, _ = ctx.counts_device(dpos, dact, P, want_probe=True)
    probe = ctx.last_probe.cpu().numpy()
    cm = cm.clone()
    _c, _i, rd, sd = ctx.counts_device(dpos[:16], dact[:16], 16, want_detail=True)
    out[mode] = {"evals_per_s": P / (ms * 1e-3), "probe_max": float(probe.max()), "probe_median": float(np.median(probe)),
                 "particles_over_1e-10": int((probe > 1e-10).sum()),
                 "max_rel_err_resid_vs_fp64": float((rd - r64).abs().max()) / rmax,
                 "max_rel_err_sd_vs_fp64": float((sd / s64 - 1).abs().max()),
                 "particles_with_count_difference": int((cm != c64).any(dim=1).sum().item())}
engine.set_precision("auto")
engine.set_check_tol(1e-10)
print(json.dumps(out, indent=1))
