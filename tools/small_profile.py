"""Phase clocks of the one-CTA loss kernel (instrumented build: libgpk built with -DGPK_SMALL_PROFILE into
gpurun_out/libgpk_prof.so, selected with GPK_LIB).  N=100, d=2, 40 particles (the reference's real PSO size)."""
import ctypes as C
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from gaussian_process_b200 import engine, pso, synthetic, _lib

lib = _lib.load()
x, y, e = synthetic.problem(int(sys.argv[1]) if len(sys.argv) > 1 else 100, 2, 0)
pos, lo, hi = synthetic.particles(x, 40, 0)
sv, ex = pso.sigma_grid()
ctx = engine.LossContext(x, y, e, sv)
for _ in range(5):
    ctx.counts_view(pos)
out = (C.c_longlong * 16)()
raw = C.CDLL(_lib.LIB_PATH)
assert raw.gpk_debug_small_prof(out) == 0
t = np.array(out[:8], dtype=np.int64)
names = ["scale points + zero hist", "covariance tile (exp)", "chol_inv_regs", "dump to smem", "z = W y", "colnorm + alpha + counts",
         "prefix + store"]
d = np.diff(t)
print(f"N={x.shape[1]}: total {d.sum()} cycles")
for n, c in zip(names, d):
    print(f"  {n:28s} {c:8d} cycles  {100.0 * c / d.sum():5.1f} %")
