"""Minimal driver for ncu: one launch each of the covariance (a), predict (c) and surface (d) kernels at
shapes large enough to be representative (inputs/outputs larger than L2) but short under --set full."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np, torch
from gaussian_process_b200 import synthetic
from gaussian_process_b200 import engine, _lib

lib = _lib.load()
st = torch.cuda.current_stream().cuda_stream
# (a) covariance: P=16 full matrices at N=4096, d=3 (2.1 GB written)
n, d, p = 4096, 3, 16
x, y, e = synthetic.problem(n, d, 0)
pos, lo, hi = synthetic.particles(x, p, 0)
dx, de, dls = engine.to_dev(x), engine.to_dev(e), engine.to_dev(pos)
out = torch.empty((p, n, n), dtype=torch.float64, device="cuda")
for _ in range(2):
    _lib.check(lib.gpk_cov_batched(dx.data_ptr(), dx.data_ptr(), de.data_ptr(), dls.data_ptr(), p, n, n, d,
                                   out.data_ptr(), n, n * n, st), "cov")
del out
# (c) predict: one chunk of 8192 grid points against a cached N=4096 fit
fit = engine.GPFit(x, y, e, np.full(d, 0.25))
grid = engine.to_dev(np.random.default_rng(0).random((d, 16384)))
mu, sd = fit.predict_device(grid, 16384)
# (d) surface: 2^20 rows x 16 experiments, 25 % gaps
t = synthetic.surface_table(1 << 18, 16, seed=3)
dt = engine.to_dev(t).repeat(4, 1).contiguous()
r = dt.shape[0]
yy = torch.empty((r, 100), dtype=torch.float64, device="cuda"); pp = torch.empty_like(yy)
valid = torch.empty(r, dtype=torch.int32, device="cuda")
for _ in range(2):
    _lib.check(lib.gpk_prob_surface(dt.data_ptr(), r, 32, 100, yy.data_ptr(), pp.data_ptr(), valid.data_ptr(), st), "surf")
torch.cuda.synchronize()
print("ok", float(mu.sum().item()), float(sd.sum().item()), int(valid.sum().item()))
