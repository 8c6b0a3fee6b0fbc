import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np, torch, bench
from gaussian_process_b200 import engine, pso, _lib
P = 256
x, y, e = bench.make_problem(0); pos, lo, hi = bench.make_particles(x, P, 0); s, ex = pso.sigma_grid()
ctx = engine.LossContext(x, y, e, s); dpos = engine.to_dev(pos); dact = torch.ones(P, dtype=torch.uint8, device="cuda")
ref = None
for G in (1, 2, 4, 8):
    _lib.load().gpk_set_groups(G, 1)
    for _ in range(2): c, info = ctx.counts_device(dpos, dact, P)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(3): c, info = ctx.counts_device(dpos, dact, P)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    c = c.cpu().numpy()
    if ref is None: ref = c
    print(f"G={G}: {ms:.1f} ms/step  {P/ms*1e3:.1f} evals/s  counts identical to G=1: {np.array_equal(c, ref)}", flush=True)
