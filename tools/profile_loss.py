"""Minimal driver for ncu: a few swarm evaluations of the bench workload (no CPU baseline, no e2e)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
import bench
from gaussian_process_b200 import engine, pso
P = int(sys.argv[1]) if len(sys.argv) > 1 else 256
reps = int(sys.argv[2]) if len(sys.argv) > 2 else 2
x, y, e = bench.make_problem(0)
pos, lo, hi = bench.make_particles(x, P, 0)
s, ex = pso.sigma_grid()
if os.environ.get("GPK_GROUPS"):
    from gaussian_process_b200 import _lib
    _lib.check(_lib.load().gpk_set_groups(int(os.environ["GPK_GROUPS"]), 1), "gpk_set_groups")
ctx = engine.LossContext(x, y, e, s)
dpos = engine.to_dev(pos); dact = torch.ones(P, dtype=torch.uint8, device="cuda")
counts, info = ctx.counts_device(dpos, dact, P)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(reps):
    counts, info = ctx.counts_device(dpos, dact, P)
e1.record()
torch.cuda.synchronize()
print("ok", int(counts.sum().item()), int(info.abs().sum().item()), f"{e0.elapsed_time(e1) / reps:.2f} ms per swarm evaluation")
