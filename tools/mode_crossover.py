"""Loss-evaluation time of the blocked path in both arithmetic modes over (N, P): where does the emulated mode pay off?"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np, torch
from gaussian_process_b200 import synthetic
from gaussian_process_b200 import engine, pso

s, ex = pso.sigma_grid()
for n, d in ((256, 2), (512, 2), (1024, 2), (2048, 3), (4096, 3), (8192, 2)):
    x, y, e = synthetic.problem(n, d, 0)
    for p in (40, 64, 256):
        if n >= 8192 and p > 64:
            continue
        pos, lo, hi = synthetic.particles(x, p, 0)
        row = []
        for mode in ("fp64", "emu"):
            engine.set_precision(mode)
            ctx = engine.LossContext(x, y, e, s)
            dpos = engine.to_dev(pos); dact = torch.ones(p, dtype=torch.uint8, device="cuda")
            for _ in range(2): ctx.counts_device(dpos, dact, p)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            reps = 3
            for _ in range(reps): ctx.counts_device(dpos, dact, p)
            e1.record(); torch.cuda.synchronize()
            row.append(e0.elapsed_time(e1) / reps)
        print(f"N={n:5d} d={d} P={p:3d}: fp64 {row[0]:8.2f} ms  emu {row[1]:8.2f} ms  ratio {row[0]/row[1]:.2f}", flush=True)
    engine.release_workspaces()
