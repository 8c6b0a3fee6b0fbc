"""C2 loss evaluation (N=1024, d=2, 64 particles): ms per swarm evaluation by arithmetic mode and number of particle groups."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gaussian_process_b200 import engine, pso, synthetic, _lib
lib = _lib.load()
N = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
P = int(sys.argv[2]) if len(sys.argv) > 2 else 64
x, y, e = synthetic.problem(N, 2, 0)
pos, lo, hi = synthetic.particles(x, P, 0)
s, ex = pso.sigma_grid()
dpos = engine.to_dev(pos); dact = torch.ones(P, dtype=torch.uint8, device="cuda")
for mode in ("fp64", "emu"):
    for graph in ("1", "0"):
        os.environ["GPK_GRAPH"] = graph
        for G in (1, 2, 4, 8):
            _lib.check(lib.gpk_set_groups(G, 1), "groups")
            engine.set_precision(mode)
            ctx = engine.LossContext(x, y, e, s)
            for _ in range(3): ctx.counts_device(dpos, dact, P)
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(20): c, i = ctx.counts_device(dpos, dact, P)
            e1.record(); torch.cuda.synchronize()
            print(f"N={N} P={P} {mode} graph={graph} groups={G}: {e0.elapsed_time(e1) / 20:.3f} ms per swarm, checksum {int(c.sum())}", flush=True)
            del ctx
