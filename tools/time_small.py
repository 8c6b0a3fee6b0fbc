import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gaussian_process_b200 import synthetic
from gaussian_process_b200 import engine, pso
for n, d in ((21, 2), (100, 2), (128, 3)):
    x, y, e = synthetic.problem(n, d, 0); pos, lo, hi = synthetic.particles(x, 40, 0); s, ex = pso.sigma_grid()
    sw = pso.SwarmLoss(x, y, e, s, ex, lo, hi)
    for _ in range(5): sw(pos)
    t = time.perf_counter()
    for _ in range(200): sw(pos)
    dt = (time.perf_counter() - t) / 200
    dpos = engine.to_dev(pos); dact = torch.ones(40, dtype=torch.uint8, device="cuda")
    torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True); e0.record()
    for _ in range(200): sw.ctx.counts_device(dpos, dact, 40)
    e1.record(); torch.cuda.synchronize()
    print(f"small path N={n} P=40: host+device {dt*1e6:.0f} us/swarm eval ({40/dt:.0f} evals/s); device only {e0.elapsed_time(e1)/200*1e3:.0f} us", flush=True)
