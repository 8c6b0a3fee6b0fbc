"""Per-kernel-class device time of one C2-sized swarm evaluation (single stream, events after every launch)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from gaussian_process_b200 import engine, pso, synthetic, _lib
N = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
P = int(sys.argv[2]) if len(sys.argv) > 2 else 64
os.environ["GPK_GRAPH"] = "0"
x, y, e = synthetic.problem(N, 2, 0)
pos, lo, hi = synthetic.particles(x, P, 0)
s, ex = pso.sigma_grid()
dpos = engine.to_dev(pos); dact = torch.ones(P, dtype=torch.uint8, device="cuda")
for mode in ("emu", "fp64"):
    engine.set_precision(mode)
    ctx = engine.LossContext(x, y, e, s)
    for _ in range(3): ctx.counts_device(dpos, dact, P)
    torch.cuda.synchronize()
    st = torch.cuda.current_stream()
    _lib.timing_begin(st.cuda_stream)
    ctx.counts_device(dpos, dact, P)
    torch.cuda.synchronize()
    classes = _lib.timing_end()
    tot = sum(v[0] for v in classes.values())
    print(f"{mode} N={N} P={P}: {tot:.3f} ms in {sum(v[1] for v in classes.values())} launches (serial, one group)")
    for k, v in sorted(classes.items(), key=lambda kv: -kv[1][0]):
        print(f"   {k:18s} {v[0]:.3f} ms / {v[1]} launches = {1e3 * v[0] / v[1]:.1f} us each")
