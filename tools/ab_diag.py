"""A/B of the diagonal-block kernel variants (GPK_DIAG_VARIANT=0: one column per step with correctly rounded pivot roots,
1: panels of 8 columns in one warp, 2: one column per step with the short reciprocal-square-root chain).  Run once per
variant (the switch is read once per process); prints a digest of the results (0 and 1 must be identical; 2 differs in the
last bits: the arrays are saved for comparison) and timings."""
import hashlib
import json
import os
import sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from gaussian_process_b200 import engine, pso, synthetic, _lib

s, ex = pso.sigma_grid()
out = {"variant": os.environ.get("GPK_DIAG_VARIANT", "0")}
h = hashlib.sha256()
keep = {}
for n, d, p, mode in ((640, 2, 6, "fp64"), (1024, 2, 64, "emu"), (300, 3, 5, "fp64")):
    x, y, e = synthetic.problem(n, d, 3)
    pos, lo, hi = synthetic.particles(x, p, 3)
    engine.set_precision(mode)
    c, i, r, sd = engine.LossContext(x, y, e, s).counts(pos, want_detail=True)
    assert not i.any()
    h.update(c.tobytes()); h.update(r.tobytes()); h.update(sd.tobytes())
    keep[f"c_{n}"], keep[f"r_{n}"], keep[f"sd_{n}"] = c, r, sd
# a non-positive-definite block must be reported at the same pivot
k = np.eye(256)[None].repeat(2, axis=0)
k[1, 200, 200] = -1.0
_, info = engine.potrf_batched(k)
out["info"] = info.tolist()
out["digest"] = h.hexdigest()
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
np.savez_compressed(os.path.join(ROOT, "gpurun_out", f"ab_diag_v{out['variant']}.npz"), **keep)
# timing: C2 swarm evaluation and the diagonal kernel alone
x, y, e = synthetic.problem(1024, 2, 0)
pos, lo, hi = synthetic.particles(x, 64, 0)
engine.set_precision("emu")
ctx = engine.LossContext(x, y, e, s)
dpos = engine.to_dev(pos)
dact = torch.ones(64, dtype=torch.uint8, device="cuda")
for _ in range(3):
    ctx.counts_device(dpos, dact, 64)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    ctx.counts_device(dpos, dact, 64)
e1.record()
torch.cuda.synchronize()
out["c2_ms_per_swarm"] = e0.elapsed_time(e1) / 20
os.environ["GPK_GRAPH"] = "0"
st = torch.cuda.current_stream()
_lib.timing_begin(st.cuda_stream)
ctx.counts_device(dpos, dact, 64)
torch.cuda.synchronize()
cls = _lib.timing_end()
out["c2_class_ms"] = {k: round(v[0], 4) for k, v in sorted(cls.items(), key=lambda kv: -kv[1][0])[:8]}
out["c2_diag_us_per_launch"] = 1e3 * cls["k_diag_block"][0] / cls["k_diag_block"][1]
print(json.dumps(out))
