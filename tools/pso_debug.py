"""Runs the seeded drop-in PSO for the golden trace cases on the GPU and stores the traces for offline diffing."""
import os, sys, io, contextlib
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-process_b200", "dropin"))
import numpy as np
from oracle import ref_shim
ref_shim.install_smt_stub()
import find_len_scales as fls
from gaussian_process_b200 import engine, pso
from oracle import gp_oracle as orc
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
for name in ("Test_file1_exp1", "syn100_d2"):
    g = np.load(os.path.join(ROOT, "tests", "golden", f"pso_trace_{name}.npz"))
    seed = int(g["seed"]); np.random.seed(seed); trace = []
    with contextlib.redirect_stdout(io.StringIO()):
        best = fls.len_scale_opt(g["x"], g["y"], g["e"], True, lhs_random_state=seed, trace=trace)
    pos = np.array([t[0] for t in trace]); sc = np.array([t[1] for t in trace])
    n = min(len(pos), len(g["positions"]))
    bad = [i for i in range(n) if not np.array_equal(pos[i], g["positions"][i])]
    print(name, "evals", len(pos), "ref", len(g["positions"]), "first position divergence", bad[:1])
    k = bad[0] - 1 if bad else n - 1
    # score differences before divergence
    dif = np.argwhere(sc[:k + 1] != g["scores"][:k + 1])
    print("  score mismatches before divergence:", len(dif), "of", (k + 1) * sc.shape[1])
    for (i, p) in dif[:10]:
        print("   eval", i, "particle", p, "gpu", repr(sc[i, p]), "ref", repr(g["scores"][i, p]), "pos", g["positions"][i, p])
    # detail at the diverging evaluation: counts from GPU vs oracle
    lo, hi = orc.pso_bounds(g["x"]); s, ex = orc.sigma_grid()
    ctx = engine.LossContext(g["x"], g["y"], g["e"], s)
    for (i, p) in dif[:3]:
        q = g["positions"][i, p]
        counts, info, resid, sd = ctx.counts(q[None, :], want_detail=True)
        np.savez(os.path.join(ROOT, "gpurun_out", f"dbg_{name}_{i}_{p}.npz"), q=q, counts=counts, resid=resid, sd=sd)
    np.savez(os.path.join(ROOT, "gpurun_out", f"trace_{name}.npz"), pos=pos, sc=sc)
