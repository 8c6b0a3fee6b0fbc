"""Emulated-FP64 (int8 tcgen05) mode vs native FP64 mode vs the oracle on the blocked loss path.
usage: python tools/oz_parity.py [big]"""
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from gaussian_process_b200 import engine  # noqa: E402
from oracle import gp_oracle as orc  # noqa: E402


def run(n, d, p, seed, mode):
    engine.set_precision(mode)  # 0 fp64, 1 emu, 2 emu28
    x, y, e = orc.synthetic_problem(n, d, seed)
    pos, lo, hi = orc.synthetic_particles(x, p, seed)
    s, ex = orc.sigma_grid()
    ctx = engine.LossContext(x, y, e, s)
    torch.cuda.synchronize()
    t0 = time.time()
    counts, info, resid, sd = ctx.counts(pos, want_detail=True)
    torch.cuda.synchronize()
    return dict(x=x, y=y, e=e, pos=pos, counts=counts, info=info, resid=resid, sd=sd, t=time.time() - t0)


def main():
    big = len(sys.argv) > 1 and sys.argv[1] == "big"
    cases = [(300, 3, 4, 0), (1024, 2, 8, 1), (2048, 3, 8, 2)]
    ok = True
    for n, d, p, seed in cases:
        a = run(n, d, p, seed, 0)
        b = run(n, d, p, seed, 1)
        rel_r = np.max(np.abs(a["resid"] - b["resid"]) / np.maximum(np.abs(a["resid"]), 1e-300))
        rel_s = np.max(np.abs(a["sd"] - b["sd"]) / a["sd"])
        dc = int(np.sum(a["counts"] != b["counts"]))
        line = f"N={n} d={d} P={p}: emu vs fp64: max rel resid {rel_r:.2e}, max rel sd {rel_s:.2e}, differing counts {dc}, info {b['info'].tolist()}"
        if n <= 1024:
            mu_r, sd_r = orc.gp(a["x"], a["y"], a["e"], a["x"], a["pos"][0])
            r_ref = mu_r - a["y"]
            e0 = np.max(np.abs(a["resid"][0] - r_ref) / np.abs(r_ref)), np.max(np.abs(a["sd"][0] - sd_r) / sd_r)
            e1 = np.max(np.abs(b["resid"][0] - r_ref) / np.abs(r_ref)), np.max(np.abs(b["sd"][0] - sd_r) / sd_r)
            line += f" | vs oracle (resid, sd): fp64 {e0[0]:.2e} {e0[1]:.2e}; emu {e1[0]:.2e} {e1[1]:.2e}"
        print(line, flush=True)
        ok &= rel_r < 1e-8 and rel_s < 1e-8
    if big:
        for mode in (0, 1):
            r = run(4096, 3, 64, 3, mode)
            r = run(4096, 3, 64, 3, mode)
            print(f"N=4096 P=64 mode {mode}: {r['t'] * 1e3:.1f} ms -> {64 / r['t']:.0f} evals/s", flush=True)
    print("OZ PARITY OK" if ok else "OZ PARITY FAILED")
    return 0 if ok else 1


if __name__ == "__main__":
    sys.exit(main())
