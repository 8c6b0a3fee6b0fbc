"""Generates tests/golden/*.npz from the UNMODIFIED Python reference -- TEST INFRASTRUCTURE.

Run in the build container only (needs /root/reference):  python -m oracle.make_golden
Every array stored here is an output of the reference's own functions
(GP_func.kernel_func / GP, find_len_scales.evaluate_loss / len_scale_opt,
calc_prob_surf.generate_prob_surf, convex_hull.fill_convex_hull, read_in.read_data) on the
inputs stored beside it.  Seeds: np.random.seed(s) + the seedable centre-LHS stub
(SURVEY F5/F6)."""
import io
import os
import sys
import tempfile
import contextlib

import numpy as np
import pandas as pd

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref_shim, gp_oracle as orc  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def shipped_experiments(ref):
    """(name, x (d,N), y, e) for the 4 shipped experiments (config C1)."""
    out = []
    rd = ref.read_in.read_data
    for path in ("Test_file1.txt", "Test_folder/Test_file2.txt", "Test_folder/Test_file3.txt"):
        xs, pairs, _ = rd(os.path.join(ref_shim.REFERENCE_DIR, path), None, [0.01, 0.01])
        for k, (x, (y, e)) in enumerate(zip(xs, pairs), start=1):
            stem = os.path.splitext(os.path.basename(path))[0]
            out.append((f"{stem}_exp{k}", np.ascontiguousarray(x, dtype=np.float64),
                        np.ascontiguousarray(y, dtype=np.float64), np.ascontiguousarray(e, dtype=np.float64)))
    return out


def gen_kernel(ref):
    d = {}
    cases = []
    rng = np.random.default_rng(11)
    for i, (dd, n1, n2) in enumerate([(2, 21, 21), (3, 200, 77), (4, 64, 130), (1, 50, 50)]):
        x1 = rng.random((dd, n1)) * 3 - 1
        x2 = x1 if n1 == n2 else rng.random((dd, n2)) * 3 - 1
        ls = rng.uniform(0.05, 2.0, dd)
        d[f"k{i}_x1"], d[f"k{i}_x2"], d[f"k{i}_ls"] = x1, x2, ls
        d[f"k{i}_K"] = ref.GP_func.kernel_func(x1, x2, ls)
        cases.append(i)
    d["n_cases"] = np.array(len(cases))
    np.savez_compressed(os.path.join(OUT, "kernel_cases.npz"), **d)


def loss_case(ref, x, y, e, pos, lower, upper):
    s, ex = orc.sigma_grid()
    losses = np.array([ref.find_len_scales.evaluate_loss(p, x, y, e, s, ex, lower, upper) for p in pos])
    n = x.shape[1]
    counts = np.zeros((len(pos), len(s)), dtype=np.int32)
    resid = np.zeros((len(pos), n))
    sd = np.zeros((len(pos), n))
    margin = np.full(len(pos), np.inf)
    for k, p in enumerate(pos):
        if orc.out_of_bounds(p, lower, upper):
            continue
        mu, sig = ref.GP_func.GP(x, y, e, x, p, batch_size=n)
        scaled = sig[:, None] * s[None, :]
        pulls = (mu[:, None] - y[:, None]) / np.maximum(scaled, 1e-12)
        counts[k] = (np.abs(pulls) <= 1).sum(axis=0)
        resid[k], sd[k] = mu - y, sig
        margin[k] = np.min(np.abs(np.abs(pulls) - 1))
    return dict(x=x, y=y, e=e, pos=pos, lower=lower, upper=upper, loss=losses, counts=counts,
                resid=resid, sd=sd, margin=margin)


def gen_loss(ref):
    d = {}
    names = []
    for name, x, y, e in shipped_experiments(ref):
        lower, upper = orc.pso_bounds(x)
        pos = orc.center_lhs(np.column_stack((lower, upper)), 12, np.random.RandomState(3))
        pos[0] = lower            # on-bound particles short-circuit (find_len_scales.py:156)
        pos[1, 0] = upper[0]
        c = loss_case(ref, x, y, e, pos, lower, upper)
        names.append(name)
        for k, v in c.items():
            d[f"{name}__{k}"] = v
    for name, n, dd, p, seed in [("syn100_d2", 100, 2, 16, 0), ("syn100_d3", 100, 3, 16, 1),
                                 ("syn128_d4", 128, 4, 8, 2), ("syn257_d3", 257, 3, 6, 3),
                                 ("syn640_d2", 640, 2, 4, 4), ("syn1024_d2", 1024, 2, 3, 5)]:
        x, y, e = orc.synthetic_problem(n, dd, seed)
        pos, lower, upper = orc.synthetic_particles(x, p, seed)
        c = loss_case(ref, x, y, e, pos, lower, upper)
        names.append(name)
        for k, v in c.items():
            d[f"{name}__{k}"] = v
        print(name, "min margin", c["margin"].min(), flush=True)
    d["names"] = np.array(names)
    np.savez_compressed(os.path.join(OUT, "loss_cases.npz"), **d)


def gen_gp(ref):
    """GP() fit + grid predict (config C1 grid from fill_convex_hull, and synthetic)."""
    d = {}
    names = []
    exps = shipped_experiments(ref)
    for name, x, y, e in (exps[0], exps[3]):
        grid = ref.convex_hull.fill_convex_hull(x.T, [0.01, 0.01])
        ls = np.array([0.35, 0.6]) if name.startswith("Test_file1") else np.array([0.21, 0.9])
        mu, sd = ref.GP_func.GP(x, y, e, grid.T, ls)
        for k, v in dict(x=x, y=y, e=e, grid=grid, ls=ls, mu=mu, sd=sd).items():
            d[f"{name}__{k}"] = v
        names.append(name)
    for name, n, dd, m, seed in [("syn300_d3", 300, 3, 700, 7), ("syn1024_d2", 1024, 2, 1500, 8)]:
        x, y, e = orc.synthetic_problem(n, dd, seed)
        rng = np.random.default_rng(seed + 100)
        grid = rng.random((m, dd))
        ls = rng.uniform(0.15, 0.6, dd)
        mu, sd = ref.GP_func.GP(x, y, e, grid.T, ls, batch_size=400)
        for k, v in dict(x=x, y=y, e=e, grid=grid, ls=ls, mu=mu, sd=sd).items():
            d[f"{name}__{k}"] = v
        names.append(name)
    d["names"] = np.array(names)
    np.savez_compressed(os.path.join(OUT, "gp_cases.npz"), **d)


def gen_pso(ref):
    """Seeded full PSO traces of the reference's len_scale_opt (40 particles, 500 iterations)."""
    fls = ref.find_len_scales
    exps = shipped_experiments(ref)
    x100, y100, e100 = orc.synthetic_problem(100, 2, 21)
    cases = [("Test_file1_exp1",) + exps[0][1:], ("Test_file3_exp2",) + exps[3][1:], ("syn100_d2", x100, y100, e100)]
    for seed, (name, x, y, e) in enumerate(cases):
        record = []
        ref_shim.set_lhs_seed(seed)
        np.random.seed(seed)
        buf = io.StringIO()
        with ref_shim.serial_pool(fls, record), contextlib.redirect_stdout(buf):
            best = fls.len_scale_opt(x, y, e, True)
        pos = np.array([r[0] for r in record])
        sc = np.array([r[1] for r in record])
        print(name, "evaluations", pos.shape, "best", best, flush=True)
        np.savez_compressed(os.path.join(OUT, f"pso_trace_{name}.npz"), x=x, y=y, e=e, seed=np.array(seed),
                            positions=pos, scores=sc, best=best, stdout=np.array(buf.getvalue()))


def gen_surface(ref):
    d = {}
    merged = pd.read_csv(os.path.join(ref_shim.REFERENCE_DIR, "output_file"), float_precision="round_trip")
    sub = pd.concat([merged.iloc[:150], merged.iloc[5000:5150], merged.iloc[-100:]]).reset_index(drop=True)
    tables = {"shipped": (sub, 2)}
    t16 = orc.synthetic_surface_table(120, 16, seed=5)
    coords = np.random.default_rng(6).random((120, 4))
    cols = [f"dim{i+1}" for i in range(4)] + [c for i in range(16) for c in (f"f_exp{i}", f"f_exp{i}_unc")]
    tables["syn_e16"] = (pd.DataFrame(np.concatenate([coords, t16], axis=1), columns=cols), 4)
    t3 = orc.synthetic_surface_table(80, 3, seed=9, p_gap=0.6)     # many rows with 0 or 1 experiment
    t3[5, 0] = np.nan                                               # odd finite count -> row dropped (:71)
    tables["syn_e3_gappy"] = (pd.DataFrame(np.concatenate([coords[:80, :1], t3], axis=1),
                                           columns=["dim1"] + [f"c{i}" for i in range(6)]), 1)
    cwd = os.getcwd()
    for name, (df, nd) in tables.items():
        with tempfile.TemporaryDirectory() as tmp:
            os.chdir(tmp)
            try:
                with contextlib.redirect_stdout(io.StringIO()):
                    ref.calc_prob_surf.generate_prob_surf(df, nd)
                out = pd.read_csv("prob_surf.txt", float_precision="round_trip")
            finally:
                os.chdir(cwd)
        d[f"{name}__table"] = df.to_numpy(dtype=np.float64)
        d[f"{name}__ndims"] = np.array(nd)
        o = out.to_numpy(dtype=np.float64)
        d[f"{name}__out_yp"] = o[:, nd:]                 # (100*Rv, 2): quantity, prob
        d[f"{name}__out_coords"] = o[::100, :nd]         # coords of each emitted row block
        print(name, df.shape, "->", out.shape, flush=True)
    d["names"] = np.array(list(tables))
    np.savez_compressed(os.path.join(OUT, "surface_cases.npz"), **d)


def gen_large(ref):
    """Reference outputs AT THE SIZES bench.py REPORTS: the C3 bench problem itself
    (N=4096, d=3, seed 0; particles 0/85/170/255 of the 256-particle bench swarm + one on-bound particle),
    a C4-sized GP() fit + 2000 grid points (N=8192, d=2), and an adversarial emulated-mode case (N=1024, e over six
    decades with isolated points).  ~10 min of CPU in the build container."""
    import time
    d = {}
    t0 = time.time()
    x, y, e = orc.synthetic_problem(4096, 3, 0)
    pos_all, lower, upper = orc.synthetic_particles(x, 256, 0)
    pos = np.concatenate([pos_all[[0, 85, 170, 255]], pos_all[[7]]])
    pos[4, 1] = upper[1]                      # on-bound -> 1e13 without a fit (find_len_scales.py:156)
    c = loss_case(ref, x, y, e, pos, lower, upper)
    c["swarm_index"] = np.array([0, 85, 170, 255, -1])
    for k, v in c.items():
        if k not in ("x", "y", "e"):          # regenerated from the seed by the test (checked by a digest)
            d[f"c3_n4096__{k}"] = v
    d["c3_n4096__xye_sum"] = np.array([x.sum(), y.sum(), e.sum()])
    print("c3_n4096 min margin", c["margin"][:4].min(), "loss", c["loss"], f"{time.time()-t0:.0f}s", flush=True)

    x, y, e = orc.adversarial_problem(1024, 2, 0)
    lower, upper = orc.pso_bounds(x)
    pos = orc.center_lhs(np.column_stack((lower, upper)), 4, np.random.RandomState(1))
    c = loss_case(ref, x, y, e, pos, lower, upper)
    for k, v in c.items():
        d[f"adv_n1024__{k}"] = v
    print("adv_n1024 min margin", c["margin"].min(), "loss", c["loss"], f"{time.time()-t0:.0f}s", flush=True)
    np.savez_compressed(os.path.join(OUT, "loss_cases_large.npz"), **d)

    g = {}
    x, y, e = orc.synthetic_problem(8192, 2, 3)
    rng = np.random.default_rng(103)
    grid = rng.random((2000, 2))
    ls = np.array([0.11, 0.17])
    mu, sd = ref.GP_func.GP(x, y, e, grid.T, ls, batch_size=1000)
    for k, v in dict(grid=grid, ls=ls, mu=mu, sd=sd).items():
        g[f"c4_n8192__{k}"] = v
    g["c4_n8192__xye_sum"] = np.array([x.sum(), y.sum(), e.sum()])
    print("c4_n8192 GP done", f"{time.time()-t0:.0f}s", "sd range", sd.min(), sd.max(), flush=True)
    x, y, e = orc.synthetic_problem(4096, 3, 0)
    grid = np.random.default_rng(104).random((1500, 3))
    ls = np.array([0.2, 0.3, 0.25])
    mu, sd = ref.GP_func.GP(x, y, e, grid.T, ls)
    for k, v in dict(grid=grid, ls=ls, mu=mu, sd=sd).items():
        g[f"c3_n4096__{k}"] = v
    print("c3_n4096 GP done", f"{time.time()-t0:.0f}s", flush=True)
    np.savez_compressed(os.path.join(OUT, "gp_cases_large.npz"), **g)


def gen_adv(ref):
    """Milder adversarial cases for the emulated arithmetic (same generator as adv_n1024 in gen_large): e log-uniform over
    1e-3 .. 1 and 1e-4 .. 1 with isolated points carrying the smallest e: cond(K) ~ 1e9 and ~ 1e11, where FP64 still delivers
    digits and the a-priori scales of the digit split are already loose by 10 and 13 bits."""
    d = {}
    for name, e_lo in (("adv3_n1024", 1e-3), ("adv4_n1024", 1e-4)):
        x, y, e = orc.adversarial_problem(1024, 2, 0, e_lo=e_lo)
        lower, upper = orc.pso_bounds(x)
        pos = orc.center_lhs(np.column_stack((lower, upper)), 4, np.random.RandomState(1))
        c = loss_case(ref, x, y, e, pos, lower, upper)
        for k, v in c.items():
            d[f"{name}__{k}"] = v
        print(name, "min margin", c["margin"].min(), "loss", c["loss"], flush=True)
    np.savez_compressed(os.path.join(OUT, "loss_cases_adv.npz"), **d)


def main():
    os.makedirs(OUT, exist_ok=True)
    ref = ref_shim.load()
    which = sys.argv[1:] or ["kernel", "loss", "gp", "surface", "pso"]
    for w in which:
        print("==", w, flush=True)
        globals()["gen_" + w](ref)


if __name__ == "__main__":
    main()
