"""Deterministic stand-ins for the numerics and the input of GP_fit.create_GP, shared by the golden generator
(tests/golden/make_gp_fit_golden.py, which runs the reference's GP_fit with them) and tests/test_gp_fit_dropin.py
(which runs the drop-in with them)."""
import numpy as np

# mode -> (write_ind, group_exps, out_file_name)
MODES = {"combined": (False, False, "out/all_results.txt"), "individual": (True, False, "out_ind/x.txt"),
         "grouped": (True, True, "out_grp/x.txt")}


def _data_list():
    rng = np.random.default_rng(5)
    out = []
    for f, n_exp in (("data/alpha.txt", 2), ("data/beta.txt", 1), ("data/gamma.txt", 3)):
        xs, pairs = [], []
        for k in range(n_exp):
            n = 0 if (f.endswith("gamma.txt") and k == 1) else int(rng.integers(4, 9))     # one empty experiment
            x = np.stack([np.around(rng.integers(0, 8, n) * 0.25, 2), rng.integers(-3, 4, n) * 1.0])
            xs.append(x)
            pairs.append((rng.normal(size=n), rng.random(n) + 0.1))
        out.append((f, xs, pairs, ["Q2", "W", "val", "err"]))
    return out


def len_scale_opt(x_known, y_known, e_known, PSO_progress):
    return np.full(x_known.shape[0], 0.5) + 0.01 * x_known.shape[1]


def fill_convex_hull(points, step):
    lo, hi = points.min(axis=0), points.max(axis=0)
    axes = [np.around(np.arange(round(l / s), round(h / s) + 1) * s, 6) for l, h, s in zip(lo, hi, step)]
    grid = np.stack(np.meshgrid(*axes, indexing="ij"), -1).reshape(-1, len(step))
    return grid


def GP(x_known, y_known, e_known, x_fit, lengths, batch_size=10000):
    w = np.exp(-0.5 * (((x_fit[:, None, :] - x_known[:, :, None]) / lengths[:, None, None]) ** 2).sum(0))
    mu = (w * y_known[:, None]).sum(0) / (w.sum(0) + 1e-3)
    sd = 1.0 / (1.0 + w.sum(0))
    return mu, sd


def install(module, mode):
    """Point `module` (the reference's GP_fit or the drop-in) at the stand-ins."""
    write_ind, group_exps, out_name = MODES[mode]
    module.len_scale_opt = len_scale_opt
    module.fill_convex_hull = fill_convex_hull
    module.GP = GP
    module.read_yaml = lambda: ([0.25, 1.0], False, out_name, ["Q2", "W", "val", "err"], _data_list(), write_ind, group_exps)
