"""The whole workflow of main.py (GP fit of every experiment -> merged table -> probability surface -> files) on the
five drop-in modules, checked piece by piece against the oracle on the shipped Test_file1 / Test_file3 points
(tests/golden/pso_trace_*.npz hold them): lattice = point-at-a-time restatement, fitted columns = oracle GP at the
length scales the GPU PSO returned, merged table = pandas chain, files = pandas to_csv of the same frames, surface =
oracle row loop.  Run on the B200 box: pytest -m gpu."""
import importlib
import os
import sys
import types
from functools import reduce

import numpy as np
import pandas as pd
import pytest

from conftest import ROOT, load_golden

pytestmark = pytest.mark.gpu


def _dropin(name):
    d = os.path.join(ROOT, "gaussian-process_b200", "dropin")
    if d not in sys.path:
        sys.path.insert(0, d)
    sys.modules.pop(name, None)
    return importlib.import_module(name)


def test_main_workflow_end_to_end(tmp_path, monkeypatch):
    from oracle import gp_oracle as orc
    from gaussian_process_b200 import lattice_seq
    g1, g3 = load_golden("pso_trace_Test_file1_exp1.npz"), load_golden("pso_trace_Test_file3_exp2.npz")
    resolution = [0.05, 0.04]
    data_list = [("data/Test_file1.txt", [g1["x"]], [(g1["y"], g1["e"])], ["energy", "angle", "pol", "err"]),
                 ("data/Test_file3.txt", [g3["x"], g3["x"][:, :0]], [(g3["y"], g3["e"]), (g3["y"][:0], g3["e"][:0])],
                  ["energy", "angle", "pol", "err"])]
    read_in = types.ModuleType("read_in")
    read_in.read_yaml = lambda: (resolution, False, "results/GP_results.txt", ["energy", "angle", "pol", "err"], data_list,
                                 False, False)
    monkeypatch.setitem(sys.modules, "read_in", read_in)
    for name in ("convex_hull", "find_len_scales", "GP_func", "calc_prob_surf"):
        _dropin(name)
    gp_fit = _dropin("GP_fit")
    found = []
    inner = gp_fit.len_scale_opt

    def recording(x, y, e, progress):
        np.random.seed(0)
        ls = inner(x, y, e, progress, lhs_random_state=0)
        found.append(ls)
        return ls

    monkeypatch.setattr(gp_fit, "len_scale_opt", recording)
    monkeypatch.chdir(tmp_path)
    df, nd = gp_fit.create_GP()
    assert nd == 2 and len(found) == 2                       # the empty experiment is skipped (GP_fit.py:26-28)

    # per experiment: lattice and fitted columns
    frames = []
    for (x, y, e), ls, cols in (((g1["x"], g1["y"], g1["e"]), found[0], ("Test_file1", "Test_file1_unc")),
                                ((g3["x"], g3["y"], g3["e"]), found[1], ("Test_file3_exp1", "Test_file3_unc1"))):
        lat = lattice_seq.fill_convex_hull(x.T.copy(), list(resolution))
        mu, sd = orc.gp(x, y, e, lat.T, ls)
        f = pd.DataFrame(lat, columns=["energy", "angle"])
        f[cols[0]], f[cols[1]] = mu, sd
        frames.append(f)
    want = reduce(lambda l, r: pd.merge(l, r, on=["energy", "angle"], how="outer"), frames).fillna(float("inf"))
    assert list(df.columns) == list(want.columns) and len(df) == len(want)
    np.testing.assert_array_equal(df[["energy", "angle"]].to_numpy(), want[["energy", "angle"]].to_numpy())
    for c in want.columns[2:]:
        a, b = df[c].to_numpy(), want[c].to_numpy()
        np.testing.assert_array_equal(np.isinf(a), np.isinf(b))
        m = np.isfinite(b)
        np.testing.assert_allclose(a[m], b[m], rtol=1e-9, atol=1e-12)        # north_star tolerance on predictions

    # the file is what pandas writes for the returned table
    df.to_csv(tmp_path / "ref.csv", index=False)
    assert open(tmp_path / "results" / "GP_results.txt", "rb").read() == open(tmp_path / "ref.csv", "rb").read()

    # probability surface of the merged table (main.py:88-90)
    with open(tmp_path / "options.yaml", "w") as fh:
        fh.write("out_file_name: results/prob.txt\n")
    sys.modules["calc_prob_surf"].generate_prob_surf(df, nd)
    surf = pd.read_csv(tmp_path / "results" / "prob.txt")
    idx, ys, ps = orc.prob_surface_rows(df.to_numpy()[:, nd:])
    assert len(surf) == len(idx) * 100 and list(surf.columns) == ["energy", "angle", "quantity", "prob"]
    np.testing.assert_array_equal(surf[["energy", "angle"]].to_numpy(), np.repeat(df.to_numpy()[idx, :nd], 100, axis=0))
    np.testing.assert_allclose(surf["quantity"].to_numpy(), ys.reshape(-1), rtol=1e-12, atol=1e-300)
    np.testing.assert_allclose(surf["prob"].to_numpy(), ps.reshape(-1), rtol=1e-9, atol=1e-15)


def test_sharded_predict_and_surface_single_rank(tmp_path):
    """The grid-tile / surface-row sharding entry points on one rank (no process group) reduce to the plain calls."""
    from oracle import gp_oracle as orc
    from gaussian_process_b200 import engine, sharding
    x, y, e = orc.synthetic_problem(200, 2, 2)
    grid = np.random.default_rng(0).random((2, 333))
    ls = np.array([0.3, 0.5])
    fit = engine.GPFit(x, y, e, ls)
    mu, sd = sharding.sharded_predict(fit.predict, grid)
    mu_r, sd_r = orc.gp(x, y, e, grid, ls)
    np.testing.assert_allclose(mu, mu_r, rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(sd, sd_r, rtol=1e-9)
    t = orc.synthetic_surface_table(500, 4, seed=1)
    off, total, idx, yy, pp = sharding.sharded_prob_surface(engine.prob_surface, t)
    idx_r, yy_r, pp_r = orc.prob_surface_rows(t)
    assert off == 0 and total == len(idx_r) and np.array_equal(idx, idx_r) and np.array_equal(yy, yy_r)
    np.testing.assert_allclose(pp, pp_r, rtol=1e-9, atol=1e-15)
