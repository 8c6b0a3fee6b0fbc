"""The oracle restatement against the golden vectors produced by the unmodified reference
(oracle/make_golden.py).  CPU only."""
import io
import contextlib

import numpy as np
import pytest

from conftest import load_golden, case
from oracle import gp_oracle as orc, ref_shim

RTOL = 1e-10     # another host CPU may pick other OpenBLAS kernels; in the generating container it is bit-exact


def test_kernel_cases():
    g = load_golden("kernel_cases.npz")
    for i in range(int(g["n_cases"])):
        K = orc.rbf_kernel(g[f"k{i}_x1"], g[f"k{i}_x2"], g[f"k{i}_ls"])
        np.testing.assert_allclose(K, g[f"k{i}_K"], rtol=1e-12, atol=1e-14)


def test_loss_cases():
    g = load_golden("loss_cases.npz")
    s, ex = orc.sigma_grid()
    for name in g["names"]:
        c = case(g, str(name))
        n = c["x"].shape[1]
        for k, p in enumerate(c["pos"]):
            loss = orc.evaluate_loss(p, c["x"], c["y"], c["e"], s, ex, c["lower"], c["upper"])
            if orc.out_of_bounds(p, c["lower"], c["upper"]):
                assert loss == 1e13 == c["loss"][k]
                continue
            counts, resid, sd, _ = orc.calibration_counts(p, c["x"], c["y"], c["e"], s)
            assert c["margin"][k] > 1e-9          # golden cases sit far from a count flip
            np.testing.assert_array_equal(counts, c["counts"][k])
            np.testing.assert_allclose(sd, c["sd"][k], rtol=1e-9)
            # the count route and the direct route give the same score
            assert orc.finish_loss(counts, n, p, s, ex, c["lower"], c["upper"]) == loss
            assert loss == pytest.approx(c["loss"][k], rel=1e-12, abs=1e-15)


def test_gp_cases():
    g = load_golden("gp_cases.npz")
    for name in g["names"]:
        c = case(g, str(name))
        mu, sd = orc.gp(c["x"], c["y"], c["e"], c["grid"].T, c["ls"])
        np.testing.assert_allclose(mu, c["mu"], rtol=RTOL, atol=1e-12)
        np.testing.assert_allclose(sd, c["sd"], rtol=RTOL, atol=1e-12)


def test_surface_cases():
    g = load_golden("surface_cases.npz")
    for name in g["names"]:
        c = case(g, str(name))
        nd = int(c["ndims"])
        idx, y, p = orc.prob_surface_rows(c["table"][:, nd:])
        np.testing.assert_array_equal(c["table"][idx, :nd], c["out_coords"])
        np.testing.assert_array_equal(y.reshape(-1), c["out_yp"][:, 0])
        np.testing.assert_allclose(p.reshape(-1), c["out_yp"][:, 1], rtol=1e-13, atol=1e-16)


@pytest.mark.parametrize("name", ["Test_file3_exp2"])
def test_pso_trace(name):
    """Seeded full trajectory of the oracle PSO == the reference's recorded trajectory."""
    g = load_golden(f"pso_trace_{name}.npz")
    seed = int(g["seed"])
    np.random.seed(seed)
    trace = []
    lines = []
    best = orc.len_scale_opt(g["x"], g["y"], g["e"], True, lhs_rng=np.random.RandomState(seed), trace=trace,
                             printer=lambda *a: lines.append(" ".join(str(v) for v in a)))
    pos = np.array([t[0] for t in trace])
    sc = np.array([t[1] for t in trace])
    assert pos.shape == g["positions"].shape
    np.testing.assert_allclose(sc, g["scores"], rtol=1e-9)
    np.testing.assert_allclose(pos, g["positions"], rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(best, g["best"], rtol=1e-9)
    assert "\n".join(lines) + "\n" == str(g["stdout"])
