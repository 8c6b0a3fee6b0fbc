"""Host-side logic of the drop-in (no GPU): score finishing, bounds, LHS, sharding, and the PSO loop
driven by the oracle loss must reproduce the reference's seeded trajectory bit for bit."""
import numpy as np
import pytest

from conftest import load_golden, case
from oracle import gp_oracle as orc
from gaussian_process_b200 import pso, lhs, sharding


def test_finish_losses_matches_scalar_expressions_bitwise():
    g = load_golden("loss_cases.npz")
    s, ex = orc.sigma_grid()
    s2, ex2 = pso.sigma_grid()
    assert np.array_equal(s, s2) and np.array_equal(ex, ex2)
    for name in g["names"]:
        c = case(g, str(name))
        n = c["x"].shape[1]
        oob = pso.out_of_bounds_mask(c["pos"], c["lower"], c["upper"])
        assert list(oob) == [orc.out_of_bounds(p, c["lower"], c["upper"]) for p in c["pos"]]
        got = pso.finish_losses(c["counts"], n, c["pos"], s, ex, c["lower"], c["upper"], oob)
        want = np.array([1e13 if oob[k] else orc.finish_loss(c["counts"][k].astype(np.int64), n, p, s, ex, c["lower"], c["upper"])
                         for k, p in enumerate(c["pos"])])
        assert np.array_equal(got, want)                     # bit-identical to the per-particle route
        np.testing.assert_allclose(got, c["loss"], rtol=1e-12)


def test_bounds_and_subsample_match_oracle():
    x, y, e = orc.synthetic_problem(300, 3, 4)
    lo, hi = pso.swarm_bounds(x)
    lo2, hi2 = orc.pso_bounds(x)
    assert np.array_equal(lo, lo2) and np.array_equal(hi, hi2)
    a = pso.kmeans_subsample(x, y, e, 100, backend="sklearn")      # the selection logic; the device backend: tests/test_gpu_kmeans.py
    b = orc.kmeans_subsample(x, y, e, 100)
    for u, v in zip(a, b):
        assert np.array_equal(u, v)


def test_builtin_lhs_is_center_lhs():
    bounds = np.array([[0.1, 2.0], [1.0, 5.0], [-1.0, 1.0]])
    a = lhs.LHS(xlimits=bounds, criterion="center", random_state=5)(40)
    b = orc.center_lhs(bounds, 40, np.random.RandomState(5))
    assert np.array_equal(a, b)
    for j in range(3):                                        # one point per cell centre
        cells = np.sort((a[:, j] - bounds[j, 0]) / (bounds[j, 1] - bounds[j, 0]) * 40 - 0.5)
        np.testing.assert_allclose(cells, np.arange(40), atol=1e-9)


def test_shard_slices_cover_everything():
    for n in (0, 1, 7, 40, 256, 257):
        for size in (1, 2, 3, 8):
            parts = [sharding.shard_slice(n, r, size) for r in range(size)]
            assert parts[0][0] == 0 and parts[-1][1] == n
            assert all(parts[i][1] == parts[i + 1][0] for i in range(size - 1))
            widths = [b - a for a, b in parts]
            assert max(widths) - min(widths) <= 1


@pytest.mark.parametrize("name", ["Test_file3_exp2"])
def test_pso_host_loop_reproduces_reference_trace(name, capsys):
    """pso.len_scale_opt with the oracle as swarm evaluator == reference trajectory and stdout."""
    g = load_golden(f"pso_trace_{name}.npz")
    seed = int(g["seed"])
    x, y, e = g["x"], g["y"], g["e"]
    lo, hi = orc.pso_bounds(x)
    s, ex = orc.sigma_grid()

    def swarm(positions):
        return np.array([orc.evaluate_loss(p, x, y, e, s, ex, lo, hi) for p in positions])

    np.random.seed(seed)
    trace = []
    best = pso.len_scale_opt(x, y, e, True, lhs_random_state=seed, swarm_fn=swarm, trace=trace, distributed=False)
    out = capsys.readouterr().out
    pos = np.array([t[0] for t in trace])
    assert pos.shape == g["positions"].shape
    np.testing.assert_allclose(pos, g["positions"], rtol=1e-9, atol=1e-12)
    np.testing.assert_allclose(best, g["best"], rtol=1e-9)
    assert out == str(g["stdout"])


def test_synthetic_generators_equal_the_oracles():
    """bench.py's measured arm generates its workload without importing oracle/: same data, value for value."""
    from oracle import gp_oracle as orc
    from gaussian_process_b200 import synthetic
    for n, d, seed in ((40, 2, 0), (300, 3, 1), (64, 1, 5)):
        for a, b in zip(synthetic.problem(n, d, seed), orc.synthetic_problem(n, d, seed)):
            np.testing.assert_array_equal(a, b)
        x = orc.synthetic_problem(n, d, seed)[0]
        for a, b in zip(synthetic.particles(x, 7, seed), orc.synthetic_particles(x, 7, seed)):
            np.testing.assert_array_equal(a, b)
    np.testing.assert_array_equal(synthetic.surface_table(500, 4, seed=3), orc.synthetic_surface_table(500, 4, seed=3))


def test_host_finish_equals_numpy_trapz():
    """libgpk's one-pass host finish (gpk_finish_wasserstein) returns the bits of np.trapz(|counts / n - expected|, sigma)
    (find_len_scales.py:163-166): random and monotone count rows, every count range the path meets, grid lengths around
    the split points of NumPy's pairwise summation."""
    import ctypes as C
    from gaussian_process_b200 import _lib, pso
    lib = _lib.load()
    vp = lambda a: a.ctypes.data_as(C.c_void_p)      # noqa: E731
    rng = np.random.default_rng(3)
    for S in (2, 3, 8, 9, 17, 128, 129, 130, 137, 257, 513, 999, 1000):
        sv = np.linspace(0.001, 3, S)
        ex = pso.sigma_to_percent(sv)
        d = np.diff(sv)
        for n in (1, 2, 21, 100, 4096, 8192):
            for mono in (False, True):
                counts = rng.integers(0, n + 1, (17, S))
                counts = (np.sort(counts, axis=1) if mono else counts).astype(np.int32)
                table = np.arange(n + 1, dtype=np.int32).astype(np.float64) / n
                out, scratch = np.empty(17), np.empty(2 * S)
                assert lib.gpk_finish_wasserstein(vp(counts), 17, S, vp(table), n, vp(ex), vp(d), vp(scratch), vp(out)) == 0
                ref = np.trapz(np.abs(counts.astype(np.float64) / n - ex[None, :]), sv, axis=1)
                np.testing.assert_array_equal(out, ref)
    bad = np.full((1, 4), 7, dtype=np.int32)
    assert lib.gpk_finish_wasserstein(vp(bad), 1, 4, vp(np.zeros(4)), 3, vp(np.zeros(4)), vp(np.zeros(3)), vp(np.empty(8)), vp(np.empty(1))) == 3
