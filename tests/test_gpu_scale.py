"""Full-size (BASELINE.json shapes) checks of the CUDA path through size-independent properties, plus
oracle comparisons on bounded samples.  Run on the B200 box: pytest -m gpu."""
import numpy as np
import pytest

from oracle import gp_oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def eng():
    from gaussian_process_b200 import engine, pso
    return engine, pso


@pytest.fixture(params=["fp64", "emu"], autouse=True)
def precision(request, eng):
    """Every full-size check runs in both arithmetic modes of the blocked path."""
    eng[0].set_precision(request.param)
    yield request.param
    eng[0].set_precision("auto")


def test_loss_n4096_two_routes_agree(eng):
    """C3 shape.  The loss kernels use mu-y = -D alpha and var = d - d^2 diag(K^-1) (blocked inverse);
    the predict kernels use the reference's form mu = K_s^T alpha, var = 1 - |L^-1 k|^2.  Two independent
    kernel chains must agree at the training points; counts must be monotone, bounded, and independent of
    the position of a particle in the swarm."""
    engine, pso = eng
    n, d = 4096, 3
    x, y, e = orc.synthetic_problem(n, d, 0)
    pos, lo, hi = orc.synthetic_particles(x, 5, 0)
    pos = np.concatenate([pos, pos[1:2]])                        # duplicate particle
    s, ex = pso.sigma_grid()
    counts, info, resid, sd = engine.LossContext(x, y, e, s).counts(pos, want_detail=True)
    assert not info.any()
    assert np.all(np.diff(counts, axis=1) >= 0) and counts.max() <= n and counts.min() >= 0
    np.testing.assert_array_equal(counts[1], counts[5])
    np.testing.assert_array_equal(resid[1], resid[5])
    for k in (0, 3):
        mu, sd2 = engine.gp(x, y, e, x, pos[k])
        np.testing.assert_allclose(mu, resid[k] + y, rtol=1e-9, atol=1e-11)
        np.testing.assert_allclose(sd2, sd[k], rtol=1e-7)         # 1 - sum v^2 loses digits by cancellation
    # the counts are what the literal formula gives for these (resid, sd)
    k = 2
    pulls = resid[k][:, None] / np.maximum(sd[k][:, None] * s[None, :], 1e-12)
    np.testing.assert_array_equal((np.abs(pulls) <= 1).sum(axis=0), counts[k])
    # wave splitting (workspace smaller than the swarm) gives identical results
    from gaussian_process_b200 import _lib
    import torch
    lib = _lib.load()
    ctx = engine.LossContext(x, y, e, s)
    mode = engine._resolve_precision(e, n)
    small = torch.empty(lib.gpk_loss_workspace_bytes(2, n, d, mode), dtype=torch.uint8, device="cuda")
    dpos = engine.to_dev(pos)
    c2 = torch.empty((6, 1000), dtype=torch.int32, device="cuda")
    inf2 = torch.empty(6, dtype=torch.int32, device="cuda")
    fb = torch.empty(6, dtype=torch.uint8, device="cuda")
    st = lib.gpk_loss_batched(ctx.x.data_ptr(), ctx.y.data_ptr(), ctx.e.data_ptr(), dpos.data_ptr(), None,
                              ctx.sig.data_ptr(), 6, n, d, 1000, mode, 0, 1e-10, c2.data_ptr(), inf2.data_ptr(), None, None,
                              None, fb.data_ptr(), small.data_ptr(), small.numel(), torch.cuda.current_stream().cuda_stream)
    assert st == 0
    np.testing.assert_array_equal(c2.cpu().numpy(), counts)
    assert not fb.cpu().numpy().any()                 # the a-posteriori check passes on the bench workload: no fallback


def test_potrf_n4096_round_trip(eng):
    engine, _ = eng
    n = 4096
    x, y, e = orc.synthetic_problem(n, 3, 2)
    k = orc.rbf_kernel(x, x, np.array([0.2, 0.3, 0.25])) + np.diag(e ** 2)
    L, info = engine.potrf_batched(k[None])
    assert not info.any()
    v = np.random.default_rng(0).standard_normal((n, 4))
    np.testing.assert_allclose(L[0] @ (L[0].T @ v), k @ v, rtol=1e-11, atol=1e-11)


def test_predict_c2_shape_vs_oracle_sample(eng):
    """C2: N=1024, d=2, 200x200 lattice (40 000 points); oracle on a 1 500-point sample."""
    engine, _ = eng
    x, y, e = orc.synthetic_problem(1024, 2, 5)
    ls = np.array([0.15, 0.22])
    g = np.linspace(0, 1, 200)
    grid = np.array(np.meshgrid(g, g, indexing="ij")).reshape(2, -1)
    mu, sd = engine.gp(x, y, e, grid, ls)
    assert mu.shape == (40000,)
    pick = np.random.default_rng(1).choice(40000, 1500, replace=False)
    mu_r, sd_r = orc.gp(x, y, e, grid[:, pick], ls)
    np.testing.assert_allclose(mu[pick], mu_r, rtol=1e-9, atol=1e-11)
    np.testing.assert_allclose(sd[pick], sd_r, rtol=1e-9)
    # chunking of the grid does not change anything
    fit = engine.GPFit(x, y, e, ls)
    mu2, sd2 = fit.predict(grid, chunk=1000)
    np.testing.assert_array_equal(mu2, mu)
    np.testing.assert_array_equal(sd2, sd)


def test_fit_n8192_inverse_property(eng):
    """C4 size (one experiment of ~8192 points): W = L^-1 satisfies W^T (W (K v)) = v."""
    engine, _ = eng
    n = 8192
    x, y, e = orc.synthetic_problem(n, 2, 7)
    ls = np.array([0.12, 0.2])
    fit = engine.GPFit(x, y, e, ls)
    W = np.tril(fit.W.cpu().numpy()[:n, :n])
    k = orc.rbf_kernel(x, x, ls) + np.diag(e ** 2)
    v = np.random.default_rng(0).standard_normal(n)
    np.testing.assert_allclose(W.T @ (W @ (k @ v)), v, rtol=0, atol=1e-7)
    np.testing.assert_allclose(k @ fit.alpha.cpu().numpy()[:n], y, rtol=0, atol=1e-8)


def test_surface_million_rows(eng):
    """C5-like: E=16, 25 % gaps.  Valid-row flags, axis end points, probability mass, oracle on a sample."""
    engine, _ = eng
    r = 1 << 20
    base = orc.synthetic_surface_table(1 << 14, 16, seed=4)
    t = np.tile(base, (r // base.shape[0], 1))
    idx, yv, p = engine.prob_surface(t)
    fin = np.isfinite(t).sum(axis=1)
    want = np.nonzero((fin >= 2) & (fin % 2 == 0))[0]
    np.testing.assert_array_equal(idx, want)
    mu = np.where(np.isfinite(t[:, 0::2]), t[:, 0::2], np.nan)
    sg = np.where(np.isfinite(t[:, 1::2]), t[:, 1::2], np.nan)
    np.testing.assert_array_equal(yv[:, 0], np.nanmin(mu - 3 * sg, axis=1)[idx])
    np.testing.assert_array_equal(yv[:, -1], np.nanmax(mu + 3 * sg, axis=1)[idx])
    assert p.min() > -1e-15 and np.all(p.sum(axis=1) < 1.0 + 1e-12) and np.all(p.sum(axis=1) > 0.5)
    np.testing.assert_array_equal(p[: base.shape[0]][:100], p[-base.shape[0]:][:100])       # tiled rows repeat exactly
    i_r, y_r, p_r = orc.prob_surface_rows(base[:300])
    np.testing.assert_array_equal(yv[:len(i_r)], y_r)
    np.testing.assert_allclose(p[:len(i_r)], p_r, rtol=1e-9, atol=1e-15)
