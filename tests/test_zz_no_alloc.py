"""gpk_predict_static: the prediction entry point that allocates nothing (scratch in a static __device__ array inside libgpk;
no counterpart in the reference, include/gpk.h).  Called through the C ABI and
checked against the existing entry point gpk_predict on two calls in a row that differ in N, d and M (the second call finds
the first call's contents in the shared scratch), within the 1e-9 bar of the other tests.  Collected last on purpose.
Passes alone in a fresh process and after the other tests.  Run on the B200 box: pytest -m gpu."""
import numpy as np
import pytest

from oracle import gp_oracle as orc

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def env():
    import torch
    from gaussian_process_b200 import _lib, engine
    return torch, _lib.load(), _lib, engine


def _fit(engine, n, d, seed):
    x, y, e = orc.synthetic_problem(n, d, seed)
    ls = np.random.default_rng(seed).uniform(0.2, 0.5, d)
    prev = engine.get_precision()
    engine.set_precision("fp64")
    try:
        fit = engine.GPFit(x, y, e, ls)
    finally:
        engine.set_precision(prev)
    return fit, (x, y, e, ls)


def _both(env, fit, grid):
    """(mu, sd) of gpk_predict (caller workspace, chunk 256) and of gpk_predict_static (no workspace) on the same fit."""
    torch, lib, _lib, engine = env
    m = grid.shape[1]
    dg = engine.to_dev(grid)
    st = torch.cuda.current_stream().cuda_stream
    out = []
    ws = torch.empty(lib.gpk_predict_workspace_bytes(fit.n, 256), dtype=torch.uint8, device="cuda")
    mu = torch.full((m,), float("nan"), dtype=torch.float64, device="cuda")
    sd = torch.full((m,), float("nan"), dtype=torch.float64, device="cuda")
    _lib.check(lib.gpk_predict(fit.W.data_ptr(), fit.alpha.data_ptr(), fit.xs.data_ptr(), fit.sq.data_ptr(), fit.ls.data_ptr(),
                               fit.n, fit.d, dg.data_ptr(), m, mu.data_ptr(), sd.data_ptr(), 256, ws.data_ptr(), ws.numel(), st),
               "gpk_predict")
    out.append((mu.cpu().numpy(), sd.cpu().numpy()))
    mu2 = torch.full((m,), float("nan"), dtype=torch.float64, device="cuda")
    sd2 = torch.full((m,), float("nan"), dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    before = (torch.cuda.memory_allocated(), torch.cuda.memory_reserved())
    _lib.check(lib.gpk_predict_static(fit.W.data_ptr(), fit.alpha.data_ptr(), fit.xs.data_ptr(), fit.sq.data_ptr(),
                                      fit.ls.data_ptr(), fit.n, fit.d, dg.data_ptr(), m, mu2.data_ptr(), sd2.data_ptr(), st),
               "gpk_predict_static")
    torch.cuda.synchronize()
    assert (torch.cuda.memory_allocated(), torch.cuda.memory_reserved()) == before      # nothing went through torch's allocator
    out.append((mu2.cpu().numpy(), sd2.cpu().numpy()))
    return out


def test_predict_static_equals_predict_on_two_calls_in_a_row(env):
    torch, lib, _lib, engine = env
    # first call: the largest problem the static scratch takes (N = 2048), several chunks with a ragged tail
    fit1, (x1, y1, e1, ls1) = _fit(engine, 2048, 3, 1)
    grid1 = np.random.default_rng(10).random((3, 1000))
    (mu_a, sd_a), (mu_b, sd_b) = _both(env, fit1, grid1)
    np.testing.assert_allclose(mu_b, mu_a, rtol=1e-9, atol=1e-9 * np.abs(mu_a).max())
    np.testing.assert_allclose(sd_b, sd_a, rtol=1e-9)
    # second call right after it: smaller, ragged N and M, other d -- the shared scratch still holds the first call's data
    fit2, (x2, y2, e2, ls2) = _fit(engine, 300, 2, 2)
    grid2 = np.random.default_rng(11).random((2, 77))
    (mu_c, sd_c), (mu_d, sd_d) = _both(env, fit2, grid2)
    np.testing.assert_allclose(mu_d, mu_c, rtol=1e-9, atol=1e-9 * np.abs(mu_c).max())
    np.testing.assert_allclose(sd_d, sd_c, rtol=1e-9)
    # and both agree with the checker (GP_func.py:26-45)
    mu_r, sd_r = orc.gp(x2, y2, e2, grid2, ls2)
    np.testing.assert_allclose(mu_d, mu_r, rtol=0, atol=1e-9 * np.abs(mu_r).max())
    np.testing.assert_allclose(sd_d, sd_r, rtol=1e-9)
    # the first problem once more: identical bits to its first evaluation (same kernels, same chunking)
    (_mu, _sd), (mu_e, sd_e) = _both(env, fit1, grid1)
    np.testing.assert_array_equal(mu_e, mu_b)
    np.testing.assert_array_equal(sd_e, sd_b)


def test_predict_static_rejects_sizes_beyond_its_scratch(env):
    torch, lib, _lib, engine = env
    dummy = torch.zeros(64, dtype=torch.float64, device="cuda")
    p = dummy.data_ptr()
    st = lib.gpk_predict_static(p, p, p, p, p, 2049, 2, p, 10, p, p, None)
    msg = lib.gpk_last_error_string().decode()
    assert st == 2 and "GPK_STATIC_MAX_N" in msg and "2049" in msg, (st, msg)
    assert lib.gpk_predict_static(p, p, p, p, p, 100, 9, p, 10, p, p, None) == 2          # d beyond GPK_MAX_D
    assert lib.gpk_predict_static(p, p, p, p, p, 100, 2, p, 0, p, p, None) == 0           # M = 0: nothing to do
