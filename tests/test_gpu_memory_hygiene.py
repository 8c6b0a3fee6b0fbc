"""Out-of-bounds and uninitialised-memory checks of our own (compute-sanitizer is closed on this GPU pool:
profiles/r02_compute_sanitizer_closed.log).  Every buffer handed to the C ABI -- the workspace and each output -- sits
between guard bands that must come back untouched (an out-of-bounds WRITE lands in one of them), and every call is made twice,
with the workspace pre-filled with 0x00 and with 0xFF bytes (NaN patterns): the outputs must be bit-identical (a READ of
memory the call did not write first would show).  Covers the one-CTA loss (both forms), the blocked path in native FP64 and in
the emulated mode with the residual probe and a forced FP64 fallback pass, fit + both prediction paths, and the surface
reduction with its compaction.  Run on the B200 box: pytest -m gpu."""
import numpy as np
import pytest

from oracle import gp_oracle as orc

pytestmark = pytest.mark.gpu

GUARD = 4096
PATTERN = 0xA5


class Guarded:
    """A device buffer of `nbytes` between two guard bands."""

    def __init__(self, torch, nbytes, fill=0):
        self.torch = torch
        self.nbytes = int(nbytes)
        pad = (-self.nbytes) % 256
        self.raw = torch.full((GUARD + self.nbytes + pad + GUARD,), PATTERN, dtype=torch.uint8, device="cuda")
        self.body = self.raw[GUARD:GUARD + self.nbytes]
        self.body.fill_(fill)

    def ptr(self):
        return self.raw.data_ptr() + GUARD

    def view(self, dtype, shape):
        return self.body.view(dtype).view(*shape)

    def intact(self):
        lo, hi = self.raw[:GUARD], self.raw[GUARD + self.nbytes:]
        return bool((lo == PATTERN).all().item()) and bool((hi == PATTERN).all().item())


@pytest.fixture(scope="module")
def env():
    import torch
    from gaussian_process_b200 import _lib, engine
    return torch, _lib.load(), _lib, engine


def _loss_call(env, x, y, e, pos, precision, flags, tol, poison):
    torch, lib, _lib, engine = env
    d, n = x.shape
    p, s = pos.shape[0], 1000
    sig, _ = orc.sigma_grid()
    dx, dy, de, dpos, dsig = (engine.to_dev(a) for a in (x, y, e, pos, sig))
    act = engine.to_dev(np.array([1] * (p - 1) + [0], dtype=np.uint8))
    ws = Guarded(torch, max(lib.gpk_loss_workspace_bytes(p, n, d, precision), 256), fill=poison)
    outs = {"counts": Guarded(torch, p * s * 4), "info": Guarded(torch, p * 4), "resid": Guarded(torch, p * n * 8),
            "sd": Guarded(torch, p * n * 8), "probe": Guarded(torch, p * 8), "fb": Guarded(torch, p)}
    st = lib.gpk_loss_batched(dx.data_ptr(), dy.data_ptr(), de.data_ptr(), dpos.data_ptr(), act.data_ptr(), dsig.data_ptr(),
                              p, n, d, s, precision, flags, tol, outs["counts"].ptr(), outs["info"].ptr(), outs["resid"].ptr(),
                              outs["sd"].ptr(), outs["probe"].ptr(), outs["fb"].ptr(), ws.ptr(), ws.nbytes,
                              torch.cuda.current_stream().cuda_stream)
    _lib.check(st, "gpk_loss_batched")
    torch.cuda.synchronize()
    for name, g in list(outs.items()) + [("workspace", ws)]:
        assert g.intact(), f"guard band of {name} overwritten"
    live = slice(0, p - 1)                                     # the inactive particle's resid / sd are not defined
    return (outs["counts"].view(torch.int32, (p, s)).cpu().numpy(), outs["info"].view(torch.int32, (p,)).cpu().numpy(),
            outs["resid"].view(torch.float64, (p, n))[live].cpu().numpy(), outs["sd"].view(torch.float64, (p, n))[live].cpu().numpy(),
            outs["fb"].view(torch.uint8, (p,)).cpu().numpy())


@pytest.mark.parametrize("n,d,p,precision,flags,tol", [
    (100, 2, 7, 0, 0, 0.0),            # one-CTA path, reduced form
    (61, 3, 4, 0, 1, 0.0),             # one-CTA path, the reference's formulae
    (300, 3, 4, 0, 0, 0.0),            # blocked, native FP64, ragged padding
    (300, 3, 4, 0, 0, 1e-10),          # ... with the probe
    (1024, 2, 5, 1, 0, 1e-10),         # emulated, probe passes
    (640, 2, 4, 1, 0, 1e-30),          # emulated, every particle fails the probe: the FP64 fallback pass runs
    (640, 2, 4, 3, 0, -1e-10),         # 21 digit products, flags only
    (385, 4, 3, 4, 0, 0.0),            # 15 digit products, no check
])
def test_loss_call_stays_in_bounds_and_reads_nothing_uninitialised(env, n, d, p, precision, flags, tol):
    x, y, e = orc.synthetic_problem(n, d, n)
    pos, lo, hi = orc.synthetic_particles(x, p, n)
    a = _loss_call(env, x, y, e, pos, precision, flags, tol, poison=0x00)
    b = _loss_call(env, x, y, e, pos, precision, flags, tol, poison=0xFF)
    for u, v in zip(a, b):
        np.testing.assert_array_equal(u, v)
    assert not a[1].any() and not a[0][p - 1].any()
    if tol == 1e-30:
        assert a[4][:p - 1].all() and not a[4][p - 1]         # every active particle took the fallback


@pytest.mark.parametrize("precision", [0, 1])
def test_fit_and_predict_stay_in_bounds(env, precision):
    torch, lib, _lib, engine = env
    n, d, m = 700, 3, 1000
    x, y, e = orc.synthetic_problem(n, d, 2)
    ls = np.array([0.3, 0.25, 0.4])
    grid = np.random.default_rng(0).random((d, m))
    n_pad = lib.gpk_padded_n(n)
    res = []
    for poison in (0x00, 0xFF):
        dx, dy, de, dls, dg = (engine.to_dev(a) for a in (x, y, e, ls, grid))
        W, alpha = Guarded(torch, n_pad * n_pad * 8), Guarded(torch, n_pad * 8)
        xs, sq = Guarded(torch, d * n_pad * 8), Guarded(torch, n_pad * 8)
        info, probe = Guarded(torch, 4), Guarded(torch, 8)
        ws = Guarded(torch, lib.gpk_fit_workspace_bytes(n, d, precision), fill=poison)
        st = lib.gpk_fit(dx.data_ptr(), dy.data_ptr(), de.data_ptr(), dls.data_ptr(), n, d, precision, W.ptr(), alpha.ptr(),
                         xs.ptr(), sq.ptr(), info.ptr(), probe.ptr(), ws.ptr(), ws.nbytes, None)
        _lib.check(st, "gpk_fit")
        mu, sd = Guarded(torch, m * 8), Guarded(torch, m * 8)
        chunk = 512
        if precision == 0:
            pw = Guarded(torch, lib.gpk_predict_workspace_bytes(n, chunk), fill=poison)
            st = lib.gpk_predict(W.ptr(), alpha.ptr(), xs.ptr(), sq.ptr(), dls.data_ptr(), n, d, dg.data_ptr(), m, mu.ptr(),
                                 sd.ptr(), chunk, pw.ptr(), pw.nbytes, None)
            _lib.check(st, "gpk_predict")
            extra = [pw]
        else:
            state = Guarded(torch, lib.gpk_predict_state_bytes(n), fill=poison)
            _lib.check(lib.gpk_predict_prepare(W.ptr(), de.data_ptr(), n, state.ptr(), state.nbytes, None), "gpk_predict_prepare")
            pw = Guarded(torch, lib.gpk_predict_emu_workspace_bytes(n, chunk), fill=poison)
            st = lib.gpk_predict_emu(state.ptr(), alpha.ptr(), xs.ptr(), sq.ptr(), dls.data_ptr(), n, d, precision, dg.data_ptr(),
                                     m, mu.ptr(), sd.ptr(), chunk, pw.ptr(), pw.nbytes, None)
            _lib.check(st, "gpk_predict_emu")
            extra = [pw, state]
        torch.cuda.synchronize()
        for g in [W, alpha, xs, sq, info, probe, ws, mu, sd] + extra:
            assert g.intact()
        assert int(info.view(torch.int32, (1,)).item()) == 0
        res.append((mu.view(torch.float64, (m,)).cpu().numpy(), sd.view(torch.float64, (m,)).cpu().numpy(),
                    float(probe.view(torch.float64, (1,)).item())))
    np.testing.assert_array_equal(res[0][0], res[1][0])
    np.testing.assert_array_equal(res[0][1], res[1][1])
    assert res[0][2] == res[1][2] and res[0][2] < 1e-10
    mu_r, sd_r = orc.gp(x, y, e, grid, ls)
    np.testing.assert_allclose(res[0][0], mu_r, rtol=0, atol=1e-9 * np.abs(mu_r).max())
    np.testing.assert_allclose(res[0][1], sd_r, rtol=1e-9)


def test_surface_and_compaction_stay_in_bounds(env):
    torch, lib, _lib, engine = env
    r, n_exp, pts = 1000, 5, 100
    t = orc.synthetic_surface_table(r, n_exp, seed=3, p_gap=0.5)
    dt = engine.to_dev(t)
    y, p, valid = Guarded(torch, r * pts * 8), Guarded(torch, r * pts * 8), Guarded(torch, r * 4)
    _lib.check(lib.gpk_prob_surface(dt.data_ptr(), r, 2 * n_exp, pts, y.ptr(), p.ptr(), valid.ptr(), None), "gpk_prob_surface")
    dst, cnt = Guarded(torch, r * 8), Guarded(torch, 8)
    _lib.check(lib.gpk_surface_scan(valid.ptr(), r, dst.ptr(), cnt.ptr(), None), "gpk_surface_scan")
    k = int(cnt.view(torch.int64, (1,)).item())
    yc, pc, idx = Guarded(torch, k * pts * 8), Guarded(torch, k * pts * 8), Guarded(torch, k * 8)
    _lib.check(lib.gpk_surface_gather(valid.ptr(), dst.ptr(), r, pts, y.ptr(), p.ptr(), yc.ptr(), pc.ptr(), idx.ptr(), None),
               "gpk_surface_gather")
    torch.cuda.synchronize()
    for g in (y, p, valid, dst, cnt, yc, pc, idx):
        assert g.intact()
    i_r, y_r, p_r = orc.prob_surface_rows(t)
    assert 0 < k == len(i_r) < r
    np.testing.assert_array_equal(idx.view(torch.int64, (k,)).cpu().numpy(), i_r)
    np.testing.assert_array_equal(yc.view(torch.float64, (k, pts)).cpu().numpy(), y_r)
    np.testing.assert_allclose(pc.view(torch.float64, (k, pts)).cpu().numpy(), p_r, rtol=1e-9, atol=1e-15)
