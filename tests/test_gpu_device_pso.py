"""The swarm loop kept on the device (gpk_pso_step, pso._device_swarm_loop) against the host loop with the reference's NumPy
update (find_len_scales.py:76-141): same positions and scores at every swarm evaluation, same result, same stdout, and the
global np.random stream left in the same state -- bit for bit.  Cases: the reference's real PSO size, several soft restarts,
many out-of-bounds particles, the blocked path (N > 128, both arithmetic modes), odd swarm sizes and dimensions.
Run on the B200 box: pytest -m gpu."""
import numpy as np
import pytest

from oracle import gp_oracle as orc

pytestmark = pytest.mark.gpu


def _run(monkeypatch, capsys, device, x, y, e, seed, **kw):
    from gaussian_process_b200 import pso
    monkeypatch.setenv("GPK_PSO_DEVICE", "1" if device else "0")
    np.random.seed(seed)
    trace = []
    best = pso.len_scale_opt(x, y, e, True, lhs_random_state=seed, trace=trace, **kw)
    out = capsys.readouterr().out
    after = np.random.random_sample(3)              # where the global stream was left
    return best, np.array([t[0] for t in trace]), np.array([t[1] for t in trace]), out, after


CASES = [
    ("default size", dict(n=100, d=2, seed=21), dict()),
    ("restarts", dict(n=60, d=2, seed=3), dict(num_particles=12, max_iter=120, patience=7)),
    ("three dims, odd swarm", dict(n=37, d=3, seed=5), dict(num_particles=9, max_iter=60, patience=20)),
    ("one dim", dict(n=50, d=1, seed=8), dict(num_particles=5, max_iter=40)),
    ("blocked path", dict(n=300, d=2, seed=9), dict(num_particles=10, max_iter=12, max_points=1000)),
    ("blocked path, emulated", dict(n=1024, d=2, seed=4), dict(num_particles=6, max_iter=5, max_points=2000)),
    ("no iterations", dict(n=30, d=2, seed=1), dict(num_particles=5, max_iter=0)),
    ("single particle", dict(n=30, d=2, seed=2), dict(num_particles=1, max_iter=25, patience=6)),
    ("four dims", dict(n=80, d=4, seed=6), dict(num_particles=16, max_iter=40, patience=9)),
    ("patience beyond the run", dict(n=45, d=2, seed=7), dict(num_particles=8, max_iter=30, patience=1000)),
    ("restart every iteration", dict(n=45, d=2, seed=7), dict(num_particles=8, max_iter=20, patience=1)),
    ("kmeans subsample first", dict(n=400, d=2, seed=8), dict(num_particles=10, max_iter=15)),
]


@pytest.mark.parametrize("name,prob,kw", CASES, ids=[c[0] for c in CASES])
def test_device_loop_equals_host_loop(monkeypatch, capsys, name, prob, kw):
    x, y, e = orc.synthetic_problem(prob["n"], prob["d"], prob["seed"])
    host = _run(monkeypatch, capsys, False, x, y, e, prob["seed"], **kw)
    dev = _run(monkeypatch, capsys, True, x, y, e, prob["seed"], **kw)
    assert host[1].shape == dev[1].shape, (host[1].shape, dev[1].shape)
    np.testing.assert_array_equal(dev[1], host[1])          # positions of every swarm evaluation
    np.testing.assert_array_equal(dev[2], host[2])          # scores of every swarm evaluation
    np.testing.assert_array_equal(dev[0], host[0])          # chosen length scales
    assert dev[3] == host[3]                                 # progress lines, restarts, final report
    np.testing.assert_array_equal(dev[4], host[4])          # np.random left where the host loop leaves it
    if name == "restarts":
        assert "soft-restarting" in host[3] and (host[2] == 1e13).any()


def test_device_loop_raises_like_the_host_loop(monkeypatch, capsys):
    """A covariance that is not positive definite (duplicate point, zero noise) raises LinAlgError from both loops."""
    from gaussian_process_b200 import pso
    x = np.array([[0.0, 0.0, 1.0, 0.3, 0.7]])
    y = np.array([0.0, 1.0, 0.5, 0.2, 0.1])
    e = np.zeros(5)
    for device in ("0", "1"):
        monkeypatch.setenv("GPK_PSO_DEVICE", device)
        np.random.seed(0)
        with pytest.raises(np.linalg.LinAlgError):
            pso.len_scale_opt(x, y, e, False, lhs_random_state=0, num_particles=4, max_iter=5)
    capsys.readouterr()


def test_pso_step_stays_inside_its_buffers():
    """Guard bands around every buffer gpk_pso_step writes (state, positions, mask) and two runs with different contents in
    the parts of the state the kernel must not depend on (logs, traces): identical results, untouched guards."""
    import ctypes as C
    import torch
    from gaussian_process_b200 import _lib, engine, pso
    from test_gpu_memory_hygiene import Guarded
    lib = _lib.load()
    P, d, S, n, max_iter, patience = 7, 2, 1000, 40, 9, 3
    max_restarts = max_iter // patience + 1
    max_evals, rng_len = 1 + max_iter + max_restarts, (2 * max_iter + max_restarts) * P * d
    offs = (C.c_ulonglong * 19)()
    total = lib.gpk_pso_layout(P, d, S, n, max_iter, max_evals, rng_len, offs)
    off = dict(zip(pso._OFFS, (int(v) for v in offs)))
    x, y, e = orc.synthetic_problem(n, d, 1)
    pos0, lo, hi = orc.synthetic_particles(x, P, 1)
    sv, ex = orc.sigma_grid()
    results = []
    for poison in (0x00, 0xFF):
        host = np.full(total, poison, dtype=np.uint8)
        host[:off["log_best"]] = 0                                   # everything the caller fills starts from zero
        hdr = host[:128].view(np.int32)
        for k, v in (("P", P), ("d", d), ("S", S), ("n_points", n), ("max_iter", max_iter), ("patience", patience),
                     ("max_evals", max_evals), ("rng_len", rng_len)):
            hdr[pso._HDR[k]] = v
        f64 = lambda name, cnt: host[off[name]:off[name] + 8 * cnt].view(np.float64)      # noqa: E731
        f64("scalars", 8)[1] = 0.002
        f64("lower", d)[:], f64("upper", d)[:], f64("vmax", d)[:] = lo, hi, hi - lo
        f64("pos", P * d)[:] = pos0.reshape(-1)
        f64("expected", S)[:] = ex
        f64("d_sigma", S)[:S - 1] = np.diff(sv)
        f64("measured_of", n + 1)[:] = np.arange(n + 1) / n
        f64("rng", rng_len)[:] = np.random.RandomState(5).random_sample(rng_len)
        state = Guarded(torch, total)
        state.body.copy_(torch.from_numpy(host))
        dpos, dact = Guarded(torch, P * d * 8), Guarded(torch, P)
        dpos.view(torch.float64, (P, d)).copy_(torch.from_numpy(pos0))
        dact.body.fill_(1)
        ctx = engine.LossContext(x, y, e, sv)
        counts = torch.zeros((P, S), dtype=torch.int32, device="cuda")
        info = torch.zeros(P, dtype=torch.int32, device="cuda")
        fb = torch.zeros(P, dtype=torch.uint8, device="cuda")
        ws = ctx._ws(P)
        for _ in range(max_evals + 2):                                # two calls more than needed: they must idle
            _lib.check(lib.gpk_loss_batched(ctx.x.data_ptr(), ctx.y.data_ptr(), ctx.e.data_ptr(), dpos.ptr(), dact.ptr(),
                                            ctx.sig.data_ptr(), P, n, d, S, 0, 0, 0.0, counts.data_ptr(), info.data_ptr(), None,
                                            None, None, fb.data_ptr(), ws.data_ptr(), ws.numel(), None), "gpk_loss_batched")
            _lib.check(lib.gpk_pso_step(state.ptr(), total, P, d, S, n, max_iter, max_evals, rng_len, counts.data_ptr(),
                                        info.data_ptr(), fb.data_ptr(), dpos.ptr(), dact.ptr(), None), "gpk_pso_step")
        torch.cuda.synchronize()
        assert state.intact() and dpos.intact() and dact.intact()
        out = state.body.cpu().numpy()
        h = out[:128].view(np.int32)
        assert h[pso._HDR["pending"]] == 3 and h[pso._HDR["error"]] == 0 and h[pso._HDR["logged"]] == max_iter
        n_ev = int(h[pso._HDR["n_evals"]])
        assert 1 + max_iter <= n_ev <= max_evals and not dact.body.cpu().numpy().any()
        results.append((out[off["gbest_pos"]:off["gbest_pos"] + 8 * d].copy(), n_ev,
                        out[off["trace_sc"]:off["trace_sc"] + 8 * n_ev * P].copy()))
    assert results[0][1] == results[1][1]
    np.testing.assert_array_equal(results[0][0], results[1][0])
    np.testing.assert_array_equal(results[0][2], results[1][2])
    assert lib.gpk_pso_step(None, 0, P, d, S, n, max_iter, max_evals, rng_len, None, None, None, None, None, None) != 0
