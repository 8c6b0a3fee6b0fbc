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
