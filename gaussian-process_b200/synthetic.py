"""Synthetic workloads of the shapes BASELINE.json names, for bench.py and the measurement tools: there
is no network for data sets.  Same generators, value for value, as the ones the parity tests feed to the CPU checker
(tests/test_host_logic.py compares them), kept inside the package so that the measured arm imports nothing else."""
import numpy as np

from . import lhs, pso


def problem(n, d, seed=0):
    """x uniform on the unit box plus its 2^d corners, y = sum sin(2 pi x) + 0.1 N(0,1), e ~ U(0.05, 0.15).
    Returns x (d,n), y (n,), e (n,)."""
    rng = np.random.default_rng(seed)
    corners = np.array(np.meshgrid(*[[0.0, 1.0]] * d, indexing="ij")).reshape(d, -1)
    k = corners.shape[1]
    x = np.concatenate([rng.random((d, n - k)), corners], axis=1)
    y = np.sin(2 * np.pi * x).sum(axis=0) + 0.1 * rng.standard_normal(n)
    e = rng.uniform(0.05, 0.15, n)
    return x, y, e


def particles(x, p, seed=0):
    """Seeded centre Latin hypercube strictly inside the PSO bounds of x (find_len_scales.py:56-70)."""
    lower, upper = pso.swarm_bounds(x)
    pos = lhs.LHS(xlimits=np.column_stack((lower, upper)), criterion="center", random_state=seed)(p)
    return pos, lower, upper


def surface_table(r, n_exp, seed=0, p_gap=0.25):
    """(R, 2E) merged-table payload: mu ~ N(0.5, 0.2), sd ~ U(0.05, 0.2), whole (mu, sd) pairs masked with +inf."""
    rng = np.random.default_rng(seed)
    mu = rng.normal(0.5, 0.2, (r, n_exp))
    sd = rng.uniform(0.05, 0.2, (r, n_exp))
    gap = rng.random((r, n_exp)) < p_gap
    mu[gap] = np.inf
    sd[gap] = np.inf
    t = np.empty((r, 2 * n_exp))
    t[:, 0::2] = mu
    t[:, 1::2] = sd
    return t


def c4_experiments(rows=8192, n_exp=8, seed=4):
    """BASELINE config C4: one 2-D file of `rows` points, `n_exp` experiments with inf-masked gaps (a row is finite in an
    experiment with probability 0.5 -> ~rows / 2 points each; the 4 box corners are always present so every hull is the
    unit box).  Returns (x_list, [(y, e), ...]) in the layout read_in.read_data hands to GP_fit (read_in.py:204-254)."""
    rng = np.random.default_rng(seed)
    x = rng.random((2, rows))
    x[:, :4] = np.array([[0, 0, 1, 1], [0, 1, 0, 1]], dtype=np.float64)
    x_list, pairs = [], []
    for k in range(n_exp):
        keep = rng.random(rows) < 0.5
        keep[:4] = True
        xk = np.ascontiguousarray(x[:, keep])
        yk = np.sin(2 * np.pi * xk[0]) + np.sin(2 * np.pi * xk[1]) + 0.3 * k + 0.1 * rng.standard_normal(xk.shape[1])
        ek = rng.uniform(0.05, 0.15, xk.shape[1])
        x_list.append(xk)
        pairs.append((yk, ek))
    return x_list, pairs
