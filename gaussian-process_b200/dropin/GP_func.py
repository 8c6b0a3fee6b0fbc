"""Drop-in replacement for the reference's GP_func module (same names, same signatures), backed by
the B200 kernels.  Put this directory ahead of the reference on sys.path: GP_fit.py's
`from GP_func import GP` (GP_fit.py:15) then resolves here."""
import _bootstrap  # noqa: F401
import numpy as np

from gaussian_process_b200 import engine, sharding

SHARD_MIN_POINTS = 1 << 16      # grid sizes from which the prediction is split over the ranks of the process group


def GP(x_known, y_known, e_known, x_fit, lengths, batch_size=10000):
    """GP regression mean and uncertainty at x_fit (reference GP_func.py:12-45).
    x_known (d,N), x_fit (d,M), lengths (d,) -> (y_fit (M,), e_fit (M,)) float64.
    Under an initialised process group in which every rank makes the same call (the swarm-sharded mode of
    GP_fit.create_GP), a grid of at least SHARD_MIN_POINTS points is predicted in contiguous tiles, one per rank
    (sharding.sharded_predict); the values are those of the single-rank call."""
    x_fit = np.asarray(x_fit, dtype=np.float64)
    if sharding.world()[1] > 1 and x_fit.ndim == 2 and x_fit.shape[1] >= SHARD_MIN_POINTS:
        fit = engine.GPFit(x_known, y_known, e_known, lengths)
        return sharding.sharded_predict(fit.predict, x_fit)
    return engine.gp(x_known, y_known, e_known, x_fit, lengths, batch_size)


def kernel_func(x1, x2, l):
    """ARD-RBF kernel matrix (reference GP_func.py:49-65).  x1 (d,N1), x2 (d,N2), l (d,) -> (N1,N2)."""
    return engine.cov_batched(x1, x2, np.asarray(l, dtype=np.float64)[None, :])[0]
