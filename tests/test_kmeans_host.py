"""Host part of the KMeans subsample (kmeans.fit_predict) against sklearn, on the CPU: the two device
steps are replaced by a NumPy stand-in that does what the kernels do (assignment by csq - 2 X C^T with first-minimum argmin,
per-cluster sums in sample order), everything else -- centring, tolerance, seeding, empty clusters, averaging, centre
shift, convergence tests, final E-step -- is the module's own code.  Labels, centres and iteration count must equal
single-threaded sklearn's bit for bit.  (The device steps themselves: tests/test_gpu_kmeans.py.)  Also: layout of the swarm
loop's state buffer (gpk_pso_layout is host code)."""
import ctypes as C
import os
import warnings

import numpy as np
import pytest


class NumpySteps:
    def __init__(self, X, k):
        self.X, self.k = X, k
        self._labels = np.full(X.shape[0], -1, dtype=np.int32)

    def step(self, centers, csq, with_sums):
        pd = csq[None, :] - 2.0 * (self.X @ centers.T)
        new_labels = np.argmin(pd, axis=1).astype(np.int32)
        changed = not np.array_equal(new_labels, self._labels)
        self._labels = new_labels
        sums = np.zeros_like(centers)
        counts = np.zeros(self.k)
        if with_sums:
            np.add.at(sums, new_labels, self.X)                 # one element after the other, in sample order
            counts = np.bincount(new_labels, minlength=self.k).astype(np.float64)
        return changed, sums, counts

    def labels(self):
        return self._labels


def _cases():
    rng = np.random.default_rng(0)
    out = []
    for rep in range(2):
        for n, d in ((120, 2), (100, 2), (104, 3), (4080, 2), (4096, 3), (1000, 4), (600, 5), (150, 2), (101, 1), (2000, 6)):
            X = rng.random((n, d))
            if rep == 1:
                X = np.round(X, 2)                               # duplicates, ties, empty clusters
            out.append((f"n={n} d={d} {'rounded' if rep else 'uniform'}", X))
    g = np.linspace(0, 1, 30)
    out.append(("lattice 30x30", np.array(np.meshgrid(g, g, indexing="ij")).reshape(2, -1).T.copy()))
    out.append(("gaussian, anisotropic", rng.normal(size=(3000, 3)) * np.array([0.1, 5.0, 1.0])))
    return out


@pytest.mark.parametrize("name,X", _cases(), ids=[n for n, _ in _cases()])
def test_kmeans_host_logic_equals_sklearn(name, X):
    from sklearn.cluster import KMeans
    from threadpoolctl import threadpool_limits
    from gaussian_process_b200 import kmeans
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        with threadpool_limits(limits=1):                        # the reference pins OMP / BLAS to one thread
            km = KMeans(n_clusters=100, n_init="auto", random_state=0)
            ref = km.fit_predict(X)
            labels, centers, n_iter = kmeans.fit_predict(X, 100, random_state=0, _backend=NumpySteps)
    assert n_iter == km.n_iter_
    np.testing.assert_array_equal(labels, ref)
    np.testing.assert_array_equal(centers, km.cluster_centers_)


def test_pso_state_layout():
    from gaussian_process_b200 import _lib, pso
    lib = _lib.load()
    P, d, S, n, max_iter, max_evals, rng_len = 40, 2, 1000, 100, 500, 507, 80480
    offs = (C.c_ulonglong * len(pso._OFFS))()
    total = lib.gpk_pso_layout(P, d, S, n, max_iter, max_evals, rng_len, offs)
    o = [int(v) for v in offs]
    assert o[0] == pso.PSO_HDR_WORDS * 4 and all(b > a for a, b in zip(o, o[1:])) and all(v % 8 == 0 for v in o)
    assert total == o[-1] + rng_len * 8
    sizes = dict(zip(pso._OFFS, [b - a for a, b in zip(o, o[1:] + [total])]))
    assert sizes["pos"] == P * d * 8 and sizes["trace_sc"] == max_evals * P * 8 and sizes["measured_of"] == (n + 1) * 8
    assert lib.gpk_pso_layout(P, 9, S, n, max_iter, max_evals, rng_len, None) == 0           # d beyond GPK_MAX_D
