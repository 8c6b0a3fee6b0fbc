"""KMeans subsample on the device (kmeans.fit_predict, gpk_kmeans_step) against sklearn itself, single
threaded as the reference runs it: labels, cluster centres and iteration count bit for bit, and the subsample
len_scale_opt keeps (find_len_scales.py:25-47).  Data sets: the C4 / C3 / C2 shapes, lattices (the reference's inputs are
kinematic grids: exact distance ties), rounded coordinates with many duplicates, fewer distinct points than clusters,
N = k + 1, 1 to 6 dimensions.  Run on the B200 box: pytest -m gpu."""
import os
import warnings

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _datasets():
    rng = np.random.default_rng(0)
    out = []
    for n, d in ((4080, 2), (8192, 2), (4096, 3), (1024, 2), (1000, 4), (600, 5), (3000, 6), (500, 1), (101, 2), (150, 3)):
        out.append((f"uniform n={n} d={d}", rng.random((n, d))))
    g = np.linspace(0, 1, 40)
    out.append(("lattice 40x40", np.array(np.meshgrid(g, g, indexing="ij")).reshape(2, -1).T.copy()))
    g3 = np.linspace(-1, 1, 12)
    out.append(("lattice 12^3", np.array(np.meshgrid(g3, g3, g3, indexing="ij")).reshape(3, -1).T.copy()))
    out.append(("rounded, duplicates", np.round(rng.random((2000, 2)), 2)))
    out.append(("few distinct points", np.round(rng.random((300, 1)), 1).repeat(2, axis=1)))
    out.append(("gaussian blobs", rng.normal(size=(5000, 3)) * np.array([0.1, 5.0, 1.0]) + rng.integers(0, 4, (5000, 1))))
    return out


@pytest.mark.parametrize("name,X", _datasets(), ids=[n for n, _ in _datasets()])
def test_device_kmeans_equals_sklearn(name, X):
    from sklearn.cluster import KMeans
    from threadpoolctl import threadpool_limits
    from gaussian_process_b200 import kmeans
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")
        with threadpool_limits(limits=1):                    # the reference pins OMP / BLAS to one thread (GP_func.py:3-7)
            km = KMeans(n_clusters=100, n_init="auto", random_state=0)
            ref_labels = km.fit_predict(X)
        labels, centers, n_iter = kmeans.fit_predict(X, 100, random_state=0)
    assert n_iter == km.n_iter_, (n_iter, km.n_iter_)
    np.testing.assert_array_equal(labels, ref_labels)
    np.testing.assert_array_equal(centers, km.cluster_centers_)


def test_subsample_equals_the_reference_selection():
    from gaussian_process_b200 import pso, synthetic
    for n, d, seed in ((4096, 3, 0), (1024, 2, 0), (8192, 2, 3)):
        x, y, e = synthetic.problem(n, d, seed)
        a = pso.kmeans_subsample(x, y, e, 100, backend="device")
        b = pso.kmeans_subsample(x, y, e, 100, backend="sklearn")
        for u, v in zip(a, b):
            np.testing.assert_array_equal(u, v)
    with pytest.raises(ValueError):
        pso.kmeans_subsample(x, y, e, 100, backend="nope")
