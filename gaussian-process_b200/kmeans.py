"""KMeans subsample of find_len_scales.len_scale_opt (find_len_scales.py:25-47), SURVEY 8f row N4.

`fit_predict(X, n_clusters, random_state)` returns what `sklearn.cluster.KMeans(n_clusters=n_clusters, n_init='auto',
random_state=random_state).fit(X)` returns as `labels_`, `cluster_centers_` and `n_iter_`, with sklearn's single-threaded
arithmetic (the reference pins OMP / BLAS to one thread, GP_func.py:3-7).  Division of labour:

* the two O(N k) steps of a Lloyd iteration -- assignment and per-cluster sums in sample order -- are libgpk kernels
  (`gpk_kmeans_step`, csrc/gpk_kmeans.cu); the points go to the device once;
* the k-means++ seeding is sklearn's own `kmeans_plusplus` on the host: it is sequential and consumes the random stream
  (like the Latin hypercube and the swarm's random numbers, which also stay the host libraries' own);
* everything on the centres' side is NumPy on the host with sklearn's expressions, statement by statement
  (`_kmeans_single_lloyd`, `_relocate_empty_clusters_dense`, `_average_centers`, `_center_shift`, `_tolerance`): k (d + 1)
  numbers per iteration.

The host part is checked bit for bit against sklearn (labels, centres, iteration count) with a NumPy stand-in for the two
device steps, including duplicated points, fewer distinct points than clusters and N = k (tests/test_kmeans_host.py, CPU);
the GPU test does the same for the whole on the B200 box (tests/test_gpu_kmeans.py)."""
import warnings

import numpy as np
import torch

from . import _lib, engine

MAX_ITER = 300            # sklearn defaults, find_len_scales.py:31 passes none of them
TOL = 1e-4
MAX_K = 1024


def _center_shift(old, new):
    """sklearn's _euclidean_dense_dense(old_j, new_j, squared=False) for every centre: groups of four squared differences
    added as ((t0 + t1) + t2) + t3 to the running sum, then the remainder one by one."""
    t = (old - new) * (old - new)
    d = old.shape[1]
    res = np.zeros(old.shape[0])
    n4 = d // 4
    for g in range(n4):
        res = res + (((t[:, 4 * g] + t[:, 4 * g + 1]) + t[:, 4 * g + 2]) + t[:, 4 * g + 3])
    for m in range(4 * n4, d):
        res = res + t[:, m]
    return np.sqrt(res)


class _DeviceSteps:
    """The two O(N k) steps of a Lloyd iteration on the device (gpk_kmeans_step): the points go up once, per iteration the
    centres and their squared norms go up and the per-cluster sums, counts and the changed flag come down."""

    def __init__(self, X, k):
        self.lib = _lib.load()
        self.n, self.d = X.shape
        self.k = k
        n, d = self.n, self.d
        dev = engine.device()
        self.stream = torch.cuda.current_stream(dev.index)
        self.d_x = torch.from_numpy(X).to(dev)
        self.d_labels = torch.full((n,), -1, dtype=torch.int32, device=dev)
        self.pin_in = torch.empty(k * (d + 1), dtype=torch.float64).pin_memory()          # [centres | squared norms]
        self.d_in = torch.empty(k * (d + 1), dtype=torch.float64, device=dev)
        self.d_out = torch.zeros(k * (d + 1) + 1, dtype=torch.float64, device=dev)        # [sums | counts | changed flag (int32)]
        self.pin_out = torch.empty(k * (d + 1) + 1, dtype=torch.float64).pin_memory()
        self.np_in, self.np_out = self.pin_in.numpy(), self.pin_out.numpy()
        self.flag_view = self.d_out[k * (d + 1):]

    def step(self, centers, csq, with_sums):
        """-> (changed, sums (k, d), counts (k,)); the labels stay on the device."""
        k, d = self.k, self.d
        self.np_in[:k * d] = centers.reshape(-1)
        self.np_in[k * d:] = csq
        self.d_in.copy_(self.pin_in, non_blocking=True)
        self.flag_view.zero_()
        base_in, base_out = self.d_in.data_ptr(), self.d_out.data_ptr()
        rc = self.lib.gpk_kmeans_step(self.d_x.data_ptr(), self.n, d, base_in, base_in + k * d * 8, k, self.d_labels.data_ptr(),
                                      base_out + k * (d + 1) * 8, base_out if with_sums else None,
                                      base_out + k * d * 8 if with_sums else None, self.stream.cuda_stream)
        _lib.check(rc, "gpk_kmeans_step")
        self.pin_out.copy_(self.d_out, non_blocking=True)
        self.stream.synchronize()
        changed = bool(self.np_out[k * (d + 1):].view(np.int32)[0])
        return changed, self.np_out[:k * d].reshape(k, d).copy(), self.np_out[k * d:k * (d + 1)].copy()

    def labels(self):
        return self.d_labels.cpu().numpy()


def fit_predict(X, n_clusters, random_state=0, _backend=None):
    """-> (labels (N,) int32, cluster_centers (k, d) float64, n_iter).  `_backend` (tests): a factory (X, k) -> object with
    step(centers, csq, with_sums) and labels() standing in for the device steps."""
    from sklearn.cluster import kmeans_plusplus
    from sklearn.exceptions import ConvergenceWarning
    X = np.array(X, dtype=np.float64, order="C", copy=True)
    n, d = X.shape
    k = int(n_clusters)
    if n < k:
        raise ValueError(f"n_samples={n} should be >= n_clusters={k}.")
    if k > MAX_K or d > engine.MAX_D:
        raise ValueError(f"KMeans on the device supports n_clusters <= {MAX_K} and <= {engine.MAX_D} dimensions")
    tol = np.mean(np.var(X, axis=0)) * TOL                       # _tolerance(X, tol), before the centring
    x_mean = X.mean(axis=0)
    X -= x_mean                                                  # "subtract of mean of x for more accurate distance computations"
    x_sq = np.einsum("ij,ij->i", X, X)                           # row_norms(X, squared=True)
    centers, _ = kmeans_plusplus(X, k, x_squared_norms=x_sq, random_state=random_state)
    centers = np.ascontiguousarray(centers, dtype=np.float64)

    backend = _backend(X, k) if _backend is not None else _DeviceSteps(X, k)

    def e_step(cur, with_sums):
        return backend.step(cur, np.einsum("ij,ij->i", cur, cur), with_sums)     # row_norms(centers_old, squared=True)

    strict = False
    n_iter = 0
    for it in range(MAX_ITER):
        changed, new, w = e_step(centers, True)
        empty = np.where(np.equal(w, 0))[0]
        if len(empty):                                           # _relocate_empty_clusters_dense
            labels = backend.labels()
            dist = ((X - centers[labels]) ** 2).sum(axis=1)
            far = np.argpartition(dist, -len(empty))[:-len(empty) - 1:-1]
            if np.max(dist) != 0:
                for idx, cid in enumerate(empty):
                    f = far[idx]
                    old_c = labels[f]
                    new[old_c] -= X[f] * 1.0
                    new[cid] = X[f] * 1.0
                    w[cid] = 1.0
                    w[old_c] -= 1.0
        if (w > 0).all():                                        # _average_centers
            new *= (1.0 / w)[:, None]
        else:
            amax = int(np.argmax(w))
            for j in range(k):                                   # sequential: an empty cluster copies the biggest one's row as
                if w[j] > 0:                                     # it is at that moment (averaged or not), as sklearn's loop does
                    new[j] *= 1.0 / w[j]
                else:
                    new[j] = new[amax]
        shift = _center_shift(centers, new)
        centers = new
        n_iter = it + 1
        if not changed:                                          # np.array_equal(labels, labels_old): strict convergence
            strict = True
            break
        if (shift ** 2).sum() <= tol:
            break
    if not strict:
        e_step(centers, False)                                   # "rerun E-step so that predicted labels match cluster centers"
    labels = backend.labels()
    if len(set(labels.tolist())) < k:
        warnings.warn("Number of distinct clusters ({}) found smaller than n_clusters ({}). Possibly due to duplicate points "
                      "in X.".format(len(set(labels.tolist())), k), ConvergenceWarning, stacklevel=2)
    return labels, centers + x_mean, n_iter
