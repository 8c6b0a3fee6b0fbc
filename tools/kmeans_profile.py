"""Where the KMeans subsample's time goes (C4-sized experiment: N=4080, d=2, k=100)."""
import cProfile, io, os, pstats, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from sklearn.cluster import kmeans_plusplus
from gaussian_process_b200 import kmeans, synthetic
x, y, e = synthetic.problem(4080, 2, 4)
X = x.T.copy()
kmeans.fit_predict(X, 100)
Xc = X - X.mean(axis=0)
xsq = np.einsum("ij,ij->i", Xc, Xc)
t0 = time.perf_counter()
for _ in range(5):
    kmeans_plusplus(Xc, 100, x_squared_norms=xsq, random_state=0)
print(f"kmeans_plusplus alone: {(time.perf_counter() - t0) / 5 * 1e3:.2f} ms")
t0 = time.perf_counter()
for _ in range(5):
    lab, cen, it = kmeans.fit_predict(X, 100)
print(f"fit_predict: {(time.perf_counter() - t0) / 5 * 1e3:.2f} ms, {it} iterations")
pr = cProfile.Profile(); pr.enable(); kmeans.fit_predict(X, 100); pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("tottime").print_stats(14); print(s.getvalue()[:3500])
# a large case: the Lloyd share grows with N
from sklearn.cluster import KMeans
Xl = np.random.default_rng(1).random((200000, 3))
kmeans.fit_predict(Xl[:5000], 100)
t0 = time.perf_counter(); lab, cen, it = kmeans.fit_predict(Xl, 100); t1 = time.perf_counter()
km = KMeans(n_clusters=100, n_init="auto", random_state=0); ref = km.fit_predict(Xl); t2 = time.perf_counter()
print(f"N=200000 d=3: device {1e3 * (t1 - t0):.0f} ms ({it} iterations), sklearn {1e3 * (t2 - t1):.0f} ms ({km.n_iter_} iterations), "
      f"labels equal {np.array_equal(lab, ref)}, centres equal {np.array_equal(cen, km.cluster_centers_)}")
