mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_kmeans.py -m gpu -q -p no:cacheprovider 2>&1 | tail -15
python - <<'PY'
import time, numpy as np, warnings
from gaussian_process_b200 import kmeans, synthetic
from sklearn.cluster import KMeans
x, y, e = synthetic.problem(4080, 2, 4)
X = x.T.copy()
for _ in range(2):
    t0 = time.perf_counter(); kmeans.fit_predict(X, 100); t1 = time.perf_counter()
    km = KMeans(n_clusters=100, n_init="auto", random_state=0); km.fit_predict(X); t2 = time.perf_counter()
print(f"N=4080 d=2 k=100: device {1e3*(t1-t0):.1f} ms, sklearn {1e3*(t2-t1):.1f} ms, iterations {km.n_iter_}")
PY
