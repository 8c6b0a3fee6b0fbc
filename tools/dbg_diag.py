import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from gaussian_process_b200 import engine
rng = np.random.default_rng(0); n = 128
A = rng.standard_normal((n, n)); A = A @ A.T + n * np.eye(n)
L, info = engine.potrf_batched(A[None])
ref = np.linalg.cholesky(A)
print("info", info, "max err", np.abs(L[0] - ref).max())
bad = np.argwhere(np.abs(L[0] - ref) > 1e-9)
print("n bad", len(bad), bad[:10].tolist())
print(np.isnan(L[0]).sum())
np.set_printoptions(precision=5, linewidth=200)
print("gpu\n", L[0][8:11, :10]); print("ref\n", ref[8:11, :10])
# what would it be if the step-1 update had been skipped / doubled?
