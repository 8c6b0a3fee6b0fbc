import os, sys
sys.path.insert(0, "/root/repo")
import numpy as np, torch
from gaussian_process_b200 import engine, _lib
lib = _lib.load()
rng = np.random.default_rng(0); n = 128; P = 64
A = rng.standard_normal((n, n)); A = A @ A.T + n * np.eye(n)
k = engine.to_dev(np.repeat(A[None], P, axis=0))
info = torch.zeros(P, dtype=torch.int32, device="cuda")
ws = engine.workspace(lib.gpk_potrf_workspace_bytes(P, n), "potrf")
st = torch.cuda.current_stream().cuda_stream
kk = k.clone()
for _ in range(3): lib.gpk_potrf_batched(kk.data_ptr(), P, n, info.data_ptr(), ws.data_ptr(), ws.numel(), st)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(50): lib.gpk_potrf_batched(kk.data_ptr(), P, n, info.data_ptr(), ws.data_ptr(), ws.numel(), st)
e1.record(); torch.cuda.synchronize()
print(os.environ.get("GPK_LIB", "default").split("/")[-1], f"potrf of {P} 128x128 blocks (memset + one k_diag_block): {e0.elapsed_time(e1) / 50 * 1e3:.1f} us")
