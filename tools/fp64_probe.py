import torch, time, os, json
print("cpu_count", os.cpu_count())
try:
    import psutil; print("mem GB", psutil.virtual_memory().total/2**30)
except Exception as e: print(e)
dev = torch.device("cuda:0")
print(torch.cuda.get_device_name(0))
for n in (4096, 8192):
    a = torch.randn(n, n, dtype=torch.float64, device=dev); b = torch.randn(n, n, dtype=torch.float64, device=dev)
    for _ in range(3): c = a @ b
    torch.cuda.synchronize()
    best = 1e9
    for _ in range(5):
        e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record(); c = a @ b; e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print(f"cublas dgemm n={n}: {best:.3f} ms  {2*n**3/best*1e-9:.2f} TFLOP/s")
    # sustained 3 s
    t0 = time.time(); k = 0
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    while time.time() - t0 < 2.0:
        for _ in range(5): c = a @ b
        k += 5
        torch.cuda.synchronize()
    e1.record(); torch.cuda.synchronize()
    print(f"  sustained: {2*n**3*k/e0.elapsed_time(e1)*1e-9:.2f} TFLOP/s")
# cholesky / triangular solve via cusolver for reference timings
n = 4096
a = torch.randn(n, n, dtype=torch.float64, device=dev); k = a @ a.T + n * torch.eye(n, dtype=torch.float64, device=dev)
for _ in range(2): L = torch.linalg.cholesky(k)
torch.cuda.synchronize(); t0 = time.time(); L = torch.linalg.cholesky(k); torch.cuda.synchronize(); print("cusolver potrf 4096 ms", (time.time()-t0)*1e3)
t0 = time.time(); V = torch.linalg.solve_triangular(L, k, upper=False); torch.cuda.synchronize(); print("trsm 4096x4096 ms", (time.time()-t0)*1e3)
