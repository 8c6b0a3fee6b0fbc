"""torchrun check of the drop-in GP() with grid tiles over the ranks (NCCL): every rank obtains the values of the single-rank
call.  torchrun --nproc-per-node 2 tools/sharded_predict_check.py"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "gaussian-process_b200", "dropin"))
import numpy as np, torch, torch.distributed as dist
local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import GP_func
from gaussian_process_b200 import engine, synthetic, sharding
x, y, e = synthetic.problem(4096, 3, 1)
grid = np.random.default_rng(0).random((3, 1_000_003))
ls = np.full(3, 0.25)
GP_func.GP(x, y, e, grid[:, :70000], ls)
dist.barrier(); t0 = time.perf_counter()
mu, sd = GP_func.GP(x, y, e, grid, ls)
dist.barrier(); t_sh = time.perf_counter() - t0
with sharding.local_swarm():
    t0 = time.perf_counter(); mu1, sd1 = GP_func.GP(x, y, e, grid, ls); t_1 = time.perf_counter() - t0
ok = np.array_equal(mu, mu1) and np.array_equal(sd, sd1)
if dist.get_rank() == 0:
    print(f"sharded GP() over {dist.get_world_size()} ranks: {t_sh:.3f} s vs single-rank {t_1:.3f} s for {grid.shape[1]} grid points at N=4096; identical values: {ok}")
dist.barrier(); dist.destroy_process_group()
sys.exit(0 if ok else 1)
