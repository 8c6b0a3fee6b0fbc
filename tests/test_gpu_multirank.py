"""SURVEY 4 tier 6 on the GPU: the N-rank sharded evaluation equals the 1-rank one bit for bit, and a sharded PSO whose
ranks are seeded differently walks exactly the single-process trajectory of rank 0.  One rank per GPU over NCCL when the box
has enough GPUs; on a one-GPU box both ranks compute on cuda:0 and exchange the losses over gloo (the kernels, the slicing and
the gather logic are the same; only the transport differs).  Run on the B200 box: pytest -m gpu."""
import os
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _worker(rank, size, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(size))
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from oracle import gp_oracle as orc, ref_shim
    from gaussian_process_b200 import pso, sharding
    ref_shim.install_smt_stub()
    nccl = torch.cuda.device_count() >= size
    torch.cuda.set_device(rank if nccl else 0)
    dist.init_process_group("nccl" if nccl else "gloo", rank=rank, world_size=size)
    s, ex = orc.sigma_grid()
    out = {"backend": np.array("nccl" if nccl else "gloo")}
    for tag, n, d, p in (("small", 100, 2, 11), ("blocked", 300, 3, 7), ("emu", 1024, 2, 6)):
        x, y, e = orc.synthetic_problem(n, d, 4)
        pos, lo, hi = orc.synthetic_particles(x, p, 4)
        pos[1] = lo                                           # an out-of-bounds particle in rank 0's slice
        fn = pso.SwarmLoss(x, y, e, s, ex, lo, hi)
        out[tag + "_sharded"] = sharding.ShardedSwarm(fn)(pos)
        out[tag + "_local"] = fn(pos)                         # the whole swarm on this rank
    # sharded PSO, ranks seeded differently: rank 0's trajectory everywhere
    x, y, e = orc.synthetic_problem(60, 2, 9)
    np.random.seed(3 + 50 * rank)
    trace = []
    best = pso.len_scale_opt(x, y, e, False, num_particles=12, max_iter=25, lhs_random_state=1 + rank, trace=trace,
                             distributed=True)
    out["pso_best"] = best
    out["pso_pos"] = np.array([t[0] for t in trace])
    out["pso_sc"] = np.array([t[1] for t in trace])
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), **out)
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_gpu_evaluation_is_bit_identical(tmp_path):
    port = 29500 + (os.getpid() * 7) % 2000
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0, r1 = np.load(tmp_path / "r0.npz"), np.load(tmp_path / "r1.npz")
    for tag in ("small", "blocked", "emu"):
        np.testing.assert_array_equal(r0[tag + "_sharded"], r0[tag + "_local"])         # N ranks == 1 rank, bit for bit
        np.testing.assert_array_equal(r0[tag + "_sharded"], r1[tag + "_sharded"])
        assert r0[tag + "_sharded"][1] == 1e13
    for k in ("pso_best", "pso_pos", "pso_sc"):
        np.testing.assert_array_equal(r0[k], r1[k])
    # ... and equals the single-process run seeded as rank 0 was
    from oracle import gp_oracle as orc, ref_shim
    from gaussian_process_b200 import pso
    ref_shim.install_smt_stub()
    x, y, e = orc.synthetic_problem(60, 2, 9)
    np.random.seed(3)
    trace = []
    best = pso.len_scale_opt(x, y, e, False, num_particles=12, max_iter=25, lhs_random_state=1, trace=trace)
    np.testing.assert_array_equal(best, r0["pso_best"])
    np.testing.assert_array_equal(np.array([t[0] for t in trace]), r0["pso_pos"])
    np.testing.assert_array_equal(np.array([t[1] for t in trace]), r0["pso_sc"])
    print("transport:", str(r0["backend"]))
