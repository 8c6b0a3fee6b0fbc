"""world_size-2 gloo test of the sharded swarm evaluation (host logic of the multi-GPU path)."""
import os
import sys

import numpy as np
import torch.multiprocessing as mp

from conftest import ROOT


def _worker(rank, size, port, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(size))
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from oracle import gp_oracle as orc, ref_shim
    from gaussian_process_b200 import pso, sharding
    ref_shim.install_smt_stub()
    dist.init_process_group("gloo", rank=rank, world_size=size)
    x, y, e = orc.synthetic_problem(24, 2, 3)
    lo, hi = orc.pso_bounds(x)
    s, ex = orc.sigma_grid()
    calls = []

    def swarm(positions):
        calls.append(len(positions))
        return np.array([orc.evaluate_loss(p, x, y, e, s, ex, lo, hi) for p in positions])

    # 1) plain gather, ragged split (7 particles over 2 ranks)
    pos = orc.center_lhs(np.column_stack((lo, hi)), 7, np.random.RandomState(1))
    full = sharding.ShardedSwarm(swarm)(pos)
    # 2) a short seeded PSO: every rank must walk the identical trajectory
    np.random.seed(11)
    trace = []
    best = pso.len_scale_opt(x, y, e, False, num_particles=9, max_iter=12, lhs_random_state=2, swarm_fn=swarm,
                             trace=trace)
    np.savez(os.path.join(out_dir, f"r{rank}.npz"), full=full, best=best, calls=np.array(calls),
             pos=np.array([t[0] for t in trace]), sc=np.array([t[1] for t in trace]))
    dist.barrier()
    dist.destroy_process_group()


def test_sharded_swarm_world2(tmp_path):
    port = 29500 + os.getpid() % 2000
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    r0 = np.load(tmp_path / "r0.npz")
    r1 = np.load(tmp_path / "r1.npz")
    for k in ("full", "best", "pos", "sc"):
        assert np.array_equal(r0[k], r1[k]), k
    # each rank evaluated only its share
    assert r0["calls"][0] == 4 and r1["calls"][0] == 3
    assert set(r0["calls"][1:]) == {5} and set(r1["calls"][1:]) == {4}

    # and the sharded result equals the single-process one
    sys.path.insert(0, ROOT)
    from oracle import gp_oracle as orc
    from gaussian_process_b200 import pso
    x, y, e = orc.synthetic_problem(24, 2, 3)
    lo, hi = orc.pso_bounds(x)
    s, ex = orc.sigma_grid()
    swarm = lambda P: np.array([orc.evaluate_loss(p, x, y, e, s, ex, lo, hi) for p in P])
    pos = orc.center_lhs(np.column_stack((lo, hi)), 7, np.random.RandomState(1))
    assert np.array_equal(swarm(pos), r0["full"])
    np.random.seed(11)
    trace = []
    best = pso.len_scale_opt(x, y, e, False, num_particles=9, max_iter=12, lhs_random_state=2, swarm_fn=swarm,
                             trace=trace, distributed=False)
    assert np.array_equal(best, r0["best"])
    assert np.array_equal(np.array([t[0] for t in trace]), r0["pos"])
