"""Particle / experiment / grid sharding over one process per GPU (torch.distributed).

The path has no data-path collective except one tiny exchange per PSO iteration: every rank
evaluates its contiguous slice of the swarm and the per-particle losses are all-gathered (P float64,
2 KB at P=256 -- latency bound; NCCL over NVLink on GPUs, gloo in the CPU tests).  Every rank then
runs the identical host PSO update from the identical seeded RNG stream, so positions never travel
(SURVEY 8e).  Experiments and grid tiles are embarrassingly parallel: `shard_slice` only."""
import numpy as np
import torch
import torch.distributed as dist


_local_depth = 0


def dist_world():
    """(rank, size) of the process group, whatever the sharding mode."""
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


def world():
    """(rank, size) the swarm shards over: the process group, or (0, 1) inside `local_swarm()`."""
    if _local_depth > 0:
        return 0, 1
    return dist_world()


class local_swarm:
    """Context in which swarm evaluations stay on this rank (no collective): used when whole experiments are the
    sharded unit (GP_fit.create_GP with at least as many experiments as ranks)."""

    def __enter__(self):
        global _local_depth
        _local_depth += 1
        return self

    def __exit__(self, *exc):
        global _local_depth
        _local_depth -= 1
        return False


def gather_objects(local, group=None):
    """List of every rank's picklable `local`, in rank order, on every rank."""
    rank, size = dist_world()
    if size == 1:
        return [local]
    out = [None] * size
    dist.all_gather_object(out, local, group=group)
    return out


def shard_slice(n, rank, size):
    """Contiguous balanced slice [lo, hi) of n units for `rank` of `size`."""
    base, rem = divmod(n, size)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def all_gather_losses(local, n_total, group=None):
    """local: float64 array holding this rank's slice (in shard_slice order) -> full (n_total,) array,
    identical on every rank."""
    rank, size = world()
    if size == 1:
        return np.asarray(local, dtype=np.float64)
    backend = dist.get_backend(group)
    width = -(-n_total // size)
    buf = torch.zeros(width, dtype=torch.float64)
    buf[:len(local)] = torch.from_numpy(np.asarray(local, dtype=np.float64))
    if backend == "nccl":
        buf = buf.cuda()
    out = torch.empty(size * width, dtype=torch.float64, device=buf.device)
    dist.all_gather_into_tensor(out, buf, group=group)
    out = out.cpu().numpy().reshape(size, width)
    parts = []
    for r in range(size):
        lo, hi = shard_slice(n_total, r, size)
        parts.append(out[r, :hi - lo])
    return np.concatenate(parts)


class ShardedSwarm:
    """Wraps a swarm evaluator `fn(positions (p,d)) -> losses (p,)` so that each rank evaluates only
    its slice and all ranks see all losses."""

    def __init__(self, fn, group=None):
        self.fn = fn
        self.group = group

    def __call__(self, positions):
        rank, size = world()
        n = positions.shape[0]
        lo, hi = shard_slice(n, rank, size)
        local = self.fn(positions[lo:hi]) if hi > lo else np.zeros(0)
        return all_gather_losses(local, n, self.group)
