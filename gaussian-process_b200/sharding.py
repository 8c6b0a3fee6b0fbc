"""Particle / experiment / grid sharding over one process per GPU (torch.distributed).

The path has no data-path collective except one tiny exchange per PSO iteration: every rank
evaluates its contiguous slice of the swarm and the per-particle losses are all-gathered (P float64,
2 KB at P=256 -- latency bound; NCCL over NVLink on GPUs, gloo in the CPU tests).  Every rank then
runs the identical host PSO update from the identical seeded RNG stream, so positions never travel
(SURVEY 8e).  Experiments, grid tiles and surface rows are embarrassingly parallel: contiguous `shard_slice` blocks per
rank; `gather_rows` reassembles per-point results (16 B per grid point), `sharded_prob_surface` keeps the surface output
sharded and exchanges only the per-rank numbers of emitting rows (one int64 each) for the compaction offsets."""
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


# A swarm is sharded only when one evaluation of it is worth more than the exchange: the all-gather of the losses costs a
# host->device copy, a latency-bound NCCL collective and a device->host copy per PSO iteration (~0.1 ms), the one-CTA path
# evaluates a 40-particle swarm of 100 points in 0.13 ms, and the blocked path below ~2000 points is bound by its chain of
# dependent launches, not by the number of particles.  Below the threshold every rank evaluates the whole
# swarm redundantly from the same (broadcast) RNG state: zero communication per iteration, identical results.
SHARD_MIN_FLOPS = 3.0e11          # P * 2/3 N^3 of one swarm evaluation (C3: 1.2e13; C2: 4.6e10; default PSO: 2.7e7)


def worth_sharding(n_particles, n_points):
    return n_points > 128 and n_particles * (2.0 / 3.0) * float(n_points) ** 3 >= SHARD_MIN_FLOPS


def broadcast_object(obj, src=0, group=None):
    """`obj` of rank `src` on every rank (pickled; used once per optimisation, not per iteration)."""
    rank, size = dist_world()
    if size == 1:
        return obj
    box = [obj if rank == src else None]
    dist.broadcast_object_list(box, src=src, group=group)
    return box[0]


class ShardedSwarm:
    """Wraps a swarm evaluator `fn(positions (p,d)) -> losses (p,)` so that each rank evaluates only
    its slice and all ranks see all losses.

    A failure on one rank (numpy.linalg.LinAlgError from a covariance that is not positive definite, GP_func.py:22, or a
    library error) must not leave the other ranks waiting in the collective: the local evaluation is caught, the slice
    travels with a status word, and the same exception type is raised on every rank after the exchange -- the reference
    aborts the whole run from its pool worker (find_len_scales.py:103)."""

    _OK, _LINALG, _OTHER = 0.0, 1.0, 2.0

    def __init__(self, fn, group=None):
        self.fn = fn
        self.group = group

    def __call__(self, positions):
        rank, size = world()
        n = positions.shape[0]
        lo, hi = shard_slice(n, rank, size)
        status, err = self._OK, None
        try:
            local = np.asarray(self.fn(positions[lo:hi]), dtype=np.float64) if hi > lo else np.zeros(0)
        except np.linalg.LinAlgError as ex:
            status, err, local = self._LINALG, ex, np.zeros(hi - lo)
        except Exception as ex:                     # noqa: BLE001 -- re-raised below, on every rank
            status, err, local = self._OTHER, ex, np.zeros(hi - lo)
        width = -(-n // size)
        padded = np.zeros(width + 1)
        padded[:hi - lo] = local
        padded[width] = status
        every = all_gather_losses(padded, size * (width + 1), self.group).reshape(size, width + 1)
        worst = every[:, width].max()
        if worst != self._OK:
            if err is not None:
                raise err
            bad = int(np.argmax(every[:, width]))
            if worst == self._LINALG:
                raise np.linalg.LinAlgError(f"Matrix is not positive definite (swarm slice of rank {bad})")
            raise RuntimeError(f"swarm evaluation failed on rank {bad}")
        return np.concatenate([every[r, :shard_slice(n, r, size)[1] - shard_slice(n, r, size)[0]] for r in range(size)])


def gather_rows(local, n_total, group=None):
    """local: this rank's contiguous block of rows (shard_slice order) of a float64 array (rows, ...) -> the full
    (n_total, ...) array on every rank.  One fixed-width all-gather (blocks padded to the widest one)."""
    rank, size = dist_world()
    local = np.ascontiguousarray(local, dtype=np.float64)
    if size == 1:
        return local
    tail = local.shape[1:]
    per_row = int(np.prod(tail)) if tail else 1
    width = -(-n_total // size)
    buf = torch.zeros(width * per_row, dtype=torch.float64)
    buf[:local.size] = torch.from_numpy(local.reshape(-1))
    if dist.get_backend(group) == "nccl":
        buf = buf.cuda()
    out = torch.empty(size * width * per_row, dtype=torch.float64, device=buf.device)
    dist.all_gather_into_tensor(out, buf, group=group)
    out = out.cpu().numpy().reshape(size, width, *tail)
    parts = []
    for r in range(size):
        lo, hi = shard_slice(n_total, r, size)
        parts.append(out[r, :hi - lo])
    return np.concatenate(parts)


def sharded_predict(predict_fn, x_fit, group=None):
    """Grid tiles over the ranks: predict_fn(x_fit_block (d,m)) -> (mean (m,), sd (m,)) is called on this
    rank's contiguous block of grid points; the full (mean, sd) come back on every rank.  The fit is replicated, not
    broadcast: refitting costs 20 ms at N=4096, less than moving the 134 MB factor."""
    rank, size = dist_world()
    x_fit = np.asarray(x_fit, dtype=np.float64)
    m = x_fit.shape[1]
    lo, hi = shard_slice(m, rank, size)
    if hi > lo:
        mu, sd = predict_fn(np.ascontiguousarray(x_fit[:, lo:hi]))
        both = np.stack([np.asarray(mu, dtype=np.float64).reshape(-1), np.asarray(sd, dtype=np.float64).reshape(-1)], axis=1)
    else:
        both = np.zeros((0, 2))
    full = gather_rows(both, m, group)
    return np.ascontiguousarray(full[:, 0]), np.ascontiguousarray(full[:, 1])


def surface_offsets(n_local, group=None):
    """Compaction of the surface output across ranks: all-gather of the per-rank numbers of emitting rows (one int64
    each) and their exclusive prefix sum.  Returns (offset of this rank's block, total)."""
    rank, size = dist_world()
    if size == 1:
        return 0, int(n_local)
    mine = torch.tensor([int(n_local)], dtype=torch.int64)
    if dist.get_backend(group) == "nccl":
        mine = mine.cuda()
    every = torch.empty(size, dtype=torch.int64, device=mine.device)
    dist.all_gather_into_tensor(every, mine, group=group)
    every = every.cpu().numpy()
    return int(every[:rank].sum()), int(every.sum())


def sharded_prob_surface(surface_fn, table, group=None):
    """Surface rows over the ranks: surface_fn(table_block) -> (row_index (k,), y (k,points), p (k,points))
    of the block's emitting rows.  This rank reduces rows [lo, hi); the outputs stay sharded (config C5 emits 27 GB).
    Returns (offset, total, global row_index, y, p): this rank's block occupies rows offset .. offset + k of the
    compacted result of `total` rows."""
    rank, size = dist_world()
    lo, hi = shard_slice(table.shape[0], rank, size)
    idx, y, p = surface_fn(table[lo:hi])
    offset, total = surface_offsets(len(idx), group)
    return offset, total, np.asarray(idx, dtype=np.int64) + lo, y, p
