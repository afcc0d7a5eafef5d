"""Host-side sharding of a batch over the GPUs of one box.

States and reactors are independent: every rank owns a slice of the batch and a full copy of the
(<= 100 KB) mechanism, and NO collective runs on the data path. The only communication is the
barrier + max-over-ranks of the measured time around a timed region (`timed_max`), over whatever
backend `torch.distributed` was initialised with (NCCL on the GPU box, gloo in the CPU tests).

For ignition batches the per-reactor cost varies ~10x with the initial temperature, so slices are
taken from a seeded random permutation (static balance); inside a GPU the integrator's slot queue
balances dynamically.
"""
import numpy as np


def slice_bounds(n, rank, world):
    """Contiguous slice [begin, end) of n items for `rank`: sizes differ by at most one."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("rank %r out of range for world size %r" % (rank, world))
    base, rem = divmod(int(n), world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def permuted_indices(n, rank, world, seed=20261104):
    """Indices of this rank's share of a seeded random permutation of range(n) (identical on every
    rank, so the shares are disjoint and cover the batch)."""
    perm = np.random.default_rng(seed).permutation(int(n))
    b, e = slice_bounds(n, rank, world)
    return np.sort(perm[b:e])


def timed_max(local_ms, device=None):
    """Max over ranks of a locally measured duration (the contract's max-over-ranks timing).
    Without an initialised process group this is the identity."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(local_ms)
    t = torch.tensor([float(local_ms)], dtype=torch.float64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def total_units(local_units, device=None):
    """Sum over ranks of the units each rank processed (whole-job throughput numerator)."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return int(local_units)
    t = torch.tensor([int(local_units)], dtype=torch.int64, device=device if device is not None else "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return int(t.item())
