"""Multi-GPU plumbing of the per-ray path: one process per GPU, torch.distributed.

Rays are independent, so rendering shards whole views (or contiguous ray ranges of a view)
across ranks with NO data-path collective; the training step has exactly one exchange, an
all-reduce(sum) of the flat fp32 gradient buffer [hash table | MLP], after which every rank
applies the same Adam step to the same sum (weights stay replicated, nothing is broadcast).
The reference has none of this (single GPU, rtxFunctions.cpp:57).
"""
import torch
import torch.distributed as dist


def view_shard(n_views, rank, world):
    """Views of rank `rank`: round-robin (view v -> rank v % world), the C3 batch render."""
    return list(range(rank, n_views, world))


def ray_shard(n_rays, rank, world):
    """Contiguous ray range [begin, end) of rank `rank`; ranges differ by at most one ray."""
    base, rem = divmod(n_rays, world)
    begin = rank * base + min(rank, rem)
    return begin, begin + base + (1 if rank < rem else 0)


def batch_first(step, rank, world, batch_rays, n_rays):
    """First position, in the seeded shuffle of a store of n_rays rays shared by all ranks, of rank
    `rank`'s batch at global step `step` (api.batch_assemble's `first`): consecutive whole batches are dealt
    round-robin, so within an epoch no ray is used twice on any rank.  Returns (epoch, first) — the epoch is
    the shuffle seed's offset; the tail that does not fill a batch for every rank is skipped (the reference
    slices past the end of its vector there, Q6)."""
    rounds_per_epoch = n_rays // (batch_rays * world)
    if rounds_per_epoch <= 0:
        raise ValueError("ray store smaller than one batch per rank")
    epoch, rnd = divmod(step, rounds_per_epoch)
    return epoch, (rnd * world + rank) * batch_rays


def allreduce_sum_(flat, group=None):
    """In-place sum over ranks of the flat gradient buffer (NCCL on GPUs, gloo in the CPU tests)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
    return flat


def grad_scale(loss_scale, group=None):
    """Factor Adam applies to the summed gradient: 1 / (loss_scale * ranks) — the mean over
    ranks folded into the optimizer's loss-scale division (optimizer->step(stream, loss_scale, ...),
    main.cu:787)."""
    world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
    return 1.0 / (float(loss_scale) * world)


def max_over_ranks(values, device, group=None):
    """Device-timed durations are reported as the max over ranks."""
    t = torch.tensor(values, dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return [float(x) for x in t]
