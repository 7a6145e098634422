"""Sharding of the cohort over ranks and the one collective of the path.

The hot path has no exchange step: a frame pair depends only on its own two feature maps / frames, and the
temporal recurrence of the reference lives inside one (patient, slice) cine
(/root/reference/nnunet/network_architecture/SegFlowGaussian.py:905-945; slices :3268; patients in the trainer
:982).  So the unit of work is one (patient, slice) cine of T frames = T-1 pairs, units are dealt round-robin
to ranks (one process per GPU), nothing is communicated while computing, and ONE gather at the end collects
the compact per-unit results (registered label maps / metrics) over NCCL (NVLink 5 / NVSwitch) -- gloo on CPU
in the tests.
"""
import torch
import torch.distributed as dist


def cohort_units(n_patients, n_slices):
    """All (patient, slice) units of a cohort, in a fixed global order."""
    return [(p, s) for p in range(n_patients) for s in range(n_slices)]


def unit_seed(patient, slice_idx, base=1005):
    """Seed of a unit's synthetic data (SURVEY.md 8(d) cfg 5): any rank can generate its shard alone."""
    return base * 10 ** 6 + patient * 100 + slice_idx


def shard_units(units, rank, world_size):
    """Round-robin: unit i goes to rank i % world_size (sizes differ by at most one)."""
    if not (0 <= rank < world_size):
        raise ValueError(f"rank {rank} outside world of {world_size}")
    return units[rank::world_size]


def gather_results(local, n_units_total, rank=None, world_size=None, group=None):
    """All-gather per-unit result rows.  ``local`` [n_local, ...] holds this rank's units in shard order; the
    return value [n_units_total, ...] is in the global unit order on every rank.  Shards are padded to the
    largest shard so one equal-size all_gather suffices (no staging copy: the padded buffer is the send buffer).
    """
    if world_size is None:
        world_size = dist.get_world_size(group) if dist.is_initialized() else 1
    if rank is None:
        rank = dist.get_rank(group) if dist.is_initialized() else 0
    if world_size == 1:
        return local[:n_units_total]
    per_rank = (n_units_total + world_size - 1) // world_size
    send = local.new_zeros((per_rank,) + tuple(local.shape[1:]))
    send[: local.shape[0]] = local
    recv = local.new_empty((world_size * per_rank,) + tuple(local.shape[1:]))
    dist.all_gather_into_tensor(recv, send, group=group)
    # rank r, slot j  ->  global unit j * world_size + r
    out = recv.view(world_size, per_rank, *local.shape[1:]).transpose(0, 1).reshape(per_rank * world_size, *local.shape[1:])
    return out[:n_units_total].contiguous()
