"""Sharding of the cohort over ranks and the one collective of the path.

The hot path has no exchange step: a frame pair depends only on its own two feature maps / frames, and the
temporal recurrence of the reference lives inside one (patient, slice) cine
(/root/reference/nnunet/network_architecture/SegFlowGaussian.py:905-945; slices :3268; patients in the trainer
:982).  So the unit of work is one (patient, slice) cine of T frames = T-1 pairs, units are dealt round-robin
to ranks (one process per GPU), nothing is communicated while computing, and ONE gather at the end collects
the per-unit results (registered label maps, uint8; metrics) over NCCL (NVLink 5 / NVSwitch) -- gloo on CPU
in the tests.  The kernels write those results straight into this rank's slice of the all-gather buffer (GatherBuffer).
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


class GatherBuffer:
    """The receive buffer of the final all-gather, allocated up front so that the kernels write their per-unit results
    STRAIGHT into this rank's slice (no staging copy precedes the collective; NCCL's in-place all-gather: the send buffer is
    ``recv[rank]``).  Layout [world, per_rank, *item_shape]; rank r, slot j holds global unit j * world + r (round-robin
    shards, padded to the largest shard).

        buf = GatherBuffer(n_units_total, (29, 1, 256, 256), torch.uint8, device)
        for j, unit in enumerate(my_units): kernel(..., out=buf.local[j])
        table = buf.gather()            # [per_rank, world, *item_shape]: a transposed VIEW of the receive buffer (no copy)
        item = buf.unit(table, k)       # result of global unit k
    """

    def __init__(self, n_units_total, item_shape, dtype, device, rank=None, world_size=None, group=None):
        if world_size is None:
            world_size = dist.get_world_size(group) if dist.is_initialized() else 1
        if rank is None:
            rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.rank, self.world, self.group, self.n_total = rank, world_size, group, n_units_total
        self.per_rank = (n_units_total + world_size - 1) // world_size
        self.recv = torch.zeros((world_size, self.per_rank) + tuple(item_shape), dtype=dtype, device=device)
        self.local = self.recv[rank]                       # this rank's slice = the send buffer

    @property
    def nbytes_sent(self):
        return self.local.numel() * self.local.element_size()

    def gather(self):
        if self.world > 1:
            dist.all_gather_into_tensor(self.recv.view(-1), self.local.reshape(-1), group=self.group)
        return self.recv.transpose(0, 1)                   # [slot j, rank r] = global unit j * world + r

    def unit(self, table, k):
        """Item of global unit k in what ``gather`` returned."""
        if not 0 <= k < self.n_total:
            raise IndexError(k)
        return table[k // self.world, k % self.world]


def gather_results(local, n_units_total, rank=None, world_size=None, group=None):
    """All-gather per-unit result rows.  ``local`` [n_local, ...] holds this rank's units in shard order; the return value
    [n_units_total, ...] is in the global unit order on every rank.  Convenience form for SMALL rows (it pads ``local`` into
    an equal-size send buffer and reorders the result: two copies of the payload); results of any size should be produced
    directly inside a ``GatherBuffer``."""
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


def sum_f64(x, out, workspace=None):
    """Deterministic float64 sum of the fp32 CUDA tensor ``x`` into the 1-element float64 tensor / view ``out`` (cf_sum_f64):
    one pass over ``x`` instead of ATen's cast copy + reduction, and a summation order that depends on ``x.numel()`` alone, so
    a unit's result row is bit-identical wherever it is recomputed.  Stream-ordered, capturable."""
    from . import _cabi
    _cabi.require_cuda(x, out)
    if x.dtype != torch.float32 or not x.is_contiguous() or out.dtype != torch.float64 or out.numel() != 1:
        raise ValueError("sum_f64: x must be a contiguous float32 tensor and out a 1-element float64 tensor")
    lib = _cabi.lib()
    nws = lib.cf_sum_f64_workspace_bytes()
    if workspace is None:
        workspace = torch.empty(nws // 8, dtype=torch.float64, device=x.device)
    with torch.cuda.device(x.device):
        _cabi.check(lib.cf_sum_f64(x.data_ptr(), x.numel(), out.data_ptr(), workspace.data_ptr(), workspace.numel() * 8,
                                   _cabi.stream()), "cf_sum_f64")
    return out


# ---------------------------------------------------------------------------------------------------------- lanes
# Inside one GPU the same independence is worth ~14 % (tools/lanes.py, profiles/r1_lanes_experiment.log): a cine unit of
# 29 pairs fills the 148 SMs for only ~3 waves per lookup, so every launch pays a latency ramp (coords -> windows -> first
# stores) and a drain during which HBM idles.  Two sub-batches of the pairs, each a serial chain build -> 12 lookups -> warp on
# its own stream, fill each other's ramps and drains.  Nothing is reordered inside a pair and results are bit-identical
# (every kernel of the path treats batch items independently).

def lane_bounds(n_items, n_lanes):
    """Contiguous, near-equal split of ``n_items`` batch items over ``n_lanes`` lanes: [(begin, end), ...]; lanes that
    would be empty are dropped."""
    if n_items < 0 or n_lanes < 1:
        raise ValueError(f"lane_bounds({n_items}, {n_lanes})")
    cuts = [(n_items * j) // n_lanes for j in range(n_lanes + 1)]
    return [(a, b) for a, b in zip(cuts[:-1], cuts[1:]) if b > a]


class Lanes:
    """Run independent sub-batches on parallel CUDA streams, fork/joined on the current stream.

        lanes = Lanes(2, device)
        lanes.run(n_pairs, lambda j, a, b: chain(f1[a:b], f2[a:b], ...))      # returns the per-lane results

    The fork and the join are event dependencies (``wait_stream``), so ``run`` is stream-ordered like a single kernel
    launch and can be captured into one CUDA graph (the graph gets parallel branches).  With one lane ``fn`` runs inline
    on the current stream.
    """

    def __init__(self, n_lanes, device=None):
        if n_lanes < 1:
            raise ValueError("n_lanes must be >= 1")
        if not torch.cuda.is_available():
            raise RuntimeError("Lanes needs a CUDA device: the path has no CPU fallback")
        self.n_lanes = int(n_lanes)
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        self.streams = [torch.cuda.Stream(device=self.device) for _ in range(self.n_lanes)] if self.n_lanes > 1 else []

    def run(self, n_items, fn):
        bounds = lane_bounds(n_items, self.n_lanes)
        if len(bounds) <= 1 or not self.streams:
            return [fn(j, a, b) for j, (a, b) in enumerate(bounds)]
        cur = torch.cuda.current_stream(self.device)
        results = []
        for (j, (a, b)), s in zip(enumerate(bounds), self.streams):
            s.wait_stream(cur)
            with torch.cuda.stream(s):
                results.append(fn(j, a, b))
        for s in self.streams[: len(bounds)]:
            cur.wait_stream(s)
        return results
