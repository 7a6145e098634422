"""Host-side sharding logic and the one collective of the path, on CPU (gloo, world_size 2 and 3).

The path shards by (patient, slice) cine with no exchange step; the only collective is the
final gather of compact per-unit results.  On the GPU box the same code runs over NCCL.
"""
import os
import socket
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _unit_result(patient, slice_idx, width):
    """Deterministic stand-in for a unit's compact result row (a metric vector), from the unit's seed."""
    from cardioflow_b200.shard import unit_seed
    g = torch.Generator().manual_seed(unit_seed(patient, slice_idx))
    return torch.randn(width, generator=g)


def _worker(rank, world, port, n_patients, n_slices, width, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from cardioflow_b200.shard import cohort_units, gather_results, shard_units
        units = cohort_units(n_patients, n_slices)
        mine = shard_units(units, rank, world)
        local = torch.stack([_unit_result(p, s, width) for p, s in mine]) if mine else torch.empty(0, width)
        out = gather_results(local, len(units))
        want = torch.stack([_unit_result(p, s, width) for p, s in units])
        q.put((rank, bool(torch.equal(out, want)), len(mine)))
    finally:
        dist.destroy_process_group()


def _worker_buffer(rank, world, port, n_units, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from cardioflow_b200.shard import GatherBuffer, cohort_units, shard_units
        units = cohort_units(n_units, 1)
        mine = shard_units(units, rank, world)
        buf = GatherBuffer(len(units), (3, 5), torch.uint8, "cpu")
        assert buf.local.data_ptr() == buf.recv[rank].data_ptr()            # results land in the send slice itself: no staging copy
        for j, (p, _) in enumerate(mine):
            buf.local[j].fill_(p + 1)                                         # "the kernel writes its label map here"
        table = buf.gather()
        ok = all(bool((buf.unit(table, k) == units[k][0] + 1).all()) for k in range(len(units)))
        ok = ok and table.data_ptr() == buf.recv.data_ptr()                  # the gathered table is a view of the receive buffer
        q.put((rank, ok, len(mine)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n_units", [(2, 7), (2, 8), (3, 10)])
def test_gather_buffer_is_in_place_and_ordered(world, n_units):
    """GatherBuffer: per-unit results (uint8 label maps on the GPU) are written straight into this rank's slice of the all-gather
    buffer; the collective is in place; global unit order is recovered by indexing, not by copying."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker_buffer, args=(r, world, port, n_units, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in results)
    assert sum(n for _, _, n in results) == n_units


@pytest.mark.parametrize("world,n_patients,n_slices", [(2, 3, 3), (2, 4, 2), (3, 5, 2)])
def test_gather_restores_global_unit_order(world, n_patients, n_slices):
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n_patients, n_slices, 6, q)) for r in range(world)]
    for p in procs:
        p.start()
    results = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert all(ok for _, ok, _ in results)
    counts = sorted(n for _, _, n in results)
    assert sum(counts) == n_patients * n_slices and counts[-1] - counts[0] <= 1    # round-robin balance


def test_shard_units_partition_and_seeds():
    from cardioflow_b200.shard import cohort_units, shard_units, unit_seed
    units = cohort_units(150, 10)
    assert len(units) == 1500
    for world in (1, 2, 4, 8):
        shards = [shard_units(units, r, world) for r in range(world)]
        assert sorted(u for s in shards for u in s) == units
        assert max(map(len, shards)) - min(map(len, shards)) <= 1
    assert len({unit_seed(p, s) for p, s in units}) == 1500                      # SURVEY.md 8(d) cfg 5 seeds
    with pytest.raises(ValueError):
        shard_units(units, 2, 2)


def test_single_process_gather_is_identity():
    from cardioflow_b200.shard import gather_results
    x = torch.arange(12.0).view(4, 3)
    assert torch.equal(gather_results(x, 4, rank=0, world_size=1), x)


def test_lane_bounds_partition():
    from cardioflow_b200 import lane_bounds
    assert lane_bounds(29, 2) == [(0, 14), (14, 29)]
    assert lane_bounds(29, 1) == [(0, 29)]
    assert lane_bounds(3, 4) == [(0, 1), (1, 2), (2, 3)]          # empty lanes are dropped
    assert lane_bounds(0, 2) == []
    for n in (1, 7, 29, 64, 300):
        for k in (1, 2, 3, 4, 8):
            b = lane_bounds(n, k)
            assert b[0][0] == 0 and b[-1][1] == n and all(x[1] == y[0] for x, y in zip(b, b[1:]))
            sizes = [e - a for a, e in b]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        lane_bounds(4, 0)


def test_lanes_refuse_to_run_without_a_gpu():
    from cardioflow_b200 import Lanes
    if torch.cuda.is_available():
        pytest.skip("CPU-only check")
    with pytest.raises(RuntimeError):
        Lanes(2)
