"""N > 1 host path on CPU: lattice sharding (edge-balanced contiguous ranges, no data-path collective) and the optional
[G,6,6] all-gather, with world_size 2 over gloo.  The per-rank compute here is the oracle (the CUDA module has no CPU
path); what is tested is that sharded execution reproduces the unsharded result in the original lattice order."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from energy_equiv_lattice_gnn_b200.data import collate, synthetic_lattice
from energy_equiv_lattice_gnn_b200.sharding import balanced_ranges, gather_stiffness, shard_batch


def test_balanced_ranges_cover_and_balance():
    costs = [120] * 10 + [864] * 3 + [6, 6, 48]
    for world in (1, 2, 3, 4, 8, 20):
        r = balanced_ranges(costs, world)
        assert len(r) == world and r[0][0] == 0 and r[-1][1] == len(costs)
        assert all(a[1] == b[0] for a, b in zip(r, r[1:])) and all(a <= b for a, b in r)
    r = balanced_ranges([120] * 1024, 8)
    assert [b - a for a, b in r] == [128] * 8
    assert balanced_ranges([], 2) == [(0, 0), (0, 0)]


def test_shard_batch_partitions_graph():
    data = [synthetic_lattice(s, n) for s, n in ((1, 3), (2, 9), (3, 4), (4, 1), (5, 6))]
    b = collate(data)
    seen_nodes = seen_edges = 0
    for rank in range(3):
        sb, (g0, g1) = shard_batch(b, rank, 3)
        ref = collate(data[g0:g1]) if g1 > g0 else None
        if ref is None:
            assert sb.num_graphs == 0
            continue
        for k in ("positions", "edge_index", "shifts", "edge_attr", "batch", "node_attrs"):
            assert torch.equal(getattr(sb, k), getattr(ref, k)), k
        seen_nodes += sb.positions.shape[0]
        seen_edges += sb.edge_index.shape[1]
    assert seen_nodes == b.positions.shape[0] and seen_edges == b.edge_index.shape[1]


def _worker(rank, world, port, q):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from tests.test_oracle_properties import small_oracle, to64
    data = [synthetic_lattice(s, n) for s, n in ((21, 4), (22, 7), (23, 3), (24, 5), (25, 2))]
    b = collate(data)
    m = small_oracle(float(b.edge_attr.max()))
    sb, (g0, g1) = shard_batch(b, rank, world)
    with torch.no_grad():
        local = m(to64(sb))["stiffness"] if g1 > g0 else torch.zeros(0, 6, 6, dtype=torch.float64)
    counts = [torch.tensor([g1 - g0]) for _ in range(world)]
    dist.all_gather(counts, torch.tensor([g1 - g0]))
    full = gather_stiffness(local, [int(c) for c in counts])
    # validation scalars are averaged over ranks with one all-reduce (Lightning sync_dist=True, train_utils.py:87-88)
    from energy_equiv_lattice_gnn_b200.metrics import sync_validation_scalars
    vs = sync_validation_scalars(torch.tensor([rank + 1.0, 10.0 * (rank + 1)]))
    assert torch.allclose(vs, torch.tensor([(world + 1) / 2.0, 5.0 * (world + 1)]))
    if rank == 0:
        with torch.no_grad():
            ref = m(to64(b))["stiffness"]
        q.put(float((full - ref).abs().max()))
    dist.destroy_process_group()


def test_sharded_equals_unsharded_world2():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    err = q.get(timeout=240)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert err < 1e-12
