"""Lattice sharding for one-process-per-GPU data parallelism.

Lattices are independent samples, so the path shards by lattice with NO data-path collective: every rank takes a
contiguous range of lattices (contiguous graph range => contiguous node / edge ranges of the collated batch), balanced
by directed-edge count (the message kernels' cost driver).  Inference needs no communication; `gather_stiffness` is the
optional [G,6,6] all-gather (144 B per lattice) for reporting.  The reference gets the same decomposition implicitly from
Lightning DDP + DistributedSampler (scripts/train_main.py:89-100, accelerator='auto').
"""
from __future__ import annotations

from typing import List, Sequence, Tuple

import torch

from .data import Batch


def balanced_ranges(costs: Sequence[int], world: int) -> List[Tuple[int, int]]:
    """Split items 0..len(costs) into `world` contiguous ranges with near-equal total cost: rank r ends at the prefix
    position closest to r/world of the total.  Every item is in exactly one range; ranges can be empty."""
    n = len(costs)
    prefix = [0]
    for c in costs:
        prefix.append(prefix[-1] + int(c))
    total = prefix[-1]
    cuts = [0]
    for r in range(1, world):
        target = total * r / world
        i = cuts[-1]
        while i < n and prefix[i + 1] <= target:
            i += 1
        if i < n and prefix[i + 1] - target < target - prefix[i]:
            i += 1
        cuts.append(max(i, cuts[-1]))
    cuts.append(n)
    return [(cuts[r], cuts[r + 1]) for r in range(world)]


def shard_batch(batch: Batch, rank: int, world: int) -> Tuple[Batch, Tuple[int, int]]:
    """The sub-batch of graphs owned by `rank` (edge-balanced contiguous graph range) and that range."""
    G = int(batch.num_graphs)
    node_graph = batch.batch
    edge_graph = node_graph[batch.edge_index[1]]
    edges_per_graph = torch.bincount(edge_graph, minlength=G).tolist()
    g0, g1 = balanced_ranges(edges_per_graph, world)[rank]
    nmask = (node_graph >= g0) & (node_graph < g1)
    emask = (edge_graph >= g0) & (edge_graph < g1)
    n_idx = torch.nonzero(nmask).flatten()
    n0 = int(n_idx[0]) if n_idx.numel() else 0
    out = Batch(positions=batch.positions[nmask], node_attrs=batch.node_attrs[nmask], edge_index=batch.edge_index[:, emask] - n0,
                shifts=batch.shifts[emask], edge_attr=batch.edge_attr[emask], batch=node_graph[nmask] - g0, num_graphs=g1 - g0)
    if hasattr(batch, "stiffness"):
        out.stiffness = batch.stiffness[g0:g1]
    return out, (g0, g1)


def gather_stiffness(local: torch.Tensor, counts: Sequence[int], group=None) -> torch.Tensor:
    """All-gather per-rank [G_r,6,6] results into the global [G,6,6] order (ranges are contiguous and ordered by rank)."""
    import torch.distributed as dist
    world = dist.get_world_size(group)
    gmax = max(counts)
    pad = torch.zeros((gmax,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    bufs = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(bufs, pad, group=group)
    return torch.cat([b[:c] for b, c in zip(bufs, counts)], 0)
