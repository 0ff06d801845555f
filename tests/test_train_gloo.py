"""World-size-2 `gloo` test (CPU) of the data-parallel training exchange: each rank back-propagates the training loss on
its shard of lattices through the oracle, `train.sync_gradients` all-reduces ONE flat buffer, and gradient / world equals
the single-process gradient on the whole batch (equal shards; scripts/train_utils.py:57-60 + Lightning DDP semantics)."""
import os
import sys

import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _flat_grad(seeds, rmax, targets):
    sys.path.insert(0, ROOT)
    from energy_equiv_lattice_gnn_b200.data import synthetic_batch
    from energy_equiv_lattice_gnn_b200.train import relative_mse_loss
    from tests.util import batch_to, make_pair
    batch, _ = synthetic_batch(seeds)
    oracle, _ = make_pair(rmax, dtype=torch.float64)
    loss = relative_mse_loss(oracle(batch_to(batch, torch.float64))["stiffness"], targets)
    loss.backward()
    return torch.cat([p.grad.reshape(-1) for p in oracle.parameters()])


def _worker(rank, world, port, rmax, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.distributed.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, ROOT)
    from energy_equiv_lattice_gnn_b200.train import random_psd_targets, sync_gradients
    torch.set_num_threads(2)
    t = random_psd_targets(4, seed=2).double()
    g = _flat_grad(range(9000 + 2 * rank, 9002 + 2 * rank), rmax, t[2 * rank:2 * rank + 2])
    w = sync_gradients(g)
    if rank == 0:
        torch.save((w, g / w), out)
    torch.distributed.destroy_process_group()


def test_flat_gradient_allreduce_equals_global_batch_gradient(tmp_path):
    sys.path.insert(0, ROOT)
    from energy_equiv_lattice_gnn_b200.data import synthetic_batch
    from energy_equiv_lattice_gnn_b200.train import random_psd_targets
    _, rmax = synthetic_batch(range(9000, 9004))
    out = str(tmp_path / "g.pt")
    mp.spawn(_worker, args=(2, 29571, rmax, out), nprocs=2, join=True)
    world, g = torch.load(out)
    ref = _flat_grad(range(9000, 9004), rmax, random_psd_targets(4, seed=2).double())
    assert world == 2
    assert float((g - ref).abs().max() / ref.abs().max()) < 1e-10
