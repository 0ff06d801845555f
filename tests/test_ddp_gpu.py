"""Multi-GPU training parity (SURVEY.md section 7 test_ddp_gpu).  Needs >= 2 CUDA devices
(`gpurun --gpus 2 -- python -m pytest tests/test_ddp_gpu.py -m gpu`); skipped on a 1-GPU box.

World-size-2 NCCL run of `train.Trainer` on the two halves of one global batch vs ONE GPU on the whole batch: after 5
optimizer steps the flat parameter vectors agree to fp32 round-off and the logged losses average to the global loss
(per-rank mean loss + DDP gradient mean == global mean for equal shards, scripts/train_utils.py:57-60).  Also runs the
model on cuda:1 without torch.cuda.set_device (device guard of native.py)."""
import os
import sys

import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pytestmark = pytest.mark.gpu
SEEDS = list(range(8400, 8416))
STEPS = 5


def _need_two():
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 CUDA devices")


def _make(device, seeds, tg):
    sys.path.insert(0, ROOT)
    from energy_equiv_lattice_gnn_b200.data import synthetic_batch
    from tests.util import cuda_model_from, make_pair
    _, rmax = synthetic_batch(SEEDS)
    oracle, params = make_pair(rmax)
    model = cuda_model_from(oracle, params, device)
    b, _ = synthetic_batch(seeds)
    b = b.to(device)
    b.stiffness = tg.to(device)
    return model, b


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    torch.cuda.set_device(rank)
    torch.distributed.init_process_group("nccl", rank=rank, world_size=world)
    sys.path.insert(0, ROOT)
    from energy_equiv_lattice_gnn_b200.train import Trainer, random_psd_targets
    tg = random_psd_targets(len(SEEDS), seed=6)
    half = len(SEEDS) // world
    model, b = _make(torch.device("cuda", rank), SEEDS[rank * half:(rank + 1) * half], tg[rank * half:(rank + 1) * half])
    tr = Trainer(model, lr=1e-3)
    losses = [tr.training_step(b) for _ in range(STEPS)]
    ls = torch.stack(losses)
    torch.distributed.all_reduce(ls)
    if rank == 0:
        torch.save((tr.opt.flat.cpu(), (ls / world).cpu()), out)
    torch.distributed.barrier()
    torch.distributed.destroy_process_group()


def test_two_gpu_trainer_matches_single_gpu(tmp_path):
    _need_two()
    sys.path.insert(0, ROOT)
    from energy_equiv_lattice_gnn_b200.train import Trainer, random_psd_targets
    out = str(tmp_path / "ddp.pt")
    mp.spawn(_worker, args=(2, 29581, out), nprocs=2, join=True)
    flat2, loss2 = torch.load(out)
    tg = random_psd_targets(len(SEEDS), seed=6)
    model, b = _make(torch.device("cuda:0"), SEEDS, tg)
    tr = Trainer(model, lr=1e-3)
    loss1 = torch.stack([tr.training_step(b) for _ in range(STEPS)]).cpu()
    flat1 = tr.opt.flat.cpu()
    assert float((flat1 - flat2).abs().max() / flat1.abs().max()) < 2e-5
    assert float((loss1 - loss2).abs().max() / loss1.abs().max()) < 1e-4


def test_model_on_second_device_without_set_device():
    _need_two()
    sys.path.insert(0, ROOT)
    from energy_equiv_lattice_gnn_b200.data import synthetic_batch
    from tests.util import cuda_model_from, make_pair
    assert torch.cuda.current_device() == 0
    batch, rmax = synthetic_batch(range(8500, 8504))
    oracle, params = make_pair(rmax)
    outs = []
    for d in (0, 1):
        model = cuda_model_from(oracle, params, torch.device("cuda", d))
        with torch.no_grad():
            outs.append(model(batch.to(torch.device("cuda", d)))["stiffness"])
        assert outs[-1].device.index == d
    assert torch.equal(outs[0].cpu(), outs[1].cpu())
    assert torch.cuda.current_device() == 0
