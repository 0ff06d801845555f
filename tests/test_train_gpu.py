"""GPU tests of the training step: fused AdamW(amsgrad) + clipping vs torch.optim.AdamW +
clip_grad_norm_ (scripts/train_utils.py:38-44, scripts/train_main.py:91), and Trainer.training_step on synthetic lattices."""
import copy

import pytest
import torch

from energy_equiv_lattice_gnn_b200.data import synthetic_batch
from energy_equiv_lattice_gnn_b200.train import FlatAdamW, Trainer, random_psd_targets
from tests.util import cuda_model_from, make_pair

pytestmark = pytest.mark.gpu


class _Tiny(torch.nn.Module):
    def __init__(self):
        super().__init__()
        g = torch.Generator().manual_seed(3)
        self.a = torch.nn.Parameter(torch.randn(37, 5, generator=g))
        self.b = torch.nn.Parameter(torch.randn(1001, generator=g))


@pytest.mark.parametrize("amsgrad,clip", [(True, 10.0), (False, 0.05), (True, None)])
def test_flat_adamw_matches_torch(cuda_device, amsgrad, clip):
    ours, ref = _Tiny().to(cuda_device), _Tiny().to(cuda_device)
    opt = FlatAdamW(ours, lr=1e-2, beta1=0.9, eps=1e-8, weight_decay=1e-3, amsgrad=amsgrad, clip_norm=clip)
    topt = torch.optim.AdamW(ref.parameters(), lr=1e-2, betas=(0.9, 0.999), eps=1e-8, amsgrad=amsgrad, weight_decay=1e-3)
    g = torch.Generator().manual_seed(9)
    for step in range(5):
        grads = [torch.randn(p.shape, generator=g) * (3.0 if step % 2 else 0.01) for p in ref.parameters()]
        opt.zero_grad()
        for p, q, gr in zip(ours.parameters(), ref.parameters(), grads):
            p.grad.add_(gr.to(cuda_device))
            q.grad = gr.to(cuda_device).clone()
        if clip:
            torch.nn.utils.clip_grad_norm_(ref.parameters(), clip)
        topt.step()
        opt.step()
        for p, q in zip(ours.parameters(), ref.parameters()):
            assert torch.allclose(p, q, rtol=2e-5, atol=2e-6), (step, float((p - q).abs().max()))


def test_training_step_reduces_loss_and_refreshes_cached_tables(cuda_device):
    n = 8
    batch, rmax = synthetic_batch(range(8000, 8000 + n))
    oracle, params = make_pair(rmax)
    model = cuda_model_from(oracle, params, cuda_device)
    batch = batch.to(cuda_device)
    batch.stiffness = random_psd_targets(n, seed=1).to(cuda_device)
    tr = Trainer(model, lr=1e-3)
    losses = [float(tr.training_step(batch)) for _ in range(12)]
    assert all(l == l for l in losses)          # finite
    assert min(losses[6:]) < losses[0], losses    # the step moves downhill on a fixed batch
    # inference after training uses the UPDATED weights (cached SC / radial tables were dropped)
    model.eval()
    with torch.no_grad():
        a = model(batch)["stiffness"]
    fresh = copy.deepcopy(model)
    fresh._cache.clear()
    with torch.no_grad():
        b = fresh(batch)["stiffness"]
    assert torch.equal(a, b)


def _trained_flat(cuda_device, seeds_list, steps, **kw):
    """Run `steps` optimizer steps; each step consumes len(seeds_list) micro-batches.  -> flat parameter vector."""
    all_seeds = [s for seeds in seeds_list for s in seeds]
    _, rmax = synthetic_batch(all_seeds)
    oracle, params = make_pair(rmax)
    model = cuda_model_from(oracle, params, cuda_device)
    tr = Trainer(model, lr=1e-3, **kw)
    tg = random_psd_targets(len(all_seeds), seed=4)
    off = 0
    batches = []
    for seeds in seeds_list:
        b, _ = synthetic_batch(seeds)
        b = b.to(cuda_device)
        b.stiffness = tg[off:off + len(seeds)].to(cuda_device)
        off += len(seeds)
        batches.append(b)
    for _ in range(steps):
        for b in batches:
            tr.training_step(b)
    return tr.opt.flat.clone(), tr.opt.t


def test_gradient_accumulation_equals_one_step_on_the_joint_batch(cuda_device):
    """accumulate_grad_batches (scripts/train_main.py:91): k equal micro-batches, loss / k each, ONE optimizer step ==
    one step on the concatenated batch (mean of equal-sized means = global mean)."""
    whole, t1 = _trained_flat(cuda_device, [list(range(8100, 8108))], steps=3)
    acc, t2 = _trained_flat(cuda_device, [list(range(8100, 8104)), list(range(8104, 8108))], steps=3, accumulate_grad_batches=2)
    assert t1 == t2 == 3
    err = float((whole - acc).abs().max() / whole.abs().max())
    assert err < 2e-5, err
    # and it is NOT what stepping on every micro-batch gives
    each, t3 = _trained_flat(cuda_device, [list(range(8100, 8104)), list(range(8104, 8108))], steps=3)
    assert t3 == 6 and float((whole - each).abs().max() / whole.abs().max()) > 1e-4


def test_flat_adamw_without_trainer_invalidates_cached_tables(cuda_device):
    """FlatAdamW.step rewrites the weights through the C ABI (no torch version bump): the SC / radial tables cached by the
    model must be dropped by the optimizer itself."""
    n = 4
    batch, rmax = synthetic_batch(range(8200, 8200 + n))
    oracle, params = make_pair(rmax)
    model = cuda_model_from(oracle, params, cuda_device).train()
    batch = batch.to(cuda_device)
    target = random_psd_targets(n, seed=2).to(cuda_device)
    opt = FlatAdamW(model, lr=1e-2)
    from energy_equiv_lattice_gnn_b200.train import relative_mse_loss
    for _ in range(2):
        opt.zero_grad()
        relative_mse_loss(model(batch)["stiffness"], target).backward()
        opt.gather_grads()
        opt.step()
    model.eval()
    with torch.no_grad():
        a = model(batch)["stiffness"]
    fresh = copy.deepcopy(model)
    fresh._cache.clear()
    with torch.no_grad():
        b = fresh(batch)["stiffness"]
    assert torch.equal(a, b)


def test_backward_is_bitwise_repeatable(cuda_device):
    n = 6
    batch, rmax = synthetic_batch(range(8300, 8300 + n))
    oracle, params = make_pair(rmax)
    model = cuda_model_from(oracle, params, cuda_device).train()
    batch = batch.to(cuda_device)
    target = random_psd_targets(n, seed=3).to(cuda_device)
    from energy_equiv_lattice_gnn_b200.train import relative_mse_loss
    runs = []
    for _ in range(2):
        model.zero_grad(set_to_none=True)
        relative_mse_loss(model(batch)["stiffness"], target).backward()
        runs.append({k: p.grad.clone() for k, p in model.named_parameters()})
    bad = [k for k in runs[0] if not torch.equal(runs[0][k], runs[1][k])]
    assert not bad, bad
