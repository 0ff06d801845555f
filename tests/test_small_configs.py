"""Constructor domain: `gnn/model.py:27-90` accepts any lmax / correlation / radial-MLP shape.
The reference itself produced tests/golden/ref_forward_small.pt (oracle/make_golden_small.py, fp64) at lmax 2 / correlation 2,
lmax 1 / correlation 1, lmax 4 / correlation 2 with a 32-wide 2-layer radial MLP and 10 edge bases, and lmax 3 with a
radial MLP of `inter_MLP_layers=1` (which the reference builds as two linears, gnn/blocks.py:541-548).  CPU: the oracle reproduces them and the CUDA module accepts the reference state_dict; GPU: forward
within 1e-5 and parameter gradients within 3e-5 of the fp64 oracle's autograd."""
import os
from argparse import Namespace

import pytest
import torch

from energy_equiv_lattice_gnn_b200 import EnergyEquivGNN
from energy_equiv_lattice_gnn_b200.data import Batch
from energy_equiv_lattice_gnn_b200.train import random_psd_targets, relative_mse_loss
from oracle.model import EnergyEquivGNN as OracleGNN

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = ["lmax2_nu2", "lmax1_nu1", "lmax4_nu2_mlp32", "lmax3_nu3_mlp1"]


def _load(name):
    return torch.load(os.path.join(GOLD, "ref_forward_small.pt"), weights_only=False)[name]


def _oracle(fx):
    old = torch.get_default_dtype()
    torch.set_default_dtype(torch.float64)
    try:
        m = OracleGNN(Namespace(**fx["params"]))
    finally:
        torch.set_default_dtype(old)
    sd = {k: v.double() if v.is_floating_point() else v for k, v in fx["state_dict"].items()}
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected and all("U_matrix" in k for k in missing), (missing, unexpected)
    b = Batch(**fx["batch"])
    for k in ("positions", "shifts", "edge_attr", "node_attrs"):
        setattr(b, k, getattr(b, k).double())
    return m, b


@pytest.mark.parametrize("name", CASES)
def test_oracle_and_state_dict(name):
    fx = _load(name)
    m, b = _oracle(fx)
    with torch.no_grad():
        out = m(b)["stiffness"]
    assert float((out - fx["stiffness"]).abs().max() / fx["stiffness"].abs().max()) < 1e-12
    ours = EnergyEquivGNN(Namespace(**fx["params"]))
    missing, unexpected = ours.load_state_dict(fx["state_dict"], strict=False)
    assert not missing and not unexpected, (missing, unexpected)


@pytest.mark.gpu
@pytest.mark.parametrize("name", CASES)
def test_forward_and_gradients_on_gpu(cuda_device, name):
    fx = _load(name)
    model = EnergyEquivGNN(Namespace(**fx["params"]))
    model.load_state_dict(fx["state_dict"], strict=False)
    model = model.to(cuda_device)
    batch = Batch(**fx["batch"]).to(cuda_device)
    model.eval()
    with torch.no_grad():
        out = model(batch)["stiffness"].cpu().double()
    ref = fx["stiffness"]
    assert float((out - ref).abs().max() / ref.abs().max()) < 1e-5
    # gradients against the fp64 oracle's autograd
    m, b = _oracle(fx)
    target = random_psd_targets(ref.shape[0], seed=8)
    relative_mse_loss(m(b)["stiffness"], target.double()).backward()
    want = {k: p.grad for k, p in m.named_parameters()}
    model.train()
    relative_mse_loss(model(batch)["stiffness"], target.to(cuda_device)).backward()
    worst = {}
    for k, p in model.named_parameters():
        assert p.grad is not None, k
        worst[k] = float((p.grad.double().cpu() - want[k]).abs().max() / max(float(want[k].abs().max()), 1e-30))
    bad = {k: v for k, v in worst.items() if not v < 3e-5}
    assert not bad, bad
