"""GPU parity at the other BASELINE.json workloads: config 4 (lmax 3, 128 channels) and config 5
(216-node lattices).  The CUDA path (through the C ABI) is compared with the fp64 CPU oracle on a few lattices (the oracle's
dense contraction makes more than that slow) and checked through size-independent properties on larger batches."""
import pytest
import torch

from energy_equiv_lattice_gnn_b200.data import collate, synthetic_batch, synthetic_lattice
from tests.util import batch_to, cuda_model_from, make_pair

pytestmark = pytest.mark.gpu
RTOL = 1e-5

CONFIG4 = dict(lmax=3, hidden_irreps="128x0e+128x1o+128x2e+128x3o", readout_irreps="64x0e+64x1o+64x2e+64x3o")


def rel_err(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max())


def properties(C):
    assert torch.isfinite(C).all()
    assert float((C - C.transpose(1, 2)).abs().max()) <= 1e-6 * float(C.abs().max())
    assert float(torch.linalg.eigvalsh(C.double().cpu()).min()) > -1e-6 * float(C.abs().max())


def test_config4_lmax3_c128(cuda_device):
    small, rmax = synthetic_batch(range(20000, 20003))
    oracle, params = make_pair(rmax, dtype=torch.float64, seed=4, **CONFIG4)
    ref = oracle(batch_to(small, torch.float64))["stiffness"].detach()
    model = cuda_model_from(oracle, params, cuda_device)
    with torch.no_grad():
        out = model(small.to(cuda_device))["stiffness"].cpu()
    assert out.shape == ref.shape == (3, 6, 6)
    assert rel_err(out, ref) < RTOL, rel_err(out, ref)
    # full config-4 size: 256 lattices per GPU
    big, _ = synthetic_batch(range(20000, 20256))
    b = big.to(cuda_device)
    with torch.no_grad():
        C1 = model(b)["stiffness"].clone()
        C2 = model(b)["stiffness"].clone()
    assert C1.shape == (256, 6, 6) and torch.equal(C1, C2)
    properties(C1)
    assert rel_err(C1[:3].cpu(), ref) < RTOL  # batch independence: the first three lattices are the small batch


def test_config5_216_node_lattices(cuda_device):
    data = [synthetic_lattice(30000 + i, 216) for i in range(6)]
    rmax = max(float(d.edge_attr.max()) for d in data)
    oracle, params = make_pair(rmax, dtype=torch.float64, seed=5)
    small = collate(data[:1])
    ref = oracle(batch_to(small, torch.float64))["stiffness"].detach()
    model = cuda_model_from(oracle, params, cuda_device)
    # 96 lattices (20 736 nodes, 82 944 edges): the six distinct lattices tiled, so every 6th result must repeat
    big = collate([data[i % 6] for i in range(96)]).to(cuda_device)
    with torch.no_grad():
        C1 = model(big)["stiffness"]
    assert C1.shape == (96, 6, 6)
    properties(C1)
    assert rel_err(C1[:1].cpu(), ref) < RTOL, rel_err(C1[:1].cpu(), ref)
    assert rel_err(C1[6:12], C1[:6]) < 1e-6 and rel_err(C1[90:96], C1[:6]) < 1e-6
