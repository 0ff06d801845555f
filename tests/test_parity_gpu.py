"""GPU parity: CUDA module vs the CPU oracle on identical synthetic lattices and weights (north-star tolerance:
fp32 relative 1e-5 on stiffness, measured against the fp64 oracle; the fp32 oracle itself is ~1e-6 from fp64)."""
import pytest
import torch

from energy_equiv_lattice_gnn_b200.data import synthetic_batch
from tests.util import batch_to, cuda_model_from, make_pair

pytestmark = pytest.mark.gpu
RTOL = 1e-5


def rel_err(a, b):
    return float((a - b).abs().max() / b.abs().max())


def test_forward_matches_oracle_config1(cuda_device):
    batch, rmax = synthetic_batch(range(1000, 1016))
    oracle, params = make_pair(rmax, dtype=torch.float64)
    ref = oracle(batch_to(batch, torch.float64))["stiffness"].detach()
    model = cuda_model_from(oracle, params, cuda_device)
    with torch.no_grad():
        out = model(batch.to(cuda_device))["stiffness"].cpu().double()
    assert out.shape == ref.shape == (16, 6, 6)
    assert rel_err(out, ref) < RTOL, rel_err(out, ref)
    # per-graph relative error too
    per = (out - ref).abs().amax(dim=(1, 2)) / ref.abs().amax(dim=(1, 2))
    assert float(per.max()) < 5 * RTOL, per


def test_bitwise_deterministic(cuda_device):
    batch, rmax = synthetic_batch(range(1000, 1004))
    oracle, params = make_pair(rmax)
    model = cuda_model_from(oracle, params, cuda_device)
    b = batch.to(cuda_device)
    with torch.no_grad():
        a1 = model(b)["stiffness"].clone()
        a2 = model(b)["stiffness"].clone()
    assert torch.equal(a1, a2)
