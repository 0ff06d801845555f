"""RotateLat (scripts/train_utils.py:114-146): oracle restatement vs the reference's own class (golden fixture made by
oracle/make_golden_rotate.py), and the device kernels vs both."""
import os

import pytest
import torch

from energy_equiv_lattice_gnn_b200.data import Batch, LatticeData, RotateLat, collate, rand_rotations
from oracle.rotate import cart4_to_mandel, rotate_batch

GOLD = os.path.join(os.path.dirname(__file__), "golden", "ref_rotate.pt")


def _cases():
    return torch.load(GOLD, weights_only=False)


def _rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max().clamp_min(1e-300))


def test_oracle_matches_reference_rotatelat():
    for c in _cases():
        n = c["positions"].shape[0]
        pos, sh, C, S = rotate_batch(c["Q"][None], c["positions"].double(), c["shifts"].double(), c["edge_index"],
                                     torch.zeros(n, dtype=torch.int64), c["stiffness_cart4"].double(), c["compliance_cart4"].double())
        assert _rel(pos, c["out_positions"]) < 1e-13
        assert _rel(sh, c["out_shifts"]) < 1e-13
        assert _rel(C, c["out_stiffness"]) < 1e-13
        assert _rel(S, c["out_compliance"]) < 1e-13
        assert _rel(cart4_to_mandel(c["stiffness_cart4"].double()), c["plain_stiffness"]) == 0.0
        assert _rel(cart4_to_mandel(c["compliance_cart4"].double()), c["plain_compliance"]) == 0.0


def test_rand_rotations_are_proper():
    Q = rand_rotations(64, torch.Generator().manual_seed(3), dtype=torch.float64)
    eye = torch.eye(3, dtype=torch.float64).expand(64, 3, 3)
    assert float((Q @ Q.transpose(1, 2) - eye).abs().max()) < 1e-12
    assert float((torch.linalg.det(Q) - 1).abs().max()) < 1e-12


def test_oracle_rotation_properties():
    g = torch.Generator().manual_seed(0)
    Q = rand_rotations(4, g, dtype=torch.float64)
    A = torch.randn(4, 6, 6, generator=g, dtype=torch.float64)
    # a symmetric Mandel matrix keeps its eigenvalues under rotation of the underlying rank-4 tensor
    c = _cases()[0]
    C4 = c["stiffness_cart4"].double().expand(4, 3, 3, 3, 3).contiguous()
    n = torch.zeros(1, dtype=torch.int64)
    _, _, C, _ = rotate_batch(Q, torch.zeros(4, 3, dtype=torch.float64), torch.zeros(0, 3, dtype=torch.float64),
                              torch.zeros(2, 0, dtype=torch.int64), torch.arange(4), C4)
    ev0 = torch.linalg.eigvalsh(cart4_to_mandel(C4))
    ev1 = torch.linalg.eigvalsh(C)
    assert float((ev0 - ev1).abs().max() / ev0.abs().max()) < 1e-12
    del A, n


def _golden_batch():
    cases = _cases()
    lats = [LatticeData(positions=c["positions"], node_attrs=torch.ones(c["positions"].shape[0], 1), edge_index=c["edge_index"],
                        shifts=c["shifts"], edge_attr=torch.ones(c["shifts"].shape[0], 1)) for c in cases]
    b = collate(lats)
    b.stiffness = torch.cat([c["stiffness_cart4"] for c in cases], 0).float()
    b.compliance = torch.cat([c["compliance_cart4"] for c in cases], 0).float()
    Q = torch.stack([c["Q"] for c in cases], 0)
    return cases, b, Q


@pytest.mark.gpu
def test_device_rotatelat_matches_reference_golden():
    cases, b, Q = _golden_batch()
    dev = torch.device("cuda:0")
    out = RotateLat(rotate=True)(b.to(dev), Q.float())
    torch.cuda.synchronize()
    tol = 1e-5  # fp32 kernel vs the reference class run in fp64 (BASELINE.json north_star tolerance)
    assert _rel(out.positions.cpu(), torch.cat([c["out_positions"] for c in cases], 0)) < tol
    assert _rel(out.shifts.cpu(), torch.cat([c["out_shifts"] for c in cases], 0)) < tol
    assert _rel(out.stiffness.cpu(), torch.cat([c["out_stiffness"] for c in cases], 0)) < tol
    assert _rel(out.compliance.cpu(), torch.cat([c["out_compliance"] for c in cases], 0)) < tol
    plain = RotateLat(rotate=False)(b.to(dev))
    # no rotation: pure index picks and the {1, sqrt2, 2} mask
    assert _rel(plain.stiffness.cpu(), torch.cat([c["plain_stiffness"] for c in cases], 0)) < 2e-7
    assert plain.positions.data_ptr() == b.to(dev).positions.data_ptr() or torch.equal(plain.positions.cpu(), b.positions)
    with pytest.raises(AssertionError):
        RotateLat(rotate=False)(b.to(dev), Q.float())


@pytest.mark.gpu
def test_device_rotatelat_large_batch_vs_oracle_and_model_equivariance():
    from energy_equiv_lattice_gnn_b200.data import synthetic_batch
    from tests.util import cuda_model_from, make_pair
    dev = torch.device("cuda:0")
    G = 64
    batch, rmax = synthetic_batch(range(5000, 5000 + G))
    g = torch.Generator().manual_seed(11)
    C4 = torch.randn(G, 3, 3, 3, 3, generator=g)
    batch.stiffness = C4
    Q = rand_rotations(G, g)
    out = RotateLat()(batch.to(dev), Q)
    pos, sh, C, _ = rotate_batch(Q.double(), batch.positions.double(), batch.shifts.double(), batch.edge_index, batch.batch, C4.double())
    assert _rel(out.positions.cpu(), pos) < 1e-5
    assert _rel(out.shifts.cpu(), sh) < 1e-5
    assert _rel(out.stiffness.cpu(), C) < 1e-5
    # size-independent property: the model is equivariant, so the prediction on the rotated batch is the rotated prediction
    oracle, params = make_pair(rmax, dtype=torch.float64)
    model = cuda_model_from(oracle, params, dev)
    with torch.no_grad():
        c0 = model(batch.to(dev))["stiffness"]
        c1 = model(out)["stiffness"]
    ev0, ev1 = torch.linalg.eigvalsh(c0.double().cpu()), torch.linalg.eigvalsh(c1.double().cpu())
    assert float((ev0 - ev1).abs().max() / ev0.abs().max()) < 1e-4
