"""GPU parity: the CUDA module (through the C ABI) vs (a) golden vectors produced by running the reference's own sources
(tests/golden, oracle/make_golden.py) and (b) the CPU oracle on seeded synthetic lattices.  North-star tolerance: fp32
relative 1e-5 on stiffness (measured against fp64 truth); graph indexing bit-exact."""
import json
import os
from argparse import Namespace

import pytest
import torch

from energy_equiv_lattice_gnn_b200 import EnergyEquivGNN, native
from energy_equiv_lattice_gnn_b200.data import Batch, LatticeData, collate, synthetic_batch, synthetic_lattice
from tests.util import batch_to, cuda_model_from, make_pair

pytestmark = pytest.mark.gpu
RTOL = 1e-5
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def rel_err(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max())


def model_from_fixture(fx, device, **over):
    m = EnergyEquivGNN(Namespace(**{**fx["params"], **over}))
    missing, unexpected = m.load_state_dict(fx["state_dict"], strict=False)
    assert not missing and not unexpected, (missing, unexpected)
    return m.to(device).eval()


@pytest.mark.parametrize("name", ["lmax4", "lmax3"])
def test_forward_matches_reference_golden(cuda_device, name):
    fx = torch.load(os.path.join(GOLD, f"ref_forward_{name}.pt"), weights_only=False)
    model = model_from_fixture(fx, cuda_device)
    batch = Batch(**fx["batch"]).to(cuda_device)
    with torch.no_grad():
        out = model(batch)["stiffness"].cpu()
    ref = fx["stiffness"]
    assert out.shape == ref.shape
    assert rel_err(out, ref) < RTOL, rel_err(out, ref)
    per = (out.double() - ref).abs().amax(dim=(1, 2)) / ref.abs().amax(dim=(1, 2))
    assert float(per.max()) < 5 * RTOL, per  # includes the 1-node lattice made only of self-loops


def test_variants_match_reference_golden(cuda_device):
    fx = torch.load(os.path.join(GOLD, "ref_forward_lmax4.pt"), weights_only=False)
    var = torch.load(os.path.join(GOLD, "ref_variants.pt"), weights_only=False)
    batch = Batch(**fx["batch"]).to(cuda_device)
    n = 0
    for key, want in var.items():
        over = json.loads(key)
        model = model_from_fixture(fx, cuda_device, **over)
        with torch.no_grad():
            out = model(batch)["stiffness"].cpu()
        assert rel_err(out, want) < RTOL, (key, rel_err(out, want))
        n += 1
    assert n >= 6 and any('matrix_exp' in k for k in var)


def test_interaction_mean_matches_reference_golden(cuda_device):
    """interaction_reduction='mean' (gnn/blocks.py:595-597): eeg_scale_rows_inv_degree after the TP scatter; the batch holds a
    1-node lattice made of self-loops and nodes of different in-degree."""
    fx = torch.load(os.path.join(GOLD, "ref_forward_lmax4.pt"), weights_only=False)
    want = torch.load(os.path.join(GOLD, "ref_interaction_mean.pt"), weights_only=False)["stiffness"]
    model = model_from_fixture(fx, cuda_device, interaction_reduction="mean")
    with torch.no_grad():
        out = model(Batch(**fx["batch"]).to(cuda_device))["stiffness"].cpu()
    assert rel_err(out, want) < RTOL, rel_err(out, want)
    with pytest.raises(NotImplementedError):
        model_from_fixture(fx, cuda_device, interaction_reduction="pna")


def test_forward_matches_oracle_config1(cuda_device):
    batch, rmax = synthetic_batch(range(1000, 1016))
    oracle, params = make_pair(rmax, dtype=torch.float64)
    ref = oracle(batch_to(batch, torch.float64))["stiffness"].detach()
    model = cuda_model_from(oracle, params, cuda_device)
    with torch.no_grad():
        out = model(batch.to(cuda_device))["stiffness"].cpu().double()
    assert out.shape == ref.shape == (16, 6, 6)
    assert rel_err(out, ref) < RTOL, rel_err(out, ref)
    per = (out - ref).abs().amax(dim=(1, 2)) / ref.abs().amax(dim=(1, 2))
    assert float(per.max()) < 5 * RTOL, per


def test_sample_cat_real_lattices_match_oracle(cuda_device):
    cat = torch.load(os.path.join(GOLD, "ref_sample_cat.pt"), weights_only=False)
    data = [LatticeData(positions=d["positions"], node_attrs=d["node_attrs"], edge_index=d["edge_index"], shifts=d["shifts"],
                        edge_attr=d["edge_attr"], name=d["name"]) for d in cat[::3]]  # one density per lattice, N in {1..48}
    batch = collate(data)
    rmax = float(batch.edge_attr.max())
    oracle, params = make_pair(rmax, dtype=torch.float64, seed=5)
    ref = oracle(batch_to(batch, torch.float64))["stiffness"].detach()
    model = cuda_model_from(oracle, params, cuda_device)
    with torch.no_grad():
        out = model(batch.to(cuda_device))["stiffness"].cpu().double()
    per = (out - ref).abs().amax(dim=(1, 2)) / ref.abs().amax(dim=(1, 2))
    assert float(per.max()) < 5 * RTOL, per


def test_csr_bit_exact(cuda_device):
    g = torch.Generator().manual_seed(0)
    for N, E in ((1, 6), (7, 0), (30, 120), (1000, 4096), (50000, 200001)):
        recv = torch.randint(0, N, (E,), generator=g)
        if E > 10:
            recv[: E // 3] = recv[0]  # a very high in-degree row (collisions)
        ei = torch.stack((torch.randint(0, N, (E,), generator=g), recv)).contiguous()
        row_ptr, perm = native.csr_build(ei.to(cuda_device), N)
        want_perm = torch.sort(recv, stable=True).indices
        want_ptr = torch.cat((torch.zeros(1, dtype=torch.long), torch.bincount(recv, minlength=N).cumsum(0)))
        assert torch.equal(perm.cpu().long(), want_perm)
        assert torch.equal(row_ptr.cpu().long(), want_ptr)


def test_graph_ptr_with_empty_graphs(cuda_device):
    batch = torch.tensor([0, 0, 2, 2, 2, 5], dtype=torch.int64)
    gp = native.graph_ptr(batch.to(cuda_device), 7).cpu().tolist()
    assert gp == [0, 2, 2, 5, 5, 5, 6, 6]


def test_empty_and_ragged_batches(cuda_device):
    oracle, params = make_pair(0.02)
    model = cuda_model_from(oracle, params, cuda_device)
    data = [synthetic_lattice(77, 1), synthetic_lattice(78, 2), synthetic_lattice(79, 33)]
    batch = collate(data)
    with torch.no_grad():
        ref = oracle(batch)["stiffness"]
        out = model(batch.to(cuda_device))["stiffness"].cpu()
    assert rel_err(out, ref) < 1e-4  # fp32 oracle on the other side
    # a graph id with no nodes pools to zero features -> C = (bias-only tail)^2, finite
    b2 = collate(data)
    b2.batch = b2.batch * 2
    b2.num_graphs = 6
    with torch.no_grad():
        out2 = model(b2.to(cuda_device))["stiffness"].cpu()
        ref2 = oracle(b2)["stiffness"]
    assert torch.isfinite(out2).all() and rel_err(out2, ref2) < 1e-4


def test_full_size_properties_config2(cuda_device):
    """BASELINE config 2 size (1024 lattices): symmetric PSD output, batch independence, bitwise determinism."""
    batch, rmax = synthetic_batch(range(2000, 3024))
    oracle, params = make_pair(rmax)
    model = cuda_model_from(oracle, params, cuda_device)
    b = batch.to(cuda_device)
    with torch.no_grad():
        C1 = model(b)["stiffness"].clone()
        C2 = model(b)["stiffness"].clone()
    assert torch.equal(C1, C2)
    assert C1.shape == (1024, 6, 6) and torch.isfinite(C1).all()
    assert float((C1 - C1.transpose(1, 2)).abs().max()) <= 1e-6 * float(C1.abs().max())
    assert float(torch.linalg.eigvalsh(C1.double().cpu()).min()) > -1e-6 * float(C1.abs().max())
    sub, _ = synthetic_batch(range(2000 + 500, 2000 + 516))
    with torch.no_grad():
        Cs = model(sub.to(cuda_device))["stiffness"]
    assert rel_err(Cs, C1[500:516]) < 1e-6
    with torch.no_grad():
        ref = oracle(sub)["stiffness"]
    assert rel_err(Cs.cpu(), ref) < 1e-4


def test_streaming_predictor_matches_direct_calls(cuda_device):
    """Public end-to-end API (pinned host batches in, pinned host stiffness out, copies overlapped with compute)."""
    from energy_equiv_lattice_gnn_b200.serve import StreamingPredictor
    oracle, params = make_pair(0.05)
    model = cuda_model_from(oracle, params, cuda_device)
    host = [synthetic_batch(range(4000 + 8 * i, 4000 + 8 * i + 5 + i))[0].pin_memory() for i in range(5)]  # ragged sizes
    with torch.no_grad():
        want = [model(b.to(cuda_device))["stiffness"].cpu() for b in host]
    got = [r.clone() for r in StreamingPredictor(model).run(host)]
    assert len(got) == len(want)
    for g, w in zip(got, want):
        assert g.shape == w.shape and torch.equal(g, w)
    assert list(StreamingPredictor(model).run([])) == []


def test_validate_inputs_rejects_malformed_batches(cuda_device):
    """Opt-in input validation: an edge index >= N, a graph id >= G or a decreasing `batch` raise instead of
    corrupting device memory; a well-formed batch passes and gives the same result as without validation."""
    batch, rmax = synthetic_batch(range(1200, 1203))
    oracle, params = make_pair(rmax)
    model = cuda_model_from(oracle, params, cuda_device)
    good = batch.to(cuda_device)
    with torch.no_grad():
        want = model(good)["stiffness"]
        model.validate_inputs = True
        assert torch.equal(model(good)["stiffness"], want)
        bad = good.to(cuda_device)
        bad.edge_index = good.edge_index.clone()
        bad.edge_index[1, 5] = good.positions.shape[0]          # receiver == N
        with pytest.raises(ValueError, match="edge_index"):
            model(bad)
        bad = good.to(cuda_device)
        bad.batch = good.batch.clone()
        bad.batch[3] = 2                                          # node of graph 0 labelled 2: not non-decreasing
        with pytest.raises(ValueError, match="non-decreasing"):
            model(bad)
        bad.batch = good.batch.clone()
        bad.batch[-1] = 3                                         # graph id == G
        with pytest.raises(ValueError, match="batch has entries"):
            model(bad)
