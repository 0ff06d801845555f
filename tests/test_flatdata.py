"""Flat binary dataset + collation: `flatdata.FlatDataset.collate` must equal the
PyG-equivalent `data.collate` (scripts/train_main.py:67-78) bit for bit, on the host and - through
`eeg_collate_gather` - on the device; `read_lat_fast` must equal `catalogue.read_lat` on the reference's sample catalogue;
the binary file round-trips."""
import os

import numpy as np
import pytest
import torch

from energy_equiv_lattice_gnn_b200 import catalogue as cat
from energy_equiv_lattice_gnn_b200.data import collate, synthetic_lattice
from energy_equiv_lattice_gnn_b200.flatdata import FlatDataset, convert_lat, read_lat_fast

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
FIELDS = ("positions", "node_attrs", "edge_index", "shifts", "edge_attr", "batch")


def _lattices():
    return [synthetic_lattice(500 + i, n, with_target=True) for i, n in enumerate((5, 30, 1, 12, 30, 7, 19))]


def _assert_same(a, b):
    assert a.num_graphs == b.num_graphs
    for k in FIELDS:
        x, y = getattr(a, k).cpu(), getattr(b, k).cpu()
        assert x.dtype == y.dtype and x.shape == y.shape, k
        assert torch.equal(x, y), k


def test_host_collate_equals_pyg_style_collate(tmp_path):
    data = _lattices()
    ds = FlatDataset.from_lattices(data)
    assert len(ds) == len(data)
    for idx in ([0, 1, 2, 3, 4, 5, 6], [6, 2, 2, 0], [3], [1, 4]):
        got = ds.collate(idx)
        want = collate([data[i] for i in idx])
        _assert_same(got, want)
        assert torch.equal(got.stiffness.reshape(len(idx), -1), want.stiffness.reshape(len(idx), -1))
    with pytest.raises(IndexError):
        ds.collate([7])
    # binary round trip (memory mapped and plain)
    path = str(tmp_path / "ds.eegflat")
    ds.save(path)
    for mm in (True, False):
        back = FlatDataset.load(path, mmap=mm)
        assert back.names == ds.names
        for k, v in ds.a.items():
            assert torch.equal(back.a[k], v), k
        _assert_same(back.collate([5, 0, 3]), collate([data[5], data[0], data[3]]))


def _write_sample(tmp_path):
    """A .lat file with the sections of the reference's sample catalogue (written by catalogue.write_lat from seeded
    synthetic cells, so the test does not need /root/reference)."""
    rng = np.random.default_rng(0)
    cells = []
    for c in range(12):
        n = int(rng.integers(2, 9))
        pos = rng.random((n, 3)).round(5)
        m = int(rng.integers(n, 2 * n + 1))
        adj = np.stack([rng.integers(0, n, m), rng.integers(0, n, m)], 1)
        adj[:n, 0] = np.arange(n)             # every node appears
        tess = rng.integers(-1, 2, (m, 3)).astype(float)
        comp = {}
        for rd in (0.001, 0.003, 0.01):
            A = rng.standard_normal((6, 6))
            comp[rd] = (A @ A.T + 6 * np.eye(6)).round(6)
        uc = dict(name=f"cell_{c}", lattice_constants=[1.0, 0.9, 0.8, 90.0, 90.0, 90.0], compliance_tensors_M=comp,
                  nodal_positions=pos.tolist(), fundamental_edge_adjacency=adj.tolist(), fundamental_tesselation_vecs=tess.tolist())
        if c % 3 == 0:
            uc["fundamental_edge_radii"] = {rd: rng.random(m).round(6).tolist() for rd in comp}
        cells.append(uc)
    path = str(tmp_path / "cat.lat")
    cat.write_lat(path, cells)
    return path


def _same_cells(a, b):
    assert len(a) == len(b)
    for x, y in zip(a, b):
        assert x["name"] == y["name"]
        assert np.array_equal(np.asarray(x["lattice_constants"], float), np.asarray(y["lattice_constants"], float))
        for k in ("nodal_positions", "fundamental_edge_adjacency", "fundamental_tesselation_vecs"):
            assert np.array_equal(np.asarray(x[k], float), np.asarray(y[k], float)), (x["name"], k)
        cx, cy = x["compliance_tensors_M"], y["compliance_tensors_M"]
        assert list(cx) == list(cy)
        for rd in cx:
            assert np.array_equal(cx[rd], cy[rd])
        assert ("fundamental_edge_radii" in x) == ("fundamental_edge_radii" in y)
        if "fundamental_edge_radii" in x:
            for rd in x["fundamental_edge_radii"]:
                assert np.array_equal(np.asarray(x["fundamental_edge_radii"][rd]), np.asarray(y["fundamental_edge_radii"][rd]))


def test_fast_reader_equals_line_reader_and_convert(tmp_path):
    path = _write_sample(tmp_path)
    slow, fast = cat.read_lat(path), read_lat_fast(path)
    _same_cells(slow, fast)
    out = str(tmp_path / "cat.eegflat")
    info = convert_lat(path, out)
    assert info["entries"] == 12 and info["lattices"] == 36 and info["lattices_per_s"] > 0
    ds = FlatDataset.load(out)
    want = [d for uc in slow for d in cat.process_one(uc)]
    assert len(ds) == len(want)
    _assert_same(ds.collate(list(range(len(want)))), collate(want))
    got = ds.collate([4, 9])
    assert got.stiffness.shape == (2, 3, 3, 3, 3) and got.compliance.shape == (2, 3, 3, 3, 3)
    assert torch.equal(got.stiffness[1], want[9].stiffness[0].to(torch.float32))


@pytest.mark.skipif(not os.path.exists("/root/reference/lattices/examples/sample_cat.lat"), reason="reference checkout not present")
def test_fast_reader_on_reference_sample_catalogue():
    path = "/root/reference/lattices/examples/sample_cat.lat"
    _same_cells([c for c in cat.read_lat(path) if "compliance_tensors_M" in c], [c for c in read_lat_fast(path) if "compliance_tensors_M" in c])


@pytest.mark.gpu
def test_device_collate_equals_host_collate(cuda_device):
    data = _lattices()
    ds = FlatDataset.from_lattices(data)
    dd = ds.to(cuda_device)
    for idx in ([0, 1, 2, 3, 4, 5, 6], [6, 2, 2, 0], [3]):
        got, want = dd.collate(idx), ds.collate(idx)
        assert got.positions.is_cuda
        _assert_same(got, want)
        assert torch.equal(got.stiffness.cpu(), want.stiffness)
    big = FlatDataset.from_lattices([synthetic_lattice(900 + i) for i in range(64)]).to(cuda_device)
    g = torch.Generator().manual_seed(1)
    idx = torch.randint(0, 64, (4096,), generator=g)
    b = big.collate(idx)
    assert b.num_graphs == 4096 and b.positions.shape[0] == 4096 * 30
    assert int(b.edge_index.max()) == b.positions.shape[0] - 1 or int(b.edge_index.max()) < b.positions.shape[0]
    assert torch.equal(b.batch.cpu(), torch.arange(4096).repeat_interleave(30))


@pytest.mark.gpu
def test_device_rotation_sampling(cuda_device):
    """eeg_rand_rotations: proper rotations with the o3.rand_matrix() distribution (scripts/train_utils.py:121): orthonormal,
    det +1, Haar moments E[R] = 0, E[R_ij^2] = 1/3; counter-based (offset continues the stream), reproducible."""
    from energy_equiv_lattice_gnn_b200 import native
    Q = native.rand_rotations(7, 0, 20000, cuda_device)
    Q2 = native.rand_rotations(7, 0, 20000, cuda_device)
    assert torch.equal(Q, Q2)
    tail = native.rand_rotations(7, 15000, 5000, cuda_device)
    assert torch.equal(tail, Q[15000:])
    Qd = Q.double().cpu()
    eye = torch.eye(3, dtype=torch.float64)
    assert float((Qd @ Qd.transpose(1, 2) - eye).abs().max()) < 2e-6
    assert float((torch.linalg.det(Qd) - 1).abs().max()) < 2e-6
    assert float(Qd.mean(0).abs().max()) < 0.02
    assert float(((Qd ** 2).mean(0) - 1 / 3).abs().max()) < 0.01
    other = native.rand_rotations(8, 0, 100, cuda_device)
    assert not torch.equal(other, Q[:100])
    # RotateLat(seed=...) consumes that stream: second call continues where the first stopped
    from energy_equiv_lattice_gnn_b200.data import RotateLat
    data = [synthetic_lattice(950 + i, 6, with_target=False) for i in range(5)]
    b = collate(data).to(cuda_device)
    rot = RotateLat(rotate=True, seed=7)
    b1, b2 = rot(b), rot(b)
    assert rot.offset == 10
    Qs = Q[:10]
    gid = b.batch
    assert torch.allclose(b1.positions, torch.einsum("nij,nj->ni", Qs[:5][gid], b.positions), atol=1e-6)
    assert torch.allclose(b2.positions, torch.einsum("nij,nj->ni", Qs[5:10][gid], b.positions), atol=1e-6)
