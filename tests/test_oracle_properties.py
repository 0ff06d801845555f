"""Properties any faithful implementation of the reference architecture satisfies: O(3)
equivariance of C (the property RotateLat relies on, scripts/train_utils.py:114-146), permutation / edge-order
invariance, supercell invariance, symmetric PSD output.  Run on the CPU oracle in fp64 with a small channel count."""
import numpy as np
import pytest
import torch

from energy_equiv_lattice_gnn_b200 import conventions as cv
from energy_equiv_lattice_gnn_b200.data import Batch, collate, lattice_from_struts, supercell, synthetic_lattice
from oracle.model import EnergyEquivGNN, default_params


def small_oracle(rmax, lmax=2, seed=0, **over):
    irr = lambda m: "+".join(f"{m}x{l}{'e' if l % 2 == 0 else 'o'}" for l in range(lmax + 1))  # noqa: E731
    p = default_params(lmax=lmax, hidden_irreps=irr(4), readout_irreps=irr(4), max_edge_radius=rmax, **over)
    torch.set_default_dtype(torch.float64)
    try:
        torch.manual_seed(seed)
        m = EnergyEquivGNN(p)
        with torch.no_grad():
            m.stiffness_head.linear.bias.normal_()
            for layer in m.stiffness_head.layers:
                layer.interaction.linear.bias.normal_()
    finally:
        torch.set_default_dtype(torch.float32)
    return m


def to64(b):
    return Batch(**{k: (v.double() if torch.is_tensor(v) and v.is_floating_point() else v) for k, v in b.__dict__.items()})


@pytest.fixture(scope="module")
def setup():
    data = [synthetic_lattice(s, 5) for s in (11, 12, 13)]
    batch = to64(collate(data))
    rmax = float(batch.edge_attr.max())
    return data, batch, small_oracle(rmax)


def test_symmetric_psd(setup):
    _, batch, m = setup
    with torch.no_grad():
        C = m(batch)["stiffness"]
    assert C.shape == (3, 6, 6)
    assert float((C - C.transpose(1, 2)).abs().max()) < 1e-14
    assert float(torch.linalg.eigvalsh(C).min()) > -1e-12


@pytest.mark.parametrize("improper", [False, True])
def test_rotation_equivariance(setup, improper):
    _, batch, m = setup
    R = cv.rotation_matrix(0.7, -1.2, 0.4)
    if improper:
        R = R @ np.diag([-1.0, 1.0, 1.0])
    Rt = torch.tensor(R)
    rb = Batch(**batch.__dict__)
    rb.positions = batch.positions @ Rt.T
    rb.shifts = batch.shifts @ Rt.T
    with torch.no_grad():
        C0 = m(batch)["stiffness"]
        C1 = m(rb)["stiffness"]
    # compare through the full rank-4 tensors: Mandel -> Cartesian, rotate, compare
    abcd, rows, cols, mask = cv.mandel_tables()

    def to_cart(Cm):
        T = np.zeros((3, 3, 3, 3))
        A = Cm / mask
        for (a, b, c, d), r, cc in zip(abcd, rows, cols):
            for (i, j) in ((a, b), (b, a)):
                for (k, l) in ((c, d), (d, c)):
                    T[i, j, k, l] = A[r, cc]
                    T[k, l, i, j] = A[r, cc]
        return T

    for g in range(C0.shape[0]):
        T0, T1 = to_cart(C0[g].numpy()), to_cart(C1[g].numpy())
        T0r = np.einsum("ijkl,ai,bj,ck,dl->abcd", T0, R, R, R, R)
        assert np.abs(T0r - T1).max() < 1e-10 * max(1.0, np.abs(T0).max())


def test_edge_permutation_and_direction_order_invariance(setup):
    _, batch, m = setup
    E = batch.edge_index.shape[1]
    perm = torch.randperm(E, generator=torch.Generator().manual_seed(1))
    pb = Batch(**batch.__dict__)
    pb.edge_index, pb.shifts, pb.edge_attr = batch.edge_index[:, perm], batch.shifts[perm], batch.edge_attr[perm]
    with torch.no_grad():
        assert float((m(batch)["stiffness"] - m(pb)["stiffness"]).abs().max()) < 1e-12


def test_node_permutation_invariance(setup):
    data, _, m = setup
    d = data[0]
    n = d.positions.shape[0]
    p = torch.randperm(n, generator=torch.Generator().manual_seed(2))
    inv = torch.empty_like(p)
    inv[p] = torch.arange(n)
    from energy_equiv_lattice_gnn_b200.data import LatticeData
    d2 = LatticeData(positions=d.positions[p], node_attrs=d.node_attrs[p], edge_index=inv[d.edge_index], shifts=d.shifts,
                     edge_attr=d.edge_attr)
    with torch.no_grad():
        a = m(to64(collate([d])))["stiffness"]
        b = m(to64(collate([d2])))["stiffness"]
    assert float((a - b).abs().max()) < 1e-12


def test_supercell_invariance():
    rng = np.random.default_rng(5)
    x = rng.random((3, 3))
    struts = np.array([[0, 1], [1, 2], [2, 0], [0, 0], [1, 2], [0, 2]])
    shifts = np.array([[0, 0, 0], [0, 1, 0], [0, 0, 0], [1, 0, 0], [0, 0, 1], [-1, 0, 1]])
    const = (1.0, 0.8, 0.9, 90.0, 90.0, 90.0)
    a = lattice_from_struts(x, struts, shifts, const, 0.01)
    x2, s2, n2, c2 = supercell(x, struts, shifts, const, (2, 1, 2))
    b = lattice_from_struts(x2, s2, n2, c2, 0.01)
    assert abs(float(a.edge_attr[0]) - float(b.edge_attr[0])) < 1e-7  # radius unchanged
    m = small_oracle(float(a.edge_attr.max()))
    with torch.no_grad():
        Ca = m(to64(collate([a])))["stiffness"]
        Cb = m(to64(collate([b])))["stiffness"]
    # geometry is stored in fp32 (datasets.py:184-185): cell and supercell coordinates round differently
    assert float((Ca - Cb).abs().max()) < 1e-6 * max(1.0, float(Ca.abs().max()))


def test_batch_independence(setup):
    data, batch, m = setup
    with torch.no_grad():
        C = m(batch)["stiffness"]
        for g, d in enumerate(data):
            Cg = m(to64(collate([d])))["stiffness"]
            assert float((C[g] - Cg[0]).abs().max()) < 1e-12
