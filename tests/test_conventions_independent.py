"""Independent pins of the angular-momentum conventions.

`conventions.py` is the single root of truth for the kernels' tables AND for the e3nn stand-ins the golden generator runs
the reference through, so these checks deliberately import NOTHING from it except the functions under test:
  * real spherical harmonics l <= 4 rebuilt from scipy.special.lpmv (associated Legendre) in e3nn's convention: polar axis
    y, azimuth measured from z towards x, component normalisation (|Y_l|^2 = 2l + 1), index m = -l..l with the sin terms
    first - e3nn `o3.spherical_harmonics(..., normalization='component')`, call site gnn/model.py:125-129,157;
  * real Wigner 3j for every tensor-product path of the model (42 for lmax 4, 23 for lmax 3) and every coupling l <= 4:
    SU(2) Clebsch-Gordan coefficients from sympy.physics.wigner (exact) + e3nn's published real <-> complex change of basis
    ((-i)^l phase), normalised to unit Frobenius norm - e3nn `o3.wigner_3j`, call sites gnn/mace.py:399, gnn/blocks.py:528-535;
  * the two are tied together without any recalled convention: each 3j tensor is invariant under the Wigner-D matrices
    that the scipy-built harmonics themselves carry (D_l from Y_l(R x) = D_l(R) Y_l(x))."""
import math

import numpy as np
import pytest
from scipy.special import lpmv
from sympy import N as sym_N
from sympy.physics.wigner import clebsch_gordan

from energy_equiv_lattice_gnn_b200 import conventions as cv
from energy_equiv_lattice_gnn_b200 import tables as tb


def sh_scipy(lmax, vec):
    v = np.asarray(vec, dtype=np.float64)
    v = v / np.linalg.norm(v, axis=-1, keepdims=True)
    x, y, z = v[..., 0], v[..., 1], v[..., 2]
    ct = y                       # polar axis: y
    phi = np.arctan2(x, z)       # azimuth from z towards x
    out = []
    for l in range(lmax + 1):
        for m in range(-l, l + 1):
            am = abs(m)
            nrm = math.sqrt((2 * l + 1) * math.factorial(l - am) / math.factorial(l + am))   # component normalisation (4 pi dropped)
            p = (-1) ** am * lpmv(am, l, ct)            # remove the Condon-Shortley phase lpmv carries
            if m == 0:
                out.append(nrm * p)
            elif m > 0:
                out.append(math.sqrt(2) * nrm * p * np.cos(am * phi))
            else:
                out.append(math.sqrt(2) * nrm * p * np.sin(am * phi))
    return np.stack(out, -1)


def test_spherical_harmonics_match_scipy_legendre():
    rng = np.random.default_rng(0)
    pts = rng.normal(size=(200, 3))
    pts = np.concatenate([pts, np.eye(3), -np.eye(3)])
    got = cv.spherical_harmonics(4, pts)
    want = sh_scipy(4, pts)
    assert np.abs(got - want).max() < 1e-12
    for l in range(5):     # component normalisation
        sl = slice(l * l, (l + 1) ** 2)
        assert np.allclose((want[:, sl] ** 2).sum(-1), 2 * l + 1)


def _q(l):
    """e3nn change_basis_real_to_complex(l): rows complex m, columns real index; overall (-i)^l makes the 3j real."""
    q = np.zeros((2 * l + 1, 2 * l + 1), dtype=np.complex128)
    s = 1 / math.sqrt(2)
    for m in range(-l, 0):
        q[l + m, l + abs(m)] = s
        q[l + m, l - abs(m)] = -1j * s
    q[l, l] = 1
    for m in range(1, l + 1):
        q[l + m, l + abs(m)] = (-1) ** m * s
        q[l + m, l - abs(m)] = 1j * (-1) ** m * s
    return (-1j) ** l * q


def w3j_sympy(l1, l2, l3):
    c = np.zeros((2 * l1 + 1, 2 * l2 + 1, 2 * l3 + 1))
    for m1 in range(-l1, l1 + 1):
        for m2 in range(-l2, l2 + 1):
            m3 = m1 + m2
            if abs(m3) <= l3:
                c[l1 + m1, l2 + m2, l3 + m3] = float(sym_N(clebsch_gordan(l1, l2, l3, m1, m2, m3), 30))
    r = np.einsum("ij,kl,mn,ikn->jlm", _q(l1), _q(l2), np.conj(_q(l3).T), c.astype(np.complex128))
    assert np.abs(r.imag).max() < 1e-12
    r = r.real
    return r / np.linalg.norm(r)


def _triples():
    seen = set()
    for lmax in (3, 4):
        seen.update(tuple(p) for p in tb.tp_plan(lmax, lmax, lmax)["paths"])     # 23 + 42 tensor-product paths
        seen.update(tuple(p) for p in tb.tp_plan(0, lmax, lmax)["paths"])
    for l1 in range(5):                                                         # every coupling the nu <= 3 contraction can use
        for l2 in range(5):
            for l3 in range(abs(l1 - l2), min(l1 + l2, 4) + 1):
                seen.add((l1, l2, l3))
    return sorted(seen)


@pytest.mark.parametrize("chunk", range(4))
def test_wigner_3j_matches_sympy_clebsch_gordan(chunk):
    tr = _triples()
    assert len(tr) >= 60
    for (l1, l2, l3) in tr[chunk::4]:
        got, want = np.asarray(cv.wigner_3j(l1, l2, l3)), w3j_sympy(l1, l2, l3)
        assert np.abs(got - want).max() < 1e-12, (l1, l2, l3)


def _D_from_scipy_sh(l, R):
    rng = np.random.default_rng(99 + l)
    pts = rng.normal(size=(12 * (2 * l + 1), 3))
    sl = slice(l * l, (l + 1) ** 2)
    A = sh_scipy(l, pts)[:, sl]
    B = sh_scipy(l, pts @ R.T)[:, sl]
    D, res, *_ = np.linalg.lstsq(A, B, rcond=None)
    assert np.abs(A @ D - B).max() < 1e-10
    return D.T


def test_3j_tensors_are_invariants_of_the_harmonics_own_representation():
    a, b, g = 0.7, 1.9, -2.4
    ry = lambda t: np.array([[math.cos(t), 0, math.sin(t)], [0, 1, 0], [-math.sin(t), 0, math.cos(t)]])      # noqa: E731
    rx = lambda t: np.array([[1, 0, 0], [0, math.cos(t), -math.sin(t)], [0, math.sin(t), math.cos(t)]])      # noqa: E731
    R = ry(a) @ rx(b) @ ry(g)
    D = [_D_from_scipy_sh(l, R) for l in range(5)]
    for l in range(5):
        assert np.abs(D[l] @ D[l].T - np.eye(2 * l + 1)).max() < 1e-9
    for (l1, l2, l3) in _triples():
        w = np.asarray(cv.wigner_3j(l1, l2, l3))
        rot = np.einsum("ai,bj,ck,ijk->abc", D[l1], D[l2], D[l3], w)
        assert np.abs(rot - w).max() < 1e-9, (l1, l2, l3)
