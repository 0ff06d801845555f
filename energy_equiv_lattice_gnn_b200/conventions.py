"""Angular-momentum conventions the lattice-GNN hot path is built on (host side, fp64 numpy).

The reference delegates all of this to e3nn 0.5.x, which is not vendored under /root/reference and is not
installable here.  This module re-derives, from the published conventions, every constant
the CUDA kernels consume:

  * real-basis Wigner 3j symbols  (e3nn `o3.wigner_3j`, call sites gnn/mace.py:399, e3nn TensorProduct in
    gnn/blocks.py:528-535)
  * component-normalised real spherical harmonics, y polar axis (e3nn `o3.SphericalHarmonics`,
    call site gnn/model.py:125-129,157)
  * Wigner-D matrices (tests only)
  * the sparse generalised Clebsch-Gordan ("U") tensors of gnn/mace.py:363-477, produced by direct path
    enumeration instead of the reference's dense recursion
  * an orthonormal equivariant basis of elasticity-symmetric rank-4 tensors (stand-in for e3nn
    `ReducedTensorProducts('ijkl=jikl=ijlk=klij')`, gnn/blocks.py:427-436)

Everything is pure numpy / python ints; nothing here touches the GPU.
"""
from __future__ import annotations

import functools
import math
from fractions import Fraction
from typing import Dict, List, Sequence, Tuple

import numpy as np

# ------------------------------------------------------------------------------------------------------------------
# SU(2) Clebsch-Gordan (Condon-Shortley), Racah's formula, exact rational arithmetic under the square root
# ------------------------------------------------------------------------------------------------------------------


def _f(n: int) -> int:
    return math.factorial(int(round(n)))


def su2_cg_coeff(j1: int, m1: int, j2: int, m2: int, j3: int, m3: int) -> float:
    """<j1 m1 j2 m2 | j3 m3> for integer spins."""
    if m3 != m1 + m2:
        return 0.0
    vmin = int(max(-j1 + j2 + m3, -j1 + m1, 0))
    vmax = int(min(j2 + j3 + m1, j3 - j1 + j2, j3 + m3))
    pref = (2.0 * j3 + 1.0) * Fraction(
        _f(j3 + j1 - j2) * _f(j3 - j1 + j2) * _f(j1 + j2 - j3) * _f(j3 + m3) * _f(j3 - m3),
        _f(j1 + j2 + j3 + 1) * _f(j1 - m1) * _f(j1 + m1) * _f(j2 - m2) * _f(j2 + m2),
    )
    s = Fraction(0)
    for v in range(vmin, vmax + 1):
        s += (-1) ** int(v + j2 + m2) * Fraction(
            _f(j2 + j3 + m1 - v) * _f(j1 - m1 + v),
            _f(v) * _f(j3 - j1 + j2 - v) * _f(j3 + m3 - v) * _f(v + j1 - j2 - m3),
        )
    return float(pref) ** 0.5 * float(s)


@functools.lru_cache(maxsize=None)
def su2_cg(j1: int, j2: int, j3: int) -> np.ndarray:
    out = np.zeros((2 * j1 + 1, 2 * j2 + 1, 2 * j3 + 1))
    if not (abs(j1 - j2) <= j3 <= j1 + j2):
        return out
    for m1 in range(-j1, j1 + 1):
        for m2 in range(-j2, j2 + 1):
            m3 = m1 + m2
            if abs(m3) <= j3:
                out[j1 + m1, j2 + m2, j3 + m3] = su2_cg_coeff(j1, m1, j2, m2, j3, m3)
    return out


@functools.lru_cache(maxsize=None)
def real_to_complex(l: int) -> np.ndarray:
    """Change of basis real -> complex harmonics, rows = complex m, cols = real index."""
    q = np.zeros((2 * l + 1, 2 * l + 1), dtype=np.complex128)
    s = 1.0 / math.sqrt(2.0)
    for m in range(-l, 0):
        q[l + m, l + abs(m)] = s
        q[l + m, l - abs(m)] = -1j * s
    q[l, l] = 1.0
    for m in range(1, l + 1):
        q[l + m, l + abs(m)] = (-1) ** m * s
        q[l + m, l - abs(m)] = 1j * (-1) ** m * s
    return (-1j) ** l * q


@functools.lru_cache(maxsize=None)
def wigner_3j(l1: int, l2: int, l3: int) -> np.ndarray:
    """Real-basis 3j tensor [2l1+1, 2l2+1, 2l3+1], Frobenius norm 1 (zeros if the triangle rule fails)."""
    if not (abs(l1 - l2) <= l3 <= l1 + l2):
        return np.zeros((2 * l1 + 1, 2 * l2 + 1, 2 * l3 + 1))
    q1, q2, q3 = real_to_complex(l1), real_to_complex(l2), real_to_complex(l3)
    c = np.einsum("ij,kl,mn,ikn->jlm", q1, q2, np.conj(q3.T), su2_cg(l1, l2, l3).astype(np.complex128))
    assert np.abs(c.imag).max() < 1e-10, (l1, l2, l3)
    c = c.real.copy()
    c[np.abs(c) < 1e-14] = 0.0
    c /= np.linalg.norm(c)
    c.setflags(write=False)
    return c


# ------------------------------------------------------------------------------------------------------------------
# Spherical harmonics (component normalisation, |Y_l|^2 = 2l+1 on the unit sphere, polar axis = y)
# ------------------------------------------------------------------------------------------------------------------


@functools.lru_cache(maxsize=None)
def sh_recursion(l: int) -> Tuple[np.ndarray, float]:
    """(w3j(1, l-1, l), scale) such that  Yhat_l[k] = scale * sum_ij w3j[i,j,k] * Yhat_1[i] * Yhat_{l-1}[j]
    where Yhat_l = Y_l / sqrt(2l+1) is the pole-value-one version (Yhat_1 = (x, y, z))."""
    assert l >= 2
    w = wigner_3j(1, l - 1, l)
    # pole: x = (0,1,0) -> Yhat_1 = e_1, Yhat_{l-1} = e_{m=0} (index l-1); demand Yhat_l[m=0] = +1
    pole = w[1, l - 1, l]
    return w, 1.0 / pole


def spherical_harmonics(lmax: int, vec: np.ndarray, normalize: bool = True) -> np.ndarray:
    """[..., 3] -> [..., (lmax+1)^2], component normalised; zero vectors give Y_0 = 1 and zeros above."""
    vec = np.asarray(vec, dtype=np.float64)
    if normalize:
        n = np.linalg.norm(vec, axis=-1, keepdims=True)
        vec = vec / np.maximum(n, 1e-12)  # F.normalize semantics
    hats = [np.ones(vec.shape[:-1] + (1,))]
    if lmax >= 1:
        hats.append(vec.copy())
    for l in range(2, lmax + 1):
        w, sc = sh_recursion(l)
        hats.append(sc * np.einsum("ijk,...i,...j->...k", w, hats[1], hats[l - 1]))
    return np.concatenate([math.sqrt(2 * l + 1) * h for l, h in enumerate(hats)], axis=-1)


def wigner_D(l: int, R: np.ndarray) -> np.ndarray:
    """D_l(R) with Y_l(R x) = D_l(R) Y_l(x); obtained by least squares on a fixed point cloud (tests only)."""
    rng = np.random.default_rng(12345 + l)
    pts = rng.normal(size=(8 * (2 * l + 1) + 8, 3))
    sl = slice(l * l, (l + 1) * (l + 1))
    A = spherical_harmonics(l, pts)[:, sl]
    B = spherical_harmonics(l, pts @ R.T)[:, sl]
    D, *_ = np.linalg.lstsq(A, B, rcond=None)
    return D.T


def rotation_matrix(alpha: float, beta: float, gamma: float) -> np.ndarray:
    """R = R_y(alpha) R_x(beta) R_y(gamma) (e3nn's YXY Euler convention)."""

    def ry(a):
        c, s = math.cos(a), math.sin(a)
        return np.array([[c, 0, s], [0, 1, 0], [-s, 0, c]])

    def rx(a):
        c, s = math.cos(a), math.sin(a)
        return np.array([[1, 0, 0], [0, c, -s], [0, s, c]])

    return ry(alpha) @ rx(beta) @ ry(gamma)


# ------------------------------------------------------------------------------------------------------------------
# Irreps bookkeeping (just what the model needs: lists of (mul, l, parity))
# ------------------------------------------------------------------------------------------------------------------


def parse_irreps(s: str) -> List[Tuple[int, int, int]]:
    """'32x0e+32x1o' -> [(32,0,1),(32,1,-1)]."""
    out = []
    for tok in str(s).replace(" ", "").split("+"):
        if not tok:
            continue
        if "x" in tok:
            mul, ir = tok.split("x")
            mul = int(mul)
        else:
            mul, ir = 1, tok
        l = int(ir[:-1])
        p = {"e": 1, "o": -1}[ir[-1]]
        out.append((mul, l, p))
    return out


def irreps_str(irreps: Sequence[Tuple[int, int, int]]) -> str:
    return "+".join(f"{m}x{l}{'e' if p == 1 else 'o'}" for m, l, p in irreps)


def irreps_dim(irreps: Sequence[Tuple[int, int, int]]) -> int:
    return sum(m * (2 * l + 1) for m, l, _ in irreps)


def tp_paths(irreps1: Sequence[Tuple[int, int, int]], lmax: int, target: Sequence[Tuple[int, int, int]]):
    """'uvu' paths of gnn/mace.py:286-314 for irreps2 = spherical harmonics up to lmax.

    Returns list of (i1, l1, l2, l3, p3) in the reference's weight order (stable sort by output irrep; generation order
    i1 asc, l2 asc, l3 asc inside), i.e. path p owns weight columns [p*mul, (p+1)*mul).
    """
    tset = {(l, p) for _, l, p in target}
    gen = []
    for i1, (_, l1, p1) in enumerate(irreps1):
        for l2 in range(lmax + 1):
            p2 = (-1) ** l2
            for l3 in range(abs(l1 - l2), l1 + l2 + 1):
                if (l3, p1 * p2) in tset:
                    gen.append((i1, l1, l2, l3, p1 * p2))
    order = sorted(range(len(gen)), key=lambda k: (gen[k][3], gen[k][4]))  # Irrep order (l, p); stable
    return [gen[k] for k in order]


# ------------------------------------------------------------------------------------------------------------------
# Generalised CG tensors of the symmetric contraction (gnn/mace.py:363-477), sparse
# ------------------------------------------------------------------------------------------------------------------


def _ir_key(l: int, p: int):
    # e3nn 0.5.x `Irrep` is a plain tuple (l, p): sorted() orders 0o < 0e < 1o < 1e ...
    return (l, p)


@functools.lru_cache(maxsize=None)
def u_matrix_paths(lmax: int, nu: int):
    """Coupling paths of `_wigner_nj([0e+1o+..+lmax]*nu)` in the order the reference generates them.

    Each entry: (l_out, p_out, ls, inter) where ls = (l_1..l_nu) are the input irreps and inter the intermediate
    (l, p) after each coupling.  The list is sorted by (l_out, p_out) (stable), like `sorted(ret, key=x[0])`.
    """
    level = [((l, (-1) ** l), (l,), ()) for l in range(lmax + 1)]  # (ir, ls, inter)
    for _ in range(1, nu):
        nxt = []
        for (ll, pl), ls, inter in level:
            for l in range(lmax + 1):
                p = (-1) ** l
                for lo in range(abs(ll - l), ll + l + 1):
                    nxt.append(((lo, pl * p), ls + (l,), inter + ((ll, pl),)))
        level = sorted(nxt, key=lambda t: _ir_key(*t[0]))
    return tuple((ir[0], ir[1], ls, inter) for ir, ls, inter in level)


@functools.lru_cache(maxsize=None)
def _coupled_tensor(ls: Tuple[int, ...], inter_l: Tuple[int, ...], l_out: int) -> np.ndarray:
    """Dense [2*l_out+1, 2l_1+1, ..., 2l_nu+1] block of one coupling path (component normalisation)."""
    if len(ls) == 1:
        return np.eye(2 * ls[0] + 1)
    left_l = inter_l[-1]
    left = _coupled_tensor(ls[:-1], inter_l[:-1], left_l)  # [2left+1, ...]
    c = wigner_3j(l_out, left_l, ls[-1]) * math.sqrt(2 * l_out + 1)  # mace.py:399-401
    out = np.einsum("j...,ijl->i...l", left, c)
    return out


def u_matrix_sparse(lmax: int, nu: int, l_out: int):
    """Sparse U_nu^L of gnn/mace.py:435-477 for coupling irreps 0e+1o+..+lmax and output irrep (L, (-1)^L).

    Returns (idx[int64, nnz x (nu+2)], val[float64, nnz], k) with columns (M, i_1..i_nu, k); i are indices into the
    (lmax+1)^2 concatenated input, k enumerates the retained coupling paths in the reference's order.
    """
    p_out = (-1) ** l_out
    rows, vals = [], []
    k = 0
    for lo, po, ls, inter in u_matrix_paths(lmax, nu):
        if lo != l_out or po != p_out:
            continue
        blk = _coupled_tensor(ls, tuple(il for il, _ in inter), lo)
        nz = np.argwhere(np.abs(blk) > 1e-12)
        offs = np.array([0] + [l * l for l in ls])
        for ix in nz:
            rows.append(tuple(int(a + b) for a, b in zip(ix, offs)) + (k,))
            vals.append(blk[tuple(ix)])
        k += 1
    idx = np.array(rows, dtype=np.int64).reshape(-1, nu + 2)
    return idx, np.array(vals, dtype=np.float64), k


def u_matrix_dense(lmax: int, nu: int, l_out: int, dtype=np.float32) -> np.ndarray:
    """Dense buffer with the reference's shape: [(2L+1), D, ..., D, k] (the M axis squeezed away when L = 0)."""
    D = (lmax + 1) ** 2
    idx, val, k = u_matrix_sparse(lmax, nu, l_out)
    out = np.zeros((2 * l_out + 1,) + (D,) * nu + (k,), dtype=dtype)
    out[tuple(idx.T)] = val
    if l_out == 0:
        out = out[0]
    return out


# ------------------------------------------------------------------------------------------------------------------
# Elasticity-symmetric rank-4 basis: 2x0e + 2x2e + 1x4e  ->  R^{3x3x3x3}
# ------------------------------------------------------------------------------------------------------------------


@functools.lru_cache(maxsize=None)
def rank4_symmetric_basis() -> np.ndarray:
    """Q[21, 3,3,3,3] with orthonormal rows spanning tensors with ijkl=jikl=ijlk=klij, rows grouped as
    2x0e (2 rows), 2x2e (2x5 rows, mul-major), 1x4e (9 rows); each block transforms with D_L.

    Construction (documented, deterministic): couple ((1 x 1)_a x (1 x 1)_b)_L for (a,b) in a fixed order, symmetrise
    over the elasticity group, Gram-Schmidt within each L on the m=first component pattern so that all 2L+1 rows of
    one multiplicity copy share coefficients.  e3nn's own choice inside the two multiplicity spaces cannot be
    reproduced without e3nn; a reference checkpoint carries its Q_flat as a buffer.
    """
    # Cartesian index (x,y,z) == l=1 real basis in e3nn's order, so w3j(1,1,a)[i,j,m] maps directly.
    def sym(T):  # T[m, i,j,k,l]
        T = T + T.transpose(0, 2, 1, 3, 4)
        T = T + T.transpose(0, 1, 2, 4, 3)
        T = T + T.transpose(0, 3, 4, 1, 2)
        return T / 8.0

    blocks = []
    for L, want in ((0, 2), (2, 2), (4, 1)):
        cands = []
        for a in (0, 2):  # symmetric pairs only (antisymmetric l=1 dies under ij=ji)
            for b in (0, 2):
                if not (abs(a - b) <= L <= a + b):
                    continue
                T = np.einsum("ijp,klq,pqm->mijkl", wigner_3j(1, 1, a), wigner_3j(1, 1, b), wigner_3j(a, b, L))
                cands.append(sym(T))
        basis = []
        for T in cands:
            T = T.copy()
            for B in basis:
                # equivariant projection: one scalar per copy (Schur), estimated with the full block inner product
                coef = np.sum(B * T) / np.sum(B * B)
                T = T - coef * B
            nrm = math.sqrt(np.sum(T * T) / (2 * L + 1))
            if nrm > 1e-9:
                basis.append(T / nrm)
            if len(basis) == want:
                break
        assert len(basis) == want, (L, len(basis))
        blocks.extend(basis)
    Q = np.concatenate(blocks, axis=0)
    assert Q.shape == (21, 3, 3, 3, 3)
    return Q


# Mandel bookkeeping of gnn/blocks.py:392-425 (index lists restated from the 6x6 upper-triangle enumeration)
MANDEL_PAIRS = ((0, 0), (1, 1), (2, 2), (1, 2), (0, 2), (0, 1))


def mandel_tables():
    """(abcd[21,4], rows[21], cols[21], mask[6,6]) such that A[r,c] = C4[a,b,c,d] * mask[r,c]."""
    rows, cols = np.triu_indices(6)
    abcd = np.array([MANDEL_PAIRS[r] + MANDEL_PAIRS[c] for r, c in zip(rows, cols)], dtype=np.int64)
    s2 = math.sqrt(2.0)
    mask = np.ones((6, 6))
    mask[:3, 3:] = s2
    mask[3:, :3] = s2
    mask[3:, 3:] = 2.0
    return abcd, rows.astype(np.int64), cols.astype(np.int64), mask
