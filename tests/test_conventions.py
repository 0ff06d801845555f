"""Self-consistency of the angular-momentum conventions every kernel constant comes from (SURVEY.md B.7 anchors)."""
import math

import numpy as np
import pytest

from energy_equiv_lattice_gnn_b200 import conventions as cv
from energy_equiv_lattice_gnn_b200 import tables as tb


def test_w3j_anchors():
    w = cv.wigner_3j(1, 2, 3)
    for (i, j, k), v in {(0, 0, 4): -0.069006555934, (0, 0, 6): -0.267261241912, (0, 1, 3): -0.169030850946,
                         (0, 1, 5): -0.218217890236, (0, 2, 2): 0.239045721867, (0, 3, 1): 0.218217890236}.items():
        assert abs(w[i, j, k] - v) < 1e-11
    assert abs(cv.wigner_3j(2, 2, 2)[0, 1, 3] - 0.207019667803) < 1e-11
    assert abs(np.abs(cv.wigner_3j(2, 2, 2)).sum() - 4.874693232301) < 1e-10
    assert abs(cv.wigner_3j(4, 4, 4)[0, 0, 4] - 0.104297703129) < 1e-11
    assert abs(cv.wigner_3j(2, 2, 4)[2, 2, 4] - 0.239045721867) < 1e-11
    w = cv.wigner_3j(3, 4, 1)
    assert abs(w.sum() - 2.580793462466) < 1e-10 and abs(np.abs(w).sum() - 5.152863745039) < 1e-10
    eps = np.zeros((3, 3, 3))
    for a, b, c in ((0, 1, 2), (1, 2, 0), (2, 0, 1)):
        eps[a, b, c], eps[a, c, b] = 1, -1
    assert np.allclose(cv.wigner_3j(1, 1, 1), eps / math.sqrt(6))
    for l in range(5):
        d = np.eye(2 * l + 1) / math.sqrt(2 * l + 1)
        assert np.allclose(cv.wigner_3j(l, l, 0)[:, :, 0], d)
        assert np.allclose(cv.wigner_3j(l, 0, l)[:, 0, :], d)
        assert np.allclose(cv.wigner_3j(0, l, l)[0], d)


def test_w3j_unit_norm_and_invariance():
    R = cv.rotation_matrix(0.3, 1.1, -0.7)
    for l1, l2, l3 in ((1, 2, 3), (2, 2, 2), (4, 4, 4), (3, 4, 1), (4, 2, 4)):
        w = cv.wigner_3j(l1, l2, l3)
        assert abs(np.linalg.norm(w) - 1) < 1e-12
        D1, D2, D3 = cv.wigner_D(l1, R), cv.wigner_D(l2, R), cv.wigner_D(l3, R)
        assert np.allclose(np.einsum("ijk,ai,bj,ck->abc", w, D1, D2, D3), w, atol=1e-9)


def test_sh_anchors_and_equivariance():
    v = np.array([0.3, -0.5, 0.8])
    Y = cv.spherical_harmonics(4, v)
    assert np.allclose(Y[1:4], (0.5248906592, -0.8748177653, 1.3997084244), atol=1e-9)
    assert np.allclose(Y[4:9], (0.9484857174, -0.5928035734, -0.2623957321, -1.5808095291, 1.0868065512), atol=1e-9)
    assert np.allclose(Y[9:16], (1.1836471578, -1.2674673539, 0.1352728248, 1.1522232113, 0.3607275328, -1.452306343, 0.6381777026), atol=1e-9)
    assert np.allclose(Y[16:25], (1.2196832539, -1.7934962824, 0.6453957136, 0.4408058327, -0.8907616618, 1.1754822206,
                                  0.7395159219, -0.9669852452, 0.1665514443), atol=1e-9)
    for l in range(5):
        sl = slice(l * l, (l + 1) ** 2)
        assert abs((Y[sl] ** 2).sum() - (2 * l + 1)) < 1e-10
        pole = cv.spherical_harmonics(4, np.array([0.0, 2.0, 0.0]))[sl]
        want = np.zeros(2 * l + 1)
        want[l] = math.sqrt(2 * l + 1)
        assert np.allclose(pole, want, atol=1e-12)
    assert np.allclose(cv.spherical_harmonics(4, np.zeros(3)), [1] + [0] * 24)  # F.normalize(0) = 0
    R = cv.rotation_matrix(0.3, 1.1, -0.7)
    assert np.allclose(cv.wigner_D(1, R), R, atol=1e-9)
    assert np.allclose(cv.wigner_D(2, R)[0], (0.3990860704, 0.7096223422, 0.3883843343, -0.2992716188, -0.3110635736), atol=1e-8)
    for l in range(5):
        D = cv.wigner_D(l, R)
        assert np.allclose(D @ D.T, np.eye(2 * l + 1), atol=1e-9)


def test_u_matrix_widths_and_norms():
    assert [cv.u_matrix_sparse(3, 3, L)[2] for L in range(4)] == [23, 51, 65, 65]
    assert [cv.u_matrix_sparse(3, 2, L)[2] for L in range(4)] == [4, 6, 7, 6]
    assert [cv.u_matrix_sparse(4, 3, L)[2] for L in range(5)] == [42, 99, 139, 155, 150]
    assert [cv.u_matrix_sparse(4, 2, L)[2] for L in range(5)] == [5, 8, 10, 10, 9]
    assert all(cv.u_matrix_sparse(4, 1, L)[2] == 1 for L in range(5))
    fro = (4.7958315233, 12.3693168769, 18.0277563773, 21.3307290077)
    sums = (42.4112413865, 155.8669849578, 263.9058020750, 300.9592377386)
    asum = (79.7576917479, 554.1920816530, 1140.0624408726, 1633.2434359647)
    for L in range(4):
        _, val, _ = cv.u_matrix_sparse(3, 3, L)
        assert abs(np.sqrt((val ** 2).sum()) - fro[L]) < 1e-8
        assert abs(val.sum() - sums[L]) < 1e-8 and abs(np.abs(val).sum() - asum[L]) < 1e-8


def test_u_matrix_equivariance():
    lmax, R = 2, cv.rotation_matrix(0.4, 0.9, 1.3)
    Dfull = np.zeros((9, 9))
    for l in range(3):
        Dfull[l * l:(l + 1) ** 2, l * l:(l + 1) ** 2] = cv.wigner_D(l, R)
    x = np.random.default_rng(0).normal(size=9)
    for L in range(3):
        U = cv.u_matrix_dense(lmax, 3, L, dtype=np.float64)
        if L == 0:
            U = U[None]
        a = np.einsum("Mijtk,i,j,t->Mk", U, Dfull @ x, Dfull @ x, Dfull @ x)
        b = np.einsum("MN,Nk->Mk", cv.wigner_D(L, R), np.einsum("Mijtk,i,j,t->Mk", U, x, x, x))
        assert np.allclose(a, b, atol=1e-9)


def test_tp_path_counts_and_cg_nnz():
    assert len(tb.tp_plan(0, 4, 4)["paths"]) == 5 and len(tb.tp_plan(4, 4, 4)["paths"]) == 42
    assert len(tb.tp_plan(0, 3, 3)["paths"]) == 4 and len(tb.tp_plan(3, 3, 3)["paths"]) == 23
    assert [p["n_paths"] for p in tb.tp_plan(4, 4, 4)["passes"]] == [5, 8, 10, 10, 9]
    assert [p["n_paths"] for p in tb.tp_plan(3, 3, 3)["passes"]] == [4, 6, 7, 6]
    assert sum(len(v) for v in tb.tp_plan(4, 4, 4)["cg"].values()) == 1158
    assert sum(len(v) for v in tb.tp_plan(3, 3, 3)["cg"].values()) == 353
    assert tb.tp_plan(4, 4, 4)["rows_total"] == 230


def test_rank4_basis_orthonormal_symmetric_equivariant():
    Q = cv.rank4_symmetric_basis()
    Qf = Q.reshape(21, 81)
    assert np.allclose(Qf @ Qf.T, np.eye(21), atol=1e-10)
    for perm in ((0, 2, 1, 3, 4), (0, 1, 2, 4, 3), (0, 3, 4, 1, 2)):
        assert np.allclose(Q, Q.transpose(perm), atol=1e-12)
    R = cv.rotation_matrix(-0.8, 0.5, 2.0)
    rot = np.einsum("mijkl,ai,bj,ck,dl->mabcd", Q, R, R, R, R)
    off = 0
    for L, mul in ((0, 2), (2, 2), (4, 1)):
        D = cv.wigner_D(L, R)
        for _ in range(mul):
            blk = slice(off, off + 2 * L + 1)
            # rows transform like Y_L: Q_m(R.T) = sum_n D[m,n]... equivalently rotating the basis tensors mixes rows by D^T
            assert np.allclose(rot[blk].reshape(2 * L + 1, -1), D.T @ Q[blk].reshape(2 * L + 1, -1), atol=1e-8)
            off += 2 * L + 1


def test_sc_plan_counts():
    plan = tb.sc_plan(4, 3)
    assert plan["n_used"] <= 7744 and plan["n_coef"] % 4 == 0
    assert plan["wshape"][(4, 3)] == 150 and plan["wshape"][(2, 2)] == 10
    wm = tb.sc_weight_map(4, 3)
    assert wm["n_wrows"] == sum(plan["wshape"].values()) and wm["row_ptr"][-1] == len(wm["col"])


@pytest.mark.parametrize("lmax", [2, 3])
def test_sc_plan_equals_reference_contraction(lmax):
    """The Horner program (whatever factorisation _factorise picks) + the weight map reproduce the reference's
    U.W.x contraction (gnn/mace.py:261-277, restated in oracle/model.py) on random inputs, in fp64."""
    import torch
    from oracle.model import SymmetricContraction
    torch.manual_seed(0)
    nu, C, B = 3, 3, 4
    D = (lmax + 1) ** 2
    irreps = [(C, L, (-1) ** L) for L in range(lmax + 1)]
    sc = SymmetricContraction(lmax, irreps, nu, C).double()
    x = torch.randn(B, C, D, dtype=torch.float64)
    ref = sc(x).detach()  # [B, sum_L C*(2L+1)], c-major per L
    plan, wm = tb.sc_plan(lmax, nu), tb.sc_weight_map(lmax, nu)
    Wflat = torch.cat([sc.contractions[k].weights[str(n)].detach() for k in sc.keys for n in range(1, nu + 1)], 0).numpy()
    A = np.zeros((wm["n_coef"], wm["n_wrows"]))
    for r in range(wm["n_coef"]):
        for q in range(wm["row_ptr"][r], wm["row_ptr"][r + 1]):
            A[r, wm["col"][q]] += wm["val"][q]
    T = A @ Wflat  # [n_coef, C]
    xn = x.numpy()
    got = np.zeros((B, C, D))
    o = 0
    for ops_L in plan["ops"]:
        for ops_M in ops_L:
            acc = np.zeros((B, C))
            for i, c1, js in ops_M:
                m1 = np.zeros((B, C)) + (T[c1] if c1 >= 0 else 0.0)
                for j, c2, ts in js:
                    m2 = np.zeros((B, C)) + (T[c2] if c2 >= 0 else 0.0)
                    for t, c3 in ts:
                        m2 = m2 + T[c3] * xn[:, :, t]
                    m1 = m1 + m2 * xn[:, :, j]
                acc = acc + m1 * xn[:, :, i]
            got[:, :, o] = acc
            o += 1
    off, refs = 0, []
    for L in range(lmax + 1):
        refs.append(ref[:, off:off + C * (2 * L + 1)].reshape(B, C, 2 * L + 1).numpy())
        off += C * (2 * L + 1)
    assert np.allclose(got, np.concatenate(refs, -1), rtol=1e-6, atol=1e-6)  # map values are stored in fp32


@pytest.mark.parametrize("s,want", [("32x0e+32x1o", [(32, 0, 1), (32, 1, -1)]), ("2x0e+2x2e+1x4e", [(2, 0, 1), (2, 2, 1), (1, 4, 1)])])
def test_parse_irreps(s, want):
    assert cv.parse_irreps(s) == want and cv.irreps_str(want) == s
