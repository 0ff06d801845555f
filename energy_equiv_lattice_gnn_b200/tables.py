"""Static tables shared by the kernel generator (tools/gen_kernels.py) and the runtime host code.

* `tp_plan`   - channelwise tensor-product paths grouped by output l3 ("passes"), CG non-zeros with the
                sqrt(2 l3 + 1) path coefficient folded in (gnn/mace.py:286-314, gnn/blocks.py:528-535).
* `sc_plan`   - symmetric-contraction "program": the generalised CG tensors of gnn/mace.py:435-477 symmetrised over
                input permutations and ordered for Horner evaluation
                    out_M = sum_i x_i (T1[M,i] + sum_j x_j (T2[M,i,j] + sum_t T3[M,i,j,t] x_t))
                (which two indices of a monomial sit on the outer levels is chosen greedily, see _factorise).
                Every coefficient T is a fixed linear combination of the trainable weights W_nu^L[k, c]; the plan
                carries that map as a sparse matrix (row = coefficient slot in consumption order).
"""
from __future__ import annotations

import functools
import math
from typing import Dict, List, Tuple

import numpy as np

from . import conventions as cv


@functools.lru_cache(maxsize=None)
def tp_plan(lmax_in: int, lmax_sh: int, lmax_out: int):
    """Input irreps = one block per l1 in 0..lmax_in with parity (-1)^l1 (lmax_in = 0: scalar-only first layer).

    Returns dict(paths=[(l1,l2,l3)], passes=[dict(l3, path_ids, n_paths)], cg={path_id: [(i,j,k,val)]},
                 row_off=[per-l3 offset into the per-node message row, in units of C floats], rows_total)
    Message row layout per node: for l3: [(2 l3 + 1), n_l3 * C]  (m-major, then path, then channel).
    """
    irreps1 = [(1, l, (-1) ** l) for l in range(lmax_in + 1)]
    target = [(1, l, (-1) ** l) for l in range(lmax_out + 1)]
    paths = [(l1, l2, l3) for (_i1, l1, l2, l3, _p) in cv.tp_paths(irreps1, lmax_sh, target)]
    passes, row_off, off = [], [], 0
    for l3 in range(lmax_out + 1):
        ids = [p for p, t in enumerate(paths) if t[2] == l3]
        passes.append(dict(l3=l3, path_ids=ids, n_paths=len(ids)))
        row_off.append(off)
        off += (2 * l3 + 1) * len(ids)
    cg = {}
    for p, (l1, l2, l3) in enumerate(paths):
        w = cv.wigner_3j(l1, l2, l3) * math.sqrt(2 * l3 + 1)
        cg[p] = [(int(i), int(j), int(k), float(w[i, j, k])) for i, j, k in np.argwhere(np.abs(w) > 1e-12)]
    return dict(paths=paths, passes=passes, cg=cg, row_off=row_off, rows_total=off)


def _sym_entries(lmax: int, nu: int, L: int):
    """{(M, sorted input tuple): np.array[k]} - U summed over all orderings that map to the sorted tuple."""
    idx, val, k = cv.u_matrix_sparse(lmax, nu, L)
    out: Dict[Tuple, np.ndarray] = {}
    for row, v in zip(idx, val):
        key = (int(row[0]),) + tuple(sorted(int(a) for a in row[1:1 + nu]))
        vec = out.get(key)
        if vec is None:
            vec = out[key] = np.zeros(k)
        vec[int(row[-1])] += v
    out = {kk: vv for kk, vv in out.items() if np.abs(vv).max() > 1e-12}
    return out, k


def _factorise(M: int, ent, correlation: int):
    """Horner tree {outer i: {middle j: set(inner t)}} for output component M.

    A symmetrised monomial x_a x_b x_c may put ANY two of its indices on the (outer, middle) levels.  The evaluation
    costs one multiply-add per coefficient, one per distinct (outer, middle) pair and one per distinct outer index, so
    pairs are chosen greedily by how many still-unassigned monomials they cover (pairs that carry a degree-2
    coefficient exist anyway and are used first).  For lmax = 4 this needs 2.2 k pairs instead of the 4.3 k of the
    fixed i <= j <= t nesting."""
    import collections
    import itertools
    monos = [kk[1:] for kk in ent.get(3, {}) if kk[0] == M] if correlation >= 3 else []
    pairs = [kk[1:] for kk in ent.get(2, {}) if kk[0] == M] if correlation >= 2 else []  # sorted (a, b)

    def pairs_of(m):
        return sorted({tuple(sorted(p)) for p in itertools.combinations(m, 2)})

    chosen = list(sorted(pairs))
    cover = {}
    have = set(chosen)
    left = []
    for m in sorted(monos):
        hit = [p for p in pairs_of(m) if p in have]
        if hit:
            cover[m] = hit[0]
        else:
            left.append(m)
    while left:
        cnt = collections.Counter()
        for m in left:
            for p in pairs_of(m):
                cnt[p] += 1
        best = max(cnt.items(), key=lambda kv: (kv[1], tuple(-a for a in kv[0])))[0]
        chosen.append(best)
        nxt = []
        for m in left:
            if best in pairs_of(m):
                cover[m] = best
            else:
                nxt.append(m)
        left = nxt
    # orientation: the index that occurs in more pairs goes outside (fewer distinct outer indices)
    freq = collections.Counter(a for p in chosen for a in set(p))
    for (m, i) in (kk for kk in ent[1] if kk[0] == M):
        freq[i] += 1000  # an outer level that exists anyway (degree-1 coefficient)
    orient = {}
    for p in chosen:
        a, b = p
        orient[p] = (a, b) if (freq[a], -a) >= (freq[b], -b) else (b, a)
    tree = {}
    for (m, i) in (kk for kk in ent[1] if kk[0] == M):
        tree.setdefault(i, {})
    for p in chosen:
        o, mid = orient[p]
        tree.setdefault(o, {}).setdefault(mid, set())
    for m, p in cover.items():
        rest = list(m)
        rest.remove(p[0])
        rest.remove(p[1])
        o, mid = orient[p]
        tree[o][mid].add(rest[0])
    return tree


@functools.lru_cache(maxsize=None)
def sc_plan(lmax: int, correlation: int):
    """Horner program + coefficient->weight map for all output irreps L = 0..lmax.

    Returns dict with
      ops      : list over L of list over M of list of (i, c1, [(j, c2, [(t, c3)])]) ; c* = coefficient slot or -1
      n_coef   : number of coefficient slots (padded to a multiple of 4)
      wshape   : {(L,nu): k}
      A        : list of (slot, L, nu, k_index, value)  - sparse map  T[slot, c] = sum A * W_{L,nu}[k, c]
    """
    assert 1 <= correlation <= 3
    slot = 0
    A: List[Tuple[int, int, int, int, float]] = []
    ops, wshape = [], {}
    for L in range(lmax + 1):
        ent = {}
        for nu in range(1, correlation + 1):
            e, k = _sym_entries(lmax, nu, L)
            ent[nu] = e
            wshape[(L, nu)] = k
        ops_L = []
        for M in range(2 * L + 1):
            tree = _factorise(M, ent, correlation)
            ops_M = []
            for i in sorted(tree):
                c1 = -1
                if (M, i) in ent[1]:
                    c1 = slot
                    A += [(slot, L, 1, kq, float(v)) for kq, v in enumerate(ent[1][(M, i)]) if abs(v) > 1e-12]
                    slot += 1
                js = []
                for j in sorted(tree[i]):
                    c2 = -1
                    k2 = (M,) + tuple(sorted((i, j)))
                    if correlation >= 2 and k2 in ent[2]:
                        c2 = slot
                        A += [(slot, L, 2, kq, float(v)) for kq, v in enumerate(ent[2][k2]) if abs(v) > 1e-12]
                        slot += 1
                    ts = []
                    for t in sorted(tree[i][j]):
                        k3 = (M,) + tuple(sorted((i, j, t)))
                        A += [(slot, L, 3, kq, float(v)) for kq, v in enumerate(ent[3][k3]) if abs(v) > 1e-12]
                        ts.append((t, slot))
                        slot += 1
                    js.append((j, c2, ts))
                ops_M.append((i, c1, js))
            ops_L.append(ops_M)
        ops.append(ops_L)
    n_coef = (slot + 3) // 4 * 4
    return dict(ops=ops, n_coef=n_coef, n_used=slot, wshape=wshape, A=A)


def sc_weight_map(lmax: int, correlation: int):
    """CSR (by coefficient slot) and CSC (by flat weight row) forms of the coefficient<-weight map.

    Flat weight rows: for L in 0..lmax, for nu in 1..correlation: k rows (the order weights are packed by the host).
    Returns dict(row_ptr[n_coef+1], col[nnz], val[nnz], t_ptr[n_wrows+1], t_row[nnz], t_val[nnz], w_off{(L,nu): first row},
                 n_wrows, n_coef)
    """
    plan = sc_plan(lmax, correlation)
    w_off, off = {}, 0
    for L in range(lmax + 1):
        for nu in range(1, correlation + 1):
            w_off[(L, nu)] = off
            off += plan["wshape"][(L, nu)]
    n_wrows, n_coef = off, plan["n_coef"]
    rows = np.array([a[0] for a in plan["A"]], dtype=np.int64)
    cols = np.array([w_off[(a[1], a[2])] + a[3] for a in plan["A"]], dtype=np.int64)
    vals = np.array([a[4] for a in plan["A"]], dtype=np.float64)
    o = np.lexsort((cols, rows))
    rows, cols, vals = rows[o], cols[o], vals[o]
    row_ptr = np.zeros(n_coef + 1, dtype=np.int32)
    np.add.at(row_ptr, rows + 1, 1)
    row_ptr = np.cumsum(row_ptr).astype(np.int32)
    ot = np.lexsort((rows, cols))
    t_ptr = np.zeros(n_wrows + 1, dtype=np.int32)
    np.add.at(t_ptr, cols + 1, 1)
    t_ptr = np.cumsum(t_ptr).astype(np.int32)
    return dict(row_ptr=row_ptr, col=cols.astype(np.int32), val=vals.astype(np.float32),
                t_ptr=t_ptr, t_row=rows[ot].astype(np.int32), t_val=vals[ot].astype(np.float32),
                w_off=w_off, n_wrows=n_wrows, n_coef=n_coef)


def sc_slot_table(lmax: int, correlation: int) -> np.ndarray:
    """int32 [n_coef, 4]: coefficient slot -> (output component, i, j, t) of the monomial it multiplies (-1 = absent);
    padding slots are (-1,-1,-1,-1).  Used by the table-driven coefficient gradient (eeg_sc_coef_grad)."""
    plan = sc_plan(lmax, correlation)
    tab = -np.ones((plan["n_coef"], 4), dtype=np.int32)
    o = 0
    for ops_L in plan["ops"]:
        for ops_M in ops_L:
            for i, c1, js in ops_M:
                if c1 >= 0:
                    tab[c1] = (o, i, -1, -1)
                for j, c2, ts in js:
                    if c2 >= 0:
                        tab[c2] = (o, i, j, -1)
                    for t, c3 in ts:
                        tab[c3] = (o, i, j, t)
            o += 1
    return tab
