"""TEST INFRASTRUCTURE - restatement of the generalised Clebsch-Gordan recursion of the reference
(gnn/mace.py:363-432 `_wigner_nj`, :435-477 `U_matrix_real`) in dense torch, as the reference does it.

It differs from the reference in one documented way: blocks whose output irrep is not requested are carried
symbolically at the LAST level only (they are discarded by `U_matrix_real` anyway, gnn/mace.py:461-476), which keeps
the dense fp64 build at seconds instead of minutes.  The product builds its tables by an independent sparse path
enumeration (energy_equiv_lattice_gnn_b200/conventions.py: u_matrix_sparse); tests compare the two.

PARITY UNPINNED (see oracle/e3nn_lite.py header).
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import numpy as np
import torch

from energy_equiv_lattice_gnn_b200 import conventions as cv

Ir = Tuple[int, int]  # (l, p); e3nn 0.5 Irrep is a tuple -> sorted() orders by (l, p)


def _ir_mul(a: Ir, b: Ir) -> List[Ir]:
    return [(l, a[1] * b[1]) for l in range(abs(a[0] - b[0]), a[0] + b[0] + 1)]


def _wigner_nj(irrepss: List[List[Ir]], dtype, keep_last: Optional[List[Ir]] = None):
    """gnn/mace.py:363-432 with mul == 1 everywhere and normalization='component'.
    Returns list of (ir, C) sorted by ir (stable), C of shape [ir.dim, D_1, ..., D_n]."""
    if len(irrepss) == 1:
        (irreps,) = irrepss
        D = sum(2 * l + 1 for l, _ in irreps)
        e = torch.eye(D, dtype=dtype)
        ret, i = [], 0
        for l, p in irreps:
            ret.append(((l, p), e[i:i + 2 * l + 1]))
            i += 2 * l + 1
        return ret
    *left, right = irrepss
    dims_left = [sum(2 * l + 1 for l, _ in irr) for irr in left]
    D_right = sum(2 * l + 1 for l, _ in right)
    ret = []
    for ir_left, C_left in _wigner_nj(left, dtype):
        i = 0
        for ir in right:
            for ir_out in _ir_mul(ir_left, ir):
                if keep_last is not None and ir_out not in keep_last:
                    continue
                C = torch.tensor(np.array(cv.wigner_3j(ir_out[0], ir_left[0], ir[0])), dtype=dtype)
                C *= (2 * ir_out[0] + 1) ** 0.5
                C = torch.einsum("jk,ijl->ikl", C_left.flatten(1), C)
                C = C.reshape(2 * ir_out[0] + 1, *dims_left, 2 * ir[0] + 1)
                E = torch.zeros(2 * ir_out[0] + 1, *dims_left, D_right, dtype=dtype)
                E[..., i:i + 2 * ir[0] + 1] = C
                ret.append((ir_out, E))
            i += 2 * ir[0] + 1
    return sorted(ret, key=lambda x: x[0])


def U_matrix_real(lmax: int, l_out: int, correlation: int, dtype=torch.float64) -> torch.Tensor:
    """gnn/mace.py:435-477 for irreps_in = 0e+1o+..+lmax, irreps_out = one irrep (L, (-1)^L).  Returns the stacked
    tensor [(2L+1), D^nu, k] with `base_o3.squeeze()` semantics (M axis dropped for L = 0)."""
    coupling = [(l, (-1) ** l) for l in range(lmax + 1)]
    target = (l_out, (-1) ** l_out)
    wig = _wigner_nj([coupling] * correlation, dtype, keep_last=[target] if correlation > 1 else None)
    stack = [C.squeeze().unsqueeze(-1) for ir, C in wig if ir == target]
    return torch.cat(stack, dim=-1)
