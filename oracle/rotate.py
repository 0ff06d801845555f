"""TEST INFRASTRUCTURE - CPU restatement of the reference's `RotateLat` transform (scripts/train_utils.py:114-146) and of
the Cartesian rank-4 -> Mandel conversion it calls (lattices/lattices/utils/elasticity_func.py:406-428,479-495), for a
collated batch with one rotation per lattice.  Pinned against the reference's own class run in the build container:
tests/golden/ref_rotate.pt (oracle/make_golden_rotate.py).  Only tests/ may import this."""
from __future__ import annotations

import math

import torch

# Mandel index -> Cartesian pair (elasticity_func.py:407-424)
_PAIRS = [(0, 0), (1, 1), (2, 2), (1, 2), (0, 2), (0, 1)]


def cart4_to_mandel(C4: torch.Tensor) -> torch.Tensor:
    """[...,3,3,3,3] -> [...,6,6]: upper triangle picked from C4[a,b,c,d], mirrored, times {1, sqrt2, 2} (:425-428)."""
    out = C4.new_zeros(C4.shape[:-4] + (6, 6))
    for i in range(6):
        a, b = _PAIRS[i]
        for j in range(i, 6):
            c, d = _PAIRS[j]
            out[..., i, j] = C4[..., a, b, c, d]
            out[..., j, i] = C4[..., a, b, c, d]
    s2 = math.sqrt(2.0)
    mask = torch.ones(6, 6, dtype=C4.dtype)
    mask[:3, 3:] = s2
    mask[3:, :3] = s2
    mask[3:, 3:] = 2.0
    return out * mask


def rotate_batch(Q, positions, shifts, edge_index, batch, stiffness4=None, compliance4=None):
    """Q [G,3,3]; returns (positions', shifts', stiffness Mandel, compliance Mandel) - train_utils.py:122-125,133-134 with
    the lattice's own Q_g selected per node / per edge (an edge belongs to its sender's lattice)."""
    Qn = Q[batch]
    Qe = Q[batch[edge_index[0]]]
    pos = torch.einsum("nij,nj->ni", Qn, positions)
    sh = torch.einsum("eij,ej->ei", Qe, shifts)
    rot4 = lambda T: torch.einsum("gijkl,gai,gbj,gck,gdl->gabcd", T, Q, Q, Q, Q)  # noqa: E731
    C = None if stiffness4 is None else cart4_to_mandel(rot4(stiffness4))
    S = None if compliance4 is None else cart4_to_mandel(rot4(compliance4))
    return pos, sh, C, S
