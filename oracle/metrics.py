"""TEST INFRASTRUCTURE - CPU restatement (torch fp64) of the reference's test metrics, scripts/train_utils.py:149-172
(`obtain_errors`) with the Mandel -> Cartesian rank-4 expansion of lattices/lattices/utils/elasticity_func.py:430-463
written out as the reference does it (81-term contraction per direction - deliberately NOT the 21-term Mandel form the
kernel uses).  Pinned by tests/golden/ref_metrics.pt (the reference's own functions, oracle/make_golden_metrics.py).
Only tests/ may import this."""
from __future__ import annotations

import math

import torch

_PAIRS = [(0, 0), (1, 1), (2, 2), (1, 2), (0, 2), (0, 1)]  # elasticity_func.py:431-449 (1-based there)


def mandel_to_cart4(C: torch.Tensor) -> torch.Tensor:
    C4 = C.new_zeros(C.shape[0], 3, 3, 3, 3)
    for i in range(6):
        a, b = _PAIRS[i]
        for j in range(i, 6):
            c, d = _PAIRS[j]
            val = C[:, i, j] / (math.sqrt(2) if i > 2 else 1.0) / (math.sqrt(2) if j > 2 else 1.0)
            for (p, q, r, s) in ((a, b, c, d), (b, a, c, d), (a, b, d, c), (c, d, a, b), (b, a, d, c), (c, d, b, a), (d, c, a, b), (d, c, b, a)):
                C4[:, p, q, r, s] = val
    return C4


def per_lattice_metrics(pred: torch.Tensor, target: torch.Tensor, directions: torch.Tensor):
    """-> dict(mseloss, loss, dir_loss, mean_stiffness [G]; predicted_eig, target_eig [G,6])  (:157-170)."""
    pred, target, d = pred.double(), target.double(), directions.double()
    c = torch.einsum("gijkl,pi,pj,pk,pl->gp", mandel_to_cart4(target), d, d, d, d)
    cp = torch.einsum("gijkl,pi,pj,pk,pl->gp", mandel_to_cart4(pred), d, d, d, d)
    return dict(mseloss=((pred - target) ** 2).mean(dim=(1, 2)), loss=(pred - target).abs().mean(dim=(1, 2)),
                dir_loss=(c - cp).abs().mean(dim=1), mean_stiffness=c.mean(dim=1),
                predicted_eig=torch.linalg.eigvalsh(pred), target_eig=torch.linalg.eigvalsh(target))
