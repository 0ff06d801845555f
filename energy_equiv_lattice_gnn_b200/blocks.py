"""Drop-in device versions of reference blocks that `EnergyEquivGNN` does not construct but `gnn/blocks.py` / `gnn/mace.py`
define: same class names, constructor arguments and forward signature."""
from __future__ import annotations

import torch

from . import native


class RadialEmbeddingBlock(torch.nn.Module):
    """gnn/blocks.py:89-102: Bessel basis (gnn/mace.py:28-71, non-trainable) x polynomial cutoff (gnn/mace.py:74-106).
    forward(edge_lengths [E,1]) -> [E, num_bessel] through `eeg_radial_embedding_fwd`.  CUDA only."""

    def __init__(self, r_max: float, num_bessel: int, num_polynomial_cutoff: int):
        super().__init__()
        self.r_max, self.num_bessel, self.p = float(r_max), int(num_bessel), int(num_polynomial_cutoff)
        self.out_dim = num_bessel

    def forward(self, edge_lengths: torch.Tensor) -> torch.Tensor:
        if not edge_lengths.is_cuda:
            raise RuntimeError("RadialEmbeddingBlock (B200) has no CPU path: move the input to a CUDA device")
        return native.radial_embedding(edge_lengths.to(torch.float32), self.num_bessel, self.r_max, self.p)
