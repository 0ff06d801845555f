"""RadialEmbeddingBlock (gnn/blocks.py:89-102: Bessel basis x polynomial cutoff, gnn/mace.py:28-106) on the device against the
reference's own class run in fp64 (tests/golden/ref_radial_embedding.pt, oracle/make_golden_radial.py).  fp32 tolerance 1e-5 of
the tensor's max (the recurrence for sin(k t) and the division by x are the only fp32 effects)."""
import os

import pytest
import torch

from energy_equiv_lattice_gnn_b200.blocks import RadialEmbeddingBlock

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_radial_embedding_matches_reference_class(cuda_device):
    cases = torch.load(os.path.join(GOLD, "ref_radial_embedding.pt"), weights_only=False)
    assert len(cases) == 3
    for c in cases:
        blk = RadialEmbeddingBlock(c["r_max"], c["num_bessel"], c["p"])
        got = blk(c["x"].to(torch.float32).to(cuda_device)).double().cpu()
        want = c["out"]
        assert got.shape == want.shape
        err = float((got - want).abs().max() / want.abs().max())
        assert err < 1e-5, (c["r_max"], c["num_bessel"], err)
        beyond = (c["x"].reshape(-1) >= c["r_max"])
        assert beyond.any() and float(got[beyond].abs().max()) == 0.0     # cutoff: exactly zero beyond r_max
    with pytest.raises(RuntimeError):
        RadialEmbeddingBlock(1.0, 8, 6)(torch.rand(4, 1))
