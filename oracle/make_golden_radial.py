#!/usr/bin/env python
"""TEST INFRASTRUCTURE - tests/golden/ref_radial_embedding.pt: the reference's own RadialEmbeddingBlock (gnn/blocks.py:89-102
over BesselBasis / PolynomialCutoff, gnn/mace.py:28-106) run in fp64 on seeded edge lengths (inside and beyond r_max).
Build container only (needs /root/reference)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import refshim  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def main():
    ref = refshim.install()
    import importlib
    blocks = importlib.import_module("gnn.blocks")
    torch.set_default_dtype(torch.float64)
    g = torch.Generator().manual_seed(11)
    cases = []
    for r_max, nb, p in ((0.6, 8, 6), (1.0, 6, 5), (2.5, 16, 3)):
        x = torch.rand(257, 1, generator=g, dtype=torch.float64) * 1.3 * r_max + 1e-3
        blk = blocks.RadialEmbeddingBlock(r_max=r_max, num_bessel=nb, num_polynomial_cutoff=p)
        cases.append(dict(r_max=r_max, num_bessel=nb, p=p, x=x, out=blk(x)))
    torch.set_default_dtype(torch.float32)
    torch.save(cases, os.path.join(GOLD, "ref_radial_embedding.pt"))
    print("ref_radial_embedding.pt:", [tuple(c["out"].shape) for c in cases])


if __name__ == "__main__":
    main()
