#!/usr/bin/env python
"""TEST INFRASTRUCTURE - tests/golden/ref_forward_small.pt: the reference EnergyEquivGNN (gnn/model.py:114-161) run in fp64 at
hyper-parameters away from the defaults (the constructor domain of gnn/model.py:27-90): lmax 2 / correlation 2, lmax 1 /
correlation 1, and lmax 4 with correlation 2, a 32-wide 2-layer radial MLP and 10 edge bases.  Build container only."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from energy_equiv_lattice_gnn_b200.data import collate, synthetic_lattice  # noqa: E402
from oracle import make_golden, refshim  # noqa: E402

CASES = {
    "lmax2_nu2": dict(lmax=2, over=dict(correlation=2)),
    "lmax1_nu1": dict(lmax=1, over=dict(correlation=1)),
    "lmax4_nu2_mlp32": dict(lmax=4, over=dict(correlation=2, inter_MLP_dim=32, inter_MLP_layers=2, num_edge_bases=10)),
    "lmax3_nu3_mlp1": dict(lmax=3, over=dict(inter_MLP_layers=1)),
}


def build(ref):
    data = [synthetic_lattice(4100, 6), synthetic_lattice(4101, 9), synthetic_lattice(4102, 3)]
    batch = collate(data)
    rmax = float(batch.edge_attr.max())
    out = {}
    for name, c in CASES.items():
        fix, _, _ = make_golden.run_reference(ref, make_golden.params_for(c["lmax"], rmax, **c["over"]), batch, seed=7)
        out[name] = dict(params=fix["params"], state_dict=fix["state_dict"], stiffness=fix["stiffness"], batch=fix["batch"])
    return out


if __name__ == "__main__":
    ref = refshim.install()
    fx = build(ref)
    torch.save(fx, os.path.join(make_golden.GOLD, "ref_forward_small.pt"))
    for k, v in fx.items():
        print(k, tuple(v["stiffness"].shape), float(v["stiffness"].abs().max()))
