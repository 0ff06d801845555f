"""Generate tests/golden/ref_interaction_mean.pt: the reference `EnergyEquivGNN` (gnn/model.py, run through oracle/refshim.py)
with `interaction_reduction='mean'` (gnn/blocks.py:595-597, torch_scatter reduce='mean') on the batch and weights of
tests/golden/ref_forward_lmax4.pt, fp64.

    python -m oracle.make_golden_mean        (build container only: needs /root/reference)
"""
from __future__ import annotations

import os
import sys
from argparse import Namespace

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import refshim  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def main():
    ref = refshim.install()
    fx = torch.load(os.path.join(GOLD, "ref_forward_lmax4.pt"), weights_only=False)
    params = Namespace(**{**fx["params"], "interaction_reduction": "mean"})
    torch.set_default_dtype(torch.float64)
    try:
        model = ref.model.EnergyEquivGNN(params)
        sd = {k: (v.double() if v.is_floating_point() else v) for k, v in fx["state_dict"].items()}
        missing, unexpected = model.load_state_dict(sd, strict=False)
        assert not unexpected and all("U_matrix" in k for k in missing), (missing, unexpected)
        import torch_geometric.data as tgd
        b = tgd.Batch(**{k: (v.double() if torch.is_tensor(v) and v.is_floating_point() else v) for k, v in fx["batch"].items()})
        with torch.no_grad():
            out = model(b)["stiffness"]
    finally:
        torch.set_default_dtype(torch.float32)
    rel = float((out - fx["stiffness"]).abs().max() / fx["stiffness"].abs().max())
    print(f"stiffness {tuple(out.shape)}; differs from the 'sum' fixture by {rel:.3f} (relative)")
    assert rel > 1e-3
    torch.save(dict(stiffness=out), os.path.join(GOLD, "ref_interaction_mean.pt"))


if __name__ == "__main__":
    main()
