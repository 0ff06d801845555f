"""Generate tests/golden/ref_metrics.pt by running the reference's OWN `obtain_errors` / `aggr_errors`
(scripts/train_utils.py:149-202), cut out of the unmodified file with `ast` (the module itself imports pytorch_lightning)
and executed against the reference's elasticity_func (oracle/refshim.py).  Targets: Mandel stiffness of the
reference-processed sample_cat.lat lattices (tests/golden/ref_sample_cat.pt); predictions: targets + seeded symmetric noise.
The 250 directions are drawn inside `obtain_errors` from torch's global CPU generator: the seed is stored so tests redraw them.

    python -m oracle.make_golden_metrics        (build container only: needs /root/reference)
"""
from __future__ import annotations

import ast
import os
import sys

import numpy as np
import pandas as pd
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import refshim  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
SEED = 1234


def reference_functions(ref):
    path = os.path.join(refshim.REFERENCE_ROOT, "scripts", "train_utils.py")
    src = open(path).read()
    ns = dict(torch=torch, np=np, pd=pd, elasticity_func=ref.elasticity_func)
    for node in ast.parse(src).body:
        if isinstance(node, ast.FunctionDef) and node.name in ("obtain_errors", "aggr_errors"):
            exec(compile(ast.get_source_segment(src, node), path, "exec"), ns)
    return ns["obtain_errors"], ns["aggr_errors"]


def main():
    ref = refshim.install()
    obtain_errors, aggr_errors = reference_functions(ref)
    cat = torch.load(os.path.join(GOLD, "ref_sample_cat.pt"), weights_only=False)
    ef = ref.elasticity_func
    target = torch.cat([ef.stiffness_cart_4_to_Mandel(d["stiffness_cart4"].double()) for d in cat], 0).float()
    g = torch.Generator().manual_seed(5)
    noise = torch.randn(target.shape, generator=g) * 0.05 * target.abs().amax(dim=(1, 2), keepdim=True)
    pred = target + 0.5 * (noise + noise.transpose(1, 2))
    names = np.array([d["name"] for d in cat])
    rel_dens = torch.tensor([d["rel_dens"] for d in cat])
    half = len(cat) // 2
    results = [({"stiffness": pred[:half]}, {"stiffness": target[:half], "name": names[:half], "rel_dens": rel_dens[:half]}),
               ({"stiffness": pred[half:]}, {"stiffness": target[half:], "name": names[half:], "rel_dens": rel_dens[half:]})]
    torch.manual_seed(SEED)
    df = obtain_errors(results, "test")
    agg = aggr_errors(df.copy())
    out = dict(seed=SEED, pred=pred, target=target, half=half, rel_dens=rel_dens,
               columns={k: torch.tensor(np.stack(df[k].to_numpy()).astype(np.float64)) for k in
                        ("mean_stiffness", "loss", "mseloss", "dir_loss", "target_eig", "predicted_eig")},
               aggregated={k: float(v) for k, v in agg.items()})
    torch.save(out, os.path.join(GOLD, "ref_metrics.pt"))
    print(len(cat), "lattices;", {k: round(v, 6) for k, v in list(out["aggregated"].items())[:6]})
    print("wrote ref_metrics.pt", os.path.getsize(os.path.join(GOLD, "ref_metrics.pt")), "bytes")


if __name__ == "__main__":
    main()
