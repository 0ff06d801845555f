"""Generate tests/golden/ref_rotate.pt by running the reference's OWN `RotateLat` class (scripts/train_utils.py:114-146).

scripts/train_utils.py imports pytorch_lightning / torch_geometric at module level (absent here), so the class source is
cut out of the unmodified file with `ast` and executed with the stand-ins of oracle/refshim.py (`Data` bag, the
reference's elasticity_func loaded from /root/reference).  Inputs: the first lattices of tests/golden/ref_sample_cat.pt
(reference-processed sample_cat.lat, Cartesian rank-4 targets) in fp64 and seeded proper rotations.

    python -m oracle.make_golden_rotate        (build container only: needs /root/reference)
"""
from __future__ import annotations

import ast
import os
import sys
from typing import Optional

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import refshim  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")


def reference_rotatelat(ref):
    path = os.path.join(refshim.REFERENCE_ROOT, "scripts", "train_utils.py")
    src = open(path).read()
    node = next(n for n in ast.parse(src).body if isinstance(n, ast.ClassDef) and n.name == "RotateLat")
    code = ast.get_source_segment(src, node)
    import torch_geometric.data as tgd  # refshim stand-in
    ns = dict(torch=torch, Tensor=torch.Tensor, Optional=Optional, Data=tgd.Data, elasticity_func=ref.elasticity_func,
              o3=None)  # Q is always passed explicitly => o3.rand_matrix is never reached
    exec(compile(code, path, "exec"), ns)
    return ns["RotateLat"]


def main():
    ref = refshim.install()
    RotateLat = reference_rotatelat(ref)
    import torch_geometric.data as tgd
    cat = torch.load(os.path.join(GOLD, "ref_sample_cat.pt"), weights_only=False)
    picks = [d for d in cat if d["positions"].shape[0] > 1][:5] + cat[:1]
    g = torch.Generator().manual_seed(7)
    cases = []
    for d in picks:
        A = torch.randn(3, 3, generator=g, dtype=torch.float64)
        Q, R = torch.linalg.qr(A)
        Q = Q * torch.sign(torch.diagonal(R))
        if torch.det(Q) < 0:
            Q[:, 0] = -Q[:, 0]
        lat = tgd.Data(node_attrs=d["node_attrs"], edge_attr=d["edge_attr"], edge_index=d["edge_index"],
                       positions=d["positions"].double(), shifts=d["shifts"].double(), rel_dens=d["rel_dens"],
                       stiffness=d["stiffness_cart4"].double(), compliance=d["compliance_cart4"].double(), name=d["name"])
        out = RotateLat(rotate=True)(lat, Q)
        plain = RotateLat(rotate=False)(lat)
        cases.append(dict(name=d["name"], Q=Q, positions=d["positions"], shifts=d["shifts"], edge_index=d["edge_index"],
                          stiffness_cart4=d["stiffness_cart4"], compliance_cart4=d["compliance_cart4"],
                          out_positions=out.positions, out_shifts=out.shifts, out_stiffness=out.stiffness,
                          out_compliance=out.compliance, plain_stiffness=plain.stiffness, plain_compliance=plain.compliance))
        print(d["name"], tuple(d["positions"].shape), tuple(out.stiffness.shape))
    torch.save(cases, os.path.join(GOLD, "ref_rotate.pt"))
    print("wrote", os.path.join(GOLD, "ref_rotate.pt"), os.path.getsize(os.path.join(GOLD, "ref_rotate.pt")), "bytes")


if __name__ == "__main__":
    main()
