#!/usr/bin/env python
"""TEST INFRASTRUCTURE - generate tests/golden/* by RUNNING THE REFERENCE'S OWN SOURCES from /root/reference
(through oracle/refshim.py, which stands in for the absent third-party packages only).

    python oracle/make_golden.py            # only works in the build container (needs /root/reference)

Fixtures written (all small; weights are rounded to fp32-representable values so fp32 and fp64 consumers see the
same model):
  ref_forward_lmax4.pt  reference EnergyEquivGNN (gnn/model.py:114-161), default hyper-parameters
                        (scripts/train_main.py:25-52), fp64, on a mixed batch: synthetic lattices +
                        real lattices of lattices/examples/sample_cat.lat processed by the reference's
                        GLAMM_Dataset.process_one (gnn/datasets.py:115-276).  Inputs, state_dict (no U buffers),
                        stiffness and intermediates (edge features, SH, per-layer node features, readout).
  ref_forward_lmax3.pt  same with lmax=3 (BASELINE config 4 irreps at 32 channels).
  ref_variants.pt       stiffness for the other positive_function / global_reduction settings on the lmax4 batch.
  ref_tables.json       reference tp_out_irreps_with_instructions (gnn/mace.py:286-314) instruction lists and
                        U_matrix_real (gnn/mace.py:435-477) shapes / norms / sums / nnz, lmax 3 and 4, nu 1..3.
  ref_mandel.pt         lattices/utils/elasticity_func.py Mandel <-> Cartesian rank-4 pairs (pins Cart_4_to_Mandel).
  ref_sample_cat.pt     the 19 usable sample_cat.lat lattices as processed by the reference (inputs + targets).
"""
from __future__ import annotations

import json
import os
import sys
import time
import warnings
from argparse import Namespace

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from energy_equiv_lattice_gnn_b200 import conventions as cv  # noqa: E402
from energy_equiv_lattice_gnn_b200.data import LatticeData, collate, synthetic_lattice  # noqa: E402
from oracle import refshim  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")
warnings.filterwarnings("ignore")


def params_for(lmax: int, rmax: float, **over) -> Namespace:
    irr = lambda m: "+".join(f"{m}x{l}{'e' if l % 2 == 0 else 'o'}" for l in range(lmax + 1))  # noqa: E731
    p = dict(lmax=lmax, hidden_irreps=irr(32), readout_irreps=irr(16), num_edge_bases=6, interaction_reduction="sum",
             interaction_bias=True, agg_norm_const=4.0, inter_MLP_dim=64, inter_MLP_layers=3, correlation=3,
             global_reduction="mean", message_passes=2, positive_function="matrix_power_2", max_edge_radius=rmax)
    p.update(over)
    return Namespace(**p)


def sample_cat(ref):
    """Reference Catalogue + process_one over lattices/examples/sample_cat.lat -> list of LatticeData-like dicts."""
    path = os.path.join(refshim.REFERENCE_ROOT, "lattices", "examples", "sample_cat.lat")
    cat = ref.Catalogue.from_file(path, 0)
    out = []
    for lat in cat:
        if "fundamental_edge_adjacency" not in lat or len(lat.get("compliance_tensors_M", lat.get("compliance_tensors_V", {}))) == 0:
            continue
        try:
            datas = ref.datasets.GLAMM_Dataset.process_one(lat, "r", "cartesian_4")
        except Exception as e:  # lattices without a fundamental representation
            print("skip", lat.get("name"), type(e).__name__, e)
            continue
        for d in datas:
            out.append(dict(name=d.name, rel_dens=float(d.rel_dens), positions=d.positions, node_attrs=d.node_attrs,
                            edge_index=d.edge_index, shifts=d.shifts, unit_shifts=d.unit_shifts, edge_attr=d.edge_attr,
                            stiffness_cart4=d.stiffness.to(torch.float64), compliance_cart4=d.compliance.to(torch.float64)))
    return out


def to_lattice(d) -> LatticeData:
    return LatticeData(positions=d["positions"], node_attrs=d["node_attrs"], edge_index=d["edge_index"], shifts=d["shifts"],
                       edge_attr=d["edge_attr"], name=d["name"], rel_dens=d["rel_dens"])


def run_reference(ref, params, batch, seed=0, extra_params=()):
    """Build the reference module in fp64 with fp32-representable weights, run it, capture intermediates."""
    torch.set_default_dtype(torch.float64)
    try:
        torch.manual_seed(seed)
        model = ref.model.EnergyEquivGNN(params)
        with torch.no_grad():
            for p in model.parameters():
                p.copy_(p.float().double())
            for name, b in model.named_buffers():  # Q_flat, Mandel mask travel in the state_dict as fp32 too
                if b.is_floating_point() and "U_matrix" not in name:
                    b.copy_(b.float().double())
            # biases are zero-initialised in e3nn; give them values so the bias paths are exercised
            g = torch.Generator("cpu").manual_seed(seed + 1)
            for name, p in model.named_parameters():
                if name.endswith(".bias") and "conv_tp_weights" not in name and "node_ft_embedding" not in name:
                    p.copy_((0.1 * torch.randn(p.shape, generator=g, dtype=torch.float64)).float().double())
        b64 = refshim._Bag(**{k: (v.double() if torch.is_tensor(v) and v.is_floating_point() else v) for k, v in batch.__dict__.items()})
        inter = {}
        head = model.stiffness_head
        hooks = []
        for i, layer in enumerate(head.layers):
            hooks.append(layer.register_forward_hook(lambda m, a, o, i=i: inter.__setitem__(f"layer{i}_out", o.detach())))
            hooks.append(layer.interaction.register_forward_hook(lambda m, a, o, i=i: inter.__setitem__(f"layer{i}_message", o[0].detach())))
        hooks.append(head.nonlin_readout.register_forward_hook(lambda m, a, o: inter.__setitem__("readout", o.detach())))
        hooks.append(head.linear.register_forward_hook(lambda m, a, o: inter.__setitem__("q", o.detach())))
        hooks.append(head.cart_to_Mandel.register_forward_hook(lambda m, a, o: inter.__setitem__("A", o.detach())))
        hooks.append(model.spherical_harmonics.register_forward_hook(lambda m, a, o: inter.__setitem__("edge_sh", o.detach())))
        t0 = time.time()
        with torch.no_grad():
            out = model(b64)["stiffness"]
        print(f"  reference forward: {time.time() - t0:.1f} s, |C|max = {float(out.abs().max()):.4e}")
        for h in hooks:
            h.remove()
        variants = {}
        for over in extra_params:
            pv = Namespace(**{**vars(params), **over})
            head.positive_layer = ref.blocks.PositiveLayer(pv)
            head.global_reduction = pv.global_reduction
            with torch.no_grad():
                variants[json.dumps(over, sort_keys=True)] = model(b64)["stiffness"]
        sd = {k: (v.float() if v.is_floating_point() else v) for k, v in model.state_dict().items() if "U_matrix" not in k}
        return dict(params=vars(params), state_dict=sd, stiffness=out, inter=inter,
                    batch={k: v for k, v in batch.__dict__.items()}), variants, model
    finally:
        torch.set_default_dtype(torch.float32)


def tables(ref):
    o3 = sys.modules["e3nn.o3"]
    out = {"tp_instructions": {}, "U": {}}
    for lmax in (3, 4):
        sh = o3.Irreps.spherical_harmonics(lmax)
        hidden = (sh * 32).sort()[0].simplify()
        for name, irr1 in (("layer0", o3.Irreps("32x0e")), ("layer1", hidden)):
            irreps_mid, ins = ref.mace.tp_out_irreps_with_instructions(irr1, sh, hidden)
            out["tp_instructions"][f"lmax{lmax}_{name}"] = dict(
                irreps_mid=str(irreps_mid), irreps_mid_simplified=str(irreps_mid.simplify()),
                instructions=[[int(a), int(b), int(c)] for a, b, c, _, _ in ins],
                l_triples=[[irr1[a].ir.l, sh[b].ir.l, irreps_mid[c].ir.l] for a, b, c, _, _ in ins])
        for nu in (1, 2, 3):
            for L in range(lmax + 1):
                t0 = time.time()
                U = ref.mace.U_matrix_real(irreps_in=sh, irreps_out=o3.Irreps(str(o3.Irrep(L, (-1) ** L))), correlation=nu,
                                           dtype=torch.float64)[-1]
                ours = torch.tensor(cv.u_matrix_dense(lmax, nu, L, dtype=np.float64))
                assert U.shape == ours.shape, (U.shape, ours.shape)
                err = float((U - ours).abs().max())
                assert err < 1e-12, (lmax, nu, L, err)
                out["U"][f"lmax{lmax}_nu{nu}_L{L}"] = dict(shape=list(U.shape), fro=float(U.norm()), sum=float(U.sum()),
                                                          abssum=float(U.abs().sum()), nnz=int((U.abs() > 1e-12).sum()),
                                                          max_abs_diff_vs_conventions=err)
                print(f"  U lmax={lmax} nu={nu} L={L}: {tuple(U.shape)} reference == conventions.u_matrix_dense ({err:.1e}) [{time.time() - t0:.1f}s]")
    return out


def mandel(ref):
    ef = ref.elasticity_func
    rng = np.random.default_rng(7)
    Ms, C4s = [], []
    for _ in range(4):
        A = rng.normal(size=(6, 6))
        M = A @ A.T
        Ms.append(M)
        C4s.append(ef.numpy_Mandel_to_cart_4(M))
    M, C4 = np.stack(Ms), np.stack(C4s)
    back = ef.numpy_cart_4_to_Mandel(C4) if hasattr(ef, "numpy_cart_4_to_Mandel") else None
    if back is not None:
        assert np.abs(back - M).max() < 1e-12
    # the reference's own torch module on the same tensors (gnn/blocks.py:392-425)
    mod = ref.blocks.Cart_4_to_Mandel()
    got = mod(torch.tensor(C4, dtype=torch.float32)).double().numpy()
    assert np.abs(got - M).max() < 1e-5 * np.abs(M).max()
    return dict(mandel=torch.tensor(M), cart4=torch.tensor(C4), a=mod.a, b=mod.b, c=mod.c, d=mod.d,
                mask=mod.mask.double(), rows=mod.rows, cols=mod.cols)


def forward_fixtures(ref, cat, verbose=True):
    """(ref_forward_lmax4, ref_variants, ref_forward_lmax3) fixtures, in memory (tests/test_golden_reproducible.py re-runs
    this against the committed files)."""
    say = print if verbose else (lambda *a, **k: None)
    # mixed batch: 2 synthetic (N=6, one cubic cell, one orthorhombic) + real lattices with N in {1, 4, 6, 12}
    real, seen = [], set()
    for d in cat:
        n = int(d["positions"].shape[0])
        if n <= 12 and n not in seen:
            seen.add(n)
            real.append(to_lattice(d))
    data = [synthetic_lattice(4000, 6), synthetic_lattice(4001, 7)] + real
    batch = collate(data)
    rmax = float(batch.edge_attr.max())
    say(f"batch: {len(data)} lattices, N={batch.positions.shape[0]}, E={batch.edge_index.shape[1]}, names={[d.name for d in data]}")

    say("reference forward lmax=4 (default hyper-parameters) ...")
    variants = [dict(positive_function=f) for f in ("matrix_power_4", "matrix_trunc_exp_2", "matrix_trunc_exp_4", "none", "matrix_exp")]
    variants.append(dict(global_reduction="sum"))
    fix4, var4, _ = run_reference(ref, params_for(4, rmax), batch, seed=0, extra_params=variants)
    fix4["lattice_names"] = [d.name for d in data]

    say("reference forward lmax=3 ...")
    fix3, _, _ = run_reference(ref, params_for(3, rmax), batch, seed=3)
    fix3["lattice_names"] = [d.name for d in data]
    return fix4, var4, fix3


def main():
    os.makedirs(GOLD, exist_ok=True)
    ref = refshim.install()
    print("sample_cat.lat through the reference Catalogue / process_one ...")
    cat = sample_cat(ref)
    print(f"  {len(cat)} (lattice, rel_dens) entries; node counts {sorted(set(int(d['positions'].shape[0]) for d in cat))}")
    torch.save(cat, os.path.join(GOLD, "ref_sample_cat.pt"))

    print("Mandel fixtures ...")
    torch.save(mandel(ref), os.path.join(GOLD, "ref_mandel.pt"))

    fix4, var4, fix3 = forward_fixtures(ref, cat)
    torch.save(fix4, os.path.join(GOLD, "ref_forward_lmax4.pt"))
    torch.save(var4, os.path.join(GOLD, "ref_variants.pt"))
    torch.save(fix3, os.path.join(GOLD, "ref_forward_lmax3.pt"))

    print("reference tables ...")
    with open(os.path.join(GOLD, "ref_tables.json"), "w") as f:
        json.dump(tables(ref), f, indent=1)
    for fn in sorted(os.listdir(GOLD)):
        print(f"  {fn}: {os.path.getsize(os.path.join(GOLD, fn)) / 1024:.0f} KiB")


if __name__ == "__main__":
    main()
