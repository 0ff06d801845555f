"""Pins the oracle (oracle/model.py) against vectors produced by RUNNING THE REFERENCE'S OWN SOURCES
(oracle/make_golden.py -> tests/golden/*; gnn/model.py, gnn/blocks.py, gnn/mace.py, gnn/datasets.py executed from
/root/reference through oracle/refshim.py).  CPU only."""
import json
import os
from argparse import Namespace

import numpy as np
import pytest
import torch

from energy_equiv_lattice_gnn_b200 import conventions as cv
from energy_equiv_lattice_gnn_b200 import tables as tb
from energy_equiv_lattice_gnn_b200.data import Batch
from oracle.model import Cart_4_to_Mandel, EnergyEquivGNN

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load(name):
    return torch.load(os.path.join(GOLD, name), weights_only=False)


def oracle_from_fixture(fx, dtype, **over):
    old = torch.get_default_dtype()
    torch.set_default_dtype(dtype)
    try:
        m = EnergyEquivGNN(Namespace(**{**fx["params"], **over}))
        sd = {k: v.to(dtype) if v.is_floating_point() else v for k, v in fx["state_dict"].items()}
        missing, unexpected = m.load_state_dict(sd, strict=False)
        assert not unexpected and all("U_matrix" in k for k in missing), (missing, unexpected)
    finally:
        torch.set_default_dtype(old)
    b = Batch(**{k: (v.to(dtype) if torch.is_tensor(v) and v.is_floating_point() else v) for k, v in fx["batch"].items()})
    return m, b


def rel(a, b):
    return float((a.double() - b.double()).abs().max() / b.double().abs().max())


@pytest.mark.parametrize("name", ["lmax4", "lmax3"])
def test_oracle_fp64_matches_reference_forward(name):
    fx = load(f"ref_forward_{name}.pt")
    m, b = oracle_from_fixture(fx, torch.float64)
    inter = {}
    head = m.stiffness_head
    for i, layer in enumerate(head.layers):
        layer.register_forward_hook(lambda mod, a, o, i=i: inter.__setitem__(f"layer{i}_out", o))
        layer.interaction.register_forward_hook(lambda mod, a, o, i=i: inter.__setitem__(f"layer{i}_message", o))
    head.nonlin_readout.register_forward_hook(lambda mod, a, o: inter.__setitem__("readout", o))
    head.linear.register_forward_hook(lambda mod, a, o: inter.__setitem__("q", o))
    head.cart_to_Mandel.register_forward_hook(lambda mod, a, o: inter.__setitem__("A", o))
    with torch.no_grad():
        out = m(b)["stiffness"]
        _, sh = m.edge_features(b)
    assert rel(out, fx["stiffness"]) < 1e-12
    assert rel(sh, fx["inter"]["edge_sh"]) < 1e-12
    for k, v in inter.items():
        assert rel(v, fx["inter"][k]) < 1e-12, k


def test_oracle_fp32_close_to_reference_fp64():
    fx = load("ref_forward_lmax4.pt")
    m, b = oracle_from_fixture(fx, torch.float32)
    with torch.no_grad():
        out = m(b)["stiffness"]
    assert rel(out, fx["stiffness"]) < 1e-5


def test_oracle_variants_match_reference():
    fx = load("ref_forward_lmax4.pt")
    var = load("ref_variants.pt")
    assert len(var) >= 6
    for key, want in var.items():
        m, b = oracle_from_fixture(fx, torch.float64, **json.loads(key))
        with torch.no_grad():
            out = m(b)["stiffness"]
        assert rel(out, want) < 1e-12, key


def test_oracle_interaction_mean_matches_reference():
    """interaction_reduction='mean' (gnn/blocks.py:595-597), fixture from oracle/make_golden_mean.py."""
    fx = load("ref_forward_lmax4.pt")
    want = load("ref_interaction_mean.pt")["stiffness"]
    m, b = oracle_from_fixture(fx, torch.float64, interaction_reduction="mean")
    with torch.no_grad():
        out = m(b)["stiffness"]
    assert rel(out, want) < 1e-12


def test_reference_tables_match_product_tables():
    with open(os.path.join(GOLD, "ref_tables.json")) as f:
        t = json.load(f)
    for lmax in (3, 4):
        for name, lin in (("layer0", 0), ("layer1", lmax)):
            ref = t["tp_instructions"][f"lmax{lmax}_{name}"]
            plan = tb.tp_plan(lin, lmax, lmax)
            assert [list(p) for p in plan["paths"]] == ref["l_triples"]
            # instruction k writes output slot k after the reference's sort => weight block order == path order
            assert [c for _, _, c in ref["instructions"]] == list(range(len(ref["instructions"])))
        for nu in (1, 2, 3):
            for L in range(lmax + 1):
                r = t["U"][f"lmax{lmax}_nu{nu}_L{L}"]
                idx, val, k = cv.u_matrix_sparse(lmax, nu, L)
                assert r["shape"][-1] == k
                assert r["nnz"] == int((np.abs(val) > 1e-12).sum())
                assert abs(r["fro"] - float(np.sqrt((val ** 2).sum()))) < 1e-9
                assert abs(r["sum"] - float(val.sum())) < 1e-9
                assert abs(r["abssum"] - float(np.abs(val).sum())) < 1e-9
                assert r["max_abs_diff_vs_conventions"] < 1e-12


def test_mandel_convention_matches_reference_elasticity_func():
    fx = load("ref_mandel.pt")
    abcd, rows, cols, mask = cv.mandel_tables()
    assert abcd[:, 0].tolist() == fx["a"] and abcd[:, 1].tolist() == fx["b"]
    assert abcd[:, 2].tolist() == fx["c"] and abcd[:, 3].tolist() == fx["d"]
    assert torch.equal(torch.tensor(rows), fx["rows"]) and torch.equal(torch.tensor(cols), fx["cols"])
    old = torch.get_default_dtype()
    torch.set_default_dtype(torch.float64)
    try:
        got = Cart_4_to_Mandel()(fx["cart4"])
    finally:
        torch.set_default_dtype(old)
    assert rel(got, fx["mandel"]) < 1e-12


def test_sample_cat_fixture_schema():
    cat = load("ref_sample_cat.pt")
    assert len(cat) == 57 and len({d["name"] for d in cat}) == 19
    for d in cat:
        n, e = d["positions"].shape[0], d["edge_index"].shape[1]
        assert d["edge_index"].dtype == torch.int64 and d["positions"].dtype == torch.float32
        assert d["shifts"].shape == (e, 3) and d["edge_attr"].shape == (e, 1) and d["node_attrs"].shape == (n, 1)
        assert e % 2 == 0
        h = e // 2  # second half = reversed struts with negated shifts (gnn/datasets.py:163-171)
        assert torch.equal(d["edge_index"][:, h:], d["edge_index"][:, :h].flip(0))
        assert torch.equal(d["shifts"][h:], -d["shifts"][:h])
