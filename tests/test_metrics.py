"""Validation / test metrics (scripts/train_utils.py:66-89,149-202): oracle restatement vs the reference's own
`obtain_errors` / `aggr_errors` (golden fixture from oracle/make_golden_metrics.py), device kernel vs both."""
import os

import numpy as np
import pytest
import torch

from energy_equiv_lattice_gnn_b200 import metrics
from oracle.metrics import mandel_to_cart4, per_lattice_metrics
from oracle.rotate import cart4_to_mandel

GOLD = os.path.join(os.path.dirname(__file__), "golden", "ref_metrics.pt")


def _fx():
    fx = torch.load(GOLD, weights_only=False)
    torch.manual_seed(fx["seed"])  # obtain_errors draws its directions from the global CPU generator (train_utils.py:155)
    d = torch.randn(250, 3)
    fx["directions"] = d / d.norm(dim=1, keepdim=True)
    return fx


def _rel(a, b):
    a, b = torch.as_tensor(a).double(), torch.as_tensor(b).double()
    return float((a - b).abs().max() / b.abs().max())


def test_oracle_matches_reference_obtain_errors():
    fx = _fx()
    got = per_lattice_metrics(10 * fx["pred"], 10 * fx["target"], fx["directions"])
    for k, ref in fx["columns"].items():
        assert _rel(got[k], ref) < 1e-5, k  # the reference ran in fp32 (10 * float32 tensors); the oracle in fp64


def test_mandel_cart4_round_trip():
    fx = _fx()
    t = fx["target"].double()
    assert _rel(cart4_to_mandel(mandel_to_cart4(t)), t) < 1e-15


def test_aggr_errors_matches_reference():
    fx = _fx()
    n = fx["target"].shape[0]
    err = {k: v.numpy() for k, v in fx["columns"].items()}
    err["tag"] = np.full(n, "test")
    got = metrics.aggr_errors(err)
    assert set(got) == set(fx["aggregated"])
    for k, v in fx["aggregated"].items():
        assert abs(got[k] - v) <= 1e-6 * abs(v) + 1e-12, k  # reference columns are fp32, pandas means in fp64


@pytest.mark.gpu
def test_device_metrics_match_reference_golden():
    fx = _fx()
    dev = torch.device("cuda:0")
    h = fx["half"]

    class B(dict):
        pass

    results = []
    for sl in (slice(0, h), slice(h, None)):
        b = B(stiffness=fx["target"][sl].to(dev))
        b.rel_dens = fx["rel_dens"][sl]
        results.append(({"stiffness": fx["pred"][sl].to(dev)}, b))
    err = metrics.obtain_errors(results, "test", directions=fx["directions"])
    for k in ("mean_stiffness", "loss", "mseloss", "dir_loss"):
        assert _rel(err[k], fx["columns"][k]) < 1e-5, k
    for k in ("target_eig", "predicted_eig"):  # fp32 Jacobi: absolute error relative to each matrix's largest eigenvalue
        ref = fx["columns"][k]
        assert float(((torch.as_tensor(err[k]).double() - ref).abs().amax(1) / ref.abs().amax(1)).max()) < 1e-5, k
    agg = metrics.aggr_errors(err)
    scale = float(fx["columns"]["target_eig"].abs().max())
    for k, v in fx["aggregated"].items():
        if k.startswith(("eig_loss", "rel_eig_loss")):
            continue  # |prod(eig_pred) - prod(eig_target)| of near-singular 6x6 matrices: ill-conditioned in fp32 on either side
        if k.startswith("min_"):
            assert abs(agg[k] - v) <= 1e-5 * scale, (k, agg[k], v)
        elif k.startswith("prop_eig_negative"):
            assert abs(agg[k] - v) <= 0.04, (k, agg[k], v)  # eigenvalues within rounding of zero may change sign
        else:
            assert abs(agg[k] - v) <= 1e-4 * abs(v) + 1e-12, (k, agg[k], v)
    # validation_step: MSE + directional L1 as 0-dim device tensors
    model = lambda batch: {"stiffness": fx["pred"].to(dev)}  # noqa: E731
    out = metrics.validation_step(model, B(stiffness=fx["target"].to(dev)), directions=fx["directions"])
    ref = per_lattice_metrics(fx["pred"], fx["target"], fx["directions"])
    assert abs(float(out["val_loss"]) - float(ref["mseloss"].mean())) < 1e-5 * float(ref["mseloss"].mean())
    assert abs(float(out["val_stiff_dir_loss"]) - float(ref["dir_loss"].mean())) < 1e-5 * float(ref["dir_loss"].mean())


@pytest.mark.gpu
def test_device_metrics_large_batch_vs_oracle():
    g = torch.Generator().manual_seed(2)
    G = 4096
    A = torch.randn(G, 6, 6, generator=g)
    target = A @ A.transpose(1, 2) / 6
    Bm = A + 0.1 * torch.randn(G, 6, 6, generator=g)
    pred = Bm @ Bm.transpose(1, 2) / 6
    d = metrics.random_directions(250, g)
    dev = torch.device("cuda:0")
    m, ep, et = metrics.per_lattice_metrics(pred.to(dev), target.to(dev), d)
    ref = per_lattice_metrics(pred, target, d)
    for i, k in enumerate(("mseloss", "loss", "dir_loss", "mean_stiffness")):
        assert _rel(m[:, i].cpu(), ref[k]) < 1e-5, k
    for got, k in ((ep, "predicted_eig"), (et, "target_eig")):
        r = ref[k]
        assert float(((got.cpu().double() - r).abs().amax(1) / r.abs().amax(1)).max()) < 1e-5, k
        assert bool((got[:, 1:] >= got[:, :-1]).all())  # ascending, like eigvalsh
    # trace = sum of eigenvalues (size-independent property)
    tr = torch.diagonal(target, dim1=1, dim2=2).sum(1)
    assert _rel(et.cpu().sum(1), tr) < 1e-5
