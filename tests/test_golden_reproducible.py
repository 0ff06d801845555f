"""Build-container-only: re-run the UNMODIFIED reference (through oracle/refshim.py) in memory and compare with every
committed forward fixture under tests/golden/ ("no test re-runs make_golden in memory").
Skipped where /root/reference does not exist (the GPU box)."""
import os
import sys

import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLD = os.path.join(ROOT, "tests", "golden")
pytestmark = pytest.mark.skipif(not os.path.exists("/root/reference/gnn/model.py"), reason="reference checkout not present")


def _same(a, b, path=""):
    if torch.is_tensor(a):
        assert torch.is_tensor(b) and a.shape == b.shape and a.dtype == b.dtype, path
        if a.is_floating_point():
            assert float((a.double() - b.double()).abs().max()) <= 1e-12 * max(float(b.double().abs().max()), 1e-300), path
        else:
            assert torch.equal(a, b), path
    elif isinstance(a, dict):
        assert set(a) == set(b), (path, set(a) ^ set(b))
        for k in a:
            _same(a[k], b[k], f"{path}/{k}")
    elif isinstance(a, (list, tuple)):
        assert len(a) == len(b), path
        for i, (x, y) in enumerate(zip(a, b)):
            _same(x, y, f"{path}[{i}]")
    else:
        assert a == b, (path, a, b)


def test_committed_forward_fixtures_are_reproducible():
    sys.path.insert(0, ROOT)
    from oracle import make_golden, refshim
    ref = refshim.install()
    cat = make_golden.sample_cat(ref)
    _same(cat, torch.load(os.path.join(GOLD, "ref_sample_cat.pt"), weights_only=False), "ref_sample_cat")
    fix4, var4, fix3 = make_golden.forward_fixtures(ref, cat, verbose=False)
    for got, name in ((fix4, "ref_forward_lmax4.pt"), (var4, "ref_variants.pt"), (fix3, "ref_forward_lmax3.pt")):
        _same(got, torch.load(os.path.join(GOLD, name), weights_only=False), name)
