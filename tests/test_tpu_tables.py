"""The tensor-core tensor-product tables (tools/gen_kernels.py: tpu_layout / gen_tpu, consumed by csrc/eeg_tp_umma.cu):
building A = G_e^T from the fill entries and multiplying by the sender row must reproduce the channelwise tensor
product defined by tables.tp_plan (gnn/blocks.py:528-535 through gnn/mace.py:286-314)."""
import os
import re
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))
import gen_kernels as gk  # noqa: E402


def tp_reference(plan, x, Y, w):
    """m[(row), u] for one edge: x [Din, C], Y [D], w [P, C] -> [rows_total, C]"""
    n_l = {ps["l3"]: ps["n_paths"] for ps in plan["passes"]}
    local = {p: pl for ps in plan["passes"] for pl, p in enumerate(ps["path_ids"])}
    out = np.zeros((plan["rows_total"], x.shape[1]))
    for p, (l1, l2, l3) in enumerate(plan["paths"]):
        for i, j, k, v in plan["cg"][p]:
            out[plan["row_off"][l3] + k * n_l[l3] + local[p]] += w[p] * v * x[l1 * l1 + i] * Y[l2 * l2 + j]
    return out


@pytest.mark.parametrize("lmax", [3, 4])
def test_layout_reproduces_tensor_product(lmax):
    lay = gk.tpu_layout(lmax, lmax)
    plan, tiles, kt = lay["plan"], lay["tiles"], lay["kt"]
    rng = np.random.default_rng(lmax)
    D, C = (lmax + 1) ** 2, 4
    x, Y, w = rng.standard_normal((D, C)), rng.standard_normal(D), rng.standard_normal((len(plan["paths"]), C))
    A = np.zeros((len(tiles), 128, kt))
    for ti, r, kk, j0, c0, j1, c1 in lay["entries"]:
        assert A[ti, r, kk] == 0.0, "two fill entries target the same operand element"
        A[ti, r, kk] = c0 * Y[j0] + c1 * Y[j1]
    got = np.zeros((plan["rows_total"], C))
    seen = set()
    for ti, t in enumerate(tiles):
        B = np.zeros((kt, C))
        B[: t["xn"]] = x[t["xoff"]: t["xoff"] + t["xn"]]
        T = A[ti] @ B
        for r, (p, _k, orow) in enumerate(t["rows"]):
            assert orow not in seen
            seen.add(orow)
            got[orow] = w[p] * T[r]
    assert seen == set(range(plan["rows_total"]))
    np.testing.assert_allclose(got, tp_reference(plan, x, Y, w), rtol=1e-12, atol=1e-12)


@pytest.mark.parametrize("lmax", [3, 4])
def test_generated_header_matches_layout(lmax):
    """The emitted schedule is a permutation of the layout's non-zeros (+ zero padding into unused lanes)."""
    lay = gk.tpu_layout(lmax, lmax)
    txt = gk.gen_tpu(lmax, lmax)
    body = txt[txt.index("gtab["):]
    ents = re.findall(r"\{(\d+), \(int\)0x([0-9a-f]{8}), \(int\)0x([0-9a-f]{8}), 0\}", body)
    nit = int(re.search(r"_NIT (\d+)", txt).group(1))
    assert len(ents) == nit * gk.TPU_FILL_THREADS
    tile_bytes = (lay["kt"] // 4) * gk.TPU_A_LBO
    want = {}
    for ti, r, kk, j0, c0, j1, c1 in lay["entries"]:
        word = (ti * tile_bytes + (kk // 4) * gk.TPU_A_LBO + r * 16 + (kk % 4) * 4) // 4
        want[word] = (j0, np.float32(c0), j1, np.float32(c1))
    seen = set()
    for packed, b0, b1 in ents:
        packed = int(packed)
        word, j0, j1 = packed & 0xffff, (packed >> 16) & 0xff, (packed >> 24) & 0xff
        c0 = np.frombuffer(bytes.fromhex(b0)[::-1], dtype=np.float32)[0]
        c1 = np.frombuffer(bytes.fromhex(b1)[::-1], dtype=np.float32)[0]
        assert word not in seen, "duplicate store target"
        seen.add(word)
        if word in want:
            assert (j0, c0, j1, c1) == want[word]
        else:
            assert c0 == 0.0 and c1 == 0.0   # padding entry: stores 0 into an unused lane
    assert set(want) <= seen
