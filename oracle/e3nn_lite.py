"""TEST INFRASTRUCTURE - CPU restatement of the e3nn 0.5.x / torch_scatter semantics the reference path uses.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this package; the
product (energy_equiv_lattice_gnn_b200/) never does.

PARITY UNPINNED: e3nn (0.5.1), torch_scatter (2.0.9), opt_einsum (3.3.0) are third-party dependencies of the reference
that are neither vendored under /root/reference nor installable here; the reference has no
tests or golden vectors for the GNN.  What pins this file instead: the numeric anchors of SURVEY.md B.7 (w3j, SH,
Wigner-D, U-matrix norms and widths), the reference's own Mandel test convention, and the O(3) equivariance /
invariance properties checked in tests/test_oracle_properties.py.

The base constants (real w3j, SH recursion coefficients) come from energy_equiv_lattice_gnn_b200.conventions - they are
e3nn constants, not reference code; everything else here follows the call sites cited per function.
"""
from __future__ import annotations

import math
from typing import List, Sequence, Tuple

import numpy as np
import torch

from energy_equiv_lattice_gnn_b200 import conventions as cv

Irreps = List[Tuple[int, int, int]]  # (mul, l, parity)

NORM2MOM_SILU = 1.6791767923989418  # e3nn normalize2mom(silu) Monte-Carlo constant


def soft_one_hot_linspace_gaussian(x: torch.Tensor, start: float, end: float, number: int) -> torch.Tensor:
    """e3nn.math.soft_one_hot_linspace(basis='gaussian', cutoff=False); call sites gnn/model.py:146-152."""
    values = torch.linspace(start, end, number, dtype=x.dtype, device=x.device)
    step = values[1] - values[0]
    diff = (x[..., None] - values) / step
    return diff.pow(2).neg().exp().div(1.12)


def spherical_harmonics(lmax: int, vec: torch.Tensor) -> torch.Tensor:
    """o3.SphericalHarmonics(normalize=True, normalization='component'); call site gnn/model.py:125-129,157."""
    v = torch.nn.functional.normalize(vec, dim=-1)
    hats = [torch.ones_like(v[..., :1])]
    if lmax >= 1:
        hats.append(v)
    for l in range(2, lmax + 1):
        w, sc = cv.sh_recursion(l)
        wt = torch.as_tensor(w * sc, dtype=v.dtype, device=v.device)
        hats.append(torch.einsum("ijk,...i,...j->...k", wt, hats[1], hats[l - 1]))
    return torch.cat([math.sqrt(2 * l + 1) * h for l, h in enumerate(hats)], dim=-1)


def scatter(src: torch.Tensor, index: torch.Tensor, dim_size: int, reduce: str) -> torch.Tensor:
    """torch_scatter.scatter(dim=0, reduce='sum'|'mean'); call sites gnn/blocks.py:595-597, gnn/model.py:100-106."""
    out = torch.zeros((dim_size,) + src.shape[1:], dtype=src.dtype, device=src.device)
    out.index_add_(0, index, src)
    if reduce == "mean":
        cnt = torch.zeros(dim_size, dtype=src.dtype, device=src.device)
        cnt.index_add_(0, index, torch.ones_like(index, dtype=src.dtype))
        out = out / cnt.clamp(min=1).view(-1, *([1] * (src.dim() - 1)))
    elif reduce != "sum":
        raise ValueError(reduce)
    return out


class Linear(torch.nn.Module):
    """o3.Linear(irreps_in, irreps_out, biases=...): one [mul_in, mul_out] block per matching irrep
    pair, flat `weight` in (i_in, i_out) order, 1/sqrt(fan_in) path normalisation, bias on 0e outputs only."""

    def __init__(self, irreps_in: Irreps, irreps_out: Irreps, biases: bool = False):
        super().__init__()
        self.irreps_in, self.irreps_out = list(irreps_in), list(irreps_out)
        self.paths = []  # (i_in, i_out, w_off)
        off = 0
        for i, (mi, li, pi) in enumerate(self.irreps_in):
            for o, (mo, lo, po) in enumerate(self.irreps_out):
                if (li, pi) == (lo, po):
                    self.paths.append((i, o, off))
                    off += mi * mo
        self.weight = torch.nn.Parameter(torch.randn(off))
        nb = sum(mo for mo, lo, po in self.irreps_out if lo == 0 and po == 1) if biases else 0
        self.bias = torch.nn.Parameter(torch.zeros(nb)) if nb else None
        self.fan_in = [0] * len(self.irreps_out)
        for i, o, _ in self.paths:
            self.fan_in[o] += self.irreps_in[i][0]

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        in_off = np.cumsum([0] + [m * (2 * l + 1) for m, l, _ in self.irreps_in])
        outs = [None] * len(self.irreps_out)
        for i, o, woff in self.paths:
            mi, li, _ = self.irreps_in[i]
            mo = self.irreps_out[o][0]
            xi = x[:, in_off[i]:in_off[i + 1]].reshape(-1, mi, 2 * li + 1)
            w = self.weight[woff:woff + mi * mo].reshape(mi, mo)
            y = torch.einsum("zui,uw->zwi", xi, w) / math.sqrt(self.fan_in[o])
            outs[o] = y if outs[o] is None else outs[o] + y
        res, boff = [], 0
        for o, (mo, lo, po) in enumerate(self.irreps_out):
            y = outs[o]
            if y is None:
                y = x.new_zeros(x.shape[0], mo, 2 * lo + 1)
            if self.bias is not None and lo == 0 and po == 1:
                y = y + self.bias[boff:boff + mo].view(1, mo, 1)
                boff += mo
            res.append(y.reshape(x.shape[0], -1))
        return torch.cat(res, dim=-1)


class TensorProductUVU(torch.nn.Module):
    """o3.TensorProduct(irreps1, sh_irreps, irreps_mid, 'uvu' instructions, shared_weights=False,
    internal_weights=False) as built in gnn/blocks.py:523-535 from gnn/mace.py:286-314.  Dense w3j einsum per path,
    path coefficient sqrt(2 l3 + 1) (component / element normalisation with one path per output slot)."""

    def __init__(self, irreps1: Irreps, lmax: int, target: Irreps):
        super().__init__()
        self.irreps1 = list(irreps1)
        self.lmax = lmax
        self.paths = cv.tp_paths(irreps1, lmax, target)
        self.weight_numel = sum(self.irreps1[i1][0] for i1, *_ in self.paths)
        self.irreps_mid = [(self.irreps1[i1][0], l3, p3) for i1, l1, l2, l3, p3 in self.paths]

    def forward(self, x1: torch.Tensor, sh: torch.Tensor, weight: torch.Tensor) -> torch.Tensor:
        in_off = np.cumsum([0] + [m * (2 * l + 1) for m, l, _ in self.irreps1])
        outs, woff = [], 0
        for i1, l1, l2, l3, _p3 in self.paths:
            mul = self.irreps1[i1][0]
            a = x1[:, in_off[i1]:in_off[i1 + 1]].reshape(-1, mul, 2 * l1 + 1)
            y = sh[:, l2 * l2:(l2 + 1) * (l2 + 1)]
            w = weight[:, woff:woff + mul]
            woff += mul
            c = torch.tensor(np.array(cv.wigner_3j(l1, l2, l3)), dtype=x1.dtype, device=x1.device)
            o = torch.einsum("zu,ijk,zui,zj->zuk", w, c, a, y) * math.sqrt(2 * l3 + 1)
            outs.append(o.reshape(x1.shape[0], -1))
        return torch.cat(outs, dim=-1)


def simplify(irreps: Irreps) -> Irreps:
    out: Irreps = []
    for m, l, p in irreps:
        if out and out[-1][1:] == (l, p):
            out[-1] = (out[-1][0] + m, l, p)
        else:
            out.append((m, l, p))
    return out


class Gate(torch.nn.Module):
    """e3nn.nn.Gate with normalize2mom(silu) on scalars and gates (gnn/blocks.py:268-274)."""

    def __init__(self, n_scalars: int, gated: Irreps):
        super().__init__()
        self.n_scalars, self.gated = n_scalars, list(gated)
        self.n_gates = sum(m for m, _, _ in gated)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        c = NORM2MOM_SILU
        s = c * torch.nn.functional.silu(x[:, :self.n_scalars])
        g = c * torch.nn.functional.silu(x[:, self.n_scalars:self.n_scalars + self.n_gates])
        off, goff, outs = self.n_scalars + self.n_gates, 0, [s]
        for m, l, _ in self.gated:
            d = 2 * l + 1
            t = x[:, off:off + m * d].reshape(-1, m, d) * g[:, goff:goff + m, None]
            outs.append(t.reshape(x.shape[0], -1))
            off += m * d
            goff += m
        return torch.cat(outs, dim=-1)
