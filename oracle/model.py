"""TEST INFRASTRUCTURE - op-for-op CPU restatement of the reference hot path in plain torch:
gnn/model.py (EnergyEquivGNN, GNN_Head), gnn/blocks.py:185-229,250-283,392-442,447-604,902-947 and
gnn/mace.py:112-277,316-352.  Same module tree and state_dict names as the reference, same
formulation (dense w3j einsum tensor product, U.W.x symmetric contraction, index_add scatter), so that it doubles as the
timed CPU baseline.  PARITY UNPINNED (see oracle/e3nn_lite.py header).
"""
from __future__ import annotations

import math
import os
from argparse import Namespace
from typing import Dict

import numpy as np
import torch

from energy_equiv_lattice_gnn_b200 import conventions as cv
from . import e3nn_lite as e3
from .mace_cg import U_matrix_real

_U_CACHE: Dict = {}


def _U(lmax: int, L: int, nu: int) -> torch.Tensor:
    key = (lmax, L, nu)
    if key not in _U_CACHE:
        _U_CACHE[key] = U_matrix_real(lmax, L, nu, dtype=torch.float64)
    return _U_CACHE[key]


class Contraction(torch.nn.Module):
    """gnn/mace.py:180-280, element_dependent=False branch."""

    def __init__(self, lmax: int, L: int, correlation: int, num_features: int):
        super().__init__()
        self.correlation, self.L = correlation, L
        for nu in range(1, correlation + 1):
            self.register_buffer(f"U_matrix_{nu}", _U(lmax, L, nu).to(torch.get_default_dtype()))
        self.weights = torch.nn.ParameterDict()
        for i in range(1, correlation + 1):
            k = self._buffers[f"U_matrix_{i}"].shape[-1]
            self.weights[str(i)] = torch.nn.Parameter(torch.randn(k, num_features) / k)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        U = lambda nu: self._buffers[f"U_matrix_{nu}"]  # noqa: E731
        # "...ik,kc,bci -> bc..." (mace.py:262-267); contraction order U.W first, then x
        uw = torch.einsum("...ik,kc->...ic", U(self.correlation), self.weights[str(self.correlation)])
        out = torch.einsum("...ic,bci->bc...", uw, x)
        for corr in range(self.correlation - 1, 0, -1):
            c_tensor = torch.einsum("...k,kc->c...", U(corr), self.weights[str(corr)])  # mace.py:269-273
            c_tensor = c_tensor + out
            out = torch.einsum("bc...i,bci->bc...", c_tensor, x)  # mace.py:275
        return out.reshape(out.shape[0], -1)


class SymmetricContraction(torch.nn.Module):
    """gnn/mace.py:112-177."""

    def __init__(self, lmax: int, irreps_out, correlation: int, num_features: int):
        super().__init__()
        self.keys = [f"{m}x{l}{'e' if p == 1 else 'o'}" for m, l, p in irreps_out]
        self.contractions = torch.nn.ModuleDict(
            {k: Contraction(lmax, l, correlation, num_features) for k, (m, l, p) in zip(self.keys, irreps_out)})

    def forward(self, x):
        return torch.cat([self.contractions[k](x) for k in self.keys], dim=-1)


def reshape_irreps(x: torch.Tensor, irreps) -> torch.Tensor:
    """gnn/mace.py:316-332."""
    ix, out = 0, []
    for mul, l, _ in irreps:
        d = 2 * l + 1
        out.append(x[:, ix:ix + mul * d].reshape(x.shape[0], mul, d))
        ix += mul * d
    return torch.cat(out, dim=-1)


class EquivariantProductBlock(torch.nn.Module):
    """gnn/blocks.py:447-490 (use_sc=False)."""

    def __init__(self, lmax, node_feats_irreps, target_irreps, correlation):
        super().__init__()
        self.in_irreps = node_feats_irreps
        mul = sum(m for m, l, p in node_feats_irreps if (l, p) == (0, 1))
        sc_out = [(mul, l, p) for _, l, p in target_irreps]
        self.symmetric_contractions = SymmetricContraction(lmax, sc_out, correlation, mul)
        self.linear = e3.Linear(sc_out, target_irreps)

    def forward(self, node_feats):
        return self.linear(self.symmetric_contractions(reshape_irreps(node_feats, self.in_irreps)))


class TensorProductInteractionBlock(torch.nn.Module):
    """gnn/blocks.py:495-604 (reduce='sum'|'mean')."""

    def __init__(self, node_feats_irreps, lmax, n_edge_feats, irreps_out, agg_norm_const, reduce, bias, MLP_dim, MLP_layers):
        super().__init__()
        self.agg_norm_const, self.reduce = agg_norm_const, reduce.lower()
        self.linear_up = e3.Linear(node_feats_irreps, node_feats_irreps)
        self.conv_tp = e3.TensorProductUVU(node_feats_irreps, lmax, irreps_out)
        layer = torch.nn.Linear(MLP_dim, self.conv_tp.weight_numel, bias=False)
        torch.nn.init.xavier_uniform_(layer.weight, gain=10)
        mods = [torch.nn.Linear(n_edge_feats, MLP_dim), torch.nn.SiLU()]
        for _ in range(MLP_layers - 2):
            mods += [torch.nn.Linear(MLP_dim, MLP_dim), torch.nn.SiLU()]
        mods.append(layer)
        self.conv_tp_weights = torch.nn.Sequential(*mods)
        self.linear = e3.Linear(e3.simplify(self.conv_tp.irreps_mid), irreps_out, biases=bias)

    def forward(self, node_feats, edge_attrs, edge_feats, edge_index):
        sender, receiver = edge_index
        node_feats = self.linear_up(node_feats)
        tp_weights = self.conv_tp_weights(edge_feats)
        mji = self.conv_tp(node_feats[sender], edge_attrs, tp_weights)
        message = e3.scatter(mji, receiver, node_feats.shape[0], self.reduce) / self.agg_norm_const
        return self.linear(message)


class MACELayer(torch.nn.Module):
    """gnn/blocks.py:902-947."""

    def __init__(self, input_irreps, lmax, n_edge_feats, interaction_irreps, output_irreps, agg, red, bias, corr, MLP_dim, MLP_layers):
        super().__init__()
        self.interaction = TensorProductInteractionBlock(input_irreps, lmax, n_edge_feats, interaction_irreps, agg, red, bias, MLP_dim, MLP_layers)
        self.product = EquivariantProductBlock(lmax, interaction_irreps, output_irreps, corr)

    def forward(self, node_ft, edge_index, edge_sh, edge_scalars):
        return self.product(self.interaction(node_ft, edge_sh, edge_scalars, edge_index))


class GeneralNonLinearReadoutBlock(torch.nn.Module):
    """gnn/blocks.py:250-283."""

    def __init__(self, irreps_in, hidden_irreps, irreps_out):
        super().__init__()
        outset = {(l, p) for _, l, p in irreps_out}
        scalars = [(m, l, p) for m, l, p in hidden_irreps if l == 0 and (l, p) in outset]
        gated = [(m, l, p) for m, l, p in hidden_irreps if l > 0 and (l, p) in outset]
        gates = [(m, 0, 1) for m, _, _ in gated]
        n_s = sum(m for m, _, _ in scalars)
        self.gate = e3.Gate(n_s, gated)
        self.linear_1 = e3.Linear(irreps_in, e3.simplify(scalars + gates + gated))
        self.linear_2 = e3.Linear(scalars + gated, irreps_out)

    def forward(self, x):
        return self.linear_2(self.gate(self.linear_1(x)))


class Cart_4_to_Mandel(torch.nn.Module):
    """gnn/blocks.py:392-425."""

    def __init__(self):
        super().__init__()
        abcd, rows, cols, mask = cv.mandel_tables()
        self.abcd = abcd
        self.register_buffer("mask", torch.tensor(mask, dtype=torch.get_default_dtype()))
        self.register_buffer("rows", torch.tensor(rows))
        self.register_buffer("cols", torch.tensor(cols))

    def forward(self, C):
        a, b, c, d = (self.abcd[:, i] for i in range(4))
        C2 = C.new_zeros((C.shape[0], 6, 6))
        C2[:, self.rows, self.cols] = C[:, a, b, c, d]
        C2[:, self.cols, self.rows] = C[:, a, b, c, d]
        return C2 * self.mask.view(1, 6, 6)


class Spherical_to_Cartesian(torch.nn.Module):
    """gnn/blocks.py:427-442; Q_flat from our documented basis."""

    def __init__(self):
        super().__init__()
        self.register_buffer("Q_flat", torch.tensor(cv.rank4_symmetric_basis().reshape(21, 81), dtype=torch.get_default_dtype()))

    def forward(self, x):
        return (x @ self.Q_flat).view(*x.shape[:-1], 3, 3, 3, 3)


def positive_function(name: str):
    """gnn/blocks.py:185-229."""
    mp = torch.linalg.matrix_power
    eye = lambda C: torch.eye(6, device=C.device, dtype=C.dtype)  # noqa: E731
    table = {
        "matrix_power_2": lambda C: mp(C, 2), "matrix_power_4": lambda C: mp(C, 4),
        "matrix_exp": torch.linalg.matrix_exp,
        "matrix_trunc_exp_2": lambda C: mp(eye(C) + C / 2, 2), "matrix_trunc_exp_4": lambda C: mp(eye(C) + C / 4, 4),
        "none": lambda C: C,
    }
    if name not in table:
        raise ValueError(f"Unknown positive function: {name}")
    return table[name]


class GNN_Head(torch.nn.Module):
    """gnn/model.py:26-112."""

    def __init__(self, params: Namespace):
        super().__init__()
        hidden = cv.parse_irreps(params.hidden_irreps)
        nfeat = sum(m for m, l, p in hidden if (l, p) == (0, 1))
        node_ft_irreps = [(nfeat, 0, 1)]
        lmax = params.lmax
        interaction_irreps = [(nfeat, l, (-1) ** l) for l in range(lmax + 1)]
        readout = cv.parse_irreps(params.readout_irreps)
        nb = 2 * params.num_edge_bases
        args = (params.agg_norm_const, params.interaction_reduction, True, params.correlation, params.inter_MLP_dim, params.inter_MLP_layers)
        self.layers = torch.nn.ModuleList([MACELayer(node_ft_irreps, lmax, nb, interaction_irreps, hidden, *args)])
        for _ in range(params.message_passes - 1):
            self.layers.append(MACELayer(hidden, lmax, nb, interaction_irreps, hidden, *args))
        self.nonlin_readout = GeneralNonLinearReadoutBlock(hidden, hidden, readout)
        self.global_reduction = params.global_reduction
        self.linear = e3.Linear(readout, [(2, 0, 1), (2, 2, 1), (1, 4, 1)], biases=True)
        self.sph_to_cart = Spherical_to_Cartesian()
        self.cart_to_Mandel = Cart_4_to_Mandel()
        self.positive = positive_function(params.positive_function)

    def forward(self, edge_index, node_ft, edge_sh, edge_feats, batch_idx, num_graphs):
        node_ft = self.layers[0](node_ft, edge_index, edge_sh, edge_feats)
        for i in range(1, len(self.layers)):
            node_ft = node_ft + self.layers[i](node_ft, edge_index, edge_sh, edge_feats)
        output = self.nonlin_readout(node_ft)
        graph_ft = e3.scatter(output, batch_idx, num_graphs, self.global_reduction)
        stiff = self.sph_to_cart(self.linear(graph_ft))
        return self.positive(self.cart_to_Mandel(stiff))


class EnergyEquivGNN(torch.nn.Module):
    """gnn/model.py:114-161."""

    def __init__(self, params: Namespace):
        super().__init__()
        self.params = params
        hidden = cv.parse_irreps(params.hidden_irreps)
        nfeat = sum(m for m, l, p in hidden if (l, p) == (0, 1))
        self.node_ft_embedding = torch.nn.Linear(1, nfeat)
        self.number_of_edge_basis = params.num_edge_bases
        self.max_edge_radius = params.max_edge_radius
        self.lmax = params.lmax
        self.stiffness_head = GNN_Head(params)

    def edge_features(self, batch):
        sender, receiver = batch.edge_index
        vectors = batch.positions[receiver] - batch.positions[sender] + batch.shifts  # mace.py:345-346
        lengths = torch.linalg.norm(vectors, dim=-1, keepdim=True)
        el = e3.soft_one_hot_linspace_gaussian(lengths.squeeze(-1), 0.0, 0.6, self.number_of_edge_basis)
        er = e3.soft_one_hot_linspace_gaussian(batch.edge_attr.squeeze(-1), 0.0, self.max_edge_radius, self.number_of_edge_basis)
        return torch.cat((el, er), dim=1), e3.spherical_harmonics(self.lmax, vectors)

    def forward(self, batch):
        node_ft = self.node_ft_embedding(batch.node_attrs)
        edge_feats, edge_sh = self.edge_features(batch)
        C = self.stiffness_head(batch.edge_index, node_ft, edge_sh, edge_feats, batch.batch, batch.num_graphs)
        return {"stiffness": C}


def default_params(**over) -> Namespace:
    """scripts/train_main.py:25-52 (network part) + max_edge_radius (train_main.py:64-65)."""
    p = dict(lmax=4, hidden_irreps="32x0e+32x1o+32x2e+32x3o+32x4e", readout_irreps="16x0e+16x1o+16x2e+16x3o+16x4e",
             num_edge_bases=6, interaction_reduction="sum", interaction_bias=True, agg_norm_const=4.0, inter_MLP_dim=64,
             inter_MLP_layers=3, correlation=3, global_reduction="mean", message_passes=2,
             positive_function="matrix_power_2", max_edge_radius=0.03,
             lr=1e-3, amsgrad=True, weight_decay=1e-8, beta1=0.9, epsilon=1e-8)
    p.update(over)
    return Namespace(**p)
