"""EnergyEquivGNN - drop-in for gnn/model.py:114-161 of the reference, running on libeeg_b200 (sm_100a kernels).

Same constructor (`EnergyEquivGNN(params: Namespace)`), same `forward(batch) -> {'stiffness': [G,6,6]}`, same
parameter / buffer names as the reference module tree so a reference `state_dict` loads with
`strict=False` (the dense `U_matrix_*` buffers are not kept: the kernels use the sparse symmetrised tables built by
`tables.sc_weight_map`).  There is no CPU path: a non-CUDA batch raises.
"""
from __future__ import annotations

import math
import os
from argparse import Namespace
from typing import Dict, List

import numpy as np
import torch

from . import conventions as cv
from . import native as nv
from . import tables as tb

NORM2MOM_SILU = 1.6791767923989418  # e3nn normalize2mom(silu), SURVEY.md B.5

_POSITIVE = {"none": 0, "matrix_power_2": 1, "matrix_power_4": 2, "matrix_trunc_exp_2": 3, "matrix_trunc_exp_4": 4,
             "matrix_exp": 5}   # gnn/blocks.py:186-208


class _Holder(torch.nn.Module):
    """Plain container so parameter paths match the reference module tree."""


def _o3_linear_param(numel: int) -> torch.nn.Parameter:
    return torch.nn.Parameter(torch.randn(numel))  # e3nn o3.Linear init: N(0,1), normalisation applied in forward


class EnergyEquivGNN(torch.nn.Module):
    def __init__(self, params: Namespace, *args, **kwargs):
        super().__init__()
        self.params = params
        hidden = cv.parse_irreps(params.hidden_irreps)
        readout = cv.parse_irreps(params.readout_irreps)
        self.lmax = lmax = int(params.lmax)
        self.C = C = sum(m for m, l, p in hidden if (l, p) == (0, 1))
        want = [(C, l, (-1) ** l) for l in range(lmax + 1)]
        if hidden != want:
            raise NotImplementedError(
                f"hidden_irreps must be {cv.irreps_str(want)} (uniform multiplicity, l = 0..lmax, parity (-1)^l); got {params.hidden_irreps}")
        if [(l, p) for _, l, p in readout] != [(l, (-1) ** l) for l in range(lmax + 1)]:
            raise NotImplementedError("readout_irreps must have one entry per l = 0..lmax with parity (-1)^l")
        if C % 32 != 0:
            raise NotImplementedError("channel count must be a multiple of 32 (one warp lane per channel)")
        if params.interaction_reduction.lower() not in ("sum", "mean"):
            raise NotImplementedError("interaction_reduction: 'sum' (reference default, train_main.py:31) and 'mean' are built; 'pna' is not")
        self.interaction_reduction = params.interaction_reduction.lower()
        if params.positive_function not in _POSITIVE:
            raise ValueError(f"Unknown positive function: {params.positive_function}")
        if params.global_reduction not in ("mean", "sum"):
            raise ValueError(f"Unknown global reduction: {params.global_reduction}")
        if not 1 <= lmax <= 4:
            raise NotImplementedError(f"lmax must be in 1..4 (generated kernels, tools/gen_kernels.py); got {lmax}")
        if not 1 <= int(params.correlation) <= 3:
            raise NotImplementedError(f"correlation must be in 1..3 (generated kernels); got {params.correlation}")
        if int(params.inter_MLP_layers) < 1 or int(params.inter_MLP_dim) % 4 != 0 or int(params.num_edge_bases) < 2 or \
                int(params.num_edge_bases) % 2 != 0:
            raise NotImplementedError("radial MLP: inter_MLP_layers >= 1, inter_MLP_dim a multiple of 4, num_edge_bases even and >= 2")
        self.correlation = int(params.correlation)
        self.number_of_edge_basis = nb = int(params.num_edge_bases)
        self.max_edge_radius = float(params.max_edge_radius)
        self.num_interactions = int(params.message_passes)
        self.agg_norm_const = float(params.agg_norm_const)
        self.rmul = [m for m, _, _ in readout]
        H, nl = int(params.inter_MLP_dim), int(params.inter_MLP_layers)
        D = (lmax + 1) ** 2

        # ---- parameters, reference names ---------------------------------------------------------------------------
        self.node_ft_embedding = torch.nn.Linear(1, C)
        head = self.stiffness_head = _Holder()
        head.layers = torch.nn.ModuleList()
        self._tp = []
        for i in range(self.num_interactions):
            lin_in = 0 if i == 0 else lmax
            plan = tb.tp_plan(lin_in, lmax, lmax)
            self._tp.append(plan)
            n_l = [ps["n_paths"] for ps in plan["passes"]]
            layer = _Holder()
            inter = layer.interaction = _Holder()
            inter.linear_up = _Holder()
            inter.linear_up.weight = _o3_linear_param((lin_in + 1) * C * C)
            wnum = len(plan["paths"]) * C
            last = torch.nn.Linear(H, wnum, bias=False)
            torch.nn.init.xavier_uniform_(last.weight, gain=10)
            mods = [torch.nn.Linear(2 * nb, H), torch.nn.SiLU()]   # gnn/blocks.py:541-548: inter_MLP_layers <= 2 gives two linears
            for _ in range(nl - 2):
                mods += [torch.nn.Linear(H, H), torch.nn.SiLU()]
            mods.append(last)
            inter.conv_tp_weights = torch.nn.Sequential(*mods)
            inter.linear = _Holder()
            inter.linear.weight = _o3_linear_param(sum(n * C * C for n in n_l))
            inter.linear.bias = torch.nn.Parameter(torch.zeros(C))
            prod = layer.product = _Holder()
            sc = prod.symmetric_contractions = _Holder()
            sc.contractions = torch.nn.ModuleDict()
            wmap = tb.sc_weight_map(lmax, self.correlation)
            plan_sc = tb.sc_plan(lmax, self.correlation)
            for L in range(lmax + 1):
                con = _Holder()
                con.weights = torch.nn.ParameterDict()
                for nu in range(1, self.correlation + 1):
                    k = plan_sc["wshape"][(L, nu)]
                    con.weights[str(nu)] = torch.nn.Parameter(torch.randn(k, C) / k)  # mace.py:235-237
                sc.contractions[f"{C}x{L}{'e' if L % 2 == 0 else 'o'}"] = con
            prod.linear = _Holder()
            prod.linear.weight = _o3_linear_param((lmax + 1) * C * C)
            head.layers.append(layer)
        self._wmap = wmap
        ro = head.nonlin_readout = _Holder()
        ro.linear_1 = _Holder()
        n_gates = C * lmax
        ro.linear_1.weight = _o3_linear_param(C * (C + n_gates) + lmax * C * C)
        ro.linear_2 = _Holder()
        ro.linear_2.weight = _o3_linear_param(sum(C * m for m in self.rmul))
        # final o3.Linear readout -> 2x0e + 2x2e + 1x4e (gnn/model.py:81-86)
        self._qmul = [{0: 2, 2: 2, 4: 1}.get(l, 0) for l in range(lmax + 1)]
        head.linear = _Holder()
        head.linear.weight = _o3_linear_param(sum(self.rmul[l] * q for l, q in enumerate(self._qmul)))
        head.linear.bias = torch.nn.Parameter(torch.zeros(2))
        head.sph_to_cart = _Holder()
        head.sph_to_cart.register_buffer("Q_flat", torch.tensor(cv.rank4_symmetric_basis().reshape(21, 81), dtype=torch.float32))
        abcd, rows, cols, mask = cv.mandel_tables()
        head.cart_to_Mandel = _Holder()
        head.cart_to_Mandel.register_buffer("mask", torch.tensor(mask, dtype=torch.float32))
        head.cart_to_Mandel.register_buffer("rows", torch.tensor(rows))
        head.cart_to_Mandel.register_buffer("cols", torch.tensor(cols))
        self._abcd = abcd

        # ---- static device tables (non-persistent buffers) ----------------------------------------------------------
        self.register_buffer("_a_row_ptr", torch.tensor(wmap["row_ptr"]), persistent=False)
        self.register_buffer("_a_col", torch.tensor(wmap["col"]), persistent=False)
        self.register_buffer("_a_val", torch.tensor(wmap["val"]), persistent=False)

        # ---- layouts ---------------------------------------------------------------------------------------------------
        rows_l = [2 * l + 1 for l in range(lmax + 1)]
        self._lay_h = nv.make_layout(rows_l, [C] * (lmax + 1))
        self._lay_h0 = nv.make_layout([1], [C])
        self._lay_f = nv.make_layout([1], [2 * nb])
        self._lay_mlp = nv.make_layout([1], [H])
        self._lay_M = [nv.make_layout(rows_l, [ps["n_paths"] * C for ps in plan["passes"]]) for plan in self._tp]
        self._lay_w = [nv.make_layout([1], [len(plan["paths"]) * C]) for plan in self._tp]
        self._lay_z = nv.make_layout(rows_l, [C + n_gates] + [C] * lmax)
        self._lay_g = nv.make_layout(rows_l, [C] * (lmax + 1))
        self._lay_o = nv.make_layout(rows_l, self.rmul)
        self._cache: Dict = {}
        # True: check edge_index / batch ranges and the ordering of `batch` on every forward (one small kernel + a host
        # synchronisation; the kernels index device memory with these tensors unchecked).  EEG_VALIDATE=1 turns it on globally.
        self.validate_inputs = os.environ.get("EEG_VALIDATE", "0") not in ("", "0")

    # ------------------------------------------------------------------------------------------------------------------
    def _sc_table(self, i: int) -> torch.Tensor:
        """T[C, n_coef] for layer i (weights -> Horner coefficients); cached on the parameters' version counters."""
        cons = self.stiffness_head.layers[i].product.symmetric_contractions.contractions
        ws = [con.weights[str(nu)] for con in cons.values() for nu in range(1, self.correlation + 1)]
        key = tuple((w.data_ptr(), w._version) for w in ws)
        hit = self._cache.get(("sc", i))
        if hit is not None and hit[0] == key:
            return hit[1]
        Wflat = torch.cat([w.detach() for w in ws], 0).contiguous()
        prev_tag = nv.get_tag()
        nv.set_tag(f'L{i}.sc_coefficients')
        T = nv.sc_coefficients(self._a_row_ptr, self._a_col, self._a_val, self._wmap["n_coef"], Wflat, self.C)
        nv.set_tag(prev_tag)   # callers evaluate this as an argument of the tagged sym_contract call
        self._cache[("sc", i)] = (key, T)
        return T

    def _radial_image(self, i: int, lins):
        """Packed hi/lo tcgen05 operand image of layer i's radial-MLP weights (cached on the weights' versions)."""
        key = tuple((l.weight.data_ptr(), l.weight._version) for l in lins[1:]) + tuple((l.bias.data_ptr(), l.bias._version) for l in lins[1:-1])
        hit = self._cache.get(("radial", i))
        if hit is not None and hit[0] == key:
            return hit[1], hit[2]
        hidden = lins[1:-1]
        Wh = torch.stack([l.weight.detach() for l in hidden]).contiguous() if hidden else None
        bh = torch.stack([l.bias.detach() for l in hidden]).contiguous() if hidden else lins[0].bias.detach().new_zeros(1, lins[0].weight.shape[0])
        prev_tag = nv.get_tag()
        nv.set_tag(f'L{i}.radial_pack')
        img = nv.radial_pack(Wh, lins[-1].weight.detach().contiguous())
        nv.set_tag(prev_tag)
        self._cache[("radial", i)] = (key, img, bh)
        return img, bh

    def _cache_bump(self):
        """Drop weight-derived tables (SC coefficients, radial operand image) after an in-place optimizer step that
        bypasses torch's version counters (train.FlatAdamW updates the flat parameter buffer through the C ABI)."""
        for k in [k for k in self._cache if isinstance(k, tuple)]:
            del self._cache[k]

    def _tail_matrix(self) -> torch.Tensor:
        """P[nq, 36]: q -> Mandel matrix A (Spherical_to_Cartesian then Cart_4_to_Mandel, gnn/blocks.py:392-442)."""
        head = self.stiffness_head
        Q = head.sph_to_cart.Q_flat
        key = (Q.data_ptr(), Q._version, head.cart_to_Mandel.mask._version)
        hit = self._cache.get("P")
        if hit is not None and hit[0] == key:
            return hit[1]
        a, b, c, d = (torch.as_tensor(self._abcd[:, k], device=Q.device) for k in range(4))
        flat = a * 27 + b * 9 + c * 3 + d
        rows, cols, mask = head.cart_to_Mandel.rows, head.cart_to_Mandel.cols, head.cart_to_Mandel.mask
        P = torch.zeros(21, 6, 6, dtype=torch.float32, device=Q.device)
        P[:, rows, cols] = Q[:, flat]
        P[:, cols, rows] = Q[:, flat]
        P = (P * mask.view(1, 6, 6)).reshape(21, 36)
        # keep only rows of q blocks that exist for this lmax (2x0e | 2x2e | 1x4e)
        keep, off = [], 0
        for l, n in ((0, 2), (2, 2), (4, 1)):
            if l <= self.lmax:
                keep += list(range(off, off + n * (2 * l + 1)))
            off += n * (2 * l + 1)
        P = P[keep].contiguous()
        self._cache["P"] = (key, P)
        return P

    # ------------------------------------------------------------------------------------------------------------------
    def forward(self, batch) -> Dict[str, torch.Tensor]:
        pos = batch.positions
        if not pos.is_cuda:
            raise RuntimeError("EnergyEquivGNN (B200) has no CPU path: move the batch to a CUDA device")
        if torch.is_grad_enabled() and any(p.requires_grad for p in self.parameters()):
            from .autograd import forward_with_grad  # explicit backward ops
            return {"stiffness": forward_with_grad(self, batch)}
        return {"stiffness": self._forward_impl(batch)}

    def _forward_impl(self, batch, save=None) -> torch.Tensor:
        C, lmax = self.C, self.lmax
        head = self.stiffness_head
        edge_index = batch.edge_index.contiguous()
        N, G = batch.node_attrs.shape[0], int(batch.num_graphs)
        f32 = torch.float32
        if self.validate_inputs:
            nv.set_tag('validate')
            nv.graph_validate(edge_index, N, batch.batch.contiguous(), G)
        nv.set_tag('csr')
        row_ptr, perm = nv.csr_build(edge_index, N)
        nv.set_tag('graph_ptr')
        gp = nv.graph_ptr(batch.batch.contiguous(), G)
        nv.set_tag('edge_featurize')
        Y, f, snd = nv.edge_featurize(batch.positions.to(f32).contiguous(), edge_index, batch.shifts.to(f32).contiguous(),
                                      batch.edge_attr.to(f32).reshape(-1).contiguous(), perm, lmax, self.number_of_edge_basis,
                                      0.6, self.max_edge_radius)
        E = edge_index.shape[1]
        emb = self.node_ft_embedding
        nv.set_tag('node_embed')
        h = nv.node_embed(batch.node_attrs.to(f32).contiguous(), emb.weight.detach().reshape(-1).contiguous(), emb.bias.detach())
        sC = 1.0 / math.sqrt(C)
        if save is not None:
            save.update(row_ptr=row_ptr, perm=perm, gp=gp, Y=Y, f=f, snd=snd, h_in=[], hup=[], acts=[], w=[], M=[], msg=[], B=[], T=[])
        o_chain = hup_next = None
        for i, layer in enumerate(head.layers):
            inter, prod = layer.interaction, layer.product
            lay_in = self._lay_h0 if i == 0 else self._lay_h
            nblk = lay_in.n_blocks
            if hup_next is not None:   # produced together with the previous layer's product.linear
                hup, hup_next = hup_next, None
            else:
                nv.set_tag(f'L{i}.linear_up')
                hup = nv.block_linear(h, lay_in, inter.linear_up.weight.detach(), [b * C * C for b in range(nblk)], [sC] * nblk, lay_in, N)
            # radial MLP (gnn/blocks.py:538-549,590): fused, last layers on tcgen05; output path-major w_t[P*C/32, E, 32]
            lins = [m for m in inter.conv_tp_weights if isinstance(m, torch.nn.Linear)]
            nv.set_tag(f'L{i}.radial_mlp')
            if nv.radial_supported(f.shape[1], int(self.params.inter_MLP_dim), len(lins)):
                img, bh = self._radial_image(i, lins)
                w = nv.radial_mlp(f, lins[0].weight.detach(), lins[0].bias.detach(), bh, len(lins) - 2, img, lins[-1].weight.shape[0])
            else:   # any other width / depth: layer-by-layer block linears (tcgen05 for K >= 128) + SiLU, then the path-major layout
                a, lay_in = f, self._lay_f
                for lin in lins[:-1]:
                    a = nv.silu_fwd(nv.block_linear(a, lay_in, lin.weight.detach().reshape(-1), [0], [1.0], self._lay_mlp, E,
                                                    bias=lin.bias.detach(), transB=True))
                    lay_in = self._lay_mlp
                w = nv.to_path_major(nv.block_linear(a, lay_in, lins[-1].weight.detach().reshape(-1), [0], [1.0], self._lay_w[i], E, transB=True),
                                     lins[-1].weight.shape[0])
            acts = None
            plan = self._tp[i]
            nv.set_tag(f'L{i}.tp_scatter')
            M = nv.tp_scatter(0 if i == 0 else lmax, lmax, C, hup, Y, w, snd, row_ptr, 1.0 / self.agg_norm_const, plan["rows_total"], N)
            if self.interaction_reduction == "mean":   # torch_scatter reduce='mean' (gnn/blocks.py:595-597)
                nv.scale_rows_inv_degree(M, row_ptr)
            n_l = [ps["n_paths"] for ps in plan["passes"]]
            woff = np.concatenate(([0], np.cumsum([n * C * C for n in n_l])))[:-1].tolist()
            nv.set_tag(f'L{i}.interaction_linear')
            msg = nv.block_linear(M, self._lay_M[i], inter.linear.weight.detach(), woff, [1.0 / math.sqrt(n * C) for n in n_l],
                                  self._lay_h, N, bias=inter.linear.bias.detach())
            nv.set_tag(f'L{i}.sym_contract')
            T = self._sc_table(i)
            if save is not None:
                save["T"].append(T)   # the backward differentiates THIS table, even if the weights change before it runs
            B = nv.sym_contract(lmax, self.correlation, C, msg, T, N)
            poff = [b * C * C for b in range(lmax + 1)]
            if save is None and i == len(head.layers) - 1 and os.environ.get("EEG_READOUT_CHAIN", "1") != "0":
                # inference, last layer: product.linear (+ residual) -> linear_1 -> Gate -> linear_2 chained in one pass;
                # h and z are never written (None when the shape is outside that kernel: the unfused calls below run)
                ro = head.nonlin_readout
                nv.set_tag('readout.chain')
                o_chain = nv.readout_chain(B, h if i > 0 else None, self._lay_h, prod.linear.weight.detach(), poff, [sC] * (lmax + 1),
                                           ro.linear_1.weight.detach(), [0] + [C * (C + C * lmax) + b * C * C for b in range(lmax)],
                                           [sC] * (lmax + 1), ro.linear_2.weight.detach(),
                                           np.concatenate(([0], np.cumsum([C * m for m in self.rmul])))[:-1].tolist(),
                                           [sC] * (lmax + 1), NORM2MOM_SILU, self._lay_o, N)
                if o_chain is not None:
                    break
            h_new = None
            if save is None and i + 1 < len(head.layers) and os.environ.get("EEG_LINEAR_PAIR", "1") != "0":
                # inference: product.linear of this layer and linear_up of the next one chained in one pass
                nv.set_tag(f'L{i}.product_linear+L{i + 1}.linear_up')
                pair = nv.linear_pair(B, h if i > 0 else None, self._lay_h, prod.linear.weight.detach(), poff, [sC] * (lmax + 1),
                                      head.layers[i + 1].interaction.linear_up.weight.detach(), poff, [sC] * (lmax + 1), N)
                if pair is not None:
                    h_new, hup_next = pair
            if h_new is None:
                nv.set_tag(f'L{i}.product_linear')
                h_new = nv.block_linear(B, self._lay_h, prod.linear.weight.detach(), poff,
                                        [sC] * (lmax + 1), self._lay_h, N, residual=h if i > 0 else None)
            if save is not None:
                save["h_in"].append(h); save["hup"].append(hup); save["acts"].append(acts); save["w"].append(w)
                save["M"].append(M); save["msg"].append(msg); save["B"].append(B)
            h = h_new
        ro = head.nonlin_readout
        n_gates = C * lmax
        w1off = [0] + [C * (C + n_gates) + b * C * C for b in range(lmax)]
        w2off = np.concatenate(([0], np.cumsum([C * m for m in self.rmul])))[:-1].tolist()
        g = z = None
        o = o_chain
        if o is None:
            nv.set_tag('readout.linear_1')
            z = nv.block_linear(h, self._lay_h, ro.linear_1.weight.detach(), w1off, [sC] * (lmax + 1), self._lay_z, N)
        if o is None and save is None:  # inference: Gate applied while linear_2's operand rows are staged (no gated tensor in HBM)
            nv.set_tag('readout.gate_linear_2')
            o = nv.gate_linear(z, self._lay_z, C, NORM2MOM_SILU, ro.linear_2.weight.detach(), w2off, [sC] * (lmax + 1), self._lay_o, N)
        if o is None:
            nv.set_tag('readout.gate')
            g = nv.gate(z, self._lay_z, self._lay_g, C, NORM2MOM_SILU, N)
            nv.set_tag('readout.linear_2')
            o = nv.block_linear(g, self._lay_g, ro.linear_2.weight.detach(), w2off, [sC] * (lmax + 1), self._lay_o, N)
        qscale = [1.0 / math.sqrt(m) for m in self.rmul]
        nv.set_tag('pool_tail')
        Cst, q, A = nv.pool_tail(o, self._lay_o, gp, head.linear.weight.detach(), self._qmul, qscale, head.linear.bias.detach(),
                                 self._tail_matrix(), _POSITIVE[self.params.positive_function],
                                 0 if self.params.global_reduction == "mean" else 1, G, want_aux=save is not None)
        if save is not None:
            save.update(h_out=h, z=z, g=g, o=o, q=q, A=A)
        return Cst
