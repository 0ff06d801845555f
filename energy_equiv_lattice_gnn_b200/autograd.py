"""Explicit backward of EnergyEquivGNN.forward: one torch.autograd.Function whose backward runs the `_bwd` / `_wgrad`
kernels of libeeg_b200 in reverse order of `EnergyEquivGNN._forward_impl` (the reference gets the same gradients from
torch autograd over e3nn / torch_scatter / opt_einsum; scripts/train_utils.py:45-64).  Positions are data, not
parameters: no gradient flows to the batch.
"""
from __future__ import annotations

import math
from typing import Dict, List

import numpy as np
import torch

from . import native as nv
from . import tables as tb


class _Forward(torch.autograd.Function):
    @staticmethod
    def forward(ctx, model, batch, *params):
        save: Dict = {}
        out = model._forward_impl(batch, save)
        ctx.model, ctx.batch, ctx.saved = model, batch, save
        return out

    @staticmethod
    def backward(ctx, dC):
        if ctx.saved is None:
            raise RuntimeError("EnergyEquivGNN: backward through the same forward a second time - the saved activations were freed "
                               "after the first backward (run the forward again; retain_graph=True is not supported by the explicit backward)")
        grads = backward_impl(ctx.model, ctx.batch, ctx.saved, dC.contiguous().to(torch.float32))
        ctx.saved = None
        return (None, None, *grads)


def forward_with_grad(model, batch) -> torch.Tensor:
    return _Forward.apply(model, batch, *list(model.parameters()))


def _sender_csr(sv, N):
    """CSR of the batch's (receiver-sorted) edges by sender, built once per backward and shared by the layers."""
    if "snd_csr" not in sv:
        sv["snd_csr"] = nv.sender_csr(sv["snd"], N)
    return sv["snd_csr"]


def _static(model, name, build):
    hit = model._cache.get(name)
    if hit is None:
        hit = model._cache[name] = build()
    return hit


def _mlp_backward(lins: List[torch.nn.Linear], f, dw, E, lay_f, lay_h, lay_w, grads):
    """conv_tp_weights MLP (gnn/blocks.py:538-549): activations are recomputed, then dW / db layer by layer."""
    F = lay_f.mul[0]
    if (len(lins) == 3 and lins[0].out_features == 64 and lins[1].in_features == 64 and lins[1].out_features == 64 and F % 4 == 0
            and F <= 32 and lins[0].in_features == F and lins[0].bias is not None and lins[1].bias is not None):
        # default shape: ds2 on tcgen05 (eeg_dx_umma.cu), the two hidden layers in one fused pass (eeg_radial_bwd.cu) which also
        # returns s2 = silu(a2), then the last layer's weight gradient on tcgen05 (eeg_wgrad_umma.cu)
        last = lins[-1]
        W2 = last.weight.detach()
        ds2 = nv.block_linear(dw, lay_w, W2.reshape(-1), [0], [1.0], lay_h, E)
        s2, dW0, db0, dW1, db1 = nv.radial_hidden_bwd(f, F, lins[0].weight.detach(), lins[0].bias.detach(), lins[1].weight.detach(),
                                                      lins[1].bias.detach(), ds2, E)
        grads[id(lins[0].weight)], grads[id(lins[0].bias)] = dW0, db0
        grads[id(lins[1].weight)], grads[id(lins[1].bias)] = dW1, db1
        grads[id(last.weight)] = nv.linear_wgrad(dw, lay_w, s2, lay_h, [0], [1.0], W2.numel(), E)[0].view_as(W2)
        return
    pre, act = [], [f]
    lay_in = lay_f
    for lin in lins[:-1]:
        a = nv.block_linear(act[-1], lay_in, lin.weight.detach().reshape(-1), [0], [1.0], lay_h, E, bias=lin.bias.detach(), transB=True)
        pre.append(a)
        act.append(nv.silu_fwd(a))
        lay_in = lay_h
    last = lins[-1]
    W2 = last.weight.detach()  # [n_out, H]
    lay_last = lay_f if len(lins) == 1 else lay_h   # single-layer MLP: the edge features feed the last linear directly
    grads[id(last.weight)] = nv.linear_wgrad(dw, lay_w, act[-1], lay_last, [0], [1.0], W2.numel(), E)[0].view_as(W2)
    if len(lins) == 1:
        return
    ds = nv.block_linear(dw, lay_w, W2.reshape(-1), [0], [1.0], lay_h, E)
    for k in range(len(lins) - 2, -1, -1):
        lin = lins[k]
        da = nv.silu_bwd(pre[k], ds)
        lay_prev = lay_f if k == 0 else lay_h
        grads[id(lin.weight)] = nv.linear_wgrad(da, lay_h, act[k], lay_prev, [0], [1.0], lin.weight.numel(), E)[0].view_as(lin.weight)
        grads[id(lin.bias)] = nv.colsum(da)
        if k > 0:
            ds = nv.block_linear(da, lay_h, lin.weight.detach().reshape(-1), [0], [1.0], lay_h, E)


def backward_impl(model, batch, sv, dC) -> List[torch.Tensor]:
    C, lmax = model.C, model.lmax
    head = model.stiffness_head
    ro = head.nonlin_readout
    N, G = batch.node_attrs.shape[0], int(batch.num_graphs)
    E = sv["snd"].shape[0]
    sC = 1.0 / math.sqrt(C)
    nblk = lmax + 1
    grads: Dict[int, torch.Tensor] = {}
    from .model import NORM2MOM_SILU, _POSITIVE

    # ---- tail: PositiveLayer, Mandel map, final linear, pooling
    qscale = [1.0 / math.sqrt(m) for m in model.rmul]
    nv.set_tag('bwd.pool_tail')
    d_o, dWq, dbq = nv.pool_tail_bwd(sv["o"], model._lay_o, sv["gp"], head.linear.weight.detach(), model._qmul, qscale,
                                     head.linear.bias.numel(), model._tail_matrix(), _POSITIVE[model.params.positive_function],
                                     0 if model.params.global_reduction == "mean" else 1, sv["A"], dC.reshape(G, 36), G)
    grads[id(head.linear.weight)] = dWq
    grads[id(head.linear.bias)] = dbq
    # ---- readout: linear_2, gate, linear_1
    w2off = np.concatenate(([0], np.cumsum([C * m for m in model.rmul])))[:-1].tolist()
    W2 = ro.linear_2.weight.detach()
    nv.set_tag('bwd.readout.linear_2')
    grads[id(ro.linear_2.weight)] = nv.linear_wgrad(sv["g"], model._lay_g, d_o, model._lay_o, w2off, [sC] * nblk, W2.numel(), N)[0]
    dg = nv.block_linear(d_o, model._lay_o, W2, w2off, [sC] * nblk, model._lay_g, N, transB=True)
    nv.set_tag('bwd.readout.gate')
    dz = nv.gate_bwd(sv["z"], model._lay_z, dg, model._lay_g, C, NORM2MOM_SILU, N)
    n_gates = C * lmax
    w1off = [0] + [C * (C + n_gates) + b * C * C for b in range(lmax)]
    W1 = ro.linear_1.weight.detach()
    nv.set_tag('bwd.readout.linear_1')
    grads[id(ro.linear_1.weight)] = nv.linear_wgrad(sv["h_out"], model._lay_h, dz, model._lay_z, w1off, [sC] * nblk, W1.numel(), N)[0]
    lay_z = model._lay_z
    if nblk > 1 and lay_z.mul[0] != 32 and all(lay_z.mul[b] == 32 for b in range(1, nblk)):
        # the scalar block (K = scalars + gates) and the multiplicity-32 blocks separately: the latter then take the warp-MMA
        # kernel for K = 32 (eeg_lin32mma.cu) instead of the mixed-shape path (0.103 -> 0.071 ms at 512 lattices)
        dh = torch.empty(N, model._lay_h.stride, dtype=torch.float32, device=dz.device)
        rest = list(range(1, nblk))
        nv.block_linear(dz, nv.sub_layout(lay_z, [0]), W1, w1off[:1], [sC], nv.sub_layout(model._lay_h, [0]), N, transB=True, out=dh)
        nv.block_linear(dz, nv.sub_layout(lay_z, rest), W1, w1off[1:], [sC] * (nblk - 1), nv.sub_layout(model._lay_h, rest), N,
                        transB=True, out=dh)
    else:
        dh = nv.block_linear(dz, lay_z, W1, w1off, [sC] * nblk, model._lay_h, N, transB=True)

    wmap = model._wmap
    dev = dC.device
    t_ptr = _static(model, "t_ptr", lambda: torch.tensor(wmap["t_ptr"], device=dev))
    t_row = _static(model, "t_row", lambda: torch.tensor(wmap["t_row"], device=dev))
    t_val = _static(model, "t_val", lambda: torch.tensor(wmap["t_val"], device=dev))
    offs = [b * C * C for b in range(nblk)]

    for i in range(len(head.layers) - 1, -1, -1):
        layer = head.layers[i]
        inter, prod = layer.interaction, layer.product
        lay_in = model._lay_h0 if i == 0 else model._lay_h
        nb_in = lay_in.n_blocks
        plan = model._tp[i]
        n_l = [ps["n_paths"] for ps in plan["passes"]]
        # h_new = product.linear(B) (+ h_in): gnn/blocks.py:488-490, gnn/model.py:95-96
        Wp = prod.linear.weight.detach()
        nv.set_tag(f'bwd.L{i}.product_linear')
        grads[id(prod.linear.weight)] = nv.linear_wgrad(sv["B"][i], model._lay_h, dh, model._lay_h, offs, [sC] * nblk, Wp.numel(), N)[0]
        dB = nv.block_linear(dh, model._lay_h, Wp, offs, [sC] * nblk, model._lay_h, N, transB=True)
        # symmetric contraction
        x = sv["msg"][i]
        T = sv["T"][i]   # the coefficient table the forward used
        nv.set_tag(f'bwd.L{i}.sym_contract_dx')
        dmsg = nv.sym_contract_bwd_x(lmax, model.correlation, C, x, dB, T, N)
        nv.set_tag(f'bwd.L{i}.sym_contract_dW')
        dWflat = nv.sc_weight_grad(lmax, model.correlation, x, dB, wmap["n_coef"], C, N, t_ptr, t_row, t_val, wmap["n_wrows"])
        row = 0
        for con in prod.symmetric_contractions.contractions.values():
            for nu in range(1, model.correlation + 1):
                w = con.weights[str(nu)]
                grads[id(w)] = dWflat[row:row + w.shape[0]]
                row += w.shape[0]
        # interaction.linear (+ bias on 0e)
        woff = np.concatenate(([0], np.cumsum([n * C * C for n in n_l])))[:-1].tolist()
        iscale = [1.0 / math.sqrt(n * C) for n in n_l]
        Wi = inter.linear.weight.detach()
        nv.set_tag(f'bwd.L{i}.interaction_linear')
        gW, gb = nv.linear_wgrad(sv["M"][i], model._lay_M[i], dmsg, model._lay_h, woff, iscale, Wi.numel(), N, want_bias=C)
        grads[id(inter.linear.weight)] = gW
        grads[id(inter.linear.bias)] = gb
        inv_norm = 1.0 / model.agg_norm_const   # folded into dM so that the TP backward needs no rescaling pass
        dM = nv.block_linear(dmsg, model._lay_h, Wi, woff, [sc * inv_norm for sc in iscale], model._lay_M[i], N, transB=True)
        if model.interaction_reduction == "mean":
            nv.scale_rows_inv_degree(dM, sv["row_ptr"])
        # tensor product + scatter
        nv.set_tag(f'bwd.L{i}.tp_scatter')
        dw, dhup = nv.tp_scatter_bwd(0 if i == 0 else lmax, lmax, C, sv["hup"][i], sv["Y"], sv["w"][i], sv["snd"], sv["row_ptr"],
                                     1.0, dM, len(plan["paths"]), N, snd_csr=_sender_csr(sv, N))
        # radial MLP
        lins = [m for m in inter.conv_tp_weights if isinstance(m, torch.nn.Linear)]
        nv.set_tag(f'bwd.L{i}.radial_mlp')
        _mlp_backward(lins, sv["f"], dw, E, model._lay_f, model._lay_mlp, model._lay_w[i], grads)
        # linear_up
        Wu = inter.linear_up.weight.detach()
        uoff = [b * C * C for b in range(nb_in)]
        nv.set_tag(f'bwd.L{i}.linear_up')
        grads[id(inter.linear_up.weight)] = nv.linear_wgrad(sv["h_in"][i], lay_in, dhup, lay_in, uoff, [sC] * nb_in, Wu.numel(), N)[0]
        dh_in = nv.block_linear(dhup, lay_in, Wu, uoff, [sC] * nb_in, lay_in, N, transB=True)
        dh = dh_in + dh if i > 0 else dh_in   # residual of layer i > 0 (gnn/model.py:95-96)
        if i == 1:
            # every parameter outside layer 0 and the node embedding has its final gradient: a data-parallel trainer can start
            # reducing those now, under the layer-0 backward (train.Trainer._early_bucket pops what it takes from `grads`)
            hook = model.__dict__.get("_grads_ready_hook")
            if hook is not None:
                hook(grads)
    # ---- node embedding: h0[n, c] = a[n] W[c] + b[c]
    emb = model.node_ft_embedding
    a = batch.node_attrs.to(torch.float32).reshape(-1).contiguous()
    nv.set_tag('bwd.node_embed')
    grads[id(emb.weight)] = nv.colsum(dh, a).view_as(emb.weight)
    grads[id(emb.bias)] = nv.colsum(dh)
    return [grads.get(id(p)) if p.requires_grad else None for p in model.parameters()]
