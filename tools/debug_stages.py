"""Stage-by-stage comparison of the CUDA path against the oracle (run on a GPU box)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from energy_equiv_lattice_gnn_b200.data import synthetic_batch
from tests.util import batch_to, cuda_model_from, make_pair
from oracle import e3nn_lite as e3
from oracle.model import reshape_irreps

dev = torch.device("cuda:0")
batch, rmax = synthetic_batch(range(1000, 1004))
oracle, params = make_pair(rmax, dtype=torch.float64)
model = cuda_model_from(oracle, params, dev)
b64 = batch_to(batch, torch.float64)
save = {}
with torch.no_grad():
    out = model._forward_impl(batch.to(dev), save=save)
torch.cuda.synchronize()
C, lmax = model.C, model.lmax


def rel(a, b, name):
    a, b = a.double().cpu(), b.double().cpu()
    print(f"{name:28s} max|ref|={b.abs().max():.4e} rel={float((a-b).abs().max()/b.abs().max().clamp(min=1e-30)):.3e}")


def planar_to_e3nn(x, muls):  # x [N, sum_l (2l+1)*mul_l] blocked m-major -> e3nn mul-major
    outs, off = [], 0
    for l, m in enumerate(muls):
        d = 2 * l + 1
        blk = x[:, off:off + d * m].reshape(-1, d, m).transpose(1, 2).reshape(x.shape[0], -1)
        outs.append(blk); off += d * m
    return torch.cat(outs, 1)


recv = batch.edge_index[1]
perm_ref = torch.sort(recv, stable=True).indices
print("perm exact:", torch.equal(save["perm"].cpu().long(), perm_ref))
rp_ref = torch.cat([torch.zeros(1, dtype=torch.long), torch.bincount(recv, minlength=batch.node_attrs.shape[0]).cumsum(0)])
print("row_ptr exact:", torch.equal(save["row_ptr"].cpu().long(), rp_ref))
ef, sh = oracle.edge_features(b64)
rel(save["f"], ef[perm_ref], "edge feats f")
rel(save["Y"][:, :sh.shape[1]], sh[perm_ref], "edge SH Y")
with torch.no_grad():
    node_ft = oracle.node_ft_embedding(b64.node_attrs)
    head = oracle.stiffness_head
    h = node_ft
    for i, layer in enumerate(head.layers):
        inter = layer.interaction
        muls_in = [C] if i == 0 else [C] * (lmax + 1)
        rel(planar_to_e3nn(save["h_in"][i], muls_in), h, f"L{i} h_in")
        hup = inter.linear_up(h)
        rel(planar_to_e3nn(save["hup"][i], muls_in), hup, f"L{i} linear_up")
        w = inter.conv_tp_weights(ef)
        rel(save["w"][i], w[perm_ref], f"L{i} tp weights")
        mji = inter.conv_tp(hup[b64.edge_index[0]], sh, w)
        Mref = e3.scatter(mji, recv, hup.shape[0], "sum") / params.agg_norm_const
        n_l = [ps["n_paths"] for ps in model._tp[i]["passes"]]
        rel(planar_to_e3nn(save["M"][i], [n * C for n in n_l]), Mref, f"L{i} message M")
        msg = inter.linear(Mref)
        rel(planar_to_e3nn(save["msg"][i], [C] * (lmax + 1)), msg, f"L{i} interaction.linear")
        Bref = layer.product.symmetric_contractions(reshape_irreps(msg, layer.product.in_irreps))
        rel(planar_to_e3nn(save["B"][i], [C] * (lmax + 1)), Bref, f"L{i} sym contraction")
        hn = layer.product.linear(Bref)
        h = hn if i == 0 else h + hn
    rel(planar_to_e3nn(save["h_out"], [C] * (lmax + 1)), h, "h_out")
    ro = head.nonlin_readout
    z = ro.linear_1(h)
    rel(planar_to_e3nn(save["z"], [C * (lmax + 1)] + [C] * lmax), z, "readout linear_1")
    g = ro.gate(z)
    rel(planar_to_e3nn(save["g"], [C] * (lmax + 1)), g, "gate")
    o = ro.linear_2(g)
    rel(planar_to_e3nn(save["o"], model.rmul), o, "readout linear_2")
    gft = e3.scatter(o, b64.batch, b64.num_graphs, "mean")
    q = head.linear(gft)
    rel(save["q"], q, "final linear q")
    A = head.cart_to_Mandel(head.sph_to_cart(q))
    rel(save["A"].reshape(-1, 6, 6), A, "Mandel A")
    rel(out, head.positive(A), "stiffness C")
