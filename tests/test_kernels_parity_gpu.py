"""Per-op / per-stage GPU parity (test_kernels_parity_gpu).

(a) eeg_tp_scatter_fwd alone (tensor-core kernel csrc/eeg_tp_umma.cu for C = 32, CUDA-core kernel otherwise) against an
    fp64 evaluation of the same sparse Clebsch-Gordan sum on a random graph with empty rows and a hub node;
(b) every stage of the model against the intermediates the UNMODIFIED reference produced (tests/golden/ref_forward_*.pt
    'inter', written by oracle/make_golden.py): edge SH, per-layer message (interaction.linear output), per-layer output
    (product.linear output), readout, q, A.  fp32 relative tolerance 1e-5 of each tensor's max (north star)."""
import os
from argparse import Namespace

import pytest
import torch

from energy_equiv_lattice_gnn_b200 import EnergyEquivGNN, native as nv, tables as tb
from energy_equiv_lattice_gnn_b200.data import Batch

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def rel(a, b):
    return float((a.double().cpu() - b.double().cpu()).abs().max() / b.double().abs().max())


def tp_reference(plan, h, Y, w_t, snd, row_ptr, inv_norm, C):
    """fp64, dense loops over the CG non-zeros: M[n, row, u] (tables.tp_plan row layout)."""
    N, E = h.shape[0], Y.shape[0]
    h, Y, w_t = h.double(), Y.double(), w_t.double()
    n_l = {ps["l3"]: ps["n_paths"] for ps in plan["passes"]}
    local = {p: pl for ps in plan["passes"] for pl, p in enumerate(ps["path_ids"])}
    recv = torch.repeat_interleave(torch.arange(N), (row_ptr[1:] - row_ptr[:-1]).long())
    msg = torch.zeros(E, plan["rows_total"], C, dtype=torch.float64)
    xs = h[snd.long()]                       # [E, Din, C]
    for p, (l1, l2, l3) in enumerate(plan["paths"]):
        for i, j, k, v in plan["cg"][p]:
            msg[:, plan["row_off"][l3] + k * n_l[l3] + local[p]] += v * w_t[p] * xs[:, l1 * l1 + i] * Y[:, l2 * l2 + j, None]
    M = torch.zeros(N, plan["rows_total"], C, dtype=torch.float64)
    M.index_add_(0, recv, msg)
    return M * inv_norm


@pytest.mark.parametrize("lmax,C,umma", [(4, 32, False), (4, 32, True), (3, 32, False), (3, 32, True), (3, 64, False)])
def test_tp_scatter_op(cuda_device, lmax, C, umma):
    plan = tb.tp_plan(lmax, lmax, lmax)
    g = torch.Generator().manual_seed(100 + lmax + C)
    N = 301
    deg = torch.randint(0, 7, (N,), generator=g)
    deg[5] = 0
    deg[N - 1] = 0
    deg[17] = 41          # hub node: more edges than ring slots
    row_ptr = torch.zeros(N + 1, dtype=torch.int32)
    row_ptr[1:] = torch.cumsum(deg, 0).int()
    E = int(row_ptr[-1])
    snd = torch.randint(0, N, (E,), generator=g, dtype=torch.int32)
    D, P = (lmax + 1) ** 2, len(plan["paths"])
    ystride = (D + 3) // 4 * 4
    h = torch.randn(N, D, C, generator=g)
    Y = torch.zeros(E, ystride)
    Y[:, :D] = torch.randn(E, D, generator=g)
    w = torch.randn(P, E, C, generator=g)                       # [path, edge, channel]
    # kernel layout: w_t[path * chunks + chunk][E][32]
    w_t = w.reshape(P, E, C // 32, 32).permute(0, 2, 1, 3).reshape(P * (C // 32), E, 32).contiguous()
    want = tp_reference(plan, h.reshape(N, D, C), Y, w, snd, row_ptr, 0.25, C)
    d = cuda_device
    got = nv.tp_scatter(lmax, lmax, C, h.reshape(N, D * C).to(d).contiguous(), Y.to(d), w_t.to(d), snd.to(d), row_ptr.to(d), 0.25,
                        plan["rows_total"], N, umma=umma)
    got2 = nv.tp_scatter(lmax, lmax, C, h.reshape(N, D * C).to(d).contiguous(), Y.to(d), w_t.to(d), snd.to(d), row_ptr.to(d), 0.25,
                         plan["rows_total"], N, umma=umma)
    torch.cuda.synchronize()
    assert torch.equal(got, got2), "tensor-product scatter is not bitwise repeatable"
    got = got.reshape(N, plan["rows_total"], C)
    assert float(got[5].abs().max()) == 0.0 and float(got[N - 1].abs().max()) == 0.0   # rows without edges are zero
    assert rel(got, want) < 2e-6, rel(got, want)


def to_e3nn(x, lmax, mul):
    """blocked m-major [N, sum_l (2l+1) mul_l] -> e3nn flat (per irrep mul-major)"""
    out, off = [], 0
    for l in range(lmax + 1):
        n = (2 * l + 1) * mul[l]
        out.append(x[:, off:off + n].reshape(-1, 2 * l + 1, mul[l]).transpose(1, 2).reshape(x.shape[0], n))
        off += n
    return torch.cat(out, 1)


@pytest.mark.parametrize("name", ["lmax4", "lmax3"])
def test_stages_match_reference_intermediates(cuda_device, name):
    fx = torch.load(os.path.join(GOLD, f"ref_forward_{name}.pt"), weights_only=False)
    model = EnergyEquivGNN(Namespace(**fx["params"]))
    missing, unexpected = model.load_state_dict(fx["state_dict"], strict=False)
    assert not missing and not unexpected
    model = model.to(cuda_device).eval()
    batch = Batch(**fx["batch"]).to(cuda_device)
    save = {}
    with torch.no_grad():
        out = model._forward_impl(batch, save)
    ref = fx["inter"]
    lmax, C = model.lmax, model.C
    D = (lmax + 1) ** 2
    perm = save["perm"].long().cpu()
    assert rel(save["Y"][:, :D], ref["edge_sh"][perm]) < 1e-5
    tol = 1e-5
    for i in range(2):
        assert rel(to_e3nn(save["msg"][i], lmax, [C] * (lmax + 1)), ref[f"layer{i}_message"]) < tol, f"layer{i}_message"
    h1 = save["h_in"][1]                      # output of layer 0 (no residual on the first layer, gnn/model.py:93-96)
    assert rel(to_e3nn(h1, lmax, [C] * (lmax + 1)), ref["layer0_out"]) < tol
    l1 = save["h_out"] - h1                   # the kernel fuses the residual h + L1(h)
    assert rel(to_e3nn(l1, lmax, [C] * (lmax + 1)), ref["layer1_out"]) < 2 * tol
    assert rel(to_e3nn(save["o"], lmax, model.rmul), ref["readout"]) < tol
    assert rel(save["A"].reshape(-1, 6, 6), ref["A"]) < tol
    assert rel(out, fx["stiffness"]) < tol
    # q: the kernel keeps only the blocks that exist for this lmax (2x0e | 2x2e | 1x4e)
    q_ref = ref["q"]
    if lmax < 4:
        q_ref = q_ref[:, :12]
    assert rel(save["q"][:, :q_ref.shape[1]], q_ref) < tol


@pytest.mark.parametrize("rows,k_in,R", [([1], [1344], 20003), ([1], [192], 30001), ([1, 3], [128, 320], 4099)])
def test_wide_weight_gradient_on_tensor_cores(cuda_device, rows, k_in, R):
    """eeg_block_linear_wgrad for 64 output columns and >= 128 weight rows runs on tcgen05 (csrc/eeg_wgrad_umma.cu; the last
    radial-MLP layer's dW = dw^T act, gnn/blocks.py:538-549): against torch fp64 at 1e-5 of the max, bitwise repeatable, with
    a row count that is not a multiple of the stage (8) or pass length, ragged last weight tile (1344 = 10.5 x 128, 192, 320),
    several rows per node, a bias gradient, and accumulation into a non-zero dW."""
    g = torch.Generator().manual_seed(7 + R)
    lin, lout = nv.make_layout(rows, k_in), nv.make_layout(rows, [64] * len(rows))
    x = torch.randn(R, lin.stride, generator=g).to(cuda_device)
    dy = torch.randn(R, lout.stride, generator=g).to(cuda_device)
    w_off, n_w = [], 0
    for k in k_in:
        w_off.append(n_w)
        n_w += k * 64
    scale = [0.5 + 0.25 * b for b in range(len(rows))]
    assert nv.lib().eeg_block_linear_wgrad_ws_floats is not None
    got, gb = nv.linear_wgrad(x, lin, dy, lout, w_off, scale, n_w, R, want_bias=64)
    again, gb2 = nv.linear_wgrad(x, lin, dy, lout, w_off, scale, n_w, R, want_bias=64)
    assert torch.equal(got, again) and torch.equal(gb, gb2)
    for b, (r, k) in enumerate(zip(rows, k_in)):
        xb = x[:, lin.off[b]:lin.off[b] + r * k].reshape(R * r, k).double()
        yb = dy[:, lout.off[b]:lout.off[b] + r * 64].reshape(R * r, 64).double()
        ref = scale[b] * xb.t() @ yb
        assert rel(got[w_off[b]:w_off[b] + k * 64].view(k, 64), ref) < 1e-5
    assert rel(gb, dy[:, :64].double().sum(0)) < 1e-5


@pytest.mark.parametrize("K,R", [(1344, 20003), (192, 30001), (2048, 4099)])
def test_long_k_input_gradient_on_tensor_cores(cuda_device, K, R):
    """eeg_block_linear_fwd with one block, one row per node, K >= 128 and 64 outputs (ds = dw W2 of the radial-MLP backward,
    gnn/blocks.py:538-549) runs on tcgen05 with TMA-loaded operands (csrc/eeg_dx_umma.cu): against torch fp64 at 1e-5 of the
    max, bitwise repeatable, row counts that are not multiples of the 512-row CTA tile, one to four accumulation passes, and
    an output row pitch wider than the 64 columns written."""
    g = torch.Generator().manual_seed(11 + K)
    lin = nv.make_layout([1], [K])
    lout = nv.make_layout([1], [64])
    x = torch.randn(R, K, generator=g).to(cuda_device)
    W = torch.randn(K, 64, generator=g).to(cuda_device)
    got = nv.block_linear(x, lin, W.reshape(-1), [0], [0.25], lout, R)
    again = nv.block_linear(x, lin, W.reshape(-1), [0], [0.25], lout, R)
    assert torch.equal(got, again)
    ref = 0.25 * x.double() @ W.double()
    assert rel(got, ref) < 1e-5


@pytest.mark.parametrize("F,E", [(20, 10007), (8, 64), (32, 4097)])
def test_fused_radial_hidden_backward(cuda_device, F, E):
    """eeg_radial_hidden_bwd (csrc/eeg_radial_bwd.cu): Linear(F,64)-SiLU-Linear(64,64)-SiLU of the radial MLP
    (gnn/blocks.py:538-549), activations recomputed and all four parameter gradients + s2 in one pass, against torch fp64
    autograd at 1e-5 of each tensor's max; edge counts that are not multiples of the 64-edge tile; bitwise repeatable."""
    g = torch.Generator().manual_seed(5 + F)
    f = torch.randn(E, F, generator=g)
    W0, b0 = torch.randn(64, F, generator=g) / F ** 0.5, 0.1 * torch.randn(64, generator=g)
    W1, b1 = torch.randn(64, 64, generator=g) / 8.0, 0.1 * torch.randn(64, generator=g)
    ds2 = torch.randn(E, 64, generator=g)
    p = [t.double().requires_grad_(True) for t in (W0, b0, W1, b1)]
    s1 = torch.nn.functional.silu(f.double() @ p[0].t() + p[1])
    s2 = torch.nn.functional.silu(s1 @ p[2].t() + p[3])
    (s2 * ds2.double()).sum().backward()
    dev = [t.to(cuda_device) for t in (f, W0, b0, W1, b1, ds2)]
    got = nv.radial_hidden_bwd(dev[0], F, dev[1], dev[2], dev[3], dev[4], dev[5], E)
    again = nv.radial_hidden_bwd(dev[0], F, dev[1], dev[2], dev[3], dev[4], dev[5], E)
    for a, b in zip(got, again):
        assert torch.equal(a, b)
    assert rel(got[0], s2.detach()) < 1e-5
    for a, ref in zip(got[1:], p):
        assert rel(a, ref.grad) < 1e-5


def test_wide_output_linear_from_multiplicity_32(cuda_device):
    """eeg_block_linear_fwd, multiplicity 32 -> 160..352 columns per irrep block (dM = dmsg W^T of the interaction.linear backward,
    gnn/blocks.py:552-559) runs on warp MMAs (csrc/eeg_lin32wide.cu): both weight layouts against torch fp64 at 1e-5, a node
    count that leaves partial 16-row tiles, bitwise repeatable."""
    g = torch.Generator().manual_seed(3)
    N, rows, nout = 1037, [1, 3, 5, 7, 9], [160, 256, 352, 32, 288]
    lin, lout = nv.make_layout(rows, [32] * 5), nv.make_layout(rows, nout)
    x = torch.randn(N, lin.stride, generator=g).to(cuda_device)
    w_off, n_w = [], 0
    for n in nout:
        w_off.append(n_w)
        n_w += 32 * n
    W = torch.randn(n_w, generator=g).to(cuda_device)
    scale = [0.3, 0.5, 0.7, 1.1, 1.3]
    for transB in (True, False):
        got = nv.block_linear(x, lin, W, w_off, scale, lout, N, transB=transB)
        assert torch.equal(got, nv.block_linear(x, lin, W, w_off, scale, lout, N, transB=transB))
        for b, (r, n) in enumerate(zip(rows, nout)):
            Wb = W[w_off[b]:w_off[b] + 32 * n].double()
            Wb = Wb.view(n, 32).t() if transB else Wb.view(32, n)
            xin = x[:, lin.off[b]:lin.off[b] + r * 32].reshape(N * r, 32).double()
            ref = scale[b] * xin @ Wb
            assert rel(got[:, lout.off[b]:lout.off[b] + r * n].reshape(N * r, n), ref) < 1e-5


def test_weight_gradient_32_outputs_on_warp_mma(cuda_device):
    """eeg_block_linear_wgrad with 32 output channels and K = 64..352 per block (path-stacked interaction.linear,
    gnn/blocks.py:552-559) runs on mma.sync (csrc/eeg_wgrad_mma.cu): against torch fp64 at 1e-5, with a bias gradient, a row
    count that leaves partial 8-row K-steps and partial 512-row slices, both weight layouts, bitwise repeatable."""
    g = torch.Generator().manual_seed(17)
    N, rows, k_in = 6007, [1, 3, 5, 7, 9], [160, 256, 352, 64, 288]
    lin, lout = nv.make_layout(rows, k_in), nv.make_layout(rows, [32] * 5)
    x = torch.randn(N, lin.stride, generator=g).to(cuda_device)
    dy = torch.randn(N, lout.stride, generator=g).to(cuda_device)
    w_off, n_w = [], 0
    for k in k_in:
        w_off.append(n_w)
        n_w += k * 32
    scale = [0.3, 0.5, 0.7, 1.1, 1.3]
    for transB in (False, True):
        got, gb = nv.linear_wgrad(x, lin, dy, lout, w_off, scale, n_w, N, transB=transB, want_bias=32)
        again, gb2 = nv.linear_wgrad(x, lin, dy, lout, w_off, scale, n_w, N, transB=transB, want_bias=32)
        assert torch.equal(got, again) and torch.equal(gb, gb2)
        for b, (r, k) in enumerate(zip(rows, k_in)):
            xb = x[:, lin.off[b]:lin.off[b] + r * k].reshape(N * r, k).double()
            yb = dy[:, lout.off[b]:lout.off[b] + r * 32].reshape(N * r, 32).double()
            ref = scale[b] * xb.t() @ yb
            gotb = got[w_off[b]:w_off[b] + k * 32]
            assert rel(gotb.view(32, k).t() if transB else gotb.view(k, 32), ref) < 1e-5
        assert rel(gb, dy[:, :32].double().sum(0)) < 1e-5
