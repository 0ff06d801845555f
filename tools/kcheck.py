#!/usr/bin/env python
"""Kernel-level numerics + timing checks for individual C-ABI entry points (run on a GPU box):

    python tools/kcheck.py radial [E] [n_out]      radial MLP (tcgen05) vs torch fp64
Each check prints max relative error (of the max magnitude) and CUDA-event time per launch.
"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from energy_equiv_lattice_gnn_b200 import native as nv  # noqa: E402

dev = torch.device("cuda:0")


def timeit(fn, n=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def check_radial(E=122880, n_out=1344, n_hidden=1, n_in=12, seed=0):
    g = torch.Generator().manual_seed(seed)
    f = torch.rand(E, n_in, generator=g)
    W0 = torch.randn(64, n_in, generator=g) / n_in ** 0.5
    b0 = 0.1 * torch.randn(64, generator=g)
    Wh = torch.randn(max(n_hidden, 1), 64, 64, generator=g) / 8.0
    bh = 0.1 * torch.randn(max(n_hidden, 1), 64, generator=g)
    W2 = torch.randn(n_out, 64, generator=g)
    x = f.double()
    x = torch.nn.functional.silu(x @ W0.double().T + b0.double())
    for l in range(n_hidden):
        x = torch.nn.functional.silu(x @ Wh[l].double().T + bh[l].double())
    ref = x @ W2.double().T  # [E, n_out]
    fd, W0d, b0d, Whd, bhd, W2d = (t.to(dev).contiguous() for t in (f, W0, b0, Wh, bh, W2))
    img = nv.radial_pack(Whd[:n_hidden] if n_hidden else None, W2d)
    w_t = nv.radial_mlp(fd, W0d, b0d, bhd, n_hidden, img, n_out)
    torch.cuda.synchronize()
    got = w_t.permute(1, 0, 2).reshape(E, n_out).double().cpu()
    err = float((got - ref).abs().max() / ref.abs().max())
    ms = timeit(lambda: nv.radial_mlp(fd, W0d, b0d, bhd, n_hidden, img, n_out))
    flops = 2.0 * E * (n_in * 64 + n_hidden * 64 * 64 + 64 * n_out)
    print(f"radial E={E} n_out={n_out} hidden={n_hidden}: rel err {err:.3e}  {ms*1e3:.1f} us  {flops/ms/1e9:.1f} TFLOP/s  "
          f"write {E*n_out*4/ms/1e6:.0f} GB/s", flush=True)
    return err


def check_sc(N=30720, C=32, lmax=4):
    """Times the symmetric-contraction variant selected by EEG_SC_VARIANT; dumps a slice for cross-variant comparison."""
    from energy_equiv_lattice_gnn_b200 import tables as tb
    n_coef = tb.sc_plan(lmax, 3)["n_coef"]
    D = (lmax + 1) ** 2
    g = torch.Generator().manual_seed(1)
    T = (0.1 * torch.randn(C, n_coef, generator=g)).to(dev)
    x = torch.randn(N, D, C, generator=g).to(dev)
    out = nv.sym_contract(lmax, 3, C, x, T, N)
    torch.cuda.synchronize()
    ms = timeit(lambda: nv.sym_contract(lmax, 3, C, x, T, N), n=10)
    v = os.environ.get("EEG_SC_VARIANT", "default")
    os.makedirs("gpurun_out", exist_ok=True)
    torch.save(torch.cat([out[:300], out[-300:]]).cpu(), f"gpurun_out/sc_{v}_l{lmax}.pt")
    print(f"sc variant={v} lmax={lmax} N={N} C={C}: {ms*1e3:.1f} us  checksum {float(out.double().sum()):.9e}", flush=True)


def check_linear(N=30720, C=32, lmax=4, npaths=(5, 8, 10, 10, 9), seed=0):
    """Block linears (tcgen05 3xTF32 unless EEG_LINEAR=cuda) vs torch fp64: the a14 shape, a 32->32 mix with residual,
    linear_1 (0e block 32->160) and the gate-fused linear_2."""
    g = torch.Generator().manual_seed(seed)
    rows = [2 * l + 1 for l in range(lmax + 1)]
    ok = True

    def run(name, lin, lout, W, woff, scale, x, bias=None, residual=None, ragged_rows=None):
        nonlocal ok
        n = x.shape[0]
        out = nv.block_linear(x, lin, W, woff, scale, lout, n, bias=bias, residual=residual)
        torch.cuda.synchronize()
        ref = torch.zeros(n, lout.stride, dtype=torch.float64)
        xd, Wd = x.double().cpu(), W.double().cpu()
        for b in range(lin.n_blocks):
            k, no, r = lin.mul[b], lout.mul[b], lin.rows[b]
            xb = xd[:, lin.off[b]:lin.off[b] + r * k].reshape(n, r, k)
            Wb = Wd[woff[b]:woff[b] + k * no].reshape(k, no)
            yb = scale[b] * xb @ Wb
            if bias is not None and b == 0:
                yb = yb + bias.double().cpu()
            ref[:, lout.off[b]:lout.off[b] + r * no] = yb.reshape(n, r * no)
        if residual is not None:
            ref += residual.double().cpu()
        err = float((out.double().cpu() - ref).abs().max() / ref.abs().max())
        ms = timeit(lambda: nv.block_linear(x, lin, W, woff, scale, lout, n, bias=bias, residual=residual), n=10)
        gb = (x.numel() + out.numel() + (residual.numel() if residual is not None else 0)) * 4 / 1e9
        print(f"linear {name:18s} N={n}: rel err {err:.2e}  {ms*1e3:7.1f} us  {gb/ms*1e3:6.0f} GB/s", flush=True)
        ok &= err < 4e-6

    for n in (N, 1000, 77):
        # a14: path-stacked messages -> hidden
        lin = nv.make_layout(rows, [p * C for p in npaths]); lout = nv.make_layout(rows, [C] * (lmax + 1))
        woff = [0]
        for p in npaths[:-1]:
            woff.append(woff[-1] + p * C * C)
        W = torch.randn(sum(p * C * C for p in npaths), generator=g).to(dev)
        x = torch.randn(n, lin.stride, generator=g).to(dev)
        run("interaction(a14)", lin, lout, W, woff, [1.0 / (p * C) ** 0.5 for p in npaths], x, bias=torch.randn(C, generator=g).to(dev))
        # 32 -> 32 with residual
        lin = nv.make_layout(rows, [C] * (lmax + 1))
        W = torch.randn((lmax + 1) * C * C, generator=g).to(dev)
        x = torch.randn(n, lin.stride, generator=g).to(dev)
        res = torch.randn(n, lin.stride, generator=g).to(dev)
        run("product+residual", lin, lin, W, [b * C * C for b in range(lmax + 1)], [C ** -0.5] * (lmax + 1), x, residual=res)
        # linear_1: 0e block -> C + lmax*C
        lz = nv.make_layout(rows, [C + lmax * C] + [C] * lmax)
        W = torch.randn(C * (C + lmax * C) + lmax * C * C, generator=g).to(dev)
        w1off = [0] + [C * (C + lmax * C) + b * C * C for b in range(lmax)]
        run("readout.linear_1", lin, lz, W, w1off, [C ** -0.5] * (lmax + 1), x)
        # gate fused into linear_2 (C -> C/2)
        lo = nv.make_layout(rows, [C // 2] * (lmax + 1)); lg = nv.make_layout(rows, [C] * (lmax + 1))
        z = torch.randn(n, lz.stride, generator=g).to(dev)
        W2 = torch.randn((lmax + 1) * C * (C // 2), generator=g).to(dev)
        w2off = [b * C * (C // 2) for b in range(lmax + 1)]
        cst = 1.6791767923989418
        o = nv.gate_linear(z, lz, C, cst, W2, w2off, [C ** -0.5] * (lmax + 1), lo, n)
        torch.cuda.synchronize()
        zd = z.double().cpu()
        sil = lambda t: t / (1 + torch.exp(-t))  # noqa: E731
        ref = torch.zeros(n, lo.stride, dtype=torch.float64)
        for b in range(lmax + 1):
            r = rows[b]
            if b == 0:
                gb_ = cst * sil(zd[:, :C]).reshape(n, 1, C)
            else:
                gate = cst * sil(zd[:, C + C * (b - 1):C + C * b])
                gb_ = zd[:, lz.off[b]:lz.off[b] + r * C].reshape(n, r, C) * gate[:, None, :]
            ref[:, lo.off[b]:lo.off[b] + r * (C // 2)] = (C ** -0.5 * gb_ @ W2.double().cpu()[w2off[b]:w2off[b] + C * C // 2].reshape(C, C // 2)).reshape(n, -1)
        err = float((o.double().cpu() - ref).abs().max() / ref.abs().max())
        ms = timeit(lambda: nv.gate_linear(z, lz, C, cst, W2, w2off, [C ** -0.5] * (lmax + 1), lo, n), n=10)
        print(f"linear {'gate+linear_2':18s} N={n}: rel err {err:.2e}  {ms*1e3:7.1f} us  {(z.numel()+o.numel())*4/1e9/ms*1e3:6.0f} GB/s", flush=True)
        ok &= err < 4e-6
    print("LINEAR", "OK" if ok else "FAIL", "(EEG_LINEAR=%s)" % os.environ.get("EEG_LINEAR", "tc"))


def check_scw(N=15360, C=32, lmax=4):
    """Symmetric-contraction weight gradient: dT program and the fold onto the weights, timed separately; dT vs fp64."""
    import ctypes as ct
    from energy_equiv_lattice_gnn_b200 import tables as tb
    wmap = tb.sc_weight_map(lmax, 3)
    n_coef, D = wmap["n_coef"], (lmax + 1) ** 2
    g = torch.Generator().manual_seed(2)
    x = torch.randn(N, D, C, generator=g).to(dev)
    dout = torch.randn(N, D, C, generator=g).to(dev)
    L = nv.lib()
    Z = L.eeg_sc_grad_slices()
    dT = torch.empty(Z, n_coef, C, device=dev)
    t_ptr, t_row, t_val = (torch.tensor(wmap[k], device=dev) for k in ("t_ptr", "t_row", "t_val"))
    dW = torch.empty(wmap["n_wrows"], C, device=dev)
    st = lambda: ct.c_void_p(torch.cuda.current_stream().cuda_stream)  # noqa: E731
    f1 = lambda: L.eeg_sc_coef_grad(ct.c_int(lmax), ct.c_int(3), ct.c_int(C), ct.c_void_p(x.data_ptr()), ct.c_void_p(dout.data_ptr()),  # noqa: E731
                                    ct.c_int(n_coef), ct.c_int64(N), ct.c_void_p(dT.data_ptr()), st())
    f2 = lambda: L.eeg_sc_coefficients_bwd(ct.c_void_p(t_ptr.data_ptr()), ct.c_void_p(t_row.data_ptr()), ct.c_void_p(t_val.data_ptr()),  # noqa: E731
                                           ct.c_int(wmap["n_wrows"]), ct.c_int(n_coef), ct.c_void_p(dT.data_ptr()), ct.c_int(C),
                                           ct.c_void_p(dW.data_ptr()), st())
    assert f1() == 0 and f2() == 0
    torch.cuda.synchronize()
    tab = torch.tensor(tb.sc_slot_table(lmax, 3))
    n_chk = min(N, 512)
    xd, gd = x[:n_chk].double().cpu(), dout[:n_chk].double().cpu()
    one = torch.ones(n_chk, 1, C, dtype=torch.float64)
    xe = torch.cat([xd, one], 1)  # index -1 -> ones row
    ok_rows = tab[:, 0] >= 0
    M, i, j, t = (tab[:, k].clamp(min=-1) for k in range(4))
    ref = gd[:, M.clamp(min=0)] * xe[:, i] * xe[:, j] * xe[:, t]
    ref = (ref * ok_rows[None, :, None]).sum(0)
    x2 = x[:n_chk].contiguous(); d2 = dout[:n_chk].contiguous()
    dT2 = torch.empty_like(dT)
    L.eeg_sc_coef_grad(ct.c_int(lmax), ct.c_int(3), ct.c_int(C), ct.c_void_p(x2.data_ptr()), ct.c_void_p(d2.data_ptr()), ct.c_int(n_coef),
                       ct.c_int64(n_chk), ct.c_void_p(dT2.data_ptr()), st())
    got = dT2.double().sum(0).cpu()
    err = float((got - ref).abs().max() / ref.abs().max())
    print(f"scw lmax={lmax} N={N} C={C} n_coef={n_coef} n_wrows={wmap['n_wrows']} nnz={len(wmap['t_row'])}: dT rel err {err:.2e}  "
          f"coef_grad {timeit(f1, n=10)*1e3:.1f} us  coefficients_bwd {timeit(f2, n=10)*1e3:.1f} us", flush=True)


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "radial"
    if what == "linear":
        check_linear()
    if what == "sc":
        check_sc(int(sys.argv[2]) if len(sys.argv) > 2 else 30720, 32, int(sys.argv[3]) if len(sys.argv) > 3 else 4)
    if what == "scw":
        check_scw(int(sys.argv[2]) if len(sys.argv) > 2 else 15360, 32, int(sys.argv[3]) if len(sys.argv) > 3 else 4)
    if what == "sc_compare":
        import glob
        for lmax in (3, 4):
            fs = sorted(glob.glob(f"gpurun_out/sc_*_l{lmax}.pt"))
            if fs:
                ref = torch.load(fs[0])
                for f in fs[1:]:
                    print(f"{f} vs {fs[0]}: max abs diff {float((torch.load(f) - ref).abs().max()):.3e}")
    if what == "radial":
        if len(sys.argv) > 2:
            check_radial(int(sys.argv[2]), int(sys.argv[3]) if len(sys.argv) > 3 else 1344)
        else:
            ok = True
            for E, n_out, nh in ((128, 32, 1), (200, 160, 1), (1000, 1344, 1), (777, 736, 0), (4096, 2944, 2), (122880, 1344, 1), (122880, 160, 1)):
                ok &= check_radial(E, n_out, nh) < 3e-6
            print("RADIAL", "OK" if ok else "FAIL")
