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


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "radial"
    if what == "radial":
        if len(sys.argv) > 2:
            check_radial(int(sys.argv[2]), int(sys.argv[3]) if len(sys.argv) > 3 else 1344)
        else:
            ok = True
            for E, n_out, nh in ((128, 32, 1), (200, 160, 1), (1000, 1344, 1), (777, 736, 0), (4096, 2944, 2), (122880, 1344, 1), (122880, 160, 1)):
                ok &= check_radial(E, n_out, nh) < 3e-6
            print("RADIAL", "OK" if ok else "FAIL")
