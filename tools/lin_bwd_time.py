#!/usr/bin/env python
"""Times and checks the dense GEMMs of the backward pass on the default workload's shapes (512 lattices):
  dM   = dmsg W^T            block_linear  K = 32  -> 160..320 per irrep block   (interaction.linear, layer 1)
  dWi  = M^T dmsg            linear_wgrad  [N x 7360] x [N x 800]
  ds   = dw W2^T             block_linear  K = 1344 -> 64                        (radial MLP, last layer)
  dW2  = dw^T act            linear_wgrad  [E x 1344] x [E x 64]
against torch fp64.  EEG_LINEAR=cuda routes block_linear to the CUDA-core GEMM (A/B).
    python tools/lin_bwd_time.py [lattices]"""
import math
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from energy_equiv_lattice_gnn_b200 import EnergyEquivGNN, native as nv  # noqa: E402
from energy_equiv_lattice_gnn_b200.params import default_params  # noqa: E402

G = int(sys.argv[1]) if len(sys.argv) > 1 else 512
dev = torch.device("cuda:0")
N, E = 30 * G, 120 * G
model = EnergyEquivGNN(default_params(max_edge_radius=1.0)).to(dev)
lay_M, lay_h = model._lay_M[1], model._lay_h
C_ = 32
n_l = [lay_M.mul[b] // C_ for b in range(lay_M.n_blocks)]
woff = np.concatenate(([0], np.cumsum([n * C_ * C_ for n in n_l])))[:-1].tolist()
iscale = [1.0 / math.sqrt(n * C_) for n in n_l]
g = torch.Generator(device=dev).manual_seed(0)
Wi = torch.randn(sum(n * C_ * C_ for n in n_l), device=dev, generator=g)
dmsg = torch.randn(N, lay_h.stride, device=dev, generator=g)
M = torch.randn(N, lay_M.stride, device=dev, generator=g)


def timeit(fn, reps=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


def rel(a, b):
    return float((a.double() - b).abs().max() / b.abs().max())


# ---- dM
dM = nv.block_linear(dmsg, lay_h, Wi, woff, iscale, lay_M, N, transB=True)
ns = min(N, 2048)
err = 0.0
for b in range(lay_M.n_blocks):
    r, k, no = lay_h.rows[b], lay_h.mul[b], lay_M.mul[b]
    Wb = Wi[woff[b]:woff[b] + no * k].view(no, k).double()
    xin = dmsg[:ns, lay_h.off[b]:lay_h.off[b] + r * k].reshape(ns, r, k).double()
    ref = iscale[b] * torch.einsum("nrk,ok->nro", xin, Wb)
    got = dM[:ns, lay_M.off[b]:lay_M.off[b] + r * no].reshape(ns, r, no)
    err = max(err, rel(got, ref))
print(f"dM   block_linear K=32 -> {[lay_M.mul[b] for b in range(lay_M.n_blocks)]}: {timeit(lambda: nv.block_linear(dmsg, lay_h, Wi, woff, iscale, lay_M, N, transB=True)):.4f} ms, rel err {err:.2e}")

# ---- dWi
gW, gb = nv.linear_wgrad(M, lay_M, dmsg, lay_h, woff, iscale, Wi.numel(), N, want_bias=C_)
err = 0.0
for b in range(lay_M.n_blocks):
    r, k, no = lay_M.rows[b], lay_M.mul[b], lay_h.mul[b]
    xin = M[:, lay_M.off[b]:lay_M.off[b] + r * k].reshape(N * r, k).double()
    dy = dmsg[:, lay_h.off[b]:lay_h.off[b] + r * no].reshape(N * r, no).double()
    ref = iscale[b] * xin.t() @ dy
    err = max(err, rel(gW[woff[b]:woff[b] + k * no].view(k, no), ref))
print(f"dWi  linear_wgrad [N x 7360]^T [N x 800]: {timeit(lambda: nv.linear_wgrad(M, lay_M, dmsg, lay_h, woff, iscale, Wi.numel(), N, want_bias=C_)):.4f} ms, rel err {err:.2e}")
del M, dM

# ---- radial MLP last layer
P = 1344
lay_w, lay_a = nv.make_layout([1], [P]), nv.make_layout([1], [64])
W2 = torch.randn(P * 64, device=dev, generator=g)     # torch.nn.Linear(64, 1344).weight: [1344, 64]
dw = torch.randn(E, P, device=dev, generator=g)
act = torch.randn(E, 64, device=dev, generator=g)
ds = nv.block_linear(dw, lay_w, W2, [0], [1.0], lay_a, E)
ref = dw[:4096].double() @ W2.view(P, 64).double()
print(f"ds   block_linear K=1344 -> 64: {timeit(lambda: nv.block_linear(dw, lay_w, W2, [0], [1.0], lay_a, E)):.4f} ms, rel err {rel(ds[:4096], ref):.2e}")
gW2, _ = nv.linear_wgrad(dw, lay_w, act, lay_a, [0], [1.0], 64 * P, E)
ref = dw.double().t() @ act.double()
print(f"dW2  linear_wgrad [E x 1344]^T [E x 64]: {timeit(lambda: nv.linear_wgrad(dw, lay_w, act, lay_a, [0], [1.0], 64 * P, E)):.4f} ms, rel err {rel(gW2.view(P, 64), ref):.2e}")

if os.environ.get("EEG_WGRAD_PROF"):
    import ctypes as C
    names = ["xform wait data", "xform wait free operand stage", "xform work", "xform drain", "mma wait operands", "mma issue", "mma wait drain", "kernel"]
    for label, fn in (("dW2", lambda: nv.linear_wgrad(dw, lay_w, act, lay_a, [0], [1.0], 64 * P, E)),):
        torch.cuda.synchronize()
        buf = (C.c_ulonglong * 8)()
        nv.lib().eeg_wgrad_umma_prof_read(buf)
        fn()
        torch.cuda.synchronize()
        nv.lib().eeg_wgrad_umma_prof_read(buf)
        print(label, "CTA 0 cycles:", {n: int(v) for n, v in zip(names, buf)})
