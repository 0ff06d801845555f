#!/usr/bin/env python
"""Stall accounting of the tensor-core tensor-product kernel (csrc/eeg_tp_umma.cu, EEG_TP_PROF=1): runs the BASELINE
config-2 layer-1 shape once and prints, per role, the share of cycles spent in each wait site (median over CTAs).

    EEG_TP_PROF=1 python tools/tp_prof.py [lattices]"""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
os.environ.setdefault("EEG_TP_PROF", "1")
import numpy as np  # noqa: E402
import torch  # noqa: E402

from energy_equiv_lattice_gnn_b200 import native as nv, tables as tb  # noqa: E402

G = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
dev = torch.device("cuda:0")
lmax, Cc = 4, 32
plan = tb.tp_plan(lmax, lmax, lmax)
N, E = 30 * G, 120 * G
g = torch.Generator().manual_seed(0)
deg = torch.full((N,), 4)
row_ptr = torch.zeros(N + 1, dtype=torch.int32)
row_ptr[1:] = torch.cumsum(deg, 0).int()
snd = (torch.arange(E) // 120 * 30 + torch.randint(0, 30, (E,), generator=g)).int()
h = torch.randn(N, 25 * Cc, generator=g).to(dev)
Y = torch.randn(E, 28, generator=g).to(dev)
w = torch.randn(len(plan["paths"]), E, 32, generator=g).to(dev)
snd, row_ptr = snd.to(dev), row_ptr.to(dev)
for _ in range(3):
    M = nv.tp_scatter(lmax, lmax, Cc, h, Y, w, snd, row_ptr, 0.25, plan["rows_total"], N, umma=True)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(10):
    M = nv.tp_scatter(lmax, lmax, Cc, h, Y, w, snd, row_ptr, 0.25, plan["rows_total"], N, umma=True)
b.record()
torch.cuda.synchronize()
print(f"tp_scatter {G} lattices: {a.elapsed_time(b) / 10 * 1e3:.1f} us per launch")
buf = (C.c_ulonglong * (148 * 16))()
rc = nv.lib().eeg_tp_umma_prof_read(buf)
p = np.array(buf[:], dtype=np.float64).reshape(148, 16)
p = p[p[:, 10] > 0]
names = ["fill.wait_slot", "fill.total", "mma.wait_filled", "mma.total", "epi.wait_w", "epi.wait_d", "epi.flush", "epi.total",
         "load.wait_free", "load.total", "edges", "nodes"]
med = np.median(p, 0)
for i, n in enumerate(names):
    print(f"  {n:16s} {med[i]:12.0f}" + (f"   {med[i] / med[10]:8.1f} clk/edge" if i < 10 else ""))
