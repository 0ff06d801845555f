#!/usr/bin/env python
"""Time eeg_sym_contract_bwd_x / fwd / coef-grad alone at a training-sized node count (CUDA events)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from energy_equiv_lattice_gnn_b200 import native as nv, tables as tb
N = int(sys.argv[1]) if len(sys.argv) > 1 else 15360
dev = torch.device("cuda:0")
n_coef = tb.sc_plan(4, 3)["n_coef"]
g = torch.Generator().manual_seed(1)
T = (0.1 * torch.randn(32, n_coef, generator=g)).to(dev)
x = torch.randn(N, 25 * 32, generator=g).to(dev)
d = torch.randn(N, 25 * 32, generator=g).to(dev)
def timeit(fn, n=20):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n
dx = nv.sym_contract_bwd_x(4, 3, 32, x, d, T, N)
print(f"N={N}: sc_bwd_x {timeit(lambda: nv.sym_contract_bwd_x(4, 3, 32, x, d, T, N))*1e3:.1f} us  sc_fwd {timeit(lambda: nv.sym_contract(4, 3, 32, x, T, N))*1e3:.1f} us  checksum {float(dx.double().sum()):.9e}")
