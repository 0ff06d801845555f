#!/usr/bin/env python
"""Time eeg_sc_coefficients alone (weights -> Horner table), CUDA events, 50 launches."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from energy_equiv_lattice_gnn_b200 import native as nv, tables as tb
dev = torch.device("cuda:0")
w = tb.sc_weight_map(4, 3)
rp, col, val = (torch.tensor(w[k], device=dev) for k in ("row_ptr", "col", "val"))
W = torch.randn(w["n_wrows"], 32, device=dev)
for _ in range(3):
    T = nv.sc_coefficients(rp, col, val, w["n_coef"], W, 32)
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
a.record()
for _ in range(50):
    T = nv.sc_coefficients(rp, col, val, w["n_coef"], W, 32)
b.record()
torch.cuda.synchronize()
print(f"sc_coefficients: {a.elapsed_time(b) / 50 * 1e3:.1f} us per launch; dtypes {rp.dtype} {col.dtype} {val.dtype}")
ref = torch.zeros(w["n_coef"], 32, dtype=torch.float64)
rows = torch.repeat_interleave(torch.arange(w["n_coef"]), torch.tensor(w["row_ptr"][1:]) - torch.tensor(w["row_ptr"][:-1]))
ref.index_add_(0, rows, torch.tensor(w["val"]).double()[:, None] * W.cpu().double()[torch.tensor(w["col"]).long()])
print("max err", float((T.t().cpu().double() - ref).abs().max()))
