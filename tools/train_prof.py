#!/usr/bin/env python
"""Where a training step's time goes: device time per tagged C-ABI call, wall time per step, and the host-side gap.
    python tools/train_prof.py [lattices]"""
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from energy_equiv_lattice_gnn_b200 import EnergyEquivGNN, native  # noqa: E402
from energy_equiv_lattice_gnn_b200.data import synthetic_batch  # noqa: E402
from energy_equiv_lattice_gnn_b200.params import default_params  # noqa: E402
from energy_equiv_lattice_gnn_b200.train import Trainer, random_psd_targets  # noqa: E402

G = int(sys.argv[1]) if len(sys.argv) > 1 else 512
dev = torch.device("cuda:0")
b, rmax = synthetic_batch(range(10000, 10000 + G), 30)
b = b.to(dev)
b.stiffness = random_psd_targets(G, seed=0).to(dev)
torch.manual_seed(0)
model = EnergyEquivGNN(default_params(max_edge_radius=rmax)).to(dev)
tr = Trainer(model)
for _ in range(3):
    tr.training_step(b)
torch.cuda.synchronize()
t0 = time.perf_counter()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
K = 10
for _ in range(K):
    tr.training_step(b)
t_host = time.perf_counter() - t0          # time to ENQUEUE K steps
e1.record()
torch.cuda.synchronize()
t_wall = time.perf_counter() - t0
print(f"{G} lattices: device {e0.elapsed_time(e1) / K:.3f} ms/step, wall {t_wall / K * 1e3:.3f} ms/step, host enqueue {t_host / K * 1e3:.3f} ms/step")
native.start_profile()
tr.training_step(b)
raw = native.stop_profile()
prof = {k: sum(v) for k, v in raw.items()}
tot = sum(prof.values())
print(f"sum of tagged device time: {tot:.3f} ms over {len(prof)} tags")
for k, v in sorted(prof.items(), key=lambda kv: -kv[1]):
    calls = " ".join(f"{x:.3f}" for x in raw[k]) if len(raw[k]) > 1 else ""
    print(f"  {k:36s} {v:8.4f}   {calls}")
