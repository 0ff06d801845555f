#!/usr/bin/env python
"""Throughput of the data path either side of the model:
  * `.lat` text -> flat binary: line-by-line reader (catalogue.read_lat, the reference's approach) vs flatdata.read_lat_fast,
    and the whole convert_lat (parse + process_one + write), on a synthetic catalogue of N entries x 3 relative densities;
  * batch construction: host `data.collate` (Python loop, PyG-equivalent) vs `FlatDataset.collate` on the host (vectorised) and
    on the GPU (one eeg_collate_gather launch) for batches of 1024 lattices.
    python tools/flatdata_bench.py [entries] [out.json]"""
import json
import os
import sys
import tempfile
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np  # noqa: E402
import torch  # noqa: E402

from energy_equiv_lattice_gnn_b200 import catalogue as cat  # noqa: E402
from energy_equiv_lattice_gnn_b200.data import collate, synthetic_lattice  # noqa: E402
from energy_equiv_lattice_gnn_b200.flatdata import FlatDataset, convert_lat, read_lat_fast  # noqa: E402

n_entries = int(sys.argv[1]) if len(sys.argv) > 1 else 2000
out_path = sys.argv[2] if len(sys.argv) > 2 else None
rng = np.random.default_rng(0)
cells = []
for c in range(n_entries):
    n = int(rng.integers(4, 13))
    m = int(rng.integers(n, 2 * n + 1))
    adj = np.stack([rng.integers(0, n, m), rng.integers(0, n, m)], 1)
    adj[:n, 0] = np.arange(n)
    comp = {}
    for rd in (0.001, 0.003, 0.01):
        A = rng.standard_normal((6, 6))
        comp[rd] = (A @ A.T + 6 * np.eye(6)).round(6)
    cells.append(dict(name=f"cell_{c}", lattice_constants=[1.0, 0.9, 0.8, 90.0, 90.0, 90.0], compliance_tensors_M=comp,
                      nodal_positions=rng.random((n, 3)).round(5).tolist(), fundamental_edge_adjacency=adj.tolist(),
                      fundamental_tesselation_vecs=rng.integers(-1, 2, (m, 3)).astype(float).tolist()))
res = {"entries": n_entries, "host_cores": os.cpu_count()}
with tempfile.TemporaryDirectory() as d:
    lat = os.path.join(d, "cat.lat")
    cat.write_lat(lat, cells)
    res["lat_bytes"] = os.path.getsize(lat)
    t = time.perf_counter(); slow = cat.read_lat(lat); res["read_lat_entries_per_s"] = n_entries / (time.perf_counter() - t)
    t = time.perf_counter(); fast = read_lat_fast(lat); res["read_lat_fast_entries_per_s"] = n_entries / (time.perf_counter() - t)
    assert len(slow) == len(fast) == n_entries
    info = convert_lat(lat, os.path.join(d, "cat.eegflat"))
    res["convert_lat"] = info
    t = time.perf_counter(); ds = FlatDataset.load(os.path.join(d, "cat.eegflat")); res["load_flat_lattices_per_s"] = len(ds) / (time.perf_counter() - t)
# batch construction
data = [synthetic_lattice(5000 + i) for i in range(256)]
ds = FlatDataset.from_lattices(data)
g = torch.Generator().manual_seed(0)
idxs = [torch.randint(0, 256, (1024,), generator=g) for _ in range(8)]
t = time.perf_counter()
for idx in idxs[:3]:
    collate([data[i] for i in idx.tolist()])
res["collate_python_loop_lattices_per_s"] = 3 * 1024 / (time.perf_counter() - t)
t = time.perf_counter()
for idx in idxs:
    ds.collate(idx)
res["collate_flat_host_lattices_per_s"] = len(idxs) * 1024 / (time.perf_counter() - t)
if torch.cuda.is_available():
    dd = ds.to("cuda")
    for idx in idxs[:2]:
        dd.collate(idx)
    torch.cuda.synchronize()
    t = time.perf_counter()
    for _ in range(10):
        for idx in idxs:
            b = dd.collate(idx)
    torch.cuda.synchronize()
    res["collate_device_lattices_per_s"] = 10 * len(idxs) * 1024 / (time.perf_counter() - t)
print(json.dumps(res, indent=1))
if out_path:
    with open(out_path, "w") as f:
        json.dump(res, f, indent=1)
