"""The rows either side of the path chained on real data: reference-processed sample_cat.lat lattices
-> collate -> upload -> RotateLat on device -> model -> validation metrics / loss, against the CPU oracle fed with a batch
rotated by the CPU restatement."""
import os

import pytest
import torch

from energy_equiv_lattice_gnn_b200 import metrics
from energy_equiv_lattice_gnn_b200.data import LatticeData, RotateLat, collate, rand_rotations
from energy_equiv_lattice_gnn_b200.train import relative_mse_loss
from oracle.metrics import per_lattice_metrics
from oracle.rotate import rotate_batch
from tests.util import batch_to, cuda_model_from, make_pair

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(__file__), "golden", "ref_sample_cat.pt")


def test_real_lattices_rotate_predict_validate(cuda_device):
    cat = torch.load(GOLD, weights_only=False)
    # stiffness of the catalogue is ~1e-4 .. 1e-2 of the base material; the reference trains on multiples of it
    # (load_datasets' multiplier, train_utils.py:150): use x100 so the targets are O(1) like the model output
    lats = []
    for d in cat:
        l = LatticeData(positions=d["positions"], node_attrs=d["node_attrs"], edge_index=d["edge_index"], shifts=d["shifts"],
                        edge_attr=d["edge_attr"], stiffness=(100 * d["stiffness_cart4"]).float(), name=d["name"], rel_dens=d["rel_dens"])
        lats.append(l)
    batch = collate(lats)
    G = batch.num_graphs
    rmax = float(batch.edge_attr.max())
    g = torch.Generator().manual_seed(4)
    Q = rand_rotations(G, g)
    dev_batch = RotateLat()(batch.to(cuda_device), Q)

    # CPU side: rotate with the restatement, run the fp64 oracle
    pos, sh, C, _ = rotate_batch(Q.double(), batch.positions.double(), batch.shifts.double(), batch.edge_index, batch.batch,
                                 batch.stiffness.double())
    cpu_batch = batch_to(batch, torch.float64)
    cpu_batch.positions, cpu_batch.shifts = pos, sh
    oracle, params = make_pair(rmax, dtype=torch.float64)
    with torch.no_grad():
        ref = oracle(cpu_batch)["stiffness"]
    model = cuda_model_from(oracle, params, cuda_device)
    with torch.no_grad():
        out = model(dev_batch)["stiffness"]
    scale = float(ref.abs().max())
    assert float((out.cpu().double() - ref).abs().max()) / scale < 2e-5   # fp32 rotation + fp32 forward vs fp64 everything

    d = metrics.random_directions(250, g)
    m, ep, et = metrics.per_lattice_metrics(out, dev_batch.stiffness, d)
    want = per_lattice_metrics(ref, C, d)
    for i, k in enumerate(("mseloss", "loss", "dir_loss", "mean_stiffness")):
        assert float((m[:, i].cpu().double() - want[k]).abs().max() / want[k].abs().max()) < 1e-4, k
    assert float((et.cpu().double() - want["target_eig"]).abs().max() / want["target_eig"].abs().max()) < 1e-5
    loss = relative_mse_loss(out, dev_batch.stiffness)
    ref_loss = relative_mse_loss(ref, C)
    assert abs(float(loss) - float(ref_loss)) < 1e-4 * abs(float(ref_loss))
