"""Serving paths on the GPU: CUDA-graph replay (`serve.GraphedPredictor`) and the stream-overlapped predictor must return
exactly what the eager forward returns."""
import pytest
import torch

from energy_equiv_lattice_gnn_b200.data import synthetic_batch
from energy_equiv_lattice_gnn_b200.serve import GraphedPredictor, StreamingPredictor
from tests.util import cuda_model_from, make_pair

pytestmark = pytest.mark.gpu


def test_graphed_predictor_equals_eager(cuda_device):
    batches = [synthetic_batch(range(8600 + 8 * i, 8608 + 8 * i))[0] for i in range(3)]
    other = synthetic_batch(range(8700, 8703), 12)[0]          # a second shape
    rmax = max(float(b.edge_attr.max()) for b in batches + [other])
    oracle, params = make_pair(rmax)
    model = cuda_model_from(oracle, params, cuda_device)
    gp = GraphedPredictor(model, max_graphs=2)
    for hb in batches + [other, batches[0]]:
        hb = hb.pin_memory()
        with torch.no_grad():
            want = model(hb.to(cuda_device))["stiffness"].cpu()
        got = gp(hb).clone()
        assert torch.equal(got, want)
    assert len(gp._graphs) == 2
    sp = StreamingPredictor(model)
    outs = [o.clone() for o in sp.run(b.pin_memory() for b in batches)]
    for o, hb in zip(outs, batches):
        with torch.no_grad():
            assert torch.equal(o, model(hb.to(cuda_device))["stiffness"].cpu())
