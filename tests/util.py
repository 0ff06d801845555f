"""Shared helpers for the parity tests (oracle <-> CUDA module)."""
import copy

import torch

from oracle.model import EnergyEquivGNN as OracleGNN
from oracle.model import default_params


def make_pair(rmax, seed=0, dtype=torch.float32, **over):
    """(oracle module [dtype, CPU], params) with weights from torch.manual_seed(seed)."""
    params = default_params(max_edge_radius=rmax, **over)
    old = torch.get_default_dtype()
    torch.set_default_dtype(dtype)
    try:
        torch.manual_seed(seed)
        oracle = OracleGNN(params)
    finally:
        torch.set_default_dtype(old)
    return oracle, params


def cuda_model_from(oracle, params, device):
    from energy_equiv_lattice_gnn_b200 import EnergyEquivGNN
    m = EnergyEquivGNN(params)
    sd = {k: v.to(torch.float32) if v.is_floating_point() else v for k, v in oracle.state_dict().items() if "U_matrix" not in k}
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not missing, missing
    assert not unexpected, unexpected
    return m.to(device).eval()


def batch_to(batch, dtype):
    b = copy.copy(batch)
    for k in ("positions", "shifts", "edge_attr", "node_attrs"):
        setattr(b, k, getattr(batch, k).to(dtype))
    return b
