"""GPU parity of the explicit backward: parameter gradients of the training loss (scripts/train_utils.py:45-64) from
the CUDA module's `_bwd` kernels vs torch autograd through the fp64 CPU oracle, same weights, same synthetic lattices.
Tolerance: 3e-5 of the largest entry of each gradient tensor (fp32 arithmetic with fixed-order reductions against fp64
autograd; measured worst case 2.3e-5 on 128 lattices, profiles/r02_gpu_tests_e.log)."""
import pytest
import torch

from energy_equiv_lattice_gnn_b200.data import synthetic_batch
from energy_equiv_lattice_gnn_b200.train import relative_mse_loss, random_psd_targets
from tests.util import batch_to, cuda_model_from, make_pair

pytestmark = pytest.mark.gpu
GTOL = 3e-5


def _grads(lmax_over, n_lat, cuda_device, positive="matrix_power_2"):
    batch, rmax = synthetic_batch(range(7000, 7000 + n_lat))
    oracle, params = make_pair(rmax, dtype=torch.float64, positive_function=positive, **lmax_over)
    target = random_psd_targets(n_lat, seed=5)
    ref_out = oracle(batch_to(batch, torch.float64))["stiffness"]
    ref_loss = relative_mse_loss(ref_out, target.double())
    ref_loss.backward()
    ref = {k: p.grad for k, p in oracle.named_parameters()}
    model = cuda_model_from(oracle, params, cuda_device).train()
    out = model(batch.to(cuda_device))["stiffness"]
    loss = relative_mse_loss(out, target.to(cuda_device))
    loss.backward()
    got = {k: p.grad for k, p in model.named_parameters()}
    return ref, got, float(ref_loss), float(loss)


@pytest.mark.parametrize("over,positive", [({}, "matrix_power_2"), ({}, "matrix_power_4"),
                                           (dict(lmax=3, hidden_irreps="32x0e+32x1o+32x2e+32x3o", readout_irreps="16x0e+16x1o+16x2e+16x3o"),
                                            "matrix_trunc_exp_2"),
                                           (dict(interaction_reduction="mean"), "matrix_power_2"),
                                           ({}, "matrix_exp"),
                                           (dict(lmax=3, hidden_irreps="128x0e+128x1o+128x2e+128x3o", readout_irreps="64x0e+64x1o+64x2e+64x3o"),
                                            "matrix_power_2")])
def test_parameter_gradients_match_oracle_autograd(cuda_device, over, positive):
    _check_grads(over, 4 if "128x0e" in over.get("hidden_irreps", "") else 6, cuda_device, positive)


def test_parameter_gradients_128_lattices(cuda_device):
    """a config-3 sized slice (BASELINE.json configs[2] trains on 512 lattices per GPU; 128 keeps the fp64 CPU autograd of the
    oracle to seconds): long reductions over nodes / edges in the weight gradients."""
    _check_grads({}, 128, cuda_device, "matrix_power_2")


def _check_grads(over, n_lat, cuda_device, positive):
    ref, got, ref_loss, loss = _grads(over, n_lat, cuda_device, positive)
    assert abs(loss - ref_loss) < 1e-5 * abs(ref_loss)
    assert set(got) == set(ref)
    worst = {}
    for k in ref:
        assert got[k] is not None, f"no gradient for {k}"
        g, r = got[k].double().cpu(), ref[k]
        assert g.shape == r.shape, k
        worst[k] = float((g - r).abs().max() / max(float(r.abs().max()), 1e-30))
    bad = {k: v for k, v in worst.items() if not v < GTOL}
    print(f"gradient parity [{positive}, {n_lat} lattices, {over}]: worst {max(worst.values()):.2e} ({max(worst, key=worst.get)})")
    assert not bad, bad
