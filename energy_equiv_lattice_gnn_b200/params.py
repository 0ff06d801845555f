"""Default hyper-parameters of the reference entry script (scripts/train_main.py:25-52) as an argparse.Namespace."""
from argparse import Namespace


def default_params(**over) -> Namespace:
    p = dict(lmax=4, hidden_irreps="32x0e+32x1o+32x2e+32x3o+32x4e", readout_irreps="16x0e+16x1o+16x2e+16x3o+16x4e",
             num_edge_bases=6, interaction_reduction="sum", interaction_bias=True, agg_norm_const=4.0, inter_MLP_dim=64,
             inter_MLP_layers=3, correlation=3, global_reduction="mean", message_passes=2,
             positive_function="matrix_power_2", max_edge_radius=0.03,
             rotate=False, batch_size=64, valid_batch_size=64, log_every_n_steps=25, optimizer="adamw",
             lr=1e-3, amsgrad=True, weight_decay=1e-8, beta1=0.9, epsilon=1e-8, num_workers=4)
    p.update(over)
    return Namespace(**p)
