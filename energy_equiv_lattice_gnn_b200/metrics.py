"""Validation / test metrics of the reference's Lightning wrapper, computed on device by `eeg_validation_metrics`
(SURVEY.md section 8(f) rank 3).

    validation_step(model, batch)  -> {'val_loss', 'val_stiff_dir_loss'}      scripts/train_utils.py:66-89
    obtain_errors(results, tag)    -> dict of per-lattice arrays              scripts/train_utils.py:149-172
    aggr_errors(errors)            -> flat dict of aggregated numbers         scripts/train_utils.py:174-202

`obtain_errors` returns a plain dict of numpy arrays with the reference DataFrame's column names (pandas is not needed on
the path); `aggr_errors` accepts that dict (or several concatenated with `concat_errors`).  CUDA only.
"""
from __future__ import annotations

from typing import Dict, Optional, Sequence

import numpy as np
import torch

from . import native

N_DIRECTIONS = 250  # train_utils.py:67,155


def random_directions(n: int = N_DIRECTIONS, generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """Unit vectors drawn like train_utils.py:67-68 / :155-156 (normalised CPU `torch.randn(n, 3)`)."""
    d = torch.randn(n, 3, generator=generator)
    return d / d.norm(dim=1, keepdim=True)


def per_lattice_metrics(pred: torch.Tensor, target: torch.Tensor, directions: torch.Tensor, eigenvalues: bool = True):
    """pred, target: CUDA Mandel [G,6,6].  -> (per_lattice [G,4] = mse, l1, dir_loss, mean_stiffness; eig_pred, eig_target [G,6])."""
    if not pred.is_cuda:
        raise RuntimeError("metrics need CUDA tensors (no CPU fallback)")
    G = pred.shape[0]
    dev = pred.device
    pred = pred.to(torch.float32).contiguous()
    target = target.to(device=dev, dtype=torch.float32).contiguous()
    directions = directions.to(device=dev, dtype=torch.float32).contiguous()
    out = torch.empty(G, 4, dtype=torch.float32, device=dev)
    ep = torch.empty(G, 6, dtype=torch.float32, device=dev) if eigenvalues else None
    et = torch.empty(G, 6, dtype=torch.float32, device=dev) if eigenvalues else None
    native.validation_metrics(pred, target, directions, out, ep, et)
    return out, ep, et


def validation_step(model, batch, directions: Optional[torch.Tensor] = None, process_group=None,
                    sync_dist: bool = True) -> Dict[str, torch.Tensor]:
    """`LightningWrappedModel.validation_step` (train_utils.py:66-89): MSE over all entries and the L1 error of the
    directional stiffness over 250 random directions.  Returns 0-dim CUDA tensors (no host sync).
    With an initialised process group and sync_dist (the reference logs with `sync_dist=True`, train_utils.py:87-88) the two
    scalars are averaged over the ranks with ONE all-reduce of a 2-float tensor."""
    if directions is None:
        directions = random_directions()
    with torch.no_grad():
        pred = model(batch)["stiffness"]
        m, _, _ = per_lattice_metrics(pred, batch["stiffness"], directions, eigenvalues=False)
    vals = torch.stack([m[:, 0].mean(), m[:, 2].mean()])
    vals = sync_validation_scalars(vals, process_group) if sync_dist else vals
    return {"val_loss": vals[0], "val_stiff_dir_loss": vals[1]}


def sync_validation_scalars(vals: torch.Tensor, process_group=None) -> torch.Tensor:
    """Lightning `self.log(..., sync_dist=True)`: mean over ranks (device-agnostic: used with gloo in the CPU tests)."""
    if torch.distributed.is_available() and torch.distributed.is_initialized():
        world = torch.distributed.get_world_size(process_group)
        if world > 1:
            vals = vals.clone()
            torch.distributed.all_reduce(vals, op=torch.distributed.ReduceOp.SUM, group=process_group)
            vals = vals / world
    return vals


def obtain_errors(results: Sequence, tag: str, directions: Optional[torch.Tensor] = None) -> Dict[str, np.ndarray]:
    """`obtain_errors` (train_utils.py:149-172).  results: [(prediction dict, batch)] as `predict_step` returns them
    (:91-93), stiffness tensors on the GPU.  The x10 of :151-152 is applied here as well."""
    target = 10 * torch.cat([x[1]["stiffness"] for x in results])
    prediction = 10 * torch.cat([x[0]["stiffness"] for x in results])
    names = np.concatenate([np.atleast_1d(np.asarray(getattr(x[1], "name", [""] * x[1]["stiffness"].shape[0]))) for x in results])
    rel = [getattr(x[1], "rel_dens", None) for x in results]
    rel_dens = (np.concatenate([np.atleast_1d(torch.as_tensor(r).cpu().numpy()) for r in rel]) if all(r is not None for r in rel)
                else np.full(target.shape[0], np.nan))
    if directions is None:
        directions = random_directions()
    m, ep, et = per_lattice_metrics(prediction, target, directions)
    m = m.cpu().numpy()
    return {"name": names, "rel_dens": rel_dens, "mean_stiffness": m[:, 3], "loss": m[:, 1], "mseloss": m[:, 0], "dir_loss": m[:, 2],
            "tag": np.full(target.shape[0], tag), "target_eig": et.cpu().numpy(), "predicted_eig": ep.cpu().numpy()}


def concat_errors(parts: Sequence[Dict[str, np.ndarray]]) -> Dict[str, np.ndarray]:
    return {k: np.concatenate([p[k] for p in parts], 0) for k in parts[0]}


def aggr_errors(err: Dict[str, np.ndarray]) -> Dict[str, float]:
    """`aggr_errors` (train_utils.py:174-202): per-tag means of the (relative) losses, minimum eigenvalues and the share of
    predictions with a negative eigenvalue; same keys (`<column>_<tag>`)."""
    params: Dict[str, float] = {}
    if np.isnan(err["loss"]).any():
        return params
    ms = err["mean_stiffness"]
    cols = {"loss": err["loss"], "rel_loss": err["loss"] / ms, "mseloss": err["mseloss"], "mse_rel_loss": np.sqrt(err["mseloss"]) / ms,
            "dir_loss": err["dir_loss"], "rel_dir_loss": err["dir_loss"] / ms}
    pv, tv = np.prod(err["predicted_eig"], axis=1), np.prod(err["target_eig"], axis=1)
    cols["eig_loss"] = np.abs(pv - tv)
    cols["rel_eig_loss"] = cols["eig_loss"] / tv
    min_pred, min_tgt = err["predicted_eig"].min(axis=1), err["target_eig"].min(axis=1)
    for tag in np.unique(err["tag"]):
        sel = err["tag"] == tag
        for k, v in cols.items():
            params[f"{k}_{tag}"] = float(np.mean(v[sel]))
        params[f"min_pred_eig_{tag}"] = float(min_pred[sel].min())
        params[f"min_target_eig_{tag}"] = float(min_tgt[sel].min())
        params[f"prop_eig_negative_{tag}"] = float((min_pred[sel] < 0).sum() / sel.sum())
    return params
