"""Training step of the reference harness (scripts/train_utils.py:38-64, scripts/train_main.py:89-100) on the B200 path:
relative-MSE loss, explicit backward (autograd.py), one flat gradient all-reduce over NCCL/NVLink, global-norm clipping
and a fused AdamW(amsgrad) kernel on one flat fp32 parameter buffer.  Lightning itself is out of scope; `Trainer` is the
60-line stand-in SURVEY.md section 8(b) allows.
"""
from __future__ import annotations

from typing import Optional

import torch

from . import native as nv


def relative_mse_loss(pred: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
    """scripts/train_utils.py:50-60: 100 * mean_g( mean_ij (pred - target)^2 / mean_ij target^2 )."""
    mean_stiffness = target.pow(2).mean(dim=(1, 2))
    stiffness_loss = torch.nn.functional.mse_loss(pred, target, reduction="none").mean(dim=(1, 2))
    return 100 * (stiffness_loss / mean_stiffness).mean()


def random_psd_targets(n: int, seed: int = 0) -> torch.Tensor:
    """A = randn(6,6); C = ((A + A^T)/2)^2 / 6 (seeded)."""
    g = torch.Generator().manual_seed(seed)
    A = torch.randn(n, 6, 6, generator=g)
    S = (A + A.transpose(1, 2)) / 2
    return S @ S / 6


def sync_gradients(flat_grad: torch.Tensor, process_group=None) -> int:
    """Data-parallel gradient exchange: ONE all-reduce (sum) of the flat gradient buffer (NCCL over NVLink on the GPU
    path, gloo in the CPU tests).  Returns the world size; the 1/world of DDP's gradient mean is folded into the
    optimizer kernel (`FlatAdamW.step(world)`).  Equal-sized shards make this identical to the single-process gradient
    of the global-mean loss (Lightning DDP + per-rank mean loss, scripts/train_utils.py:57-60)."""
    if not (torch.distributed.is_available() and torch.distributed.is_initialized()):
        return 1
    world = torch.distributed.get_world_size(process_group)
    if world > 1:
        torch.distributed.all_reduce(flat_grad, op=torch.distributed.ReduceOp.SUM, group=process_group)
    return world


class FlatAdamW:
    """torch.optim.AdamW(lr, betas=(beta1, 0.999), eps, amsgrad, weight_decay) semantics (scripts/train_utils.py:38-44) +
    Lightning's gradient_clip_val (clip_grad_norm_, scripts/train_main.py:91), fused over ONE flat fp32 buffer.

    The model's parameters become views into `self.flat`; gradients are gathered into `self.grad` (also flat), which is
    what the data-parallel all-reduce sends: one 0.89 MB message per step for the default model."""

    def __init__(self, model: torch.nn.Module, lr=1e-3, beta1=0.9, beta2=0.999, eps=1e-8, weight_decay=1e-2, amsgrad=True,
                 clip_norm: Optional[float] = 10.0):
        self.model = model
        self.params = [p for p in model.parameters() if p.requires_grad]
        dev = self.params[0].device
        if dev.type != "cuda":
            raise RuntimeError("FlatAdamW runs on the CUDA path only")
        n = sum(p.numel() for p in self.params)
        self.flat = torch.empty(n, dtype=torch.float32, device=dev)
        self.grad = torch.zeros(n, dtype=torch.float32, device=dev)
        off = 0
        self.slices = []
        for p in self.params:
            k = p.numel()
            self.flat[off:off + k].copy_(p.detach().reshape(-1))
            p.data = self.flat[off:off + k].view_as(p)
            p.grad = self.grad[off:off + k].view_as(p)
            self.slices.append((off, k))
            off += k
        self.m = torch.zeros_like(self.flat)
        self.v = torch.zeros_like(self.flat)
        self.vmax = torch.zeros_like(self.flat) if amsgrad else None
        self.gnorm_sq = torch.zeros(1, dtype=torch.float32, device=dev)
        self.lr, self.b1, self.b2, self.eps, self.wd, self.clip = lr, beta1, beta2, eps, weight_decay, clip_norm
        self.t = 0

    def zero_grad(self):
        self.grad.zero_()

    def gather_grads(self):
        """Copy gradients that autograd left in separate tensors into the flat buffer (views are already in place)."""
        for p, (off, k) in zip(self.params, self.slices):
            if p.grad is not None and p.grad.data_ptr() != self.grad[off:off + k].data_ptr():
                self.grad[off:off + k].copy_(p.grad.reshape(-1))
                p.grad = self.grad[off:off + k].view_as(p)

    def step(self, world_size: int = 1):
        """Gradients in `self.grad` hold the SUM over ranks when world_size > 1 (all-reduce sum): scaled by 1/world here."""
        self.t += 1
        nv.set_tag('optimizer')
        scale = 1.0 / world_size
        gn = None
        if self.clip is not None and self.clip > 0:
            self.gnorm_sq.zero_()
            nv.sumsq(self.grad, self.gnorm_sq)
            gn = self.gnorm_sq
        nv.adamw_step(self.flat, self.grad, self.m, self.v, self.vmax, self.lr, self.b1, self.b2, self.eps, self.wd, self.t, gn,
                      self.clip or 0.0, scale)
        # the kernel rewrote the parameters behind torch's version counters: drop the model's weight-derived tables
        # (SC Horner coefficients, radial tcgen05 operand image) here, so FlatAdamW is safe to use without Trainer
        bump = getattr(self.model, "_cache_bump", None)
        if bump is not None:
            bump()
        return self


class Trainer:
    """training_step of scripts/train_utils.py:45-64 + optimizer step; DDP = one flat all-reduce (sum) per optimizer step.

    accumulate_grad_batches = Lightning's Trainer(accumulate_grad_batches=k) (scripts/train_main.py:91, k = 4 there): the
    loss of each micro-batch is scaled by 1/k before the backward, gradients add up in the flat buffer, and the all-reduce,
    the clip and the AdamW step run after every k-th micro-batch.
    global_mean = False reproduces the reference exactly (per-rank mean loss, DDP gradient MEAN over ranks: identical to the
    global-batch gradient only for equal-sized shards); True weights each rank by its graph count, so edge-balanced shards
    of unequal size (sharding.shard_batch) still give the gradient of the global-mean loss."""

    def __init__(self, model, lr=1e-3, beta1=0.9, eps=1e-8, weight_decay=1e-8, amsgrad=True, clip_norm=10.0, process_group=None,
                 accumulate_grad_batches: int = 1, global_mean: bool = False, overlap_allreduce: bool = True):
        self.model = model
        self.opt = FlatAdamW(model, lr=lr, beta1=beta1, eps=eps, weight_decay=weight_decay, amsgrad=amsgrad, clip_norm=clip_norm)
        self.pg = process_group
        self.accumulate = max(int(accumulate_grad_batches), 1)
        self.global_mean = bool(global_mean)
        self._micro = 0
        self._count = None   # graphs accumulated since the last optimizer step (global_mean only)
        # Overlap of the gradient all-reduce with the backward: the parameters are laid out [node embedding, layer 0, layer 1 ...,
        # readout, final linear]; everything from layer 1 on (79 % of the default model's gradient) is final when the backward
        # enters layer 0, so that tail of the flat buffer is all-reduced asynchronously under the layer-0 backward and only
        # the head is reduced after it.
        self.overlap_allreduce = bool(overlap_allreduce)
        names = [n for n, p in model.named_parameters() if p.requires_grad]
        head = [n.startswith("node_ft_embedding.") or n.startswith("stiffness_head.layers.0.") for n in names]
        self._tail_idx = head.index(False) if False in head else len(head)
        if any(head[self._tail_idx:]) or self._tail_idx == 0:
            self._tail_idx = len(head)          # unexpected parameter order: no overlap
        self._tail_off = self.opt.slices[self._tail_idx][0] if self._tail_idx < len(head) else self.opt.grad.numel()
        self._pending = None

    def _distributed(self) -> bool:
        return torch.distributed.is_available() and torch.distributed.is_initialized() and \
            torch.distributed.get_world_size(self.pg) > 1

    def _early_bucket(self, grads) -> None:
        """Called by the explicit backward (autograd.backward_impl) once every gradient outside layer 0 / the node embedding is
        final: add them into the flat buffer (what autograd's accumulation would do after the backward), take them out of
        `grads` so that they are not added twice, and start the all-reduce of that part of the buffer."""
        for p, (off, k) in zip(self.opt.params[self._tail_idx:], self.opt.slices[self._tail_idx:]):
            g = grads.pop(id(p), None)
            if g is not None:
                self.opt.grad[off:off + k].add_(g.reshape(-1))
        self._pending = torch.distributed.all_reduce(self.opt.grad[self._tail_off:], op=torch.distributed.ReduceOp.SUM, group=self.pg,
                                                     async_op=True)

    def training_step(self, batch) -> torch.Tensor:
        self.model.train()
        if self._micro == 0:
            self.opt.zero_grad()
        out = self.model(batch)["stiffness"]
        loss = relative_mse_loss(out, batch.stiffness)
        n_local = float(out.shape[0])
        # global_mean: back-propagate the SUM over this rank's lattices; divided by the global count at the step
        scaled = loss * (n_local if self.global_mean else 1.0 / self.accumulate)
        early = (self.overlap_allreduce and not self.global_mean and self._micro == self.accumulate - 1 and self._distributed()
                 and self._tail_off < self.opt.grad.numel())
        if early:
            self.model._grads_ready_hook = self._early_bucket
        try:
            scaled.backward()
        finally:
            self.model.__dict__.pop("_grads_ready_hook", None)
        self.opt.gather_grads()
        if self.global_mean:
            c = torch.tensor([n_local], dtype=torch.float64, device=out.device)
            self._count = c if self._count is None else self._count + c
        self._micro += 1
        if self._micro < self.accumulate:
            return loss.detach()
        self._micro = 0
        if self._pending is not None:           # the tail is already in flight: reduce the head, then join
            world = sync_gradients(self.opt.grad[:self._tail_off], self.pg)
            self._pending.wait()
            self._pending = None
        else:
            world = sync_gradients(self.opt.grad, self.pg)
        if self.global_mean:
            if self._distributed():
                torch.distributed.all_reduce(self._count, op=torch.distributed.ReduceOp.SUM, group=self.pg)
            self.opt.grad.mul_(float(world) / float(self._count.item()))   # step() divides by world again
            self._count = None
        self.opt.step(world)
        return loss.detach()
