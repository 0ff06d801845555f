"""Host-side streaming inference: the counterpart of `LightningWrappedModel.predict_step` driven by a DataLoader
(scripts/train_utils.py:91-99, scripts/train_main.py:120-135 of the reference), where every batch arrives in pinned host
memory and the predicted stiffness goes back to the host.

`StreamingPredictor` overlaps the three legs of a step on separate CUDA streams:
    copy stream   : batch i+1 host -> device          (cudaMemcpyAsync from pinned memory)
    compute stream: forward of batch i                (the C-ABI kernels, on torch's current stream)
    copy stream   : stiffness of batch i-1 device -> pinned host ring
Every step still moves its own inputs and outputs; only the waiting is hidden.  No CPU compute path exists here either."""
from __future__ import annotations

from typing import Iterable, Iterator, Optional

import torch


class StreamingPredictor:
    def __init__(self, model: torch.nn.Module, depth: int = 2):
        p = next(model.parameters())
        if p.device.type != "cuda":
            raise RuntimeError("StreamingPredictor needs the model on a CUDA device (there is no CPU path)")
        self.model = model
        self.device = p.device
        self.depth = max(2, int(depth))
        self.copy_stream = torch.cuda.Stream(device=self.device)
        self._host_out: list = [None] * self.depth
        self.last_done: Optional[torch.cuda.Event] = None   # completion of the most recent device -> host copy

    def _upload(self, hb, compute):
        with torch.cuda.stream(self.copy_stream):
            db = hb.to(self.device, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
        for v in db.__dict__.values():
            if torch.is_tensor(v):
                v.record_stream(compute)   # allocated on the copy stream, consumed on the compute stream
        return db, ev

    @torch.no_grad()
    def run(self, host_batches: Iterable) -> Iterator[torch.Tensor]:
        """Yields, in order, the [G,6,6] stiffness of every batch as a pinned host tensor.  The tensor is a slot of a ring:
        it stays valid until the generator is advanced again (copy it to keep it)."""
        compute = torch.cuda.current_stream(self.device)
        it = iter(host_batches)
        try:
            nxt = self._upload(next(it), compute)
        except StopIteration:
            return
        pending = []   # (slot, event) of device -> host copies in flight
        i = 0
        while nxt is not None:
            db, ev_in = nxt
            try:
                nxt = self._upload(next(it), compute)   # overlaps with the forward below
            except StopIteration:
                nxt = None
            compute.wait_event(ev_in)
            out = self.model(db)["stiffness"]
            ev_out = torch.cuda.Event()
            ev_out.record(compute)
            slot = i % self.depth
            if len(pending) == self.depth:       # the ring slot about to be reused must have been handed out
                s0, e0 = pending.pop(0)
                e0.synchronize()
                yield self._host_out[s0]
            buf = self._host_out[slot]
            if buf is None or buf.shape != out.shape:
                buf = torch.empty(out.shape, dtype=out.dtype, pin_memory=True)
                self._host_out[slot] = buf
            with torch.cuda.stream(self.copy_stream):
                self.copy_stream.wait_event(ev_out)
                buf.copy_(out, non_blocking=True)
                ev_done = torch.cuda.Event()
                ev_done.record(self.copy_stream)
            out.record_stream(self.copy_stream)
            self.last_done = ev_done
            pending.append((slot, ev_done))
            i += 1
        for s0, e0 in pending:
            e0.synchronize()
            yield self._host_out[s0]


class GraphedPredictor:
    """Small-batch inference through CUDA graphs (the reference trains and predicts at batch 64, scripts/train_main.py:42,
    where the ~18 C-ABI launches of a forward are launch bound).

    One graph per batch shape (N nodes, E edges, G graphs): the first call with a new shape runs the forward eagerly twice
    on a side stream (lazy shared-memory opt-ins, weight-derived tables) and then captures it on static device buffers; later
    calls copy the pinned host batch into those buffers, replay the graph and copy the [G,6,6] result to a pinned host
    tensor.  At most `max_graphs` shapes are kept (least recently used evicted); weights must not be replaced between calls
    (in-place updates are fine only if the model's cached tables are refreshed: call `clear()` after an optimizer step)."""

    _FIELDS = ("positions", "node_attrs", "edge_index", "shifts", "edge_attr", "batch")

    def __init__(self, model: torch.nn.Module, max_graphs: int = 8):
        p = next(model.parameters())
        if p.device.type != "cuda":
            raise RuntimeError("GraphedPredictor needs the model on a CUDA device (there is no CPU path)")
        self.model = model
        self.device = p.device
        self.max_graphs = int(max_graphs)
        self._graphs: dict = {}
        self.launches_per_replay = 1

    def clear(self):
        self._graphs.clear()

    def _capture(self, hb):
        from .data import Batch
        dev = self.device
        static = Batch(**{k: torch.empty_like(getattr(hb, k), device=dev) for k in self._FIELDS}, num_graphs=int(hb.num_graphs))
        for k in self._FIELDS:
            getattr(static, k).copy_(getattr(hb, k), non_blocking=True)
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(2):
                self.model(static)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        g = torch.cuda.CUDAGraph()
        with torch.no_grad(), torch.cuda.graph(g):
            out = self.model(static)["stiffness"]
        host = torch.empty(out.shape, dtype=out.dtype, pin_memory=True)
        return dict(graph=g, static=static, out=out, host=host)

    @torch.no_grad()
    def __call__(self, host_batch, sync: bool = True) -> torch.Tensor:
        """-> pinned host [G,6,6] (valid until the next call with the same shape).  sync=False returns right after
        enqueueing; wait on `self.last_done` before reading."""
        key = (int(host_batch.positions.shape[0]), int(host_batch.edge_index.shape[1]), int(host_batch.num_graphs))
        ent = self._graphs.pop(key, None)
        if ent is None:
            if len(self._graphs) >= self.max_graphs:
                self._graphs.pop(next(iter(self._graphs)))
            ent = self._capture(host_batch)
        self._graphs[key] = ent   # most recently used last
        st = ent["static"]
        for k in self._FIELDS:
            getattr(st, k).copy_(getattr(host_batch, k), non_blocking=True)
        ent["graph"].replay()
        ent["host"].copy_(ent["out"], non_blocking=True)
        self.last_done = torch.cuda.Event()
        self.last_done.record(torch.cuda.current_stream(self.device))
        if sync:
            self.last_done.synchronize()
        return ent["host"]
