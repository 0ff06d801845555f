"""Flat binary lattice dataset + batch construction on the device.

The reference turns a `.lat` catalogue into a list of PyG `Data` objects, `torch.save`s the collated list
(`gnn/datasets.py:278-307`) and lets a `DataLoader` re-collate `batch_size` of them per step on the host
(`scripts/train_main.py:67-78`).  Here the processed dataset is ONE flat file of concatenated arrays that can be memory
mapped, uploaded to HBM once (a 10^6-lattice catalogue is a few GB of the 180 GB) and batched by a single gather kernel:

    FlatDataset.from_lattices(list[LatticeData])        concatenate (host)
    FlatDataset.save(path) / FlatDataset.load(path)     flat binary: magic, JSON header, 64-byte aligned raw arrays
    FlatDataset.to(device)                              dataset resident on the GPU
    FlatDataset.collate(indices) -> data.Batch          PyG `Batch.from_data_list` semantics.  On a CUDA
                                                        dataset: host computes the 4 x (G+1) offset table from the host
                                                        copy of node_ptr / edge_ptr (no device sync), one small H2D
                                                        copy, ONE `eeg_collate_gather` launch.
    convert_lat(lat_path, out_path)                     `.lat` text -> flat binary (fast parser + process_one), with the
                                                        measured lattices/s

Targets are stored as the full Cartesian rank-4 tensors `[81]` (stiffness, compliance): what `GLAMM_Dataset.process_one`
stores for graph format 'cartesian_4' and what `RotateLat` consumes.  Throughput: tools/flatdata_bench.py
(profiles/r02_flatdata_throughput.json).
"""
from __future__ import annotations

import json
import os
import time
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from .data import Batch, LatticeData

MAGIC = b"EEGFLAT1"
_ALIGN = 64
_NODE = ("positions", "node_attrs")
_EDGE = ("edge_snd", "edge_rcv", "shifts", "edge_attr")


class FlatDataset:
    """Concatenated lattices.  Arrays (torch tensors, host or device):
    node_ptr, edge_ptr int64 [L+1] (always also kept on the host), positions f32 [SN,3], node_attrs f32 [SN],
    edge_snd / edge_rcv int32 [SE] (local node ids), shifts f32 [SE,3], edge_attr f32 [SE], rel_dens f32 [L],
    optional stiffness / compliance f32 [L,81] (Cartesian rank 4) or [L,36] (Mandel); names: list[str]."""

    def __init__(self, arrays: Dict[str, torch.Tensor], names: Optional[List[str]] = None):
        self.a = arrays
        self.names = names or []
        self.node_ptr_host = arrays["node_ptr"].cpu().numpy()
        self.edge_ptr_host = arrays["edge_ptr"].cpu().numpy()

    def __len__(self) -> int:
        return len(self.node_ptr_host) - 1

    @property
    def device(self):
        return self.a["positions"].device

    # ---------------------------------------------------------------------------------------------------------------
    @staticmethod
    def from_lattices(data: Sequence[LatticeData]) -> "FlatDataset":
        n = np.array([d.positions.shape[0] for d in data], dtype=np.int64)
        e = np.array([d.edge_index.shape[1] for d in data], dtype=np.int64)
        arr = {
            "node_ptr": torch.from_numpy(np.concatenate(([0], np.cumsum(n)))),
            "edge_ptr": torch.from_numpy(np.concatenate(([0], np.cumsum(e)))),
            "positions": torch.cat([d.positions for d in data], 0).to(torch.float32).contiguous(),
            "node_attrs": torch.cat([d.node_attrs.reshape(-1) for d in data], 0).to(torch.float32).contiguous(),
            "edge_snd": torch.cat([d.edge_index[0] for d in data], 0).to(torch.int32).contiguous(),
            "edge_rcv": torch.cat([d.edge_index[1] for d in data], 0).to(torch.int32).contiguous(),
            "shifts": torch.cat([d.shifts for d in data], 0).to(torch.float32).contiguous(),
            "edge_attr": torch.cat([d.edge_attr.reshape(-1) for d in data], 0).to(torch.float32).contiguous(),
            "rel_dens": torch.tensor([float(getattr(d, "rel_dens", 0.0)) for d in data], dtype=torch.float32),
        }
        for key in ("stiffness", "compliance"):
            ts = [getattr(d, key, None) for d in data]
            if all(t is not None for t in ts):
                arr[key] = torch.cat([t.reshape(1, -1) for t in ts], 0).to(torch.float32).contiguous()
        return FlatDataset(arr, [getattr(d, "name", "") for d in data])

    # ---------------------------------------------------------------------------------------------------------------
    def save(self, path: str) -> None:
        specs, off = [], 0
        blobs = []
        for k, t in self.a.items():
            x = t.detach().cpu().contiguous().numpy()
            off = (off + _ALIGN - 1) // _ALIGN * _ALIGN
            specs.append(dict(name=k, dtype=str(x.dtype), shape=list(x.shape), offset=off))
            blobs.append((off, x))
            off += x.nbytes
        header = json.dumps(dict(version=1, arrays=specs, names=self.names)).encode()
        base = (len(MAGIC) + 8 + len(header) + _ALIGN - 1) // _ALIGN * _ALIGN
        with open(path, "wb") as f:
            f.write(MAGIC)
            f.write(np.array([len(header)], dtype="<u8").tobytes())
            f.write(header)
            for o, x in blobs:
                f.seek(base + o)
                f.write(x.tobytes())

    @staticmethod
    def load(path: str, device=None, mmap: bool = True) -> "FlatDataset":
        with open(path, "rb") as f:
            if f.read(len(MAGIC)) != MAGIC:
                raise ValueError(f"{path}: not an EEGFLAT1 file")
            hlen = int(np.frombuffer(f.read(8), dtype="<u8")[0])
            header = json.loads(f.read(hlen).decode())
        base = (len(MAGIC) + 8 + hlen + _ALIGN - 1) // _ALIGN * _ALIGN
        arr = {}
        for s in header["arrays"]:
            n = int(np.prod(s["shape"])) if s["shape"] else 1
            if mmap and n > 0:
                x = np.memmap(path, dtype=np.dtype(s["dtype"]), mode="r", offset=base + s["offset"], shape=tuple(s["shape"]))
                t = torch.from_numpy(np.ascontiguousarray(x))
            else:
                x = np.fromfile(path, dtype=np.dtype(s["dtype"]), count=n, offset=base + s["offset"]).reshape(s["shape"])
                t = torch.from_numpy(x)
            arr[s["name"]] = t
        ds = FlatDataset(arr, header.get("names", []))
        return ds.to(device) if device is not None else ds

    def to(self, device, non_blocking: bool = False) -> "FlatDataset":
        return FlatDataset({k: v.to(device, non_blocking=non_blocking) for k, v in self.a.items()}, self.names)

    # ---------------------------------------------------------------------------------------------------------------
    def _offsets(self, idx: np.ndarray):
        """[4, G+1] int64 table: src_node_off, src_edge_off, out_node_ptr, out_edge_ptr; and (N, E)."""
        G = len(idx)
        n0, e0 = self.node_ptr_host[idx], self.edge_ptr_host[idx]
        nn, ne = self.node_ptr_host[idx + 1] - n0, self.edge_ptr_host[idx + 1] - e0
        meta = np.zeros((4, G + 1), dtype=np.int64)
        meta[0, :G], meta[1, :G] = n0, e0
        np.cumsum(nn, out=meta[2, 1:])
        np.cumsum(ne, out=meta[3, 1:])
        return meta, int(meta[2, G]), int(meta[3, G])

    def collate(self, indices, targets: Sequence[str] = ("stiffness", "compliance")) -> Batch:
        """PyG `Batch.from_data_list([dataset[i] for i in indices])` (fields of gnn/datasets.py:256-269 the path reads)."""
        idx = np.asarray(indices.cpu() if torch.is_tensor(indices) else indices, dtype=np.int64).reshape(-1)
        if idx.size and (idx.min() < 0 or idx.max() >= len(self)):
            raise IndexError("lattice index out of range")
        G = len(idx)
        meta, N, E = self._offsets(idx)
        a = self.a
        dev = self.device
        present = [k for k in targets if k in a]
        if dev.type == "cuda":
            from . import native
            meta_d = torch.from_numpy(meta).pin_memory().to(dev, non_blocking=True)
            idx_d = torch.from_numpy(idx).pin_memory().to(dev, non_blocking=True) if present else None
            first = present[0] if present else None
            pos, na, ei, sh, ea, bvec, g0 = native.collate_gather(a["positions"], a["node_attrs"], a["edge_snd"], a["edge_rcv"], a["shifts"],
                                                                 a["edge_attr"], meta_d, G, N, E,
                                                                 a[first] if first else None, idx_d)
            out = Batch(positions=pos, node_attrs=na, edge_index=ei, shifts=sh, edge_attr=ea, batch=bvec, num_graphs=G)
            for k in present:
                setattr(out, k, g0 if k == first else a[k].index_select(0, idx_d))
        else:
            # host collation (the reference's own DataLoader path), vectorised in numpy: repeat-interleave of the offsets
            gid_n = np.repeat(np.arange(G), np.diff(meta[2]))
            gid_e = np.repeat(np.arange(G), np.diff(meta[3]))
            src_n = meta[0, gid_n] + (np.arange(N) - meta[2, gid_n])
            src_e = meta[1, gid_e] + (np.arange(E) - meta[3, gid_e])
            noff = meta[2, gid_e]
            h = {k: v.numpy() for k, v in a.items()}
            ei = np.stack([h["edge_snd"][src_e].astype(np.int64) + noff, h["edge_rcv"][src_e].astype(np.int64) + noff])
            out = Batch(positions=torch.from_numpy(h["positions"][src_n]), node_attrs=torch.from_numpy(h["node_attrs"][src_n]).reshape(-1, 1),
                        edge_index=torch.from_numpy(ei), shifts=torch.from_numpy(h["shifts"][src_e]),
                        edge_attr=torch.from_numpy(h["edge_attr"][src_e]).reshape(-1, 1), batch=torch.from_numpy(gid_n), num_graphs=G)
            for k in present:
                setattr(out, k, torch.from_numpy(h[k][idx]))
        for k in present:   # Cartesian rank-4 targets keep the reference's [G,3,3,3,3] shape
            t = getattr(out, k)
            if t.shape[1] == 81:
                setattr(out, k, t.reshape(G, 3, 3, 3, 3))
            elif t.shape[1] == 36:
                setattr(out, k, t.reshape(G, 6, 6))
        out.rel_dens = a["rel_dens"][torch.from_numpy(idx).to(a["rel_dens"].device)]
        if self.names:
            out.name = [self.names[i] for i in idx]
        return out


# ----------------------------------------------------------------------------------------------------------------------
# fast `.lat` reader: one pass of str.find over the whole text + np.fromstring per numeric block
# ----------------------------------------------------------------------------------------------------------------------
def _num_block(text: str, start: int) -> np.ndarray:
    """Numbers of the rows after the header line that ends at `start`, up to the first blank line."""
    end = text.find("\n\n", start)
    if end < 0:
        end = len(text)
    return np.array(text[start:end].split(), dtype=np.float64)


def read_lat_fast(path: str) -> List[dict]:
    """Same result as `catalogue.read_lat(path)` for the sections `catalogue.process_one` reads (name, lattice constants,
    Mandel compliance per relative density, nodal positions, fundamental adjacency / tessellation vectors, fundamental
    radii) - lattices/lattices/data/catalogue.py:316-479 - without the per-line Python loop."""
    from .catalogue import SEPARATOR
    with open(path) as f:
        text = f.read()
    cells = []
    for chunk in text.split(SEPARATOR):
        if not chunk.strip():
            continue
        uc: dict = {}
        i = chunk.find("Name:")
        if i < 0:
            raise ValueError("catalogue entry without a name or nodal positions")
        uc["name"] = chunk[i + 5: chunk.find("\n", i)].strip()
        i = chunk.find("unit cell parameters")
        if i >= 0:
            j = chunk.find("\n", i) + 1
            uc["lattice_constants"] = [float(w) for w in chunk[j: chunk.find("\n", j)].split(",")]

        def per_density(lo_key, hi_key):
            lo, hi = chunk.find(lo_key), chunk.find(hi_key)
            out = {}
            if lo < 0:
                return None, ""
            head = chunk[lo: chunk.find("\n", lo)]
            pos = lo
            while True:
                pos = chunk.find("at relative density", pos, hi)
                if pos < 0:
                    break
                eol = chunk.find("\n", pos)
                rd = float(chunk[pos + len("at relative density"): eol].strip().rstrip(":"))
                eol2 = chunk.find("\n", eol + 1)
                out[rd] = np.array(chunk[eol + 1: eol2 if eol2 >= 0 else len(chunk)].split(","), dtype=np.float64)
                pos = eol
            return out, head
        comp, head = per_density("Compliance tensors", "Compliance tensors (Mandel) end" if "Compliance tensors (Mandel) end" in chunk
                                 else "Compliance tensors (Voigt) end")
        if comp is not None:
            tens = {}
            iu = np.triu_indices(6)
            for rd, nums in comp.items():
                if len(nums) != 21:
                    raise ValueError(f"{uc['name']}: compliance at relative density {rd} has {len(nums)} numbers, expected 21")
                S = np.zeros((6, 6))
                S[iu] = nums
                S[iu[::-1]] = nums
                tens[rd] = S
            uc["compliance_tensors_M" if "Mandel" in head else "compliance_tensors_V"] = tens
        i = chunk.find("Nodal positions")
        if i < 0:
            raise ValueError("catalogue entry without a name or nodal positions")
        uc["nodal_positions"] = _num_block(chunk, chunk.find("\n", i) + 1).reshape(-1, 3)
        i = chunk.find("Fundamental edge adjacency")
        if i >= 0:
            uc["fundamental_edge_adjacency"] = _num_block(chunk, chunk.find("\n", i) + 1).astype(np.int64).reshape(-1, 2)
            i = chunk.find("Fundamental tesselation vectors")
            t = _num_block(chunk, chunk.find("\n", i) + 1)
            uc["fundamental_tesselation_vecs"] = t.reshape(len(uc["fundamental_edge_adjacency"]), -1)
        rad, _ = per_density("Fundamental edge radii start", "Fundamental edge radii end")
        if rad is not None:
            uc["fundamental_edge_radii"] = {k: list(v) for k, v in rad.items()}
        cells.append(uc)
    return cells


def convert_lat(lat_path: str, out_path: str, reldens_slice: slice = slice(None, None, 1)) -> dict:
    """`.lat` catalogue -> flat binary dataset (what `GLAMM_Dataset.process` + `torch.save(collate(...))` do in
    gnn/datasets.py:278-307).  Returns timing: catalogue entries / s of the parser, processed lattices / s overall."""
    from .catalogue import process_one
    t0 = time.perf_counter()
    cells = read_lat_fast(lat_path)
    t1 = time.perf_counter()
    data: List[LatticeData] = []
    skipped = 0
    for uc in cells:
        try:
            data.extend(process_one(uc, reldens_slice))
        except AssertionError:      # gnn/datasets.py:292-296: entries without enough data are skipped
            skipped += 1
    t2 = time.perf_counter()
    ds = FlatDataset.from_lattices(data)
    ds.save(out_path)
    t3 = time.perf_counter()
    return dict(entries=len(cells), lattices=len(data), skipped=skipped, parse_s=t1 - t0, process_s=t2 - t1, write_s=t3 - t2,
                entries_per_s=len(cells) / max(t1 - t0, 1e-9), lattices_per_s=len(data) / max(t3 - t0, 1e-9),
                bytes=os.path.getsize(out_path))
