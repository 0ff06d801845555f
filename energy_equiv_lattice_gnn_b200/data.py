"""Lattice graphs: the duck-typed `Batch` the model consumes, PyG-equivalent collation, and the synthetic
periodic-lattice recipe used by tests and bench (mirrors gnn/datasets.py:147-269).

The reference gets its batches from torch_geometric (`Batch.from_data_list`, scripts/train_main.py:67-78); that package
is absent here, so `collate` restates the only behaviour the model depends on: node/edge tensors concatenated along
dim 0, `edge_index` offset by the cumulative node count, `batch` = graph id per node.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import List, Optional, Sequence

import numpy as np
import torch


@dataclass
class LatticeData:
    """One lattice in the schema of gnn/datasets.py:256-269 (fields the model reads)."""
    positions: torch.Tensor   # f32 [N,3]
    node_attrs: torch.Tensor  # f32 [N,1] (all ones, datasets.py:182)
    edge_index: torch.Tensor  # i64 [2,E]  row 0 sender, row 1 receiver
    shifts: torch.Tensor      # f32 [E,3]
    edge_attr: torch.Tensor   # f32 [E,1]  strut radius
    stiffness: Optional[torch.Tensor] = None  # f32 [1,6,6] (training target, Mandel)
    name: str = ""
    rel_dens: float = 0.0


class Batch:
    """Attribute bag with the PyG `Batch` fields used by `EnergyEquivGNN.forward` (gnn/model.py:137-161)."""

    def __init__(self, **kw):
        for k, v in kw.items():
            setattr(self, k, v)

    def to(self, device, non_blocking: bool = False):
        out = Batch()
        for k, v in self.__dict__.items():
            setattr(out, k, v.to(device, non_blocking=non_blocking) if torch.is_tensor(v) else v)
        return out

    def pin_memory(self):
        out = Batch()
        for k, v in self.__dict__.items():
            setattr(out, k, v.pin_memory() if torch.is_tensor(v) else v)
        return out

    def __getitem__(self, k):  # batch['stiffness'] as in scripts/train_utils.py:49
        return getattr(self, k)


def collate(data_list: Sequence[LatticeData]) -> Batch:
    n_off = 0
    ei, bidx = [], []
    for g, d in enumerate(data_list):
        ei.append(d.edge_index + n_off)
        n = d.node_attrs.shape[0]
        bidx.append(torch.full((n,), g, dtype=torch.int64))
        n_off += n
    b = Batch(
        positions=torch.cat([d.positions for d in data_list], 0),
        node_attrs=torch.cat([d.node_attrs for d in data_list], 0),
        edge_index=torch.cat(ei, 1),
        shifts=torch.cat([d.shifts for d in data_list], 0),
        edge_attr=torch.cat([d.edge_attr for d in data_list], 0),
        batch=torch.cat(bidx, 0),
        num_graphs=len(data_list),
    )
    if all(d.stiffness is not None for d in data_list):
        b.stiffness = torch.cat([d.stiffness for d in data_list], 0)
    if all(getattr(d, "compliance", None) is not None for d in data_list):  # catalogue.process_one output
        b.compliance = torch.cat([d.compliance for d in data_list], 0)
    return b


# ------------------------------------------------------------------------------------------------------------------
# geometry helpers restated from lattices/lattices/data/lattice.py:280-295,343-368
# ------------------------------------------------------------------------------------------------------------------


def uc_volume(c) -> float:
    a, b, cc = c[0], c[1], c[2]
    al, be, ga = (math.radians(x) for x in c[3:6])
    t = 1.0 - math.cos(al) ** 2 - math.cos(be) ** 2 - math.cos(ga) ** 2 + 2.0 * math.cos(al) * math.cos(be) * math.cos(ga)
    return a * b * cc * math.sqrt(t)


def transform_matrix(c) -> np.ndarray:
    a, b, cc = c[0], c[1], c[2]
    al, be, ga = (math.radians(x) for x in c[3:6])
    om = uc_volume(c)
    T = np.zeros((3, 3))
    T[0, 0] = a
    T[0, 1] = b * math.cos(ga)
    T[0, 2] = cc * math.cos(be)
    T[1, 1] = b * math.sin(ga)
    T[1, 2] = cc * (math.cos(al) - math.cos(be) * math.cos(ga)) / math.sin(ga)
    T[2, 2] = om / (a * b * math.sin(ga))
    return T


def lattice_from_struts(frac_pos: np.ndarray, struts: np.ndarray, unit_shifts: np.ndarray, constants, rel_dens: float,
                        stiffness: Optional[np.ndarray] = None, name: str = "") -> LatticeData:
    """Steps of gnn/datasets.py:159-204: fractional -> Cartesian, doubled edge list, radius from relative density
    (sum of lengths runs over the DOUBLED list - quirk kept)."""
    T = transform_matrix(constants)
    pos = frac_pos @ T.T
    tess = unit_shifts.astype(np.float64) @ T.T
    adj = np.vstack((struts, struts[:, ::-1]))
    tess = np.vstack((tess, -tess))
    vec = pos[adj[:, 1]] - pos[adj[:, 0]] + tess
    lens = np.linalg.norm(vec, axis=1)
    rad = math.sqrt(rel_dens * uc_volume(constants) / (lens.sum() * math.pi))
    E = adj.shape[0]
    return LatticeData(
        positions=torch.tensor(pos, dtype=torch.float32),
        node_attrs=torch.ones((pos.shape[0], 1), dtype=torch.float32),
        edge_index=torch.tensor(adj.T.copy(), dtype=torch.int64),
        shifts=torch.tensor(tess, dtype=torch.float32),
        edge_attr=torch.full((E, 1), rad, dtype=torch.float32),
        stiffness=None if stiffness is None else torch.tensor(stiffness, dtype=torch.float32).reshape(1, 6, 6),
        name=name, rel_dens=rel_dens,
    )


def synthetic_lattice(seed: int, n_nodes: int = 30, n_struts: Optional[int] = None, with_target: bool = False) -> LatticeData:
    """Deterministic random periodic lattice."""
    g = torch.Generator(device="cpu").manual_seed(int(seed))
    N = n_nodes
    S = 2 * N if n_struts is None else n_struts
    x = torch.rand(N, 3, generator=g, dtype=torch.float64).numpy()
    if seed % 2 == 0:
        const = (1.0, 1.0, 1.0, 90.0, 90.0, 90.0)
    else:
        abc = (0.6 + 0.4 * torch.rand(3, generator=g, dtype=torch.float64)).numpy()
        abc = abc / abc.max()
        const = (float(abc[0]), float(abc[1]), float(abc[2]), 90.0, 90.0, 90.0)
    T = transform_matrix(const)
    # candidate struts i<=j, n in {-1,0,1}^3 (for i==j only n lexicographically > 0)
    ns = np.array([(a, b, c) for a in (-1, 0, 1) for b in (-1, 0, 1) for c in (-1, 0, 1)], dtype=np.int64)
    ii, jj = np.triu_indices(N)
    I = np.repeat(ii, 27)
    J = np.repeat(jj, 27)
    NS = np.tile(ns, (len(ii), 1))
    lexpos = (NS[:, 0] > 0) | ((NS[:, 0] == 0) & ((NS[:, 1] > 0) | ((NS[:, 1] == 0) & (NS[:, 2] > 0))))
    keep = (I != J) | lexpos
    I, J, NS = I[keep], J[keep], NS[keep]
    L = np.linalg.norm((x[J] + NS - x[I]) @ T.T, axis=1)
    order = np.lexsort((NS[:, 2], NS[:, 1], NS[:, 0], J, I, L))  # by (length, i, j, n)
    rank = np.empty_like(order)
    rank[order] = np.arange(len(order))
    chosen: List[int] = []
    chosen_set = set()
    best_rank = np.full(N, len(order), dtype=np.int64)  # each node's shortest candidate
    np.minimum.at(best_rank, I, rank)
    np.minimum.at(best_rank, J, rank)
    for n in range(N):
        best = order[best_rank[n]]
        if best not in chosen_set:
            chosen_set.add(int(best))
            chosen.append(int(best))
    for c in order:
        if len(chosen) >= S:
            break
        if int(c) not in chosen_set:
            chosen_set.add(int(c))
            chosen.append(int(c))
    chosen = np.array(chosen[:max(S, len(chosen))], dtype=np.int64)
    struts = np.stack((I[chosen], J[chosen]), 1)
    rho = (0.001, 0.003, 0.01)[seed % 3]
    target = None
    if with_target:
        A = torch.randn(6, 6, generator=g, dtype=torch.float64).numpy()
        A = 0.5 * (A + A.T)
        target = (A @ A) / 6.0
    return lattice_from_struts(x, struts, NS[chosen], const, rho, stiffness=target, name=f"syn{seed}")


def supercell(frac_pos: np.ndarray, struts: np.ndarray, unit_shifts: np.ndarray, constants, reps):
    """n_x x n_y x n_z replication without rescaling: C(supercell) == C(unit cell)."""
    reps = np.asarray(reps, dtype=np.int64)
    N = frac_pos.shape[0]
    cells = [(a, b, c) for a in range(reps[0]) for b in range(reps[1]) for c in range(reps[2])]
    cid = {c: k for k, c in enumerate(cells)}
    pos = np.concatenate([(frac_pos + np.array(c)) / reps for c in cells], 0)
    new_s, new_n = [], []
    for c in cells:
        for (i, j), n in zip(struts, unit_shifts):
            tgt = np.array(c) + n
            wrap = np.floor_divide(tgt, reps)
            tc = tuple(int(v) for v in (tgt - wrap * reps))
            new_s.append((cid[c] * N + i, cid[tc] * N + j))
            new_n.append(wrap)
    const = (constants[0] * reps[0], constants[1] * reps[1], constants[2] * reps[2]) + tuple(constants[3:6])
    return pos, np.array(new_s, dtype=np.int64), np.array(new_n, dtype=np.int64), const


def synthetic_batch(seeds: Sequence[int], n_nodes: int = 30, with_target: bool = False):
    """(Batch, max_edge_radius) for a list of seeds."""
    data = [synthetic_lattice(s, n_nodes, with_target=with_target) for s in seeds]
    rmax = max(float(d.edge_attr.max()) for d in data)
    return collate(data), rmax


# ------------------------------------------------------------------------------------------------------------------
# RotateLat on a collated device batch (scripts/train_utils.py:114-146)
# ------------------------------------------------------------------------------------------------------------------
def rand_rotations(n: int, generator: Optional[torch.Generator] = None, dtype=torch.float32) -> torch.Tensor:
    """[n,3,3] random proper rotations, the distribution of e3nn `o3.rand_matrix()` (train_utils.py:121): alpha, gamma
    uniform on [0, 2pi), cos(beta) uniform on [-1, 1], R = Ry(alpha) Rx(beta) Ry(gamma)."""
    u = torch.rand(3, n, generator=generator, dtype=torch.float64)
    alpha, gamma = 2 * math.pi * u[0], 2 * math.pi * u[2]
    beta = torch.acos(2 * u[1] - 1)

    def ry(a):
        c, s, o, z = a.cos(), a.sin(), torch.ones_like(a), torch.zeros_like(a)
        return torch.stack([c, z, s, z, o, z, -s, z, c], -1).reshape(-1, 3, 3)

    def rx(a):
        c, s, o, z = a.cos(), a.sin(), torch.ones_like(a), torch.zeros_like(a)
        return torch.stack([o, z, z, z, c, -s, z, s, c], -1).reshape(-1, 3, 3)

    return (ry(alpha) @ rx(beta) @ ry(gamma)).to(dtype)


class RotateLat:
    """Batched, on-device counterpart of the reference's per-lattice `RotateLat` transform (scripts/train_utils.py:114-146).

    `__call__(batch, Q=None)` takes a collated CUDA `Batch` whose targets are still Cartesian rank-4 tensors
    (`stiffness`, and optionally `compliance`, `[G,3,3,3,3]`, as `GLAMM_Dataset.process_one` stores them for
    'cartesian_4', gnn/datasets.py:256-269) and returns a new `Batch` with rotated `positions` / `shifts` and Mandel
    `[G,6,6]` targets.  `Q` is `[G,3,3]` (one rotation per lattice, drawn like `o3.rand_matrix()` when omitted);
    `rotate=False` only converts the targets, and then `Q` must be None, as in the reference (:126-127).
    `seed` (int) draws the rotations ON THE DEVICE (`eeg_rand_rotations`, counter-based: the stream position advances by G per
    call, so epochs are reproducible from (seed, number of lattices seen)); `generator` draws them on the host like round 1.
    CUDA only: the kernels live in libeeg_b200 (`eeg_rotate_vec3`, `eeg_rotate_cart4_to_mandel`)."""

    def __init__(self, rotate: bool = True, generator: Optional[torch.Generator] = None, seed: Optional[int] = None):
        self.rotate = rotate
        self.generator = generator
        self.seed = seed
        self.offset = 0

    def __call__(self, batch: Batch, Q: Optional[torch.Tensor] = None) -> Batch:
        from . import native
        G = int(batch.num_graphs)
        dev = batch.positions.device
        out = Batch(**batch.__dict__)
        if self.rotate:
            if Q is None and self.seed is not None:
                Q = native.rand_rotations(int(self.seed), self.offset, G, dev)
                self.offset += G
            elif Q is None:
                Q = rand_rotations(G, self.generator)
            Q = Q.to(device=dev, dtype=torch.float32).contiguous()
            if Q.shape != (G, 3, 3):
                raise ValueError(f"Q must be [{G},3,3], got {tuple(Q.shape)}")
            out.positions = torch.empty_like(batch.positions)
            out.shifts = torch.empty_like(batch.shifts)
            native.rotate_vec3(Q, batch.batch, None, batch.positions.contiguous(), out.positions)
            native.rotate_vec3(Q, batch.batch, batch.edge_index[0].contiguous(), batch.shifts.contiguous(), out.shifts)
        else:
            assert Q is None, "Q should be None if instance initialized with rotate=False"
        for key in ("stiffness", "compliance"):
            t = getattr(batch, key, None)
            if t is None:
                continue
            if tuple(t.shape[1:]) != (3, 3, 3, 3):
                raise ValueError(f"batch.{key} must be Cartesian rank-4 [G,3,3,3,3], got {tuple(t.shape)}")
            m = torch.empty(G, 6, 6, dtype=torch.float32, device=dev)
            native.rotate_cart4_to_mandel(Q if self.rotate else None, t.to(torch.float32).contiguous(), m)
            setattr(out, key, m)
        return out
