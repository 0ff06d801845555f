"""`.lat` catalogue -> model-ready lattices: the on-disk format one step upstream of the hot path (SURVEY.md section 8(f)
rank 2).

    read_lat(path)          text catalogue -> list of unit-cell dicts      lattices/lattices/data/catalogue.py:316-479
                            (`Catalogue.from_file` + `get_unit_cell`; same keys)
    process_one(uc)         unit-cell dict -> one `LatticeData` per relative density
                                                                             gnn/datasets.py:115-276 (edge format 'r',
                                                                             graph format 'cartesian_4')
    write_lat(path, cells)  the inverse of read_lat for the sections process_one reads (synthetic catalogues, tests)

Host-side numpy, as in the reference; what it produces feeds `data.collate` -> `RotateLat` (device) -> the model.
Validated against the reference's own Catalogue / process_one output on lattices/examples/sample_cat.lat
(tests/golden/ref_sample_cat.pt, tests/test_catalogue.py).
"""
from __future__ import annotations

import math
from typing import Dict, Iterable, List, Optional

import numpy as np
import torch

from .data import LatticeData, transform_matrix, uc_volume

SEPARATOR = "----- lattice_transition -----"
_PAIRS = [(0, 0), (1, 1), (2, 2), (1, 2), (0, 2), (0, 1)]  # Mandel index -> Cartesian pair (elasticity_func.py:430-449)


def _block(lines: List[str], start: int, conv) -> list:
    """Rows after the header line `start` up to the first blank line (catalogue.py:405-447)."""
    rows = []
    for line in lines[start + 1:]:
        if len(line.strip()) == 0:
            break
        rows.append([conv(w) for w in line.split()])
    return rows


def _per_density(lines: List[str], lo: int, hi: int) -> Dict[float, List[float]]:
    out: Dict[float, List[float]] = {}
    for i in range(lo + 1, hi):
        if "at relative density" in lines[i]:
            rd = float(lines[i].strip().rstrip(":").split("density")[1])
            nums = []
            for w in lines[i + 1].split(","):
                try:
                    nums.append(float(w))
                except ValueError:
                    continue
            out[rd] = nums
    return out


def parse_unit_cell(lines: List[str], indexing: int = 0) -> dict:
    """One catalogue entry (the lines after a separator) -> dict with the keys of `Catalogue.get_unit_cell`."""
    uc: dict = {}
    mark: Dict[str, int] = {}
    compl_format = "V"
    for i, line in enumerate(lines):
        if line.startswith("Name:"):
            uc["name"] = line.split(":", 1)[1].strip()
        elif "unit cell parameters" in line:
            uc["lattice_constants"] = [float(w) for w in lines[i + 1].split(",")]
        elif "connectivity" in line and "Z_avg = " in line:
            uc["average_connectivity"] = float(line.split("Z_avg = ")[1])
        elif "Base name" in line:
            uc["base_name"] = line.split(":")[1].strip()
        elif "Imperfection level" in line:
            uc["imperfection_level"] = float(line.split(":")[1])
        elif "Imperfection kind" in line:
            uc["imperfection_kind"] = line.split(":")[1].strip()
        elif "Nodal hash" in line:
            uc["nodal_hash"] = line.split(":")[1].strip()
        elif "Compliance tensors" in line:
            mark["compl_start" if "start" in line else "compl_end"] = i
            compl_format = "M" if "Mandel" in line else "V"
        elif "Nodal positions" in line:
            mark["nodes"] = i
        elif "Bar connectivities" in line:
            mark["bars"] = i
        elif "Fundamental edge adjacency" in line:
            mark["fund_adj"] = i
        elif "Fundamental tesselation vectors" in line:
            mark["fund_tess"] = i
        elif "Fundamental edge radii start" in line:
            mark["rad_start"] = i
        elif "Fundamental edge radii end" in line:
            mark["rad_end"] = i
    if "name" not in uc or "nodes" not in mark:
        raise ValueError("catalogue entry without a name or nodal positions")
    uc["nodal_positions"] = _block(lines, mark["nodes"], float)
    if "bars" in mark:
        uc["edge_adjacency"] = _block(lines, mark["bars"], lambda w: int(w) - indexing)
    if "fund_adj" in mark:
        uc["fundamental_edge_adjacency"] = _block(lines, mark["fund_adj"], lambda w: int(w) - indexing)
        uc["fundamental_tesselation_vecs"] = _block(lines, mark["fund_tess"], float)
    if "compl_start" in mark:
        tens = {}
        for rd, nums in _per_density(lines, mark["compl_start"], mark["compl_end"]).items():
            if len(nums) != 21:
                raise ValueError(f"{uc['name']}: compliance at relative density {rd} has {len(nums)} numbers, expected 21")
            S = np.zeros((6, 6))
            S[np.triu_indices(6)] = nums
            S[np.triu_indices(6)[::-1]] = nums
            tens[rd] = S
        uc[f"compliance_tensors_{compl_format}"] = tens
    if "rad_start" in mark:
        uc["fundamental_edge_radii"] = _per_density(lines, mark["rad_start"], mark["rad_end"])
    return uc


def read_lat(path: str, indexing: int = 0) -> List[dict]:
    """`Catalogue.from_file(path, indexing)` + `get_unit_cell` for every entry, in file order."""
    with open(path) as f:
        text = f.read().split("\n")
    cells, cur = [], None
    for line in text:
        if line.strip() == SEPARATOR:
            if cur:
                cells.append(cur)
            cur = []
        elif cur is not None:
            cur.append(line)
    if cur:
        cells.append(cur)
    return [parse_unit_cell(c, indexing) for c in cells if any(l.strip() for l in c)]


def mandel_to_cart4(C: np.ndarray) -> np.ndarray:
    """`numpy_Mandel_to_cart_4` (elasticity_func.py:367-378, :430-463): upper triangle, fully symmetrised."""
    C4 = np.zeros((3, 3, 3, 3))
    r2 = math.sqrt(2.0)
    for i in range(6):
        a, b = _PAIRS[i]
        for j in range(i, 6):
            c, d = _PAIRS[j]
            val = C[i, j]
            if i > 2:
                val = val / r2
            if j > 2:
                val = val / r2
            for (p, q, r, s) in ((a, b, c, d), (b, a, c, d), (a, b, d, c), (c, d, a, b), (b, a, d, c), (c, d, b, a), (d, c, a, b), (d, c, b, a)):
                C4[p, q, r, s] = val
    return C4


def process_one(uc: dict, reldens_slice: slice = slice(None, None, 1)) -> List[LatticeData]:
    """`GLAMM_Dataset.process_one(lat, 'r', 'cartesian_4')` (gnn/datasets.py:115-276): fundamental representation ->
    Cartesian positions / shifts, doubled edge list (forward struts, then reversed), strut radius from the relative density
    (or the stored fundamental radii), stiffness = inv(compliance) in Mandel form, both expanded to rank 4 (fp64).
    The returned `LatticeData` carry `stiffness` / `compliance` as `[1,3,3,3,3]` and `unit_shifts` - what `RotateLat` takes."""
    name = uc["name"]
    if "nodal_positions" in uc:
        pos = np.atleast_2d(np.asarray(uc["nodal_positions"], dtype=np.float64))
    elif "reduced_node_coordinates" in uc:
        pos = np.atleast_2d(np.asarray(uc["reduced_node_coordinates"], dtype=np.float64))
    else:
        raise ValueError("No nodal positions found")
    fadj = np.atleast_2d(np.asarray(uc["fundamental_edge_adjacency"]))
    ftess = np.atleast_2d(np.asarray(uc["fundamental_tesselation_vecs"], dtype=np.float64))
    const = np.asarray(uc["lattice_constants"], dtype=np.float64)
    if "compliance_tensors_M" in uc:
        compl = uc["compliance_tensors_M"]
    elif "compliance_tensors_V" in uc:
        raise NotImplementedError("Voigt compliance catalogues are not supported (the GLAMM catalogues store Mandel)")
    else:
        compl = {}
    uq = np.unique(fadj)
    pos = pos[uq]
    adj = np.searchsorted(uq, fadj)
    if ftess.shape[1] == 6:
        tess = ftess[:, 3:] - ftess[:, :3]
    elif ftess.shape[1] == 3:
        tess = ftess
    else:
        raise ValueError(f"Fundamental tessellation vectors shape {ftess.shape} not recognised")
    unit_shifts = tess.astype(int)
    T = transform_matrix(const)
    pos = pos @ T.T
    tess = tess @ T.T
    adj = np.vstack((adj, adj[:, ::-1]))
    unit_shifts = np.vstack((unit_shifts, -unit_shifts))
    tess = np.vstack((tess, -tess))
    vec = pos[adj[:, 1]] - pos[adj[:, 0]] + tess
    lens = np.linalg.norm(vec, axis=1)
    vol = uc_volume(const)
    n_nodes = len(np.unique(adj))
    if len(compl) == 0:
        raise AssertionError(f"Lattice {name} does not have enough data")
    out = []
    for rd in list(compl.keys())[reldens_slice]:
        if "fundamental_edge_radii" in uc:
            keys = np.array(list(uc["fundamental_edge_radii"].keys()))
            k = keys[np.argmin(np.abs(keys - rd))]
            if abs(k - rd) >= 1e-4:
                raise AssertionError(f"Closest relative density {k} is not close enough to {rd}")
            r = np.asarray(uc["fundamental_edge_radii"][k], dtype=np.float64).reshape(-1, 1)
            radii = np.concatenate((r, r), axis=0)
            if radii.shape[0] != adj.shape[0]:
                raise AssertionError(f"Edge radii shape {radii.shape} does not match edge adjacency shape {adj.shape}")
        else:
            radii = math.sqrt(rd * vol / (lens.sum() * math.pi)) * np.ones((adj.shape[0], 1))
        S = compl[rd]
        C4 = S4 = None
        if S is not None:
            C4 = torch.from_numpy(mandel_to_cart4(np.linalg.inv(S))).unsqueeze(0)
            S4 = torch.from_numpy(mandel_to_cart4(S)).unsqueeze(0)
        d = LatticeData(positions=torch.tensor(pos, dtype=torch.float32), node_attrs=torch.ones((n_nodes, 1), dtype=torch.float32),
                        edge_index=torch.tensor(adj.T.copy(), dtype=torch.int64), shifts=torch.tensor(tess, dtype=torch.float32),
                        edge_attr=torch.tensor(radii, dtype=torch.float32), stiffness=C4, name=name, rel_dens=float(rd))
        d.compliance = S4
        d.unit_shifts = torch.tensor(unit_shifts, dtype=torch.int64)
        out.append(d)
    return out


def write_lat(path: str, cells: Iterable[dict]) -> None:
    """Write the sections `read_lat` / `process_one` consume (name, lattice constants, Mandel compliance per relative
    density, nodal positions, fundamental adjacency + tessellation vectors, optional fundamental radii)."""
    with open(path, "w") as f:
        for uc in cells:
            f.write(f"{SEPARATOR}\nName: {uc['name']}\n\n")
            f.write("Normalized unit cell parameters (a,b,c,alpha,beta,gamma):\n")
            f.write(", ".join(repr(float(x)) for x in uc["lattice_constants"]) + "\n\n\n")
            if uc.get("compliance_tensors_M"):
                f.write("Compliance tensors (Mandel) start (flattened upper triangular)\n")
                for rd, S in uc["compliance_tensors_M"].items():
                    f.write(f"-> at relative density {rd!r}:\n")
                    f.write(",".join(repr(float(x)) for x in np.asarray(S)[np.triu_indices(6)]) + "\n")
                f.write("Compliance tensors (Mandel) end\n\n")
            f.write("Nodal positions:\n")
            for p in uc["nodal_positions"]:
                f.write(" ".join(repr(float(x)) for x in p) + "\n")
            f.write("\n")
            if "edge_adjacency" in uc:
                f.write("Bar connectivities:\n")
                for e in uc["edge_adjacency"]:
                    f.write(" ".join(str(int(x)) for x in e) + "\n")
                f.write("\n")
            f.write("Fundamental edge adjacency\n")
            for e in uc["fundamental_edge_adjacency"]:
                f.write(" ".join(str(int(x)) for x in e) + "\n")
            f.write("\nFundamental tesselation vectors\n")
            for t in uc["fundamental_tesselation_vecs"]:
                f.write(" ".join(repr(float(x)) for x in t) + "\n")
            f.write("\n")
            if uc.get("fundamental_edge_radii"):
                f.write("Fundamental edge radii start\n")
                for rd, r in uc["fundamental_edge_radii"].items():
                    f.write(f"-> at relative density {rd!r}:\n" + ",".join(repr(float(x)) for x in r) + "\n")
                f.write("Fundamental edge radii end\n\n")
