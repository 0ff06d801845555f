"""TEST INFRASTRUCTURE - lets the UNMODIFIED reference sources under /root/reference be imported and run in this
container, where their third-party dependencies (e3nn 0.5.1, torch_scatter 2.0.9, torch_geometric 2.1, opt_einsum 3.3,
opt_einsum_fx, pytorch_lightning, matplotlib) are absent and cannot be installed.

`install()` registers minimal stand-ins for exactly the third-party names the reference hot path touches
(gnn/model.py:1-24, gnn/blocks.py:10-28, gnn/mace.py:12-22, gnn/datasets.py:1-24) and exposes the reference
packages `gnn` and `lattices.lattices` WITHOUT running their `__init__.py` (those import plotting / Lightning-era
modules).  The reference's own code - model wiring, MACE blocks, `_wigner_nj`/`U_matrix_real`, `Contraction`,
`tp_out_irreps_with_instructions`, `Cart_4_to_Mandel`, `GLAMM_Dataset.process_one`, `Catalogue`, `Lattice`,
`elasticity_func` - then executes as written.  oracle/make_golden.py uses this to produce tests/golden/*.

What is still a restatement: the e3nn primitives themselves (o3.Linear, o3.TensorProduct, SphericalHarmonics,
wigner_3j, Gate, soft_one_hot_linspace, ReducedTensorProducts) - written here a SECOND time, generically from the e3nn
semantics of SURVEY.md App. B and independently of oracle/e3nn_lite.py where practical (generic instruction-driven
TensorProduct, generic Gate, Irreps algebra), so the golden vectors cross-check the oracle's specialisations.

Only oracle/make_golden.py and tests may import this file.  /root/reference does not exist on the GPU box: nothing
that runs there imports it.
"""
from __future__ import annotations

import collections
import importlib.util
import math
import os
import sys
import types
from typing import List

import numpy as np
import torch

from energy_equiv_lattice_gnn_b200 import conventions as cv

REFERENCE_ROOT = "/root/reference"


# ----------------------------------------------------------------------------------------------------------------
# e3nn.o3 irreps algebra (e3nn 0.5.x: Irrep is a tuple (l, p); Irreps a tuple of (mul, Irrep))
# ----------------------------------------------------------------------------------------------------------------
class Irrep(tuple):
    def __new__(cls, l, p=None):
        if p is None:
            if isinstance(l, Irrep):
                return l
            if isinstance(l, str):
                s = l.strip()
                l, p = int(s[:-1]), {"e": 1, "o": -1}[s[-1]]
            elif isinstance(l, tuple):
                l, p = l
        return super().__new__(cls, (int(l), int(p)))

    l = property(lambda self: self[0])
    p = property(lambda self: self[1])
    dim = property(lambda self: 2 * self[0] + 1)

    def __repr__(self):
        return f"{self.l}{'e' if self.p == 1 else 'o'}"

    __str__ = __repr__

    def __mul__(self, other):
        other = Irrep(other)
        for l in range(abs(self.l - other.l), self.l + other.l + 1):
            yield Irrep(l, self.p * other.p)


class _MulIr(tuple):
    def __new__(cls, mul, ir):
        return super().__new__(cls, (int(mul), Irrep(ir)))

    mul = property(lambda self: self[0])
    ir = property(lambda self: self[1])
    dim = property(lambda self: self[0] * self[1].dim)

    def __repr__(self):
        return f"{self.mul}x{self.ir}"

    __str__ = __repr__


class Irreps(tuple):
    def __new__(cls, irreps=None):
        if isinstance(irreps, Irreps):
            return super().__new__(cls, irreps)
        out = []
        if isinstance(irreps, Irrep):
            out.append(_MulIr(1, irreps))
        elif isinstance(irreps, str):
            for tok in irreps.replace(" ", "").split("+"):
                if not tok:
                    continue
                if "x" in tok:
                    m, ir = tok.split("x")
                    out.append(_MulIr(int(m), Irrep(ir)))
                else:
                    out.append(_MulIr(1, Irrep(tok)))
        elif irreps is not None:
            for it in irreps:
                if isinstance(it, _MulIr):
                    out.append(it)
                elif isinstance(it, (Irrep, str)):
                    out.append(_MulIr(1, Irrep(it)))
                else:
                    m, ir = it
                    out.append(_MulIr(m, Irrep(ir)))
        return super().__new__(cls, out)

    @staticmethod
    def spherical_harmonics(lmax, p=-1):
        return Irreps([(1, (l, p ** l)) for l in range(lmax + 1)])

    dim = property(lambda self: sum(mi.dim for mi in self))
    num_irreps = property(lambda self: sum(mi.mul for mi in self))
    ls = property(lambda self: [mi.ir.l for mi in self for _ in range(mi.mul)])
    lmax = property(lambda self: max(mi.ir.l for mi in self))

    def count(self, ir):
        ir = Irrep(ir)
        return sum(mi.mul for mi in self if mi.ir == ir)

    def __contains__(self, ir):
        ir = Irrep(ir)
        return any(mi.ir == ir for mi in self)

    def __mul__(self, n):
        return Irreps(tuple.__mul__(self, n))

    __rmul__ = __mul__

    def __add__(self, other):
        return Irreps(tuple.__add__(self, Irreps(other)))

    def __getitem__(self, i):
        x = tuple.__getitem__(self, i)
        return Irreps(x) if isinstance(i, slice) else x

    def simplify(self):
        out = []
        for mi in self:
            if out and out[-1][1] == mi.ir:
                out[-1] = (out[-1][0] + mi.mul, mi.ir)
            elif mi.mul > 0:
                out.append((mi.mul, mi.ir))
        return Irreps(out)

    def sort(self):
        Ret = collections.namedtuple("sort", ["irreps", "p", "inv"])
        out = sorted((mi.ir, i, mi.mul) for i, mi in enumerate(self))  # stable: ties broken by original index
        inv = tuple(i for _, i, _ in out)
        p = [0] * len(inv)
        for k, i in enumerate(inv):
            p[i] = k
        return Ret(Irreps([(mul, ir) for ir, _, mul in out]), tuple(p), inv)

    def slices(self):
        s, i = [], 0
        for mi in self:
            s.append(slice(i, i + mi.dim))
            i += mi.dim
        return s

    def __repr__(self):
        return "+".join(repr(mi) for mi in self)

    __str__ = __repr__


def wigner_3j(l1, l2, l3, dtype=None, device=None):
    """o3.wigner_3j: fresh tensor each call (the reference scales it in place, gnn/mace.py:399-403)."""
    return torch.tensor(np.array(cv.wigner_3j(int(l1), int(l2), int(l3))), dtype=dtype or torch.get_default_dtype(), device=device)


class Linear(torch.nn.Module):
    """o3.Linear: flat weight, blocks in (i_in, i_out) order, 1/sqrt(fan_in), bias on 0e outputs."""

    def __init__(self, irreps_in, irreps_out, internal_weights=True, shared_weights=True, biases=False, **_):
        super().__init__()
        self.irreps_in, self.irreps_out = Irreps(irreps_in), Irreps(irreps_out)
        self.instr = []
        n = 0
        for i, mi in enumerate(self.irreps_in):
            for o, mo in enumerate(self.irreps_out):
                if mi.ir == mo.ir:
                    self.instr.append((i, o, n, mi.mul, mo.mul))
                    n += mi.mul * mo.mul
        self.weight_numel = n
        self.weight = torch.nn.Parameter(torch.randn(n))
        nb = sum(mo.mul for mo in self.irreps_out if mo.ir == Irrep(0, 1)) if biases else 0
        self.bias = torch.nn.Parameter(torch.zeros(nb)) if nb else None

    def forward(self, x):
        sin, z = self.irreps_in.slices(), x.shape[0]
        fan = [sum(mi for (_, o2, _, mi, _) in self.instr if o2 == o) for o in range(len(self.irreps_out))]
        outs = [x.new_zeros(z, mo.mul, mo.ir.dim) for mo in self.irreps_out]
        for i, o, off, mi, mo in self.instr:
            w = self.weight[off:off + mi * mo].view(mi, mo)
            xi = x[:, sin[i]].reshape(z, mi, -1)
            outs[o] = outs[o] + torch.einsum("zui,uw->zwi", xi, w) / math.sqrt(fan[o])
        b = 0
        for o, mo in enumerate(self.irreps_out):
            if self.bias is not None and mo.ir == Irrep(0, 1):
                outs[o] = outs[o] + self.bias[b:b + mo.mul].view(1, -1, 1)
                b += mo.mul
        return torch.cat([t.reshape(z, -1) for t in outs], dim=-1)


class TensorProduct(torch.nn.Module):
    """o3.TensorProduct, instruction driven, modes 'uvu' and 'uvw', external or internal weights;
    irrep_normalization='component', path_normalization='element' (e3nn defaults)."""

    def __init__(self, irreps_in1, irreps_in2, irreps_out, instructions, shared_weights=None, internal_weights=None, **_):
        super().__init__()
        self.irreps_in1, self.irreps_in2, self.irreps_out = Irreps(irreps_in1), Irreps(irreps_in2), Irreps(irreps_out)
        self.instructions = []
        off = 0
        for ins in instructions:
            i1, i2, io, mode, train = ins[:5]
            m1, m2, mo = self.irreps_in1[i1].mul, self.irreps_in2[i2].mul, self.irreps_out[io].mul
            shape = {"uvu": (m1, m2), "uvw": (m1, m2, mo)}[mode]
            self.instructions.append((i1, i2, io, mode, shape, off))
            off += int(np.prod(shape))
        self.weight_numel = off
        # element normalisation: alpha = dim_out / sum over paths into the same output of num_elements
        nel = {"uvu": lambda m1, m2: m2, "uvw": lambda m1, m2: m1 * m2}
        tot = collections.defaultdict(int)
        for i1, i2, io, mode, shape, _ in self.instructions:
            tot[io] += nel[mode](self.irreps_in1[i1].mul, self.irreps_in2[i2].mul)
        self.coeff = [math.sqrt(self.irreps_out[io].ir.dim / tot[io]) for (_, _, io, _, _, _) in self.instructions]
        if internal_weights:
            self.weight = torch.nn.Parameter(torch.randn(off))

    def forward(self, x1, x2, weight=None):
        z = x1.shape[0]
        s1, s2 = self.irreps_in1.slices(), self.irreps_in2.slices()
        outs = [x1.new_zeros(z, mo.mul, mo.ir.dim) for mo in self.irreps_out]
        if weight is None:
            weight = self.weight
        for (i1, i2, io, mode, shape, off), a in zip(self.instructions, self.coeff):
            mi1, mi2, mo = self.irreps_in1[i1], self.irreps_in2[i2], self.irreps_out[io]
            c = wigner_3j(mi1.ir.l, mi2.ir.l, mo.ir.l, dtype=x1.dtype, device=x1.device)
            a1 = x1[:, s1[i1]].reshape(z, mi1.mul, mi1.ir.dim)
            a2 = x2[:, s2[i2]].reshape(z, mi2.mul, mi2.ir.dim)
            w = weight[..., off:off + int(np.prod(shape))]
            w = w.reshape(z, *shape) if w.dim() == 2 else w.reshape(1, *shape).expand(z, *shape)
            eq = {"uvu": "zuv,ijk,zui,zvj->zuk", "uvw": "zuvw,ijk,zui,zvj->zwk"}[mode]
            outs[io] = outs[io] + a * torch.einsum(eq, w, c, a1, a2)
        return torch.cat([t.reshape(z, -1) for t in outs], dim=-1)


class SphericalHarmonics(torch.nn.Module):
    """o3.SphericalHarmonics(irreps_out, normalize, normalization) for irreps 0e+1o+..+lmax."""

    def __init__(self, irreps_out, normalize, normalization="integral", **_):
        super().__init__()
        self.irreps_out = Irreps(irreps_out)
        assert normalization == "component" and [mi.ir.l for mi in self.irreps_out] == list(range(len(self.irreps_out)))
        self.normalize, self.lmax = normalize, len(self.irreps_out) - 1

    def forward(self, x):
        if self.normalize:
            x = torch.nn.functional.normalize(x, dim=-1)
        hats = [torch.ones_like(x[..., :1]), x]
        for l in range(2, self.lmax + 1):
            w, sc = cv.sh_recursion(l)
            hats.append(torch.einsum("ijk,...i,...j->...k", torch.as_tensor(w * sc, dtype=x.dtype), hats[1], hats[l - 1]))
        return torch.cat([math.sqrt(2 * l + 1) * h for l, h in enumerate(hats[:self.lmax + 1])], dim=-1)


class ReducedTensorProducts:
    """Only the instance the reference builds (gnn/blocks.py:430-433): 'ijkl=jikl=ijlk=klij' on four 1o vectors.
    The basis inside the 2x0e / 2x2e multiplicity spaces is OURS, not e3nn's."""

    def __init__(self, formula, **irreps):
        assert formula == "ijkl=jikl=ijlk=klij" and all(str(v) == "1o" for v in irreps.values()), formula
        self.irreps_out = Irreps("2x0e+2x2e+1x4e")
        self.change_of_basis = torch.tensor(cv.rank4_symmetric_basis(), dtype=torch.get_default_dtype())


_NORM2MOM_CACHE = {}


def _normalize2mom_const(act):
    """e3nn.math.normalize2mom: second moment of act on a fixed 1e6-sample N(0,1) draw."""
    if act not in _NORM2MOM_CACHE:
        z = torch.randn(1_000_000, generator=torch.Generator("cpu").manual_seed(0), dtype=torch.float64)
        _NORM2MOM_CACHE[act] = float(act(z).pow(2).mean().pow(-0.5))
    return _NORM2MOM_CACHE[act]


class Gate(torch.nn.Module):
    """e3nn.nn.Gate(irreps_scalars, act_scalars, irreps_gates, act_gates, irreps_gated)."""

    def __init__(self, irreps_scalars, act_scalars, irreps_gates, act_gates, irreps_gated):
        super().__init__()
        self.irreps_scalars, self.irreps_gates, self.irreps_gated = Irreps(irreps_scalars), Irreps(irreps_gates), Irreps(irreps_gated)
        assert self.irreps_gates.num_irreps == self.irreps_gated.num_irreps
        self.act_scalars, self.act_gates = list(act_scalars), list(act_gates)
        self.irreps_in = (self.irreps_scalars + self.irreps_gates + self.irreps_gated)
        self.irreps_out = self.irreps_scalars + self.irreps_gated

    @staticmethod
    def _activate(x, irreps, acts):
        out, i = [], 0
        for mi, act in zip(irreps, acts):
            seg = x[:, i:i + mi.dim]
            out.append(seg if act is None else _normalize2mom_const(act) * act(seg))
            i += mi.dim
        return torch.cat(out, dim=-1) if out else x[:, :0]

    def forward(self, x):
        ns, ng = self.irreps_scalars.dim, self.irreps_gates.dim
        s = self._activate(x[:, :ns], self.irreps_scalars, self.act_scalars)
        g = self._activate(x[:, ns:ns + ng], self.irreps_gates, self.act_gates)
        outs, off, goff = [s], ns + ng, 0
        for mi in self.irreps_gated:
            t = x[:, off:off + mi.dim].reshape(x.shape[0], mi.mul, mi.ir.dim) * g[:, goff:goff + mi.mul, None]
            outs.append(t.reshape(x.shape[0], -1))
            off += mi.dim
            goff += mi.mul
        return torch.cat(outs, dim=-1)


def soft_one_hot_linspace(x, start, end, number, basis=None, cutoff=None):
    assert basis == "gaussian" and cutoff is False
    values = torch.linspace(start, end, number, dtype=x.dtype, device=x.device)
    step = values[1] - values[0]
    diff = (x[..., None] - values) / step
    return diff.pow(2).neg().exp().div(1.12)


def scatter(src, index, dim=-1, out=None, dim_size=None, reduce="sum"):
    """torch_scatter.scatter for dim=0."""
    assert dim in (0, -2) and out is None
    n = int(index.max()) + 1 if dim_size is None else dim_size
    res = torch.zeros((n,) + src.shape[1:], dtype=src.dtype, device=src.device)
    res.index_add_(0, index, src)
    if reduce == "mean":
        cnt = torch.zeros(n, dtype=src.dtype, device=src.device).index_add_(0, index, torch.ones_like(index, dtype=src.dtype))
        res = res / cnt.clamp(min=1).view(-1, *([1] * (src.dim() - 1)))
    else:
        assert reduce in ("sum", "add"), reduce
    return res


class _Bag:
    """torch_geometric.data.Data / Batch stand-in: attribute bag."""

    def __init__(self, **kw):
        self.__dict__.update(kw)

    def __getitem__(self, k):
        return getattr(self, k)


def _module(name, **attrs):
    m = types.ModuleType(name)
    m.__dict__.update(attrs)
    sys.modules[name] = m
    return m


def _load(name, path, package=None):
    spec = importlib.util.spec_from_file_location(name, path, submodule_search_locations=None)
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


_installed = False


def install(root: str = REFERENCE_ROOT):
    """Register the stand-ins and expose the reference packages.  Returns a namespace with the reference modules."""
    global _installed
    if not os.path.isdir(root):
        raise RuntimeError(f"{root} not present (the reference only exists in the build container)")
    if not _installed:
        if not hasattr(np, "float_"):
            np.float_ = np.float64  # removed in numpy 2; used as an annotation/astype target in lattice.py
        ident = lambda *a, **k: (lambda cls: cls)  # noqa: E731
        o3 = _module("e3nn.o3", Irrep=Irrep, Irreps=Irreps, Linear=Linear, TensorProduct=TensorProduct,
                     SphericalHarmonics=SphericalHarmonics, wigner_3j=wigner_3j, ReducedTensorProducts=ReducedTensorProducts)
        nn = _module("e3nn.nn", Gate=Gate, FullyConnectedNet=type("FullyConnectedNet", (torch.nn.Module,), {}))
        mth = _module("e3nn.math", soft_one_hot_linspace=soft_one_hot_linspace)
        io = _module("e3nn.io", CartesianTensor=type("CartesianTensor", (), {}))
        jit = _module("e3nn.util.jit", compile_mode=ident)
        cg = _module("e3nn.util.codegen", CodeGenMixin=type("CodeGenMixin", (), {}))
        util = _module("e3nn.util", jit=jit, codegen=cg)
        _module("e3nn", o3=o3, nn=nn, math=mth, io=io, util=util)
        stub = lambda *a, **k: (_ for _ in ()).throw(NotImplementedError("not on the hot path"))  # noqa: E731
        _module("torch_scatter", scatter=scatter, scatter_min=stub, scatter_max=stub)
        tgd = _module("torch_geometric.data", Data=_Bag, Batch=_Bag, InMemoryDataset=type("InMemoryDataset", (), {}))
        tgu = _module("torch_geometric.utils", degree=stub)
        _module("torch_geometric", data=tgd, utils=tgu)
        _module("opt_einsum", contract=lambda eq, *ops, **k: torch.einsum(eq.replace(" ", ""), *ops))
        _module("opt_einsum_fx")
        # reference packages, without executing their __init__.py
        lat = _module("lattices")
        lat.__path__ = [os.path.join(root, "lattices")]
        ll = _module("lattices.lattices")
        ll.__path__ = [os.path.join(root, "lattices", "lattices")]
        lu = _module("lattices.lattices.utils")
        lu.__path__ = [os.path.join(root, "lattices", "lattices", "utils")]
        ld = _module("lattices.lattices.data")
        ld.__path__ = [os.path.join(root, "lattices", "lattices", "data")]
        ef = _load("lattices.lattices.utils.elasticity_func", os.path.join(root, "lattices/lattices/utils/elasticity_func.py"))
        lu.elasticity_func = ef
        cat = _load("lattices.lattices.data.catalogue", os.path.join(root, "lattices/lattices/data/catalogue.py"))
        lt = _load("lattices.lattices.data.lattice", os.path.join(root, "lattices/lattices/data/lattice.py"))
        ll.Lattice, ll.Catalogue, ll.elasticity_func = lt.Lattice, cat.Catalogue, ef
        g = _module("gnn")
        g.__path__ = [os.path.join(root, "gnn")]
        _installed = True
    import gnn.blocks as blocks  # noqa: E402
    import gnn.datasets as datasets  # noqa: E402
    import gnn.mace as mace  # noqa: E402
    import gnn.model as model  # noqa: E402
    ll = sys.modules["lattices.lattices"]
    return types.SimpleNamespace(model=model, blocks=blocks, mace=mace, datasets=datasets, Lattice=ll.Lattice,
                                 Catalogue=ll.Catalogue, elasticity_func=ll.elasticity_func)
