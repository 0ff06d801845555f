"""ctypes binding of libeeg_b200.so (the C ABI declared in include/eeg_b200.h).

There is no CPU fallback: if the library is missing, or a tensor is not a contiguous CUDA tensor of the stated dtype,
the call raises.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import torch

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libeeg_b200.so")
EEG_MAX_L = 8

SYMBOLS = [
    "eeg_last_error", "eeg_abi_version", "eeg_csr_workspace_bytes", "eeg_csr_build", "eeg_graph_ptr",
    "eeg_edge_featurize", "eeg_block_linear_fwd", "eeg_gate_linear_fwd", "eeg_readout_chain_fwd", "eeg_linear_pair_fwd", "eeg_tp_scatter_fwd", "eeg_sc_coefficients", "eeg_sym_contract_fwd",
    "eeg_gate_fwd", "eeg_pool_tail_fwd", "eeg_node_embed_fwd", "eeg_radial_image_floats", "eeg_radial_pack", "eeg_radial_mlp_fwd",
    # explicit backward ops + optimizer
    "eeg_block_linear_wgrad", "eeg_colsum", "eeg_silu_fwd", "eeg_silu_bwd", "eeg_gate_bwd", "eeg_pool_tail_bwd",
    "eeg_tp_scatter_bwd", "eeg_sym_contract_bwd_x", "eeg_sc_coef_grad", "eeg_sc_coefficients_bwd", "eeg_sc_grad_slices",
    "eeg_sumsq", "eeg_adamw_step", "eeg_rotate_vec3", "eeg_rotate_cart4_to_mandel", "eeg_validation_metrics", "eeg_scale_rows_inv_degree",
    "eeg_collate_gather", "eeg_rand_rotations", "eeg_tp_scatter_umma_fwd", "eeg_segment_sum_rows",
    "eeg_block_linear_wgrad_ws_floats", "eeg_radial_embedding_fwd", "eeg_blocks32_transpose", "eeg_graph_validate", "eeg_radial_hidden_bwd", "eeg_radial_hidden_bwd_ws_floats",
]


class EegLayout(C.Structure):
    _fields_ = [("n_blocks", C.c_int32), ("rows", C.c_int32 * EEG_MAX_L), ("mul", C.c_int32 * EEG_MAX_L),
                ("off", C.c_int32 * EEG_MAX_L), ("stride", C.c_int32)]


def make_layout(rows: Sequence[int], mul: Sequence[int]) -> EegLayout:
    """Blocked m-major layout: block b = [rows_b, mul_b] at the running offset."""
    assert len(rows) == len(mul) <= EEG_MAX_L
    lay = EegLayout()
    lay.n_blocks = len(rows)
    off = 0
    for b, (r, m) in enumerate(zip(rows, mul)):
        lay.rows[b], lay.mul[b], lay.off[b] = r, m, off
        off += r * m
    lay.stride = off
    return lay


def sub_layout(lay: EegLayout, blocks: Sequence[int]) -> EegLayout:
    """The given irrep blocks of `lay` as a layout of their own (same node stride, same offsets inside the node row)."""
    sub = EegLayout()
    sub.n_blocks = len(blocks)
    for q, b in enumerate(blocks):
        sub.rows[q], sub.mul[q], sub.off[q] = lay.rows[b], lay.mul[b], lay.off[b]
    sub.stride = lay.stride
    return sub


_lib = None


def lib() -> C.CDLL:
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m energy_equiv_lattice_gnn_b200.build` "
                "(or __graft_entry__.build()); there is no CPU/PyTorch fallback for the hot path.")
        _lib = C.CDLL(LIB_PATH)
        _lib.eeg_last_error.restype = C.c_char_p
        _lib.eeg_csr_workspace_bytes.restype = C.c_size_t
        _lib.eeg_csr_workspace_bytes.argtypes = [C.c_int64, C.c_int64]
        for s in SYMBOLS:
            getattr(_lib, s)  # fail loudly on a stale library
    return _lib


# Optional per-entry-point CUDA-event timing (bench.py roofline leg) and launch accounting.
# KERNELS_PER_CALL = kernels one C-ABI call launches (memsets excluded).
KERNELS_PER_CALL = {"eeg_csr_build": 6, "eeg_graph_ptr": 1, "eeg_node_embed_fwd": 1, "eeg_edge_featurize": 1,
                    "eeg_block_linear_fwd": 1, "eeg_gate_linear_fwd": 1, "eeg_tp_scatter_fwd": 1, "eeg_sc_coefficients": 1,
                    "eeg_sym_contract_fwd": 1, "eeg_gate_fwd": 1, "eeg_pool_tail_fwd": 1,
                    "eeg_radial_pack": 1, "eeg_radial_mlp_fwd": 1, "eeg_pool_tail_bwd": 2, "eeg_readout_chain_fwd": 2, "eeg_linear_pair_fwd": 1}
launch_count = 0
_events = None  # list of (name, tag, start_event, end_event) while profiling
_tag = ""


def set_tag(tag: str):
    """Label subsequent calls (e.g. 'L1.mlp3') for the per-kernel timing table."""
    global _tag
    _tag = tag


def get_tag() -> str:
    return _tag


def start_profile():
    global _events
    _events = []


def stop_profile():
    """-> {tag or name: [ms, ...]} (synchronises)."""
    global _events
    torch.cuda.synchronize()
    out = {}
    for name, tag, a, b in _events:
        out.setdefault(tag or name, []).append(a.elapsed_time(b))
    _events = None
    return out


def _check(rc: int, what: str):
    global launch_count
    if rc != 0:
        raise RuntimeError(f"{what} failed (code {rc}): {lib().eeg_last_error().decode()}")
    launch_count += KERNELS_PER_CALL.get(what, 1)
    if _events is not None and _events and _events[-1][3] is None:
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        _events[-1] = _events[-1][:3] + (ev,)


def _begin(what: str):
    if _events is not None:
        ev = torch.cuda.Event(enable_timing=True)
        ev.record()
        _events.append((what, _tag, ev, None))


def _p(t: Optional[torch.Tensor], dtype=None):
    if t is None:
        return C.c_void_p(0)
    if not t.is_cuda:
        raise RuntimeError("libeeg_b200 needs CUDA tensors (no CPU fallback)")
    if not t.is_contiguous():
        raise RuntimeError("libeeg_b200 needs contiguous tensors")
    if dtype is not None and t.dtype != dtype:
        raise RuntimeError(f"expected dtype {dtype}, got {t.dtype}")
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


f32, i32, i64 = torch.float32, torch.int32, torch.int64


def csr_build(edge_index: torch.Tensor, n_nodes: int):
    E = edge_index.shape[1]
    dev = edge_index.device
    row_ptr = torch.empty(n_nodes + 1, dtype=i32, device=dev)
    perm = torch.empty(E, dtype=i32, device=dev)
    nbytes = lib().eeg_csr_workspace_bytes(E, n_nodes)
    ws = torch.empty(nbytes, dtype=torch.uint8, device=dev)
    _begin("eeg_csr_build")
    _check(lib().eeg_csr_build(_p(edge_index, i64), C.c_int64(E), C.c_int64(n_nodes), _p(row_ptr), _p(perm), _p(ws),
                               C.c_size_t(nbytes), _stream()), "eeg_csr_build")
    return row_ptr, perm


def graph_validate(edge_index: torch.Tensor, n_nodes: int, batch: Optional[torch.Tensor], num_graphs: int) -> None:
    """Raise ValueError on an out-of-range edge_index / batch entry or a decreasing `batch` (one small kernel + a host read)."""
    flags = torch.empty(1, dtype=i32, device=edge_index.device)
    _begin("eeg_graph_validate")
    _check(lib().eeg_graph_validate(_p(edge_index, i64), C.c_int64(edge_index.shape[1]), C.c_int64(n_nodes), _p(batch, i64) if batch is not None
                                    else C.c_void_p(0), C.c_int64(num_graphs), _p(flags), _stream()), "eeg_graph_validate")
    f = int(flags.item())
    if f:
        what = [m for bit, m in ((1, f"edge_index has entries outside [0, {n_nodes})"), (2, f"batch has entries outside [0, {num_graphs})"),
                                 (4, "batch is not non-decreasing")) if f & bit]
        raise ValueError("malformed graph batch: " + "; ".join(what))


def graph_ptr(batch: torch.Tensor, num_graphs: int):
    gp = torch.empty(num_graphs + 1, dtype=i32, device=batch.device)
    _begin("eeg_graph_ptr")
    _check(lib().eeg_graph_ptr(_p(batch, i64), C.c_int64(batch.shape[0]), C.c_int64(num_graphs), _p(gp), _stream()), "eeg_graph_ptr")
    return gp


def node_embed(node_attrs, W, b):
    N, Cch = node_attrs.shape[0], W.shape[0]
    out = torch.empty(N, Cch, dtype=f32, device=node_attrs.device)
    _begin("eeg_node_embed_fwd")
    _check(lib().eeg_node_embed_fwd(_p(node_attrs, f32), _p(W, f32), _p(b, f32), _p(out), C.c_int64(N), C.c_int(Cch), _stream()),
           "eeg_node_embed_fwd")
    return out


def edge_featurize(pos, edge_index, shifts, radius, perm, lmax, nbasis, len_end, rmax):
    E = edge_index.shape[1]
    D = (lmax + 1) ** 2
    ystride = (D + 3) // 4 * 4
    dev = pos.device
    Y = torch.empty(E, ystride, dtype=f32, device=dev)
    f = torch.empty(E, 2 * nbasis, dtype=f32, device=dev)
    snd = torch.empty(E, dtype=i32, device=dev)
    _begin("eeg_edge_featurize")
    _check(lib().eeg_edge_featurize(_p(pos, f32), _p(edge_index, i64), _p(shifts, f32), _p(radius, f32), _p(perm, i32),
                                    C.c_int64(E), C.c_int(lmax), C.c_int(nbasis), C.c_float(len_end), C.c_float(rmax),
                                    _p(Y), C.c_int(ystride), _p(f), _p(snd), _stream()), "eeg_edge_featurize")
    return Y, f, snd


def block_linear(x, lin: EegLayout, W, w_off: Sequence[int], scale: Sequence[float], lout: EegLayout, n_rows: int,
                 bias=None, transB=False, act=0, residual=None, out=None):
    nb = lin.n_blocks
    if out is None:
        out = torch.empty(n_rows, lout.stride, dtype=f32, device=x.device)
    woff = (C.c_int64 * EEG_MAX_L)(*list(w_off) + [0] * (EEG_MAX_L - nb))
    sc = (C.c_float * EEG_MAX_L)(*list(scale) + [0.0] * (EEG_MAX_L - nb))
    _begin("eeg_block_linear_fwd")
    _check(lib().eeg_block_linear_fwd(_p(x, f32), C.byref(lin), _p(W, f32), woff, sc, _p(bias, f32), C.c_int(int(transB)),
                                      C.c_int(act), _p(residual, f32), _p(out, f32), C.byref(lout), C.c_int64(n_rows),
                                      _stream()), "eeg_block_linear_fwd")
    return out


def gate_linear(z, lz: EegLayout, n_scalars: int, act_const: float, W, w_off: Sequence[int], scale: Sequence[float],
                lout: EegLayout, n_rows: int):
    """Gate fused into the following o3.Linear (tensor-core path). Returns None when the shape is not supported."""
    nb = lz.n_blocks
    if any(lz.mul[b] % 32 for b in range(nb)) or n_scalars % 32 or any(lout.mul[b] % 16 for b in range(nb)):
        return None
    out = torch.empty(n_rows, lout.stride, dtype=f32, device=z.device)
    woff = (C.c_int64 * EEG_MAX_L)(*list(w_off) + [0] * (EEG_MAX_L - nb))
    sc = (C.c_float * EEG_MAX_L)(*list(scale) + [0.0] * (EEG_MAX_L - nb))
    _begin("eeg_gate_linear_fwd")
    _check(lib().eeg_gate_linear_fwd(_p(z, f32), C.byref(lz), C.c_int(n_scalars), C.c_float(act_const), _p(W, f32), woff, sc,
                                     _p(out, f32), C.byref(lout), C.c_int64(n_rows), _stream()), "eeg_gate_linear_fwd")
    return out


def linear_pair(x, residual, lh: EegLayout, Wa, wa_off, sa, Wb, wb_off, sb, n_rows: int):
    """y1 = Wa x (+ residual), y2 = Wb y1 in one pass (multiplicity-32 blocks).  Returns (y1, y2) or None if unsupported."""
    nb = lh.n_blocks
    if any(lh.mul[b] != 32 for b in range(nb)) or nb > 5:
        return None
    y1, y2 = torch.empty_like(x), torch.empty_like(x)
    arr_i = lambda v: (C.c_int64 * EEG_MAX_L)(*list(v) + [0] * (EEG_MAX_L - nb))  # noqa: E731
    arr_f = lambda v: (C.c_float * EEG_MAX_L)(*list(v) + [0.0] * (EEG_MAX_L - nb))  # noqa: E731
    _begin("eeg_linear_pair_fwd")
    _check(lib().eeg_linear_pair_fwd(_p(x, f32), _p(residual, f32) if residual is not None else None, C.byref(lh), _p(Wa, f32),
                                     arr_i(wa_off), arr_f(sa), _p(Wb, f32), arr_i(wb_off), arr_f(sb), _p(y1), _p(y2),
                                     C.c_int64(n_rows), _stream()), "eeg_linear_pair_fwd")
    return y1, y2


def readout_chain(B, residual, lh: EegLayout, Wp, wp_off, sp, W1, w1_off, s1, W2, w2_off, s2, act_const: float, lout: EegLayout,
                  n_rows: int):
    """Inference tail product.linear(+residual) -> linear_1 -> Gate -> linear_2 in one pass (multiplicity-32 models).
    Returns o [N, lout.stride], or None when the shape is outside the fused kernel."""
    nb = lh.n_blocks
    if any(lh.mul[b] != 32 for b in range(nb)) or any(not (1 <= lout.mul[b] <= 32) for b in range(nb)) or nb > 5:
        return None
    out = torch.empty(n_rows, lout.stride, dtype=f32, device=B.device)
    gates = torch.empty(n_rows, max(1, 32 * (nb - 1)), dtype=f32, device=B.device)
    arr_i = lambda v: (C.c_int64 * EEG_MAX_L)(*list(v) + [0] * (EEG_MAX_L - nb))  # noqa: E731
    arr_f = lambda v: (C.c_float * EEG_MAX_L)(*list(v) + [0.0] * (EEG_MAX_L - nb))  # noqa: E731
    _begin("eeg_readout_chain_fwd")
    _check(lib().eeg_readout_chain_fwd(_p(B, f32), _p(residual, f32) if residual is not None else None, C.byref(lh), _p(Wp, f32),
                                       arr_i(wp_off), arr_f(sp), _p(W1, f32), arr_i(w1_off), arr_f(s1), _p(W2, f32), arr_i(w2_off),
                                       arr_f(s2), C.c_float(act_const), _p(gates), _p(out), C.byref(lout), C.c_int64(n_rows), _stream()),
           "eeg_readout_chain_fwd")
    return out


def radial_pack(Wh: Optional[torch.Tensor], W2: torch.Tensor) -> torch.Tensor:
    """Hidden 64x64 weights [n_hidden,64,64] (or None) + last-layer weight [n_out,64] -> staged hi/lo operand image."""
    n_hidden = 0 if Wh is None else Wh.shape[0]
    n_out, H = W2.shape
    L = lib()
    L.eeg_radial_image_floats.restype = C.c_int64
    img = torch.empty(L.eeg_radial_image_floats(C.c_int(n_hidden), C.c_int(n_out)), dtype=f32, device=W2.device)
    _begin("eeg_radial_pack")
    _check(L.eeg_radial_pack(_p(Wh, f32), C.c_int(n_hidden), C.c_int(H), _p(W2, f32), C.c_int(n_out), _p(img), _stream()),
           "eeg_radial_pack")
    return img


def radial_mlp(f, W0, b0, bh, n_hidden, img, n_out):
    """-> w_t [n_out/32, E, 32] path-major per-edge tensor-product weights."""
    E = f.shape[0]
    w_t = torch.empty(n_out // 32, E, 32, dtype=f32, device=f.device)
    _begin("eeg_radial_mlp_fwd")
    _check(lib().eeg_radial_mlp_fwd(_p(f, f32), C.c_int(f.shape[1]), _p(W0, f32), _p(b0, f32), _p(bh, f32), C.c_int(n_hidden),
                                    C.c_int(W0.shape[0]), _p(img, f32), C.c_int(n_out), _p(w_t), C.c_int64(E), _stream()),
           "eeg_radial_mlp_fwd")
    return w_t


def radial_supported(n_in: int, hidden_dim: int, n_linear: int) -> bool:
    """Shape domain of the fused tcgen05 radial MLP (eeg_radial_mlp_fwd): inter_MLP_dim 64, >= 2 linears, <= 16 inputs."""
    return hidden_dim == 64 and 2 <= n_linear <= 6 and 0 < n_in <= 16


def to_path_major(w_edge_major, n_out: int):
    """[E, n_out] -> w_t [n_out/32, E, 32]"""
    E = w_edge_major.shape[0]
    out = torch.empty(n_out // 32, E, 32, dtype=f32, device=w_edge_major.device)
    _begin("eeg_blocks32_transpose")
    _check(lib().eeg_blocks32_transpose(_p(w_edge_major, f32), _p(out), C.c_int64(E), C.c_int(n_out // 32), C.c_int(1), _stream()),
           "eeg_blocks32_transpose")
    return out


def tp_scatter(lmax_in, lmax, Cch, h, Y, w_t, sender_sorted, row_ptr, inv_norm, rows_total, n_nodes, umma: bool = False):
    """umma=True calls the tensor-core variant (eeg_tp_scatter_umma_fwd: C = 32, lmax_in = lmax only)."""
    M = torch.empty(n_nodes, rows_total * Cch, dtype=f32, device=h.device)
    E = w_t.shape[1]
    name = "eeg_tp_scatter_umma_fwd" if umma else "eeg_tp_scatter_fwd"
    _begin(name)
    if umma:
        _check(lib().eeg_tp_scatter_umma_fwd(C.c_int(lmax_in), C.c_int(lmax), C.c_int(Cch), _p(h, f32), _p(Y, f32), C.c_int(Y.shape[1]),
                                             _p(w_t, f32), C.c_int64(E), _p(sender_sorted, i32), _p(row_ptr, i32), C.c_float(inv_norm),
                                             _p(M), C.c_int64(n_nodes), _stream()), name)
        return M
    _check(lib().eeg_tp_scatter_fwd(C.c_int(lmax_in), C.c_int(lmax), C.c_int(Cch), _p(h, f32), _p(Y, f32), C.c_int(Y.shape[1]),
                                    _p(w_t, f32), C.c_int64(E), _p(sender_sorted, i32), _p(row_ptr, i32), C.c_float(inv_norm),
                                    _p(M), C.c_int64(n_nodes), _stream()), "eeg_tp_scatter_fwd")
    return M


def sc_coefficients(a_row_ptr, a_col, a_val, n_coef, Wflat, Cch):
    T = torch.empty(Cch, n_coef, dtype=f32, device=Wflat.device)
    _begin("eeg_sc_coefficients")
    _check(lib().eeg_sc_coefficients(_p(a_row_ptr, i32), _p(a_col, i32), _p(a_val, f32), C.c_int(n_coef), _p(Wflat, f32),
                                     C.c_int(Cch), _p(T), _stream()), "eeg_sc_coefficients")
    return T


def sym_contract(lmax, nu, Cch, x, T, n_nodes):
    out = torch.empty_like(x)
    _begin("eeg_sym_contract_fwd")
    _check(lib().eeg_sym_contract_fwd(C.c_int(lmax), C.c_int(nu), C.c_int(Cch), _p(x, f32), _p(T, f32), _p(out),
                                      C.c_int64(n_nodes), _stream()), "eeg_sym_contract_fwd")
    return out


def gate(z, lz: EegLayout, lout: EegLayout, n_scalars: int, act_const: float, n_nodes: int):
    out = torch.empty(n_nodes, lout.stride, dtype=f32, device=z.device)
    _begin("eeg_gate_fwd")
    _check(lib().eeg_gate_fwd(_p(z, f32), C.byref(lz), _p(out), C.byref(lout), C.c_int(n_scalars), C.c_float(act_const),
                              C.c_int64(n_nodes), _stream()), "eeg_gate_fwd")
    return out


def pool_tail(o, lo: EegLayout, gp, Wq, qmul: Sequence[int], qscale: Sequence[float], bias, P, positive: int, reduce: int,
              num_graphs: int, want_aux: bool = False):
    dev = o.device
    nq = sum(q * lo.rows[b] for b, q in enumerate(qmul))
    Cout = torch.empty(num_graphs, 6, 6, dtype=f32, device=dev)
    q_out = torch.empty(num_graphs, nq, dtype=f32, device=dev) if want_aux else None
    A_out = torch.empty(num_graphs, 6, 6, dtype=f32, device=dev) if want_aux else None
    nb = lo.n_blocks
    qm = (C.c_int32 * EEG_MAX_L)(*list(qmul) + [0] * (EEG_MAX_L - nb))
    qs = (C.c_float * EEG_MAX_L)(*list(qscale) + [0.0] * (EEG_MAX_L - nb))
    _begin("eeg_pool_tail_fwd")
    _check(lib().eeg_pool_tail_fwd(_p(o, f32), C.byref(lo), _p(gp, i32), _p(Wq, f32), qm, qs, _p(bias, f32), _p(P, f32),
                                   C.c_int(positive), C.c_int(reduce), _p(q_out), _p(A_out), _p(Cout), C.c_int64(num_graphs),
                                   _stream()), "eeg_pool_tail_fwd")
    return Cout, q_out, A_out


# ---------------------------------------------------------------------------------------------------------------------
# explicit backward ops (include/eeg_b200.h, "backward" section)
def _arr(ctype, vals, n=EEG_MAX_L, fill=0):
    vals = list(vals)
    return (ctype * n)(*(vals + [fill] * (n - len(vals))))


def linear_wgrad(x, lin: EegLayout, dy, lout: EegLayout, w_off, scale, n_weights: int, n_rows: int, transB=False, want_bias=0):
    """-> (dW [n_weights], dbias [want_bias] or None); dW_b[u,w] = scale_b sum_rows x[.,u] dy[.,w]."""
    dW = torch.zeros(n_weights, dtype=f32, device=x.device)
    db = torch.zeros(want_bias, dtype=f32, device=x.device) if want_bias else None
    L = lib()
    L.eeg_block_linear_wgrad_ws_floats.restype = C.c_int64
    nws = int(L.eeg_block_linear_wgrad_ws_floats(C.byref(lin), C.byref(lout), C.c_int64(n_weights), C.c_int(want_bias), C.c_int64(n_rows)))
    ws = torch.empty(max(nws, 1), dtype=f32, device=x.device)   # deterministic mode: row slices are added in fixed order
    _begin("eeg_block_linear_wgrad")
    _check(L.eeg_block_linear_wgrad(_p(x, f32), C.byref(lin), _p(dy, f32), C.byref(lout), _arr(C.c_int64, w_off), _arr(C.c_float, scale, fill=0.0),
                                    C.c_int(int(transB)), C.c_void_p(0), C.c_void_p(0), _p(dW), _p(db), C.c_int64(n_rows),
                                    C.c_int64(n_weights), _p(ws), C.c_int64(ws.numel()), _stream()),
           "eeg_block_linear_wgrad")
    return dW, db


def radial_hidden_bwd(f, F: int, W0, b0, W1, b1, ds2, E: int):
    """Fused backward of Linear(F,64)-SiLU-Linear(64,64)-SiLU: -> (s2 [E,64], dW0 [64,F], db0 [64], dW1 [64,64], db1 [64])."""
    dev = f.device
    s2 = torch.empty(E, 64, dtype=f32, device=dev)
    dW0 = torch.zeros(64, F, dtype=f32, device=dev)
    db0 = torch.zeros(64, dtype=f32, device=dev)
    dW1 = torch.zeros(64, 64, dtype=f32, device=dev)
    db1 = torch.zeros(64, dtype=f32, device=dev)
    L = lib()
    L.eeg_radial_hidden_bwd_ws_floats.restype = C.c_int64
    ws = torch.empty(max(int(L.eeg_radial_hidden_bwd_ws_floats(C.c_int(F), C.c_int64(E))), 1), dtype=f32, device=dev)
    _begin("eeg_radial_hidden_bwd")
    _check(L.eeg_radial_hidden_bwd(_p(f, f32), C.c_int(F), C.c_int(f.shape[1]), _p(W0, f32), _p(b0, f32), _p(W1, f32), _p(b1, f32),
                                   _p(ds2, f32), C.c_int64(E), _p(s2), _p(dW0), _p(db0), _p(dW1), _p(db1), _p(ws), C.c_int64(ws.numel()),
                                   _stream()), "eeg_radial_hidden_bwd")
    return s2, dW0, db0, dW1, db1


def colsum(x, row_scale=None):
    """out[c] = sum_r row_scale[r] * x[r, c] for a 2-D contiguous x."""
    R, cols = x.shape
    out = torch.zeros(cols, dtype=f32, device=x.device)
    ws = torch.empty(((R + 255) // 256) * cols + 1, dtype=f32, device=x.device)
    _begin("eeg_colsum")
    _check(lib().eeg_colsum(_p(x, f32), C.c_int64(R), C.c_int(cols), C.c_int(cols), _p(row_scale, f32), _p(out), _p(ws),
                            C.c_int64(ws.numel()), _stream()), "eeg_colsum")
    return out


def silu_fwd(a):
    s = torch.empty_like(a)
    _begin("eeg_silu_fwd")
    _check(lib().eeg_silu_fwd(_p(a, f32), _p(s), C.c_int64(a.numel()), _stream()), "eeg_silu_fwd")
    return s


def silu_bwd(a, ds):
    da = torch.empty_like(a)
    _begin("eeg_silu_bwd")
    _check(lib().eeg_silu_bwd(_p(a, f32), _p(ds, f32), _p(da), C.c_int64(a.numel()), _stream()), "eeg_silu_bwd")
    return da


def gate_bwd(z, lz: EegLayout, dg, lg: EegLayout, n_scalars: int, act_const: float, n_nodes: int):
    dz = torch.empty_like(z)
    _begin("eeg_gate_bwd")
    _check(lib().eeg_gate_bwd(_p(z, f32), C.byref(lz), _p(dg, f32), C.byref(lg), C.c_int(n_scalars), C.c_float(act_const), _p(dz),
                              C.c_int64(n_nodes), _stream()), "eeg_gate_bwd")
    return dz


def pool_tail_bwd(o, lo: EegLayout, gp, Wq, qmul, qscale, bias_len: int, P, positive: int, reduce: int, A, dC, num_graphs: int):
    """-> (d_o [N, lo.stride], dWq, dbias)"""
    dev = o.device
    nb = lo.n_blocks
    nq = sum(q * lo.rows[b] for b, q in enumerate(qmul))
    npool = sum(lo.rows[b] * lo.mul[b] for b, q in enumerate(qmul) if q > 0)
    d_o = torch.empty_like(o)
    dq = torch.empty(num_graphs, nq, dtype=f32, device=dev)
    pooled = torch.empty(num_graphs, npool, dtype=f32, device=dev)
    dWq = torch.empty_like(Wq)
    db = torch.empty(bias_len, dtype=f32, device=dev) if bias_len else None
    _begin("eeg_pool_tail_bwd")
    _check(lib().eeg_pool_tail_bwd(_p(o, f32), C.byref(lo), _p(gp, i32), _p(Wq, f32), _arr(C.c_int32, qmul), _arr(C.c_float, qscale, fill=0.0),
                                   _p(P, f32), C.c_int(positive), C.c_int(reduce), _p(A, f32), _p(dC, f32), _p(d_o), _p(dq), _p(pooled),
                                   _p(dWq), _p(db), C.c_int64(num_graphs), _stream()), "eeg_pool_tail_bwd")
    return d_o, dWq, db


def sender_csr(sender_sorted, n_nodes: int):
    """CSR of the receiver-sorted edge list by SENDER: (ptr int32 [N+1], perm int32 [E]), stable -> fixed summation order."""
    E = sender_sorted.shape[0]
    ei = torch.stack([torch.zeros(E, dtype=i64, device=sender_sorted.device), sender_sorted.to(i64)]).contiguous()
    return csr_build(ei, n_nodes)


def tp_scatter_bwd(lmax_in, lmax, Cch, h, Y, w_t, sender_sorted, row_ptr, inv_norm, dM, n_paths, n_nodes, snd_csr=None):
    """-> (dw [E, n_paths*C] edge-major CSR order, dh [N, (lmax_in+1)^2 * C]).  No atomics: per-edge input gradients are
    written edge-major and summed per sender by a CSR walk (snd_csr = sender_csr(sender_sorted, N), built here if omitted)."""
    E = w_t.shape[1]
    dw = torch.empty(E, n_paths * Cch, dtype=f32, device=h.device)
    width = h.shape[1]
    dxe = torch.empty(E, width, dtype=f32, device=h.device)
    _begin("eeg_tp_scatter_bwd")
    _check(lib().eeg_tp_scatter_bwd(C.c_int(lmax_in), C.c_int(lmax), C.c_int(Cch), _p(h, f32), _p(Y, f32), C.c_int(Y.shape[1]), _p(w_t, f32),
                                    C.c_int64(E), _p(sender_sorted, i32), _p(row_ptr, i32), C.c_float(inv_norm), _p(dM, f32), _p(dw), _p(dxe),
                                    C.c_int64(n_nodes), _stream()), "eeg_tp_scatter_bwd")
    if snd_csr is None:
        snd_csr = sender_csr(sender_sorted, n_nodes)
    ptr, perm = snd_csr
    dh = torch.empty_like(h)
    _begin("eeg_segment_sum_rows")
    _check(lib().eeg_segment_sum_rows(_p(dxe, f32), _p(perm, i32), _p(ptr, i32), C.c_int64(n_nodes), C.c_int64(width), _p(dh), _stream()),
           "eeg_segment_sum_rows")
    return dw, dh


def sym_contract_bwd_x(lmax, nu, Cch, x, dout, T, n_nodes):
    dx = torch.empty_like(x)
    _begin("eeg_sym_contract_bwd_x")
    _check(lib().eeg_sym_contract_bwd_x(C.c_int(lmax), C.c_int(nu), C.c_int(Cch), _p(x, f32), _p(dout, f32), _p(T, f32), _p(dx),
                                        C.c_int64(n_nodes), _stream()), "eeg_sym_contract_bwd_x")
    return dx


def sc_weight_grad(lmax, nu, x, dout, n_coef, Cch, n_nodes, t_ptr, t_row, t_val, n_wrows):
    """-> dWflat [n_wrows, C]: coefficient gradient (generated program, sliced over nodes) folded back onto the weights."""
    L = lib()
    Z = L.eeg_sc_grad_slices()
    dT = torch.empty(Z, n_coef, Cch, dtype=f32, device=x.device)
    _begin("eeg_sc_coef_grad")
    _check(L.eeg_sc_coef_grad(C.c_int(lmax), C.c_int(nu), C.c_int(Cch), _p(x, f32), _p(dout, f32), C.c_int(n_coef), C.c_int64(n_nodes),
                              _p(dT), _stream()), "eeg_sc_coef_grad")
    dW = torch.empty(n_wrows, Cch, dtype=f32, device=x.device)
    _begin("eeg_sc_coefficients_bwd")
    _check(L.eeg_sc_coefficients_bwd(_p(t_ptr, i32), _p(t_row, i32), _p(t_val, f32), C.c_int(n_wrows), C.c_int(n_coef), _p(dT), C.c_int(Cch),
                                     _p(dW), _stream()), "eeg_sc_coefficients_bwd")
    return dW


def sumsq(x, out):
    """out[0] += sum(x^2) (global gradient norm for clipping)."""
    _begin("eeg_sumsq")
    _check(lib().eeg_sumsq(_p(x, f32), C.c_int64(x.numel()), _p(out, f32), _stream()), "eeg_sumsq")


def adamw_step(p, g, m, v, vmax, lr, beta1, beta2, eps, weight_decay, step, gnorm_sq, max_norm, grad_scale):
    """Fused AdamW (amsgrad when vmax is given) on flat fp32 buffers; clips by the global norm sqrt(gnorm_sq[0])."""
    _begin("eeg_adamw_step")
    _check(lib().eeg_adamw_step(_p(p, f32), _p(g, f32), _p(m, f32), _p(v, f32), _p(vmax, f32), C.c_int64(p.numel()), C.c_float(lr), C.c_float(beta1),
                                C.c_float(beta2), C.c_float(eps), C.c_float(weight_decay), C.c_int(step), _p(gnorm_sq, f32), C.c_float(max_norm),
                                C.c_float(grad_scale), _stream()), "eeg_adamw_step")


def rotate_vec3(Q, graph_of, via, x, out):
    """out[i] = Q[graph_of[via[i] if via is not None else i]] @ x[i]  (RotateLat, scripts/train_utils.py:124-125)."""
    _begin("eeg_rotate_vec3")
    _check(lib().eeg_rotate_vec3(_p(Q, f32), _p(graph_of, i64), _p(via, i64), _p(x, f32), _p(out, f32), C.c_int64(x.shape[0]), _stream()),
           "eeg_rotate_vec3")


def rotate_cart4_to_mandel(Q, C4, out):
    """out[g] = Mandel(Q_g x4 C4[g])  (scripts/train_utils.py:122-123,133-134); Q None = no rotation."""
    _begin("eeg_rotate_cart4_to_mandel")
    _check(lib().eeg_rotate_cart4_to_mandel(_p(Q, f32), _p(C4, f32), _p(out, f32), C.c_int64(C4.shape[0]), _stream()),
           "eeg_rotate_cart4_to_mandel")


def validation_metrics(pred, target, directions, per_lattice, eig_pred=None, eig_target=None):
    """per_lattice[g] = (mse, l1, dir_loss, mean_stiffness); optional ascending eigenvalues (scripts/train_utils.py:149-172)."""
    _begin("eeg_validation_metrics")
    _check(lib().eeg_validation_metrics(_p(pred, f32), _p(target, f32), _p(directions, f32), C.c_int(directions.shape[0]),
                                        C.c_int64(pred.shape[0]), _p(per_lattice, f32), _p(eig_pred, f32), _p(eig_target, f32),
                                        _stream()), "eeg_validation_metrics")


def scale_rows_inv_degree(x, row_ptr):
    """x[n, :] /= max(in-degree(n), 1) in place (interaction_reduction='mean', gnn/blocks.py:595-597)."""
    _begin("eeg_scale_rows_inv_degree")
    _check(lib().eeg_scale_rows_inv_degree(_p(x, f32), _p(row_ptr, i32), C.c_int64(x.shape[0]), C.c_int64(x.numel() // max(x.shape[0], 1)),
                                           _stream()), "eeg_scale_rows_inv_degree")


def collate_gather(ds_pos, ds_nattr, ds_snd, ds_rcv, ds_shifts, ds_eattr, meta, G: int, N: int, E: int, graph_rows=None, graph_idx=None):
    """meta: int64 device [4, G+1] = src_node_off, src_edge_off, out_node_ptr, out_edge_ptr (rows 0-1 use the first G entries).
    -> (positions, node_attrs [N,1], edge_index [2,E], shifts, edge_attr [E,1], batch, gathered graph rows or None)"""
    dev = ds_pos.device
    o_pos = torch.empty(N, 3, dtype=f32, device=dev)
    o_na = torch.empty(N, 1, dtype=f32, device=dev)
    o_ei = torch.empty(2, E, dtype=i64, device=dev)
    o_sh = torch.empty(E, 3, dtype=f32, device=dev)
    o_ea = torch.empty(E, 1, dtype=f32, device=dev)
    o_b = torch.empty(N, dtype=i64, device=dev)
    width = 0 if graph_rows is None else graph_rows.shape[1]
    o_g = torch.empty(G, width, dtype=f32, device=dev) if graph_rows is not None else None
    _begin("eeg_collate_gather")
    _check(lib().eeg_collate_gather(_p(ds_pos, f32), _p(ds_nattr, f32), _p(ds_snd, i32), _p(ds_rcv, i32), _p(ds_shifts, f32),
                                    _p(ds_eattr, f32), _p(meta[0], i64), _p(meta[1], i64), _p(meta[2], i64), _p(meta[3], i64),
                                    C.c_int64(G), C.c_int64(N), C.c_int64(E), _p(o_pos), _p(o_na), _p(o_ei), _p(o_sh), _p(o_ea), _p(o_b),
                                    _p(graph_rows, f32) if graph_rows is not None else C.c_void_p(0),
                                    _p(graph_idx, i64) if graph_idx is not None else C.c_void_p(0), C.c_int64(width), _p(o_g), _stream()),
           "eeg_collate_gather")
    return o_pos, o_na, o_ei, o_sh, o_ea, o_b, o_g


def radial_embedding(edge_lengths, num_bessel: int, r_max: float, num_polynomial_cutoff: int) -> torch.Tensor:
    """RadialEmbeddingBlock.forward (gnn/blocks.py:96-102): [E,1] or [E] -> [E, num_bessel]."""
    x = edge_lengths.reshape(-1).contiguous()
    out = torch.empty(x.shape[0], num_bessel, dtype=f32, device=x.device)
    _begin("eeg_radial_embedding_fwd")
    _check(lib().eeg_radial_embedding_fwd(_p(x, f32), C.c_int64(x.shape[0]), C.c_int(num_bessel), C.c_float(r_max),
                                          C.c_int(num_polynomial_cutoff), _p(out), _stream()), "eeg_radial_embedding_fwd")
    return out


def rand_rotations(seed: int, offset: int, G: int, device) -> torch.Tensor:
    Q = torch.empty(G, 3, 3, dtype=f32, device=device)
    _begin("eeg_rand_rotations")
    _check(lib().eeg_rand_rotations(C.c_uint64(seed & (2 ** 64 - 1)), C.c_uint64(offset), C.c_int64(G), _p(Q), _stream()), "eeg_rand_rotations")
    return Q


# ---- device guard -----------------------------------------------------------------------------------------------------
# The C-ABI launchers run on the CUDA device that is current in the calling thread and take the stream they are given.
# Every binding above therefore runs with the device of its first CUDA tensor argument made current (so `_stream()` is
# that device's current stream and the launch, its >48 KB shared-memory opt-ins included, lands on the tensors' device
# even when the caller never called torch.cuda.set_device).
def _first_cuda_device(args, kwargs):
    for a in list(args) + list(kwargs.values()):
        if isinstance(a, torch.Tensor) and a.is_cuda:
            return a.device
        if isinstance(a, (list, tuple)):
            for b in a:
                if isinstance(b, torch.Tensor) and b.is_cuda:
                    return b.device
    return None


def _guarded(fn):
    import functools

    @functools.wraps(fn)
    def wrapper(*args, **kwargs):
        dev = _first_cuda_device(args, kwargs)
        if dev is None or dev.index == torch.cuda.current_device():
            return fn(*args, **kwargs)
        with torch.cuda.device(dev):
            return fn(*args, **kwargs)
    return wrapper


def _install_guards():
    import inspect
    g = globals()
    skip = {"lib", "make_layout", "set_tag", "get_tag", "start_profile", "stop_profile"}
    for name, fn in list(g.items()):
        if name.startswith("_") or name in skip or not inspect.isfunction(fn) or fn.__module__ != __name__:
            continue
        g[name] = _guarded(fn)


_install_guards()
