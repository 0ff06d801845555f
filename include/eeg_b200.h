/* eeg_b200.h - C ABI of the B200-native EnergyEquivGNN hot path (libeeg_b200.so).
 *
 * The reference (igrega348/energy-equiv-lattice-gnn) has no native code and no FFI: its arithmetic lives in
 * e3nn / torch_scatter / opt_einsum calls made from gnn/model.py, gnn/blocks.py and gnn/mace.py.  Each entry point
 * below replaces one such call site (cited per function).  Conventions:
 *   - every pointer is a DEVICE pointer unless named host_*; tensors are dense row-major fp32, indices as stated;
 *   - no allocation inside: caller passes outputs and workspace; `stream` is a cudaStream_t passed as void*;
 *   - return 0 on success, non-zero error code otherwise; text via eeg_last_error() (thread-local);
 *   - node-feature layout ("blocked, m-major"): per node a row  concat_l [ (2l+1) x mul_l ]  with element
 *     (l, m, u) at  off_l + m*mul_l + u.  With mul_l == C for all l this is [N, (lmax+1)^2, C].
 */
#ifndef EEG_B200_H
#define EEG_B200_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define EEG_MAX_L 8 /* max number of irrep blocks in one layout */

const char* eeg_last_error(void);
int eeg_abi_version(void);

/* Blocked layout descriptor (host struct, passed by pointer). */
typedef struct EegLayout {
  int32_t n_blocks;            /* number of irrep blocks */
  int32_t rows[EEG_MAX_L];     /* 2l+1 per block */
  int32_t mul[EEG_MAX_L];      /* multiplicity per block */
  int32_t off[EEG_MAX_L];      /* float offset of the block inside a node row */
  int32_t stride;              /* floats per node row */
} EegLayout;

/* ---- graph indexing: replaces torch_scatter's unsorted atomics (gnn/blocks.py:595-597) with a CSR ------------------
 * edge_index int64 [2,E] (row 0 sender, row 1 receiver; gnn/mace.py:345).  perm = edge ids stably sorted by receiver
 * (== torch.sort(receiver, stable=True).indices, bit-exact), row_ptr = exclusive cumsum of bincount(receiver).
 * ws: at least eeg_csr_workspace_bytes(E, N) bytes. */
size_t eeg_csr_workspace_bytes(int64_t E, int64_t N);
int eeg_csr_build(const int64_t* edge_index, int64_t E, int64_t N, int32_t* row_ptr /*[N+1]*/, int32_t* perm /*[E]*/,
                  void* ws, size_t ws_bytes, void* stream);
/* batch int64 [N] non-decreasing graph id (PyG Batch.batch; gnn/model.py:100-106) -> graph_ptr int32 [G+1] */
int eeg_graph_ptr(const int64_t* batch, int64_t N, int64_t G, int32_t* graph_ptr, void* stream);

/* Optional validation of the graph inputs the kernels index with (torch_scatter / PyG would assert on the device; these kernels
 * do not check).  Writes one int32 to `flags` (device): bit 0 = an edge_index entry outside [0, N), bit 1 = a batch entry outside
 * [0, G), bit 2 = batch is not non-decreasing (required by eeg_graph_ptr; gnn/model.py:100-106 relies on it through
 * scatter(mean)).  batch may be NULL.  PRECONDITION of eeg_csr_build / eeg_edge_featurize / eeg_graph_ptr when this is skipped:
 * all three conditions hold. */
int eeg_graph_validate(const int64_t* edge_index, int64_t E, int64_t N, const int64_t* batch, int64_t G, int32_t* flags,
                       void* stream);

/* ---- node embedding: torch.nn.Linear(1, C) on node_attrs [N,1] (gnn/model.py:122,142): out[n,c] = a[n]*W[c] + b[c] */
int eeg_node_embed_fwd(const float* node_attrs, const float* W /*[C,1]*/, const float* b /*[C]*/, float* out /*[N,C]*/,
                       int64_t N, int C, void* stream);

/* ---- edge featurisation: gnn/mace.py:338-352 + gnn/model.py:146-157 (e3nn soft_one_hot_linspace x2, SphericalHarmonics)
 * Outputs are written in CSR order (slot j <-> edge perm[j]):  Y [E, ystride] (ystride = (lmax+1)^2 rounded up to 4),
 * f [E, 2*nbasis] (length basis first), sender_sorted int32 [E]. */
int eeg_edge_featurize(const float* pos /*[N,3]*/, const int64_t* edge_index, const float* shifts /*[E,3]*/,
                       const float* radius /*[E]*/, const int32_t* perm, int64_t E, int lmax, int nbasis, float len_end,
                       float rmax, float* Y, int ystride, float* f, int32_t* sender_sorted, void* stream);

/* ---- generic fp32 block GEMM over blocked layouts: e3nn o3.Linear (gnn/blocks.py:516-521,552-559,471-476,276-279,
 * gnn/model.py:82-86) and torch.nn.Linear of the radial MLP (gnn/blocks.py:538-549).
 * For block b: out[n, b, m, w] = act( scale[b] * sum_u in[n, b, m, u] * W_b[u, w] + bias_b[w] ) (+ residual[n,b,m,w])
 * W_b at W + w_off[b]; row-major [mul_in, mul_out] (transB=0) or [mul_out, mul_in] (transB=1, torch.nn.Linear).
 * bias may be NULL; bias applies to block 0 only (0e scalars). act: 0 none, 1 SiLU.  residual may be NULL (layout = out). */
int eeg_block_linear_fwd(const float* in, const EegLayout* lin, const float* W, const int64_t* host_w_off,
                         const float* host_scale, const float* bias, int transB, int act, const float* residual,
                         float* out, const EegLayout* lout, int64_t N, void* stream);

/* Shapes with every mul_in a multiple of 32 and mul_out a multiple of 16 (act == 0) run on the tcgen05 tensor cores
 * (3xTF32 split, fp32 accumulation in TMEM; eeg_tclinear.cu); anything else on the CUDA-core GEMM.  Same results to
 * fp32 round-off.  EEG_LINEAR=cuda in the environment forces the CUDA-core path. */

/* ---- e3nn.nn.Gate (gnn/blocks.py:268-282) fused into the following o3.Linear (readout linear_2, gnn/blocks.py:276-283):
 * z layout lz: block 0 = [n_scalars scalars | gates of block 1 | gates of block 2 ...], blocks b >= 1 = gated features.
 * out[n,b,m,w] = scale[b] * sum_u g[n,b,m,u] W_b[u,w],  g_0 = c*silu(scalars),  g_b = z_b * c*silu(gates_b).
 * Tensor-core path only (mul multiples of 32, mul_out multiples of 16); returns 3 when the shape is not supported so
 * the caller can run eeg_gate_fwd + eeg_block_linear_fwd instead. */
int eeg_gate_linear_fwd(const float* z, const EegLayout* lz, int n_scalars, float act_const, const float* W,
                        const int64_t* host_w_off, const float* host_scale, float* out, const EegLayout* lout, int64_t N,
                        void* stream);

/* ---- radial MLP -> per-edge tensor-product weights: `conv_tp_weights` Sequential (gnn/blocks.py:538-549,590) -----
 * Linear(n_in,64)+SiLU on CUDA cores; n_hidden x [Linear(64,64)+SiLU] and the last Linear(64,n_out,bias=False) on
 * tcgen05 tensor cores (3xTF32 split, fp32 accumulate in TMEM), weights streamed by the TMA engine (cp.async.bulk).
 * eeg_radial_pack: one-off (per weight version) repack of Wh [n_hidden,64,64] and W2 [n_out,64] (torch.nn.Linear
 *   layout) into the staged hi/lo operand image the kernel streams; img has eeg_radial_image_floats() floats.
 * eeg_radial_mlp_fwd: f [E,n_in] (CSR order), W0 [64,n_in], b0 [64], bh [n_hidden,64], n_out = n_paths*C.
 *   Output is PATH-MAJOR: w_t [n_out/32][E][32] (tile = path*(C/32) + channel chunk), as eeg_tp_scatter_fwd reads it. */
int64_t eeg_radial_image_floats(int n_hidden, int n_out);
int eeg_radial_pack(const float* Wh, int n_hidden, int hidden_dim, const float* W2, int n_out, float* img, void* stream);
int eeg_radial_mlp_fwd(const float* f, int n_in, const float* W0, const float* b0, const float* bh, int n_hidden,
                       int hidden_dim, const float* img, int n_out, float* w_t, int64_t E, void* stream);

/* ---- channelwise tensor product + scatter-sum: e3nn TensorProduct 'uvu' + torch_scatter (gnn/blocks.py:591-597)
 * h [N, (lmax_in+1)^2, C] sender features (after linear_up); Y / sender in CSR order;
 * w_t [n_paths * C/32][E][32] PATH-MAJOR per-edge weights in CSR order (eeg_radial_mlp_fwd output).
 * M [N, rows_total, C] with rows laid out per l3 as [(2l3+1), n_paths(l3)] (EegLayout with mul = n_paths*C).
 * (lmax_in, lmax) must be a generated combination: (0,4),(4,4),(0,3),(3,3).  C multiple of 32. */
int eeg_tp_scatter_fwd(int lmax_in, int lmax, int C, const float* h, const float* Y, int ystride, const float* w_t,
                       int64_t E, const int32_t* sender_sorted, const int32_t* row_ptr, float inv_norm, float* M,
                       int64_t N, void* stream);
/* Same contract on the tensor cores (csrc/eeg_tp_umma.cu: one [230 x 25] x [25 x 32] 3xTF32 tcgen05 product per edge, TMEM
 * accumulators, TMA weight stream); C = 32 and lmax_in = lmax in {3, 4} only, otherwise returns 1. */
int eeg_tp_scatter_umma_fwd(int lmax_in, int lmax, int C, const float* h, const float* Y, int ystride, const float* w_t,
                            int64_t E, const int32_t* sender_sorted, const int32_t* row_ptr, float inv_norm, float* M,
                            int64_t N, void* stream);

/* ---- symmetric contraction (product basis): gnn/mace.py:112-277
 * eeg_sc_coefficients: T[C, n_coef] (channel-major) = (A_csr @ Wflat[n_wrows, C])^T   (A = symmetrised U tensors, built by the host)
 * eeg_sym_contract_fwd: x [N, D, C] -> out [N, D, C]; (lmax, nu) must be generated: (4,3), (3,3). */
int eeg_sc_coefficients(const int32_t* a_row_ptr, const int32_t* a_col, const float* a_val, int n_coef,
                        const float* Wflat, int C, float* T, void* stream);
int eeg_sym_contract_fwd(int lmax, int nu, int C, const float* x, const float* T, float* out, int64_t N, void* stream);

/* ---- gated non-linearity: e3nn.nn.Gate (gnn/blocks.py:268-282).  z layout: block0 mul = S + G (scalars then gates),
 * blocks l>=1 mul = g_l with sum g_l = G; out layout: block0 mul = S, blocks l>=1 mul g_l.  act = c * silu. */
int eeg_gate_fwd(const float* z, const EegLayout* lz, float* out, const EegLayout* lout, int n_scalars, float act_const,
                 int64_t N, void* stream);

/* ---- pooling + tail: scatter(mean) (gnn/model.py:100-106), final o3.Linear (:108), Spherical_to_Cartesian +
 * Cart_4_to_Mandel (gnn/blocks.py:392-442) folded into P [n_q, 36], PositiveLayer (gnn/blocks.py:185-229).
 * o: node features (layout lo); Wq [sum_b mul_b*q_b] per-block [mul_b, q_b] (q_b = 0 for unused blocks), bias [q_0];
 * q layout = concat_b [q_b, rows_b] (mul-major, e3nn);  positive: 0 none, 1 A^2, 2 A^4, 3 (I+A/2)^2, 4 (I+A/4)^4.
 * reduce: 0 mean, 1 sum. */
int eeg_pool_tail_fwd(const float* o, const EegLayout* lo, const int32_t* graph_ptr, const float* Wq,
                      const int32_t* host_qmul, const float* host_qscale, const float* bias, const float* P /*[nq,36]*/,
                      int positive, int reduce, float* q_out /*[G,nq] or NULL*/, float* A_out /*[G,36] or NULL*/,
                      float* C_out /*[G,36]*/, int64_t G, void* stream);

/* ====================================================================================================================
 * Explicit backward ops (the reference differentiates the same graph with torch autograd over e3nn / torch_scatter /
 * opt_einsum; scripts/train_utils.py:45-64 calls loss.backward()).  Input gradients of a linear are
 * eeg_block_linear_fwd with transB = 1 (dX = dY W^T).  Gradient buffers marked "+=" must be zeroed by the caller.
 * ==================================================================================================================== */

/* dW_b[u,w] += scale_b * sum_{n,m} x[n,b,m,u] dy[n,b,m,w]  (transB = 1: stored as dW_b[w,u], torch.nn.Linear layout);
 * dbias[w] += sum_n dy[n,0,0,w] (may be NULL).  in_ld / k_override may be NULL (row stride / K = mul of the layout).
 * n_weights = floats of dW.  ws: workspace of eeg_block_linear_wgrad_ws_floats(lin, lout, n_weights, bias columns, N) floats ->
 * DETERMINISTIC (row slices are written to ws and added in slice order); ws NULL -> atomicAdd per slice (order varies).
 * With a workspace, 64 output columns and >= 128 weight rows per block (the last radial-MLP layer) the sums run on tcgen05
 * (3xTF32, accumulators drained every <= 512 rows, operands by tensor-map TMA loads when x and dy are 2-D matrices:
 * csrc/eeg_wgrad_umma.cu); EEG_WGRAD=cuda keeps the CUDA-core kernel, EEG_WGRAD_TMA=0 the cp.async loads. */
int eeg_block_linear_wgrad(const float* x, const EegLayout* lin, const float* dy, const EegLayout* lout,
                           const int64_t* host_w_off, const float* host_scale, int transB, const int32_t* in_ld,
                           const int32_t* k_override, float* dW, float* dbias, int64_t N, int64_t n_weights, float* ws,
                           int64_t ws_floats, void* stream);
int64_t eeg_block_linear_wgrad_ws_floats(const EegLayout* lin, const EegLayout* lout, int64_t n_weights, int bias_cols, int64_t N);

/* Backward of the two hidden layers of the radial MLP (gnn/blocks.py:538-549: Linear(F,64) - SiLU - Linear(64,64) - SiLU) in one
 * pass over the edges: given the edge features f [E, f_stride] (first F columns), the torch.nn.Linear weights W0 [64,F], b0,
 * W1 [64,64], b1 and ds2 = dL/d(silu(a2)) [E,64], recomputes the activations and adds dW0 [64,F], db0, dW1 [64,64], db1 (+=:
 * zero them first); s2_out [E,64] (may be NULL) receives silu(a2), the input of the last linear, for its weight gradient.
 * F a multiple of 4, <= 32.  Deterministic: per-CTA partials in ws (eeg_radial_hidden_bwd_ws_floats floats), fixed-order sum. */
int eeg_radial_hidden_bwd(const float* f, int F, int f_stride, const float* W0, const float* b0, const float* W1, const float* b1,
                          const float* ds2, int64_t E, float* s2_out, float* dW0, float* db0, float* dW1, float* db1, float* ws,
                          int64_t ws_floats, void* stream);
int64_t eeg_radial_hidden_bwd_ws_floats(int F, int64_t E);
/* out[c] += sum_r row_scale[r] * x[r*ld + c]   (row_scale may be NULL = 1): bias / embedding gradients.
 * ws: ceil(R / 256) * cols floats -> deterministic (CTA partials added in order); NULL -> atomicAdd. */
int eeg_colsum(const float* x, int64_t R, int ld, int cols, const float* row_scale, float* out, float* ws, int64_t ws_floats,
               void* stream);
/* Same result as eeg_colsum in its deterministic mode, with NO caller workspace and no allocation of any kind (no cudaMalloc /
 * cudaFree / allocating library call).  This entry point has no counterpart in the reference: it was added
 * for callers that cannot allocate (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid cudaMalloc).  The CTA partials go to a file-scope device array inside the library of eeg_colsum_static_max_floats()
 * = 2^20 floats (4 MiB, held by every GPU that loads the library until exit); a call needing more than that, i.e.
 * ceil(R / 256) * cols > 2^20, is rejected with return code 1 and a message - use eeg_colsum with a workspace then.
 * out[c] += ... as for eeg_colsum (the caller zeroes out).  Every call on a device shares the one array: calls that overlap
 * on the GPU (different streams) race; calls ordered on one stream are safe, and nothing has to be cleared between calls. */
int eeg_colsum_static(const float* x, int64_t R, int ld, int cols, const float* row_scale, float* out, void* stream);
int64_t eeg_colsum_static_max_floats(void);
/* out[n, :] = sum_{j in [ptr[n], ptr[n+1])} in[perm[j], :]  (rows of `width` floats, width % 4 == 0; fixed order): the
 * sender-side scatter of the tensor-product backward as a CSR walk instead of atomics. */
int eeg_segment_sum_rows(const float* in, const int32_t* perm, const int32_t* ptr, int64_t N, int64_t width, float* out,
                         void* stream);
/* SiLU of the radial MLP (gnn/blocks.py:538-549): s = silu(a);  da = ds * silu'(a) */
int eeg_silu_fwd(const float* a, float* s, int64_t n, void* stream);
int eeg_silu_bwd(const float* a, const float* ds, float* da, int64_t n, void* stream);
/* e3nn.nn.Gate backward (gnn/blocks.py:268-282): dz from z (layout lz) and dg (layout lg) */
int eeg_gate_bwd(const float* z, const EegLayout* lz, const float* dg, const EegLayout* lg, int n_scalars, float act_const,
                 float* dz, int64_t N, void* stream);
/* Backward of eeg_pool_tail_fwd: A [G,36] is the forward's A_out; dC [G,36]; writes d_o [N, lo.stride] (all columns),
 * dWq (same shape as Wq), dbias [q_0] (may be NULL); dq_ws [G,nq] and pooled_ws [G,npool] are workspaces. */
int eeg_pool_tail_bwd(const float* o, const EegLayout* lo, const int32_t* graph_ptr, const float* Wq,
                      const int32_t* host_qmul, const float* host_qscale, const float* P, int positive, int reduce,
                      const float* A, const float* dC, float* d_o, float* dq_ws, float* pooled_ws, float* dWq, float* dbias,
                      int64_t G, void* stream);
/* Backward of eeg_tp_scatter_fwd: dw [E, n_paths*C] and dx_edge [E,(lmax_in+1)^2*C] (edge-major, CSR order; fully written, no
 * atomics).  dh[s] = sum of dx_edge over the edges sent by s: eeg_segment_sum_rows with a CSR of sender_sorted. */
int eeg_tp_scatter_bwd(int lmax_in, int lmax, int C, const float* h, const float* Y, int ystride, const float* w_t,
                       int64_t E, const int32_t* sender_sorted, const int32_t* row_ptr, float inv_norm, const float* dM,
                       float* dw, float* dx_edge, int64_t N, void* stream);
/* Backward of eeg_sym_contract_fwd w.r.t. x (same coefficient table T as the forward) */
int eeg_sym_contract_bwd_x(int lmax, int nu, int C, const float* x, const float* dout, const float* T, float* dx, int64_t N,
                           void* stream);
/* Coefficient gradient: dT [eeg_sc_grad_slices()][n_coef][C] partial sums over node slices (fixed order: deterministic);
 * n_coef must equal the slot count of the compiled (lmax, nu) program.  eeg_sc_coefficients_bwd folds them onto the
 * weights through the transposed (CSC) coefficient map: dW [n_wrows, C]; it sums the slices into slice 0 in place first
 * (dT is scratch). */
int eeg_sc_grad_slices(void);
int eeg_sc_coef_grad(int lmax, int nu, int C, const float* x, const float* dout, int n_coef, int64_t N, float* dT, void* stream);
int eeg_sc_coefficients_bwd(const int32_t* t_ptr, const int32_t* t_row, const float* t_val, int n_wrows, int n_coef,
                            float* dT, int C, float* dW, void* stream);

/* ---- two chained multiplicity-32 per-irrep mixes in one pass (inference): y1 = Wa x (+ residual) = product.linear of layer i
 * (gnn/blocks.py:471-476,488-490, gnn/model.py:95-96), y2 = Wb y1 = linear_up of layer i+1 (gnn/blocks.py:516-521,589).
 * x, residual (may be NULL), y1, y2 share the layout lh (every block multiplicity 32); weights [32][32] per block. */
int eeg_linear_pair_fwd(const float* x, const float* residual, const EegLayout* lh, const float* Wa, const int64_t* host_wa_off,
                        const float* host_sa, const float* Wb, const int64_t* host_wb_off, const float* host_sb, float* y1,
                        float* y2, int64_t N, void* stream);

/* ---- fused inference tail for multiplicity-32 models: product.linear (+ residual) -> readout linear_1 -> Gate -> linear_2
 * (gnn/blocks.py:471-476,488-490, gnn/model.py:95-96, gnn/blocks.py:250-283) in one streaming pass per irrep block; only
 * out [N, lout] and the gate activations (scratch `gates` [N, 32 (n_blocks - 1)]) are written.  lh: layout of B / residual
 * (every block multiplicity 32, rows 2l+1, l <= 4); residual may be NULL.  Weight offsets / scales per block as in
 * eeg_block_linear_fwd; W1's 0e block is [32][32 + 32 (n_blocks - 1)] (scalars first, then the gates of l = 1, 2, ...). */
int eeg_readout_chain_fwd(const float* B, const float* residual, const EegLayout* lh, const float* Wp, const int64_t* host_wp_off,
                          const float* host_sp, const float* W1, const int64_t* host_w1_off, const float* host_s1, const float* W2,
                          const int64_t* host_w2_off, const float* host_s2, float act_const, float* gates, float* out,
                          const EegLayout* lout, int64_t N, void* stream);

/* ---- optimizer step (scripts/train_utils.py:38-44 AdamW(amsgrad), scripts/train_main.py:91 gradient_clip_val) on flat
 * fp32 buffers.  eeg_sumsq: out[0] += sum x^2.  eeg_adamw_step: g' = grad_scale * g clipped to max_norm by the global norm
 * sqrt(gnorm_sq[0]) * grad_scale (gnorm_sq NULL or max_norm <= 0: no clipping); vmax NULL: plain AdamW. step counts from 1. */
int eeg_sumsq(const float* x, int64_t n, float* out, void* stream);
int eeg_adamw_step(float* p, const float* g, float* m, float* v, float* vmax, int64_t n, float lr, float beta1, float beta2,
                   float eps, float weight_decay, int step, const float* gnorm_sq, float max_norm, float grad_scale,
                   void* stream);

/* ---- interaction_reduction = 'mean' (gnn/blocks.py:595-597, torch_scatter reduce='mean'): x[n, :] /= max(in-degree(n), 1)
 * in place, x = M[N, row_floats] after eeg_tp_scatter_fwd, or dM before eeg_tp_scatter_bwd (the scaling is its own adjoint).
 * row_ptr = CSR by receiver (eeg_csr_build). */
int eeg_scale_rows_inv_degree(float* x, const int32_t* row_ptr, int64_t N, int64_t row_floats, void* stream);

/* ---- RotateLat on a collated batch (scripts/train_utils.py:114-146).  Q = [G,3,3] row-major, one
 * proper rotation per lattice.  eeg_rotate_vec3: out[i] = Q[graph_of[via ? via[i] : i]] . in[i] for n rows of 3 floats
 * (positions: graph_of = batch, via = NULL (:124); shifts: via = edge_index row 0, the sender (:125)); in == out is allowed.
 * eeg_rotate_cart4_to_mandel: C4 [G,3,3,3,3] -> Q_ai Q_bj Q_ck Q_dl C_ijkl (:122-123) -> Mandel [G,6,6]
 * (lattices/lattices/utils/elasticity_func.py:406-428, the conversion RotateLat applies at :133-134); Q NULL = no
 * rotation (RotateLat(rotate=False), :126-131). */
int eeg_rotate_vec3(const float* Q, const int64_t* graph_of, const int64_t* via, const float* in, float* out, int64_t n,
                    void* stream);
int eeg_rotate_cart4_to_mandel(const float* Q, const float* C4, float* out, int64_t G, void* stream);

/* ---- batch construction on the device (scripts/train_main.py:67-78: PyG DataLoader -> Batch.from_data_list).
 * The dataset lives in HBM in flat form (flatdata.FlatDataset): node arrays pos [SN,3], node_attrs [SN]; edge arrays
 * edge_snd / edge_rcv int32 [SE] (LOCAL node ids of the lattice), shifts [SE,3], edge_attr [SE].  A batch of G lattices is
 * described by four int64 device arrays: src_node_off / src_edge_off [G] (first node / edge of each chosen lattice in the
 * dataset) and out_node_ptr / out_edge_ptr [G+1] (exclusive cumsum of their node / edge counts; [G] = N / E).
 * Writes the PyG batch: positions, node_attrs, edge_index int64 [2,E] (local id + node offset of the graph), shifts, edge_attr,
 * batch int64 [N]; optionally gathers per-lattice rows (targets): o_graph_rows[g, :] = graph_rows[graph_idx[g], :]. */
int eeg_collate_gather(const float* pos, const float* node_attrs, const int32_t* edge_snd, const int32_t* edge_rcv,
                       const float* shifts, const float* edge_attr, const int64_t* src_node_off, const int64_t* src_edge_off,
                       const int64_t* out_node_ptr, const int64_t* out_edge_ptr, int64_t G, int64_t N, int64_t E, float* o_pos,
                       float* o_node_attrs, int64_t* o_edge_index, float* o_shifts, float* o_edge_attr, int64_t* o_batch,
                       const float* graph_rows /*[L,width] or NULL*/, const int64_t* graph_idx /*[G]*/, int64_t graph_width,
                       float* o_graph_rows /*[G,width]*/, void* stream);
/* G proper rotations drawn like e3nn o3.rand_matrix() (scripts/train_utils.py:121): alpha, gamma ~ U[0,2pi), cos(beta) ~ U[-1,1],
 * Q = Ry(alpha) Rx(beta) Ry(gamma); counter-based generator: lattice g of the call uses counters 3*(offset+g) .. +2. */
int eeg_rand_rotations(uint64_t seed, uint64_t offset, int64_t G, float* Q /*[G,3,3]*/, void* stream);

/* [E, B, 32] <-> [B, E, 32] (to_path_major = 1 / 0): edge-major per-edge weights of a generic radial MLP <-> the path-major
 * w_t layout of eeg_tp_scatter_fwd (used when conv_tp_weights is outside eeg_radial_mlp_fwd's fused shape). */
int eeg_blocks32_transpose(const float* in, float* out, int64_t E, int B, int to_path_major, void* stream);

/* ---- RadialEmbeddingBlock (gnn/blocks.py:89-102): BesselBasis (gnn/mace.py:28-71) x PolynomialCutoff (gnn/mace.py:74-106).
 * edge_lengths [E] -> out [E, num_bessel].  Defined by the reference, not used by EnergyEquivGNN. */
int eeg_radial_embedding_fwd(const float* edge_lengths, int64_t E, int num_bessel, float r_max, int num_polynomial_cutoff,
                             float* out, void* stream);

/* ---- validation / test metrics (scripts/train_utils.py:66-89 validation_step, :149-172 obtain_errors; SURVEY.md 8(f)
 * rank 3).  pred, target: Mandel [G,6,6]; directions: [P,3] unit vectors.  per_lattice [G,4] = (mean squared error and mean
 * absolute error over the 36 entries (:157-158), mean_p |c(d_p) - c_pred(d_p)| (:161-163), mean_p c(d_p) of the target
 * (:164)) with c(d) = C_ijkl d_i d_j d_k d_l of the Mandel -> Cartesian expansion of
 * lattices/lattices/utils/elasticity_func.py:430-463.  eig_pred / eig_target [G,6] (either may be NULL): ascending
 * eigenvalues as torch.linalg.eigvalsh returns them (:166-170). */
int eeg_validation_metrics(const float* pred, const float* target, const float* directions, int P, int64_t G,
                           float* per_lattice, float* eig_pred, float* eig_target, void* stream);

#ifdef __cplusplus
}
#endif
#endif
