/*
 * gdr.h — C ABI of libgdr.so: the B200 (sm_100a) propagation path of
 * bowang-lab/gcn-drug-repurposing.
 *
 * The reference has no FFI: its "operator API" is a set of Python call sites that
 * end in torch / numpy / scipy library calls.  Every entry point below replaces one
 * of those library calls (cited as reference file:line).  The host-side mirror of the
 * reference classes (package `gcn-drug-repurposing_b200`) binds these with ctypes.
 *
 * Conventions
 *   - plain C types only: device pointers, int64 sizes, a cudaStream_t passed as void*.
 *   - every pointer is a DEVICE pointer unless the name ends in `_host`.
 *   - the caller owns every buffer, including workspaces (sizes from *_workspace_bytes).
 *   - return value: 0 = ok, >0 = cudaError_t, <0 = argument error (GDR_E_*).
 *     gdr_last_error_string() gives the text of the last failure on this thread.
 *   - all kernels are stream-ordered on `stream`; nothing synchronises the device
 *     unless the entry point's comment says so.
 *   - feature matrices are dense row-major, row stride == d (number of columns),
 *     d % 4 == 0 and 16-byte aligned base pointers.
 *   - CSR: int32 rowptr[n_rows+1], int32 col[nnz], values fp32 or fp64.
 */
#ifndef GDR_H_
#define GDR_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define GDR_OK 0
#define GDR_E_BADARG (-1)   /* null pointer, negative size */
#define GDR_E_SHAPE (-2)    /* unsupported width / alignment */
#define GDR_E_WORKSPACE (-3) /* workspace too small */
#define GDR_E_INTERNAL (-5) /* driver entry point missing / tensor-map encode failed */
#define GDR_E_NOCONV (-4)   /* power iteration hit max_iter (reference raises, diffusion_profiles.py:90) */

typedef void *gdr_stream_t; /* cudaStream_t */

/* ---- library ------------------------------------------------------------------ */
int gdr_version(void);
const char *gdr_last_error_string(void);
/* number of kernels this library has launched in this process (bench "gpu_launches") */
int64_t gdr_launch_count(void);
void gdr_launch_count_reset(void);

/* ---- CSR construction ----------------------------------------------------------- *
 * replaces scipy's csr/coo conversions around helpers/helper.py:53,89,92-96 and the
 * torch sparse COO the model receives (modules/model.py:163).                        */

/* COO (int64 row/col as torch gives them, any order, duplicates kept) -> CSR sorted by
 * (row, col), stable.  perm_out (optional, int64[nnz]) receives the source position of
 * every CSR entry so that values in another dtype can be permuted by the caller.     */
int64_t gdr_csr_from_coo_workspace_bytes(int64_t nnz, int64_t n_rows, int64_t n_cols);
int gdr_csr_from_coo(const int64_t *coo_row, const int64_t *coo_col, int64_t nnz,
                     int64_t n_rows, int64_t n_cols, int32_t *rowptr_out,
                     int32_t *col_out, int64_t *perm_out, void *workspace,
                     int64_t workspace_bytes, gdr_stream_t stream);
/* values_out[i] = values_in[perm[i]] (elem_bytes 1, 4 or 8) */
int gdr_permute_values(const void *values_in, const int64_t *perm, int64_t nnz,
                       int elem_bytes, void *values_out, gdr_stream_t stream);

/* CSR -> CSR of the transpose, entries of every transposed row ordered by source row
 * (stable), i.e. exactly scipy's A.T.tocsr().  perm_out as above.                    */
int64_t gdr_csr_transpose_workspace_bytes(int64_t nnz, int64_t n_rows, int64_t n_cols);
int gdr_csr_transpose(const int32_t *rowptr, const int32_t *col, int64_t nnz,
                      int64_t n_rows, int64_t n_cols, int32_t *t_rowptr_out,
                      int32_t *t_col_out, int64_t *perm_out, void *workspace,
                      int64_t workspace_bytes, gdr_stream_t stream);

/* preprocess_graph (helpers/helper.py:82-89):  Â = D^-1/2 (A+I) D^-1/2 in fp64, entry
 * value (a_ij * d_j) * d_i with d = rowsum(A+I)^-1/2.  Input CSR must have sorted rows.
 * Output nnz = nnz + (rows without a stored diagonal).  Call the _count entry first
 * (writes out_rowptr and *nnz_out_host after synchronising the stream).              */
int gdr_csr_add_identity_count(const int32_t *rowptr, const int32_t *col, int64_t n,
                               int32_t *out_rowptr, int64_t *nnz_out_host,
                               void *workspace, int64_t workspace_bytes,
                               gdr_stream_t stream);
int64_t gdr_csr_add_identity_workspace_bytes(int64_t n);
int gdr_csr_normalize_f64(const int32_t *rowptr, const int32_t *col, const double *val,
                          int64_t n, const int32_t *out_rowptr, int32_t *out_col,
                          double *out_val, double *deg_inv_sqrt_out,
                          gdr_stream_t stream);

/* ---- SpMM: the aggregation  Y = Â·X  ---------------------------------------------- *
 * replaces torch.sparse.mm (modules/model.py:163,169), its autograd Âᵀ·G (pass the
 * transposed CSR) and scipy's x*M (multiscale/diff_prof/diffusion_profiles.py:85).
 *   Y[r,:] = sum_e val[e] * X[col[e],:] (* P[col[e],:] if mul != NULL)
 *            (+ E1[r,:] * E2[r,:]       if e1 != NULL; e2 == NULL means + E1[r,:], and e1 may be y itself:
 *                                        in-place accumulation  Y += Â·X  for column-split operands)
 * mul: the reference's torch.mul(Ax, features) (model.py:168) folded into the gather.
 * e1,e2: the backward's element-wise addend  dX = Âᵀ·dP + dR ⊙ P.                     */
int gdr_spmm_csr_f32(const int32_t *rowptr, const int32_t *col, const float *val,
                     int64_t n_rows, int64_t d, const float *x, const float *mul,
                     const float *e1, const float *e2, float *y, gdr_stream_t stream);
/* Same product with a degree-aware schedule (the "plan"): rows with more than hub_threshold entries
 * are cut into chunks; chunk_info[c] = {hub index, first entry, end entry, 0} and
 * hub_info[h] = {row, first chunk, #chunks, 0} (int32 x4 each, 16-byte aligned).  One warp reduces one
 * chunk into partial[c][0..d) and the last warp of a hub (tickets[h], zero on entry, reset on exit)
 * adds the partials in chunk order, so results are independent of scheduling.  n_chunks == 0 means
 * "no hubs" and is what gdr_spmm_csr_f32 passes.  partial: fp32 [n_chunks, d], caller-owned.        */
int gdr_spmm_csr_planned_f32(const int32_t *rowptr, const int32_t *col, const float *val,
                             int64_t n_rows, int64_t d, const float *x, const float *mul,
                             const float *e1, const float *e2, float *y, int32_t hub_threshold,
                             int64_t n_hubs, const int32_t *hub_info, int64_t n_chunks,
                             const int32_t *chunk_info, int32_t *tickets, float *partial,
                             gdr_stream_t stream);
/* y = Â·x (+ e1 ⊙ e2) and, from the same launch, y2 = y ⊙ m2 (m2, y2 row-aligned with y).  The layer
 * needs both P = Â·X and torch.mul(P, X) (modules/model.py:163,168): writing the product here makes the
 * following T = Â·(P ⊙ X) a one-operand gather.  Plan arguments as in gdr_spmm_csr_planned_f32.       */
int gdr_spmm_csr_dual_f32(const int32_t *rowptr, const int32_t *col, const float *val, int64_t n_rows,
                          int64_t d, const float *x, const float *e1, const float *e2, float *y,
                          const float *m2, float *y2, int32_t hub_threshold, int64_t n_hubs,
                          const int32_t *hub_info, int64_t n_chunks, const int32_t *chunk_info,
                          int32_t *tickets, float *partial, gdr_stream_t stream);
int gdr_spmm_csr_f64(const int32_t *rowptr, const int32_t *col, const double *val,
                     int64_t n_rows, int64_t d, const double *x, double *y,
                     gdr_stream_t stream);

/* ---- dense transform -------------------------------------------------------------- *
 * GSS_GNNLayer.forward (modules/model.py:163-175) after the two aggregations:
 *   Z = P·W1ᵀ + b1 + T·W2ᵀ + b2 ;  Y = ELU(Z) ;  Xnext = residual ? residual + decay*Y : Y
 * (the residual update is ResidualGraphConvolutionalNetwork.forward, model.py:201-203).
 * W row-major [d_out=d, d_in=d] as nn.Linear stores it.  z_out, xnext_out: [n, d].   */
int gdr_gcn_layer_fwd_f32(const float *p, const float *t, const float *w1,
                          const float *b1, const float *w2, const float *b2,
                          const float *residual, float decay, int64_t n, int64_t d,
                          float *z_out, float *xnext_out, gdr_stream_t stream);

/* F.normalize(x, dim=1) (model.py:205): h = x / max(||x||_2, 1e-12); norm_out[n]. */
int gdr_row_normalize_fwd_f32(const float *x, int64_t n, int64_t d, float *h_out,
                              float *norm_out, gdr_stream_t stream);
/* backward of the above: gx = (gh - h * <h, gh>) / max(norm, eps).
 * rows (optional int64[n_rows_sel]) restricts the work to those rows of gh/h (the
 * batch rows of the loss); gx must then be zero-filled by the caller.               */
int gdr_row_normalize_bwd_f32(const float *gh, const float *h, const float *norm,
                              int64_t n, int64_t d, float *gx_out, gdr_stream_t stream);

/* gZ = scale * gXnext * ELU'(Z) (+ gres if not NULL): autograd of F.elu + residual.  */
int gdr_elu_bwd_f32(const float *gxnext, const float *z, const float *gres, float scale,
                    int64_t count, float *gz_out, gdr_stream_t stream);

/* C[n,d_in] = G[n,d_out]·W[d_out,d_in] (+ E1 ⊙ E2): autograd of nn.Linear w.r.t. its
 * input (dP = dZ·W1, dT = dZ·W2), with the dP += dR ⊙ X addend folded in.           */
int gdr_linear_bwd_input_f32(const float *g, const float *w, const float *e1,
                             const float *e2, int64_t n, int64_t d, float *c_out,
                             gdr_stream_t stream);

/* dW[d,d] (+)= Gᵀ·A  and  db[d] (+)= colsum(G): autograd of nn.Linear w.r.t. weight
 * and bias (dW1 = dZᵀ·P, modules/model.py:165).  accumulate != 0 adds into dw/db.    */
int64_t gdr_linear_bwd_weight_workspace_bytes(int64_t n, int64_t d);
int gdr_linear_bwd_weight_f32(const float *g, const float *a, int64_t n, int64_t d,
                              int accumulate, float *dw, float *db, void *workspace,
                              int64_t workspace_bytes, gdr_stream_t stream);

/* ---- the same contractions on tcgen05 tensor cores (3xTF32: hi/lo TF32 split of both operands,
 * three kind::tf32 MMAs per K step, corrections and K chunks in separate TMEM accumulators, summed in
 * fp32 RN in the epilogue; TMA-fed, csrc/gemm_tc.cu).  Same arguments and results (to fp32 rounding
 * noise) as the functions above, plus a caller-owned workspace for the split operands.               */
int64_t gdr_gemm_nt_tc_workspace_bytes(int64_t m, int64_t n, int64_t k);
/* C[m,n] = A[m,k]·B[n,k]ᵀ, row-major, K contiguous: torch.mm(A, B.T) / np.dot(H, H.T) (train.py:166) */
int gdr_gemm_nt_tc_f32(const float *a, int64_t lda, const float *b, int64_t ldb, int64_t m, int64_t n,
                       int64_t k, float *c, int64_t ldc, void *workspace, int64_t workspace_bytes,
                       gdr_stream_t stream);
int64_t gdr_gcn_layer_fwd_tc_workspace_bytes(int64_t n, int64_t d);
int gdr_gcn_layer_fwd_tc_f32(const float *p, const float *t, const float *w1, const float *b1,
                             const float *w2, const float *b2, const float *residual, float decay,
                             int64_t n, int64_t d, float *z_out, float *xnext_out, void *workspace,
                             int64_t workspace_bytes, gdr_stream_t stream);
int64_t gdr_linear_bwd_input_tc_workspace_bytes(int64_t n, int64_t d);
int gdr_linear_bwd_input_tc_f32(const float *g, const float *w, const float *e1, const float *e2,
                                int64_t n, int64_t d, float *c_out, void *workspace,
                                int64_t workspace_bytes, gdr_stream_t stream);
int64_t gdr_linear_bwd_weight_tc_workspace_bytes(int64_t n, int64_t d);
int gdr_linear_bwd_weight_tc_f32(const float *g, const float *a, int64_t n, int64_t d, int accumulate,
                                 float *dw, float *db, void *workspace, int64_t workspace_bytes,
                                 gdr_stream_t stream);

/* Both weight gradients of a layer from one split-K product [dW1 | dW2] = Gᵀ·[A1 | A2], db1 = db2 = colsum(G). */
int64_t gdr_linear_bwd_weight2_tc_workspace_bytes(int64_t n, int64_t d);
int gdr_linear_bwd_weight2_tc_f32(const float *g, const float *a1, const float *a2, int64_t n, int64_t d,
                                  int accumulate, float *dw1, float *db1, float *dw2, float *db2,
                                  void *workspace, int64_t workspace_bytes, gdr_stream_t stream);
int64_t gdr_gss_loss_tc_workspace_bytes(int64_t b, int64_t d);
int gdr_gss_loss_fwd_bwd_tc_f32(const float *embs, const int64_t *index, int64_t b, int64_t n, int64_t d,
                                float alpha, float beta, float *loss_out, float *grad_e_out,
                                void *workspace, int64_t workspace_bytes, gdr_stream_t stream);
int64_t gdr_gram_percentile_tc_workspace_bytes(int64_t n, int64_t d);
int gdr_gram_percentile_tc_f32(const float *h, int64_t n, int64_t d, double q_percent,
                               float *beta_out_host, void *workspace, int64_t workspace_bytes,
                               gdr_stream_t stream);

/* ---- GSS loss ----------------------------------------------------------------------- *
 * GSS_loss.gss_loss (modules/model.py:214-221):  E = embs[index]; G = E·Eᵀ;
 * loss = mean(-0.5*alpha*(relu(G)-beta)^2).  One call gives the loss and dLoss/dE
 * (B x d, un-scattered); index == NULL means all n rows (B = n).                      */
int64_t gdr_gss_loss_workspace_bytes(int64_t b, int64_t d);
int gdr_gss_loss_fwd_bwd_f32(const float *embs, const int64_t *index, int64_t b,
                             int64_t n, int64_t d, float alpha, float beta,
                             float *loss_out, float *grad_e_out, void *workspace,
                             int64_t workspace_bytes, gdr_stream_t stream);
/* The same loss and gradient for callers that cannot allocate device memory.  NO COUNTERPART IN THE REFERENCE
 * (torch allocates freely there); it exists for hosts that cannot give this library a workspace: a plug-in inside a
 * process whose device-memory pool belongs to someone else, a latency-critical loop where allocator calls
 * (cudaMalloc takes a device-wide lock and may synchronise) are banned after start-up, or a plain C / Fortran
 * program without a device allocator.  Inputs and outputs still come from the caller; the B x d + B x B workspace
 * is a fixed 18.1 MiB array inside libgdr.so, held on every GPU that loads the library until the process exits.
 *   - limits: B <= 2048 and B*d <= 2048*256 (the reference's training batch, train.py --batch-size 2048 at
 *     hidden 256); larger sizes return GDR_E_SHAPE without launching anything.
 *   - makes no cudaMalloc / cudaFree / allocating library call.
 *   - the first call on each device resolves the array's device address (cudaGetSymbolAddress): do it once at
 *     set-up, with that device current and before any stream capture begins.
 *   - one array per GPU, shared by every call: calls that may overlap on the GPU (two streams, two threads) race
 *     on it, so keep them ordered.  Stale contents from the previous call are harmless (each call writes every
 *     byte it later reads).
 * Results equal gdr_gss_loss_fwd_bwd_f32 on the same inputs.                                                  */
int gdr_gss_loss_fwd_bwd_static_f32(const float *embs, const int64_t *index, int64_t b,
                                    int64_t n, int64_t d, float alpha, float beta,
                                    float *loss_out, float *grad_e_out, gdr_stream_t stream);
/* dst[i,:] = src[index[i],:]  (embs[index], model.py:216; also the send-buffer pack of the partitioned SpMM) */
int gdr_gather_rows_f32(const float *src, const int64_t *index, int64_t b, int64_t n, int64_t d,
                        float *dst, gdr_stream_t stream);
/* The exchange step of the partitioned SpMM as direct stores into peer memory (NVLink P2P): for every
 * destination k, dst[k][j,:] = src[idx[k][j],:] (idx[k] == NULL: rows src_off[k]+j).  dst[k] may be a peer-GPU
 * address obtained from a symmetric-memory rendezvous.  Tables are HOST arrays (n_dst <= 16).  Replaces
 * gdr_gather_rows_f32 + ncclSend of the packed buffer.  n_ctas <= 0: default.                            */
int gdr_push_rows_f32(const float *src, int64_t d, int n_dst, float *const *dst, const int64_t *const *idx,
                      const int64_t *src_off, const int64_t *counts, int n_ctas, gdr_stream_t stream);
/* NVSwitch multicast (NVLS) forms of the two hub exchanges; mc_* are multicast addresses of a symmetric
 * allocation (multimem.st / multimem.ld_reduce).  broadcast: the hub block of X reaches every GPU while leaving
 * this one once.  reduce_add: y += the sum over all GPUs of their partial-row buffers, reduced in the switch.  */
int gdr_mc_broadcast_f32(const float *src, float *mc_dst, int64_t count, int n_ctas, gdr_stream_t stream);
int gdr_mc_reduce_add_f32(float *y, const float *mc_src, int64_t count, int n_ctas, gdr_stream_t stream);
/* gh[index[i],:] += scale_dev[0] * ge[i,:]  (autograd of embs[index]); gh pre-zeroed. */
int gdr_scatter_rows_f32(const float *ge, const int64_t *index, int64_t b, int64_t n,
                         int64_t d, const float *scale_dev, float *gh, gdr_stream_t stream);

/* beta selection (train.py:163-167): np.percentile(flatten(H·Hᵀ), q), linear
 * interpolation, computed by radix select over recomputed Gram tiles (N×N never
 * stored).  Synchronises the stream; result to *beta_out_host.                        */
int64_t gdr_gram_percentile_workspace_bytes(int64_t n, int64_t d);
int gdr_gram_percentile_f32(const float *h, int64_t n, int64_t d, double q_percent,
                            float *beta_out_host, void *workspace,
                            int64_t workspace_bytes, gdr_stream_t stream);

/* ---- kNN graph (gen_graph, helpers/helper.py:39-44) ---------------------------------- *
 * S = X·Xᵀ in fp64 (x: [n,d] row-major = the reference's X.T) and, per row, the k
 * largest entries (np.argpartition(S,-k,1)[:,-k:], an unordered set): indices int32
 * [n,k] sorted by descending similarity, values fp64 [n,k].  N×N is never stored.     */
int gdr_gram_topk_f64(const double *x, int64_t n, int64_t d, int64_t k,
                      int32_t *idx_out, double *val_out, gdr_stream_t stream);
/* The same result with the N×N contraction on the tcgen05 tensor cores and the selection fused into its
 * epilogue (gram_topk_tc.cu): TF32 Gram tiles (operand A resident in tensor memory, B streamed by TMA) keep
 * the C largest approximate similarities per row and column range, the candidates are re-ranked with exact
 * fp64 inner products, and a row whose k-th exact value is not provably above everything that was dropped
 * (margin 1.1e-3·|x_i|·max|x_j|) is recomputed by the exact fp64 kernel — so idx_out/val_out equal
 * gdr_gram_topk_f64's (order: value descending, then column ascending), fp64 ties excepted.  d <= 256,
 * k <= 16.  n_fallback_out (device int32, may be NULL) receives the number of rows that took the fp64 path. */
int64_t gdr_gram_topk_tc_workspace_bytes(int64_t n, int64_t d, int64_t k);
int gdr_gram_topk_tc_f64(const double *x, int64_t n, int64_t d, int64_t k, int32_t *idx_out,
                         double *val_out, int32_t *n_fallback_out, void *workspace,
                         int64_t workspace_bytes, gdr_stream_t stream);

/* ---- Kipf & Welling graph convolution (multiscale/openne/gcn/layers.py:133-190, metrics.py:4-21) ----------- *
 * act(sum_i A_i·(X·W_i) + b): the products are gdr_spmm_csr_f32 (supports, sparse X) and gdr_gemm_nt_tc_f32
 * (dense X·W, dW, dX); these are the row-local pieces.                                                     */
/* y = act(z + bias)   (bias may be NULL; relu != 0: ReLU, else identity; y may alias z) */
int gdr_bias_act_f32(const float *z, const float *bias, int64_t n, int64_t d, int relu, float *y,
                     gdr_stream_t stream);
/* dz = dy ⊙ [y > 0] (relu != 0) or dy; db = column sums of dz.  dz and db may be NULL (not wanted). */
int gdr_act_bwd_colsum_f32(const float *dy, const float *y, int64_t n, int64_t d, int relu, float *dz,
                           float *db, gdr_stream_t stream);
/* y = x * floor(keep_prob + u) / keep_prob  (layers.py:22-29, tf.nn.dropout); u: uniform [0,1) numbers from the
 * caller's generator; mask_scale_out (may be NULL) keeps the factor for the backward pass.                */
int gdr_dropout_apply_f32(const float *x, const float *uniform, int64_t count, float keep_prob, float *y,
                          float *mask_scale_out, gdr_stream_t stream);
/* masked_softmax_cross_entropy + masked_accuracy (metrics.py:4-21) and dLoss/dLogits in one pass.
 * logits [n, ld] (ld >= c: padded rows allowed), labels [n, c], mask [n] (0/1 as fp32).                   */
int64_t gdr_masked_softmax_ce_workspace_bytes(int64_t n);
int gdr_masked_softmax_ce_f32(const float *logits, int64_t ld, const float *labels, const float *mask,
                              int64_t n, int64_t c, float *loss_out, float *accuracy_out,
                              float *dlogits_out, void *workspace, int64_t workspace_bytes,
                              gdr_stream_t stream);

/* ---- node2vec / DeepWalk walk generation (multiscale/openne/walker.py:11-161) ------------------------------- *
 * CSR of the directed graph with ascending columns per row; cum = per-row running sums of the edge weights
 * (gdr_walk_row_cumsum_f32; weight == NULL: all ones = BasicWalker's uniform choice).  One thread per walk:
 * first-order draws by inverse CDF, second-order steps (second_order != 0) by rejection against max(1/p,1,1/q)
 * with the reference's factors (1/p back to the previous node, 1 if next -> previous is an edge, 1/q otherwise).
 * walks_out [n_walks, walk_length] int32 node indices, -1 after a walk ended at a node without out-neighbours.
 * Random numbers: counter-based, a pure function of (seed, walk_id0 + walk, draw).                            */
int gdr_walk_row_cumsum_f32(const int32_t *rowptr, const float *weight, int64_t n, float *cum_out,
                            gdr_stream_t stream);
int gdr_node2vec_walks(const int32_t *rowptr, const int32_t *col, const float *cum, int64_t n,
                       const int32_t *start, int64_t n_walks, int64_t walk_length, double p, double q,
                       int second_order, uint64_t seed, int64_t walk_id0, int32_t *walks_out,
                       gdr_stream_t stream);

/* ---- optimizer ------------------------------------------------------------------------ *
 * torch.optim.Adam single step (train.py:139,184), one launch for all tensors laid out
 * back to back in flat buffers of `count` floats; step_dev: device int32 counter that
 * the kernel increments (so the launch is CUDA-graph replayable).                       */
int gdr_adam_step_f32(float *param, const float *grad, float *exp_avg, float *exp_avg_sq,
                      int64_t count, float lr, float beta1, float beta2, float eps,
                      int32_t *step_dev, gdr_stream_t stream);
/* The same update for up to 16 parameter tensors in one launch (train.py:139 hands Adam four).  The tables
 * are HOST arrays of device pointers / element counts; `step` is the 1-based step number of this update.   */
int gdr_adam_step_multi_f32(int n_tensors, float *const *params, const float *const *grads,
                            float *const *exp_avg, float *const *exp_avg_sq, const int64_t *counts,
                            int64_t step, double lr, double beta1, double beta2, double eps,
                            gdr_stream_t stream);

/* ---- diffusion profiles (multiscale/diff_prof/diffusion_profiles.py:29-90) ----------- *
 * Batched personalised-PageRank power iteration for `n_seeds` seeds at once, each with the
 * reference's per-seed matrix surgery applied on the fly (never materialised):
 *   convert_M_to_make_all_drugs_indications_sinks_except_selected (:29-47), refine_M_S (:49-56),
 *   power_iteration (:65-90).
 *   M   : CSR of the weighted MSI adjacency, M[u,v] = w(u->v) (get_initial_M, :21-27), fp64,
 *         sorted rows.  Mt: CSR of its transpose (gdr_csr_transpose: sources ascending).
 *   *_flag[e] (uint8 per stored entry, same order as the matching val array):
 *         bit0: entry (t,p), t a drug/indication and p in drug_or_indication2proteins[t]
 *               -> zero for every seed except t itself                        (:39-46)
 *         bit1: entry (p,t), same pairs -> zero for seed t only               (:33-36)
 *   seed_node[s]     : node index of seed s (personalisation = one-hot, :58-63).
 *   x_out [n_seeds, n]: diffusion profiles (row s = what the reference np.save()s for seed s).
 *   iters_out[n_seeds]: iteration (1-based) at which each seed met ||dx||_1 < n*tol.
 * Synchronises the stream once per iteration (convergence count).  Returns GDR_E_NOCONV if any
 * seed hit max_iter (the reference raises, :90).                                           */
int64_t gdr_rwr_workspace_bytes(int64_t n, int64_t nnz, int64_t n_seeds);
int gdr_rwr_solve_f64(const int32_t *m_rowptr, const int32_t *m_col, const double *m_val,
                      const uint8_t *m_flag, const int32_t *mt_rowptr, const int32_t *mt_col,
                      const double *mt_val, const uint8_t *mt_flag, int64_t n, int64_t nnz,
                      const int32_t *seed_node, int64_t n_seeds, double alpha, double tol,
                      int64_t max_iter, double *x_out, int32_t *iters_out, void *workspace,
                      int64_t workspace_bytes, gdr_stream_t stream);

/* The reference's per-seed methods with a caller-supplied matrix (one vector at a time):
 *   refine_M_S(M) (diffusion_profiles.py:49-56): s_out[n] = 1/rowsum(M) in numpy's reduceat order (0 stays 0),
 *       val_out[nnz] = s_out[row]*val (scipy's Q*M).  CSR rows sorted.
 *   power_iteration(M, S, nodelist, per_dict) (:65-90): Mt = CSR of the transpose of the row-scaled M (sources
 *       ascending, gdr_csr_transpose), s = the S above (only S == 0 matters), p[n] = per_dict values in nodelist
 *       order (normalised on the device, p / p.sum()).  x_out[n] device; *iters_out_host = iterations used.
 *       Synchronises the stream every iteration; GDR_E_NOCONV after max_iter iterations (the reference raises, :90). */
int gdr_rwr_refine_f64(const int32_t *rowptr, const int32_t *col, const double *val, int64_t n, int64_t nnz,
                       double *s_out, double *val_out, gdr_stream_t stream);
int64_t gdr_rwr_power_iteration_workspace_bytes(int64_t n);
int gdr_rwr_power_iteration_f64(const int32_t *mt_rowptr, const int32_t *mt_col, const double *mt_val, int64_t n,
                                const double *s, const double *p, double alpha, double tol, int64_t max_iter,
                                double *x_out, int32_t *iters_out_host, void *workspace, int64_t workspace_bytes,
                                gdr_stream_t stream);

/* ---- consumer ranking step (predict_drug.py:49-74,111-124; run_covid.py:294-306) ----------------------------- *
 * fp64 like the reference (np.loadtxt).  embs: [n, d] row-major.  prox_out[i] = <normalize(embs[sel[i]]),
 * normalize(embs[query_row])> (sklearn normalize: a zero norm divides by 1) when normalize != 0, else the raw dot
 * product (np.matmul(drug_embs, covid_emb)).  order_out[j] = position i into sel of the j-th largest score
 * (np.argsort(proximities)[::-1]; ties keep ascending position), sorted_out[j] (optional) = that score.
 * sel == NULL: rows 0..m-1.  gdr_rank_values_f64: values_out[i] = values[sel[i]] and its descending order.        */
int64_t gdr_rank_workspace_bytes(int64_t m);
int gdr_rank_embedding_f64(const double *embs, int64_t n, int64_t d, int normalize, int64_t query_row,
                           const int32_t *sel, int64_t m, double *prox_out, int32_t *order_out, double *sorted_out,
                           void *workspace, int64_t workspace_bytes, gdr_stream_t stream);
int gdr_rank_values_f64(const double *values, const int32_t *sel, int64_t m, double *values_out, int32_t *order_out,
                        double *sorted_out, void *workspace, int64_t workspace_bytes, gdr_stream_t stream);

/* ---- host-side text formats around the path (no device work) ------------------------------------------ *
 * graph_embs.txt exactly as np.savetxt(fname, hidden_emb) writes it (train.py:193; '%.18e', ' ', '\n', no header):
 * returns the bytes written or a negative GDR_E_* code.  x is a HOST array [n, d] of fp32 (4) or fp64 (8).      */
int64_t gdr_text_savetxt_bytes(int64_t n, int64_t d);
int64_t gdr_text_savetxt_e18(const void *x, int64_t n, int64_t d, int elem_bytes, char *out, int64_t out_cap,
                             int n_threads);
/* node2vec .embs.txt as train.py:79-80 consumes it (skip_rows header lines; first column = node name, the rest
 * float()).  values == NULL: returns the number of data lines, *d_out = value columns.  Otherwise fills
 * values[N, d] (fp64) and the byte span of every name; returns N or a negative code.                           */
int64_t gdr_text_parse_embs(const char *text, int64_t len, int skip_rows, double *values, int64_t d,
                            int64_t *name_off, int32_t *name_len, int64_t *d_out, int n_threads);

/* whole_graph.weighted.edgelist as nx.write_weighted_edgelist / nx.read_weighted_edgelist handle it
 * (predict_drug.py:225-226, run_covid.py:150-151): one line "u v w" per edge, w = Python's repr(float).
 * format: names = one blob + per-node (offset, length); returns bytes written, or -needed when out is NULL / too
 * small.  parse: pass 1 (w_out == NULL) returns the edge count ('#' comments and blank lines skipped), pass 2 fills
 * the byte spans of both node names and the weights.                                                             */
int64_t gdr_text_format_edgelist(const char *names, const int64_t *name_off, const int32_t *name_len, int64_t n_nodes,
                                 const int32_t *src, const int32_t *dst, const double *w, int64_t n_edges, char *out,
                                 int64_t out_cap, int n_threads);
int64_t gdr_text_parse_edgelist(const char *text, int64_t len, int64_t *u_off, int32_t *u_len, int64_t *v_off,
                                int32_t *v_len, double *w_out, int n_threads);

#ifdef __cplusplus
}
#endif
#endif /* GDR_H_ */
