/*
 * nichecompass_b200 -- C ABI of the B200-native VGPGAE hot path (libnichecompass_b200.so).
 *
 * The reference (Lotfollahi-lab/nichecompass v0.3.3) has no FFI of its own: every device op on
 * its hot path is reached through PyTorch / PyG.  Each entry point below names
 * the reference call site (path:line relative to /root/reference/src/nichecompass) whose
 * arithmetic it replaces.  The Python classes in nichecompass_b200/{nn,modules,data,train}
 * bind these through ctypes (nichecompass_b200/_lib.py); INTEGRATION.md shows the same binding
 * as a patch against the reference.
 *
 * Conventions
 *   - plain pointers + sizes, no torch types.  Every pointer is a DEVICE pointer unless the
 *     name ends in _host.  The caller owns every buffer (incl. workspaces); kernels never
 *     allocate, free or synchronise.  `stream` is a cudaStream_t passed as void*.
 *   - all launches are asynchronous on `stream` and CUDA-graph capturable.
 *   - return value: 0 (NC_OK) or a negative nc_status; nc_last_error() returns the message
 *     of the last failure on the calling thread.
 *   - element types of activations are selected by an nc_dtype argument (fp32 parity mode /
 *     bf16 tensor-core mode); reductions, losses and gradients of parameters are fp32
 *     (the NB likelihood is evaluated in fp64 internally).
 *   - local node / edge ids are int32; feature-store row pointers are int64.
 *   - "dst-CSR": rowptr[i]..rowptr[i+1] lists the LOCAL source ids (col) of the edges that
 *     point at local node i, i.e. the induced subgraph grouped by target (PyG order,
 *     SURVEY.md App. A.5).
 */
#ifndef NICHECOMPASS_B200_H
#define NICHECOMPASS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum { NC_OK = 0, NC_EINVAL = -1, NC_ESHAPE = -2, NC_ECUDA = -3, NC_EUNSUPPORTED = -4 } nc_status;
typedef enum { NC_F32 = 0, NC_BF16 = 1 } nc_dtype;

const char* nc_last_error(void);
int nc_version(void);
/* number of kernels launched by this library in this process so far (bench.py gpu_launches) */
int64_t nc_launch_count(void);
/* adds the launches of a replayed CUDA graph (the caller counted them at capture time); returns the new total */
int64_t nc_launch_count_add(int64_t n);

/* Page-locked, device-mapped host memory for a count store that stays in host memory
 * (data/store.py, location="host"; the reference keeps the dense count tensor in pageable host memory and
 * slices it on the CPU, data/datasets.py:59-71).  The block is owned by the caller until nc_host_free, which
 * waits for the device before unmapping; blocks are never recycled through a caching allocator.           */
int nc_host_alloc(int64_t bytes, void** out_host);
int nc_host_free(void* host);

/* ------------------------------------------------------------------------------------------
 * Graph helpers on a dst-CSR (replaces torch_sparse SparseStorage sort / fill_diag / sum and
 * PyG add_remaining_self_loops + degree scatter: nn/aggregators.py:205-209, PyG gcn_norm)
 * ---------------------------------------------------------------------------------------- */

/* Per-entry and per-row weights of the symmetric GCN normalisation with exactly one self
 * loop per node.  deg[v] = 1 + #{non-self entries whose SOURCE is v} if use_out_degree
 * (torch_sparse row-sum, aggregator A.3) else 1 + #{non-self entries of row v} (GCNConv A.2).
 * w[e] = deg[col[e]]^-1/2 * deg[i]^-1/2 (0 for self entries, which are replaced by the
 * single self term wself[i] = 1/deg[i]).  deg_ws: int32[n_nodes] scratch.               */
int nc_gcn_norm_weights(const int32_t* rowptr, const int32_t* col, int32_t n_nodes,
                        int32_t n_rows, int use_out_degree, int32_t* deg_ws, float* w,
                        float* wself, void* stream);

/* nc_gcn_norm_weights without the deg_ws argument and without any allocation: the degree scratch
 * is a static device array inside the library (4 MiB per GPU, held until exit), for callers
 * that cannot allocate (stream capture, serving loops).  NO COUNTERPART IN THE REFERENCE.
 * Limit: n_nodes <= 1 048 576 (NC_EUNSUPPORTED beyond
 * it).  Same w / wself as nc_gcn_norm_weights.  All calls on a device share the scratch: it is
 * cleared at the start of every call, and calls that overlap on the GPU (different streams
 * without ordering) race on it.                                                              */
int nc_gcn_norm_weights_noalloc(const int32_t* rowptr, const int32_t* col, int32_t n_nodes,
                                int32_t n_rows, int use_out_degree, float* w, float* wself,
                                void* stream);

/* Reverse (by-source) view of the first n_rows rows of a dst-CSR.  add_self == 1: self entries
 * dropped and one self edge i->i appended for every i < n_rows (GATv2Conv); 0: self entries
 * dropped; 2: every entry kept as given, nothing appended (attention aggregator).
 * rptr[n_nodes+1], rdst[*], rslot[*]: target row and "slot" of each out-edge, where the slot
 * of entry e of row i is e + i and the slot of the appended self edge is rowptr[i+1] + i
 * (the indexing used by nc_gatv2_*).  Entries of one source are sorted by slot.
 * cursor_ws: int32[n_nodes] scratch.                                                      */
int nc_reverse_csr(const int32_t* rowptr, const int32_t* col, int32_t n_nodes, int32_t n_rows,
                   int add_self, int32_t* rptr, int32_t* rdst, int32_t* rslot,
                   int32_t* cursor_ws, void* stream);

/* ------------------------------------------------------------------------------------------
 * One-hop node-label aggregators = weighted CSR SpMM with 128-bit gathers
 * (nn/aggregators.py:205-213 gcn-norm, :260-264 sum; torch_sparse spmm_sum)
 * out[i, :] = sum_e w[e] * x[col[e], :]  (+ wself[i] * x[i, :]),  i < n_rows.
 * w == NULL -> all ones; wself == NULL -> no extra self term; skip_self != 0 drops entries
 * with col[e] == i; bias (fp32 [n_cols], nullable) is added to every output row.
 * x: [*, ldx] of x_dtype; out: [n_rows, ldo] of out_dtype.
 * Also used for GCNConv message passing (PyG GCNConv.propagate) and its backward.        */
int nc_spmm_csr(const void* x, int x_dtype, int64_t ldx, int32_t n_cols,
                const int32_t* rowptr, const int32_t* col, const float* w, const float* wself,
                const float* bias, int skip_self, int32_t n_rows, void* out, int out_dtype,
                int64_t ldo, void* stream);

/* Same aggregation read straight from the HBM-resident CSR count store (no dense x):
 * out[i, :] = sum_e w[e] * X[n_id[col[e]], :] (+ wself[i] * X[n_id[i], :]), fp32 out.
 * (replaces data/datasets.py:59-71 dense x + loader row slicing + the SpMM above)        */
int nc_spmm_csr_store(const int64_t* x_rowptr, const int32_t* x_col, const float* x_val,
                      int32_t n_cols, const int32_t* n_id, const int32_t* rowptr,
                      const int32_t* col, const float* w, const float* wself, int skip_self,
                      int32_t n_rows, float* out, int64_t ldo, void* stream);

/* ------------------------------------------------------------------------------------------
 * Feature store -> dense encoder input tile
 * (data/datasets.py:59-71 + loader x[n_id] slicing + modules/vgpgae.py:598-601 log(1+x))
 * out[r, :] = f(X[n_id[r], :]) with f = log1p if apply_log1p; columns [n_cols, ldo) zeroed.
 * n_id == NULL -> identity.                                                               */
int nc_csr_gather_rows(const int64_t* x_rowptr, const int32_t* x_col, const float* x_val,
                       int32_t n_cols, const int32_t* n_id, int32_t n_rows, int apply_log1p,
                       void* out, int out_dtype, int64_t ldo, void* stream);
/* Same result as nc_csr_gather_rows from a kernel that can run BESIDE the compute stream: no shared memory,
 * at most max_ctas CTAs (<= 0: one per two SMs) striding over the rows -- used to stage the next iteration's
 * encoder input from a store in pinned host memory while the current iteration computes.      */
int nc_csr_gather_rows_bg(const int64_t* x_rowptr, const int32_t* x_col, const float* x_val,
                          int32_t n_cols, const int32_t* n_id, int32_t n_rows, int apply_log1p,
                          void* out, int out_dtype, int64_t ldo, int32_t max_ctas, void* stream);
int nc_dense_gather_rows(const float* x, int64_t ldx, int32_t n_cols, const int32_t* n_id,
                         int32_t n_rows, int apply_log1p, void* out, int out_dtype,
                         int64_t ldo, void* stream);

/* ------------------------------------------------------------------------------------------
 * Dense contractions (ATen addmm / cuBLAS call sites: nn/encoders.py:179 fc_l1, PyG
 * GATv2Conv lin_l / lin_r, GCNConv lin, nn/layercomponents.py:77-80 MaskedLinear, and all
 * their backward GEMMs)
 *   C[M,N] (+)= act( opA(A)[M,K] * opB(B)[K,N] + bias[N] ) (* mask[M,N])
 * A is stored [M,K] (lda) or, if trans_a, [K,M]; B is stored as the weight matrix [N,K]
 * (ldb) or, if trans_b, [K,N].  bias / mask may be NULL; mask is fp32 [M, ldmask];
 * relu != 0 applies max(0, .); accumulate != 0 adds to the existing C.
 * nc_gemm_f32: fp32 SIMT tiles (parity mode).
 * nc_gemm_bf16: bf16 operands, fp32 accumulation in TMEM (tcgen05.mma + TMA), C fp32 or
 * bf16.  Requires 16-byte aligned base pointers and lda/ldb multiples of 8 elements.      */
int nc_gemm_f32(const float* A, int64_t lda, int trans_a, const float* B, int64_t ldb,
                int trans_b, float* C, int64_t ldc, int32_t M, int32_t N, int32_t K,
                const float* bias, int relu, const float* mask, int64_t ldmask,
                int accumulate, void* stream);
int nc_gemm_bf16(const void* A, int64_t lda, int trans_a, const void* B, int64_t ldb,
                 int trans_b, void* C, int c_dtype, int64_t ldc, int32_t M, int32_t N,
                 int32_t K, const float* bias, int relu, const float* mask, int64_t ldmask,
                 int accumulate, void* stream);

/* Weight gradient of a MaskedLinear whose mask is sparse (nn/layercomponents.py:77-80 backward):
 * dW[f, p] += mval[e] * sum_b dlT[f, b] * zT[p, b] for every stored entry e = (f, p = mcol[e]) of
 * the mask in CSR form by feature (mrowptr int32[F+1]); mval == NULL means 1.  dlT [F, nb] and
 * zT [P, nb] are the transposed fp32 logit gradients / decoder inputs.  dW is ACCUMULATED.        */
int nc_masked_sddmm(const float* dlT, int64_t lddl, const float* zT, int64_t ldz, int32_t nb,
                    const int32_t* mrowptr, const int32_t* mcol, const float* mval, int32_t F,
                    float* dW, int64_t lddw, void* stream);

/* ------------------------------------------------------------------------------------------
 * Fused GATv2 edge kernels (PyG GATv2Conv.forward message passing; nn/encoders.py:124-140,
 * 197-207): gather x_l[j] + x_r[i] -> LeakyReLU -> att dot -> segment
 * softmax over the incoming edges of i -> alpha-weighted sum of x_l[j] -> mean over heads
 * -> + bias, for all convs of the encoder in one launch.
 * xl [n_src, ld], xr [n_dst, ld]: lin_l / lin_r outputs of all convs side by side; conv c
 * head h occupies columns [off_c + h*C_c, off_c + (h+1)*C_c); att uses the same indexing.
 * seg: int32[4*n_conv] = {off_c, C_c, out_off_c, 0} on the device, seg_host the same on the
 * host (alignment checks).  Self loops are re-normalised as PyG does: CSR entries with
 * col == i are ignored and one self edge is appended per target.  max_deg >= the longest
 * CSR row among the n_dst targets.
 * alpha [(rowptr[n_dst] + n_dst) * n_conv * H]: attention coefficient of "slot" s (entry e of
 * row i -> s = e + i, appended self edge -> s = rowptr[i+1] + i), saved for the backward.
 * out [n_dst, ldo] fp32, conv c written at columns [out_off_c, out_off_c + C_c).
 * head_ws (nullable): fp32 [n_dst, ld] workspace.  When it is given, the rows are bf16 and every
 * C_c / off_c / out_off_c is a multiple of 4 (n_conv * H <= 16 segments of <= 1536 columns), the
 * ring kernels of csrc/gatv2_ring.cu run: persistent CTAs that stream whole x_l rows through a
 * shared-memory ring (cp.async.bulk + mbarrier), one warp per (conv, head) segment.  head_ws then
 * receives the per-head outputs sum_j alpha_ijh x_l[j, h, :] (before the head mean) and must be
 * handed to nc_gatv2_bwd unchanged.  Otherwise the generic kernels run and head_ws is untouched. */
int nc_gatv2_fwd(const void* xl, const void* xr, int dtype, int64_t ld, const float* att,
                 const float* bias, const int32_t* rowptr, const int32_t* col, int32_t n_dst,
                 const int32_t* seg, const int32_t* seg_host, int32_t n_conv, int32_t H,
                 int32_t max_deg, float neg_slope, float* out, int64_t ldo, float* alpha,
                 float* head_ws, void* stream);
/* Forward only, over a TILE of targets of a large resident graph (full-graph latent inference,
 * models/nichecompass.py:1264-1381 / modules/vgpgae.py:1368-1437 without the loader's per-batch
 * recomputation): targets are rows [dst_off, dst_off + n_dst) of rowptr and -- for the self edge --
 * of xl; xr [n_dst, ld] and out [n_dst, ldo] hold the tile only.  col indexes xl rows.  No
 * attention coefficients are stored.  head_ws: as in nc_gatv2_fwd (scratch of the tile).      */
int nc_gatv2_fwd_tile(const void* xl, const void* xr, int dtype, int64_t ld, const float* att,
                      const float* bias, const int32_t* rowptr, const int32_t* col, int32_t dst_off,
                      int32_t n_dst, const int32_t* seg, const int32_t* seg_host, int32_t n_conv,
                      int32_t H, int32_t max_deg, float neg_slope, float* out, int64_t ldo,
                      float* head_ws, void* stream);
/* Backward of nc_gatv2_fwd.  rptr/rdst/rslot: nc_reverse_csr(..., add_self=1) of the same graph.
 * dxl [n_src, ld] and dxr [n_dst, ld] (same dtype as xl) are fully overwritten; datt (indexed
 * like att) and dbias (indexed like out columns, may be NULL) are ACCUMULATED with atomics.
 * ds_ws: fp32 workspace of the size of alpha.
 * head_ws: what nc_gatv2_fwd wrote (NULL: generic kernels).  bwd_ws / bwd_ws_bytes: scratch of
 * nc_gatv2_bwd_ws_bytes(n_slots, n_dst, n_conv, H) bytes, 256-byte aligned (LeakyReLU sign bits of
 * every (edge, element) + per-target softmax-backward constants); n_slots >= rowptr[n_dst] + n_dst.
 * With both given (and the shape conditions of nc_gatv2_fwd, n_conv * H a multiple of 4, alpha and
 * ds_ws 16-byte aligned) the ring kernels run: one streaming pass per target (d score, d x_r,
 * d att, sign bits) and one pass per source that reads the sign bits instead of
 * x_l[j] + x_r[i]; otherwise the generic kernels.                                             */
int nc_gatv2_bwd(const void* xl, const void* xr, int dtype, int64_t ld, const float* att,
                 const int32_t* rowptr, const int32_t* col, int32_t n_dst, int32_t n_src,
                 const int32_t* rptr, const int32_t* rdst, const int32_t* rslot,
                 const int32_t* seg, const int32_t* seg_host, int32_t n_conv, int32_t H,
                 int32_t max_deg, float neg_slope, const float* dout, int64_t ldo,
                 const float* alpha, float* ds_ws, void* dxl, void* dxr, float* datt,
                 float* dbias, const float* head_ws, void* bwd_ws, int64_t bwd_ws_bytes,
                 int64_t n_slots, void* stream);
int64_t nc_gatv2_bwd_ws_bytes(int64_t n_slots, int64_t n_dst, int32_t n_conv, int32_t H);

/* ------------------------------------------------------------------------------------------
 * Attention node-label aggregator (nn/aggregators.py:113-156).
 * gl [n_src, ld] = linear_l_l(x), gr [n_dst, ld] = linear_r_l(x) (head h in columns [h*F, (h+1)*F)),
 * attn [H*F].  Edges exactly as sampled (no self-loop re-normalisation).  The messages are the raw
 * rows x[j], identical for every head, so out[i] = sum_j w_ij x[j] with w = mean_h alpha:
 * nc_attn_agg_fwd writes alpha [(rowptr[n_dst] + n_dst) * H] (slot of entry e of row i = e + i)
 * and the per-entry SpMM weights w [rowptr[n_dst]]; the aggregation itself is nc_spmm_csr.
 * seg: device int32[4] = {0, F, 0, 0}.
 * nc_attn_agg_bwd: dout [n_dst, ldo] (gradient of out), x [n_src, ldx] fp32 raw rows;
 * rptr/rdst/rslot = nc_reverse_csr(..., add_self = 2); dalpha_ws / ds_ws: fp32 workspaces of the
 * size of alpha; dgl [n_src, ld] / dgr [n_dst, ld] overwritten, dattn ACCUMULATED.               */
int nc_attn_agg_fwd(const void* gl, const void* gr, int dtype, int64_t ld, const float* attn,
                    const int32_t* rowptr, const int32_t* col, int32_t n_dst, const int32_t* seg,
                    int32_t H, int32_t F, int32_t max_deg, float neg_slope, float* alpha, float* w,
                    void* stream);
int nc_attn_agg_bwd(const void* gl, const void* gr, int dtype, int64_t ld, const float* attn,
                    const int32_t* rowptr, const int32_t* col, int32_t n_dst, int32_t n_src,
                    const int32_t* rptr, const int32_t* rdst, const int32_t* rslot, const int32_t* seg,
                    int32_t H, int32_t F, int32_t max_deg, float neg_slope, const float* dout,
                    int64_t ldo, const float* x, int64_t ldx, const float* alpha, float* dalpha_ws,
                    float* ds_ws, void* dgl, void* dgr, float* dattn, void* stream);

/* ------------------------------------------------------------------------------------------
 * Decoder epilogue + negative-binomial loss, forward value and gradients in one pass
 * (nn/layers.py:143 softmax over ALL features, nn/decoders.py:168 exp(log_lib)*softmax,
 * modules/vgpgae.py:714-749 column selection + library size over selected columns,
 * modules/losses.py:169-212 NB log-likelihood, theta = exp(theta_param)).
 * logits [n_rows, ldl], x [n_rows, ldx] raw labels over all F features, sel[F] != 0 marks the
 * reconstructed columns.  loss_sum += scale * sum_rows sum_sel -ll (fp64; `scale` = lambda / n_rows); dlogits = scale * d(sum -ll)/dlogits; dtheta += scale * ...;
 * means (nullable) receives exp(log_lib)*softmax over all F columns; dx (nullable) the gradient
 * w.r.t. the labels x (direct term + library size; labels produced by the attention aggregator
 * are differentiable, modules/vgpgae.py:704-726).                       */
int nc_nb_loss_fused(const float* logits, int64_t ldl, const float* x, int64_t ldx,
                     const uint8_t* sel, const float* theta_param, int32_t n_rows, int32_t F,
                     float scale, float* means, int64_t ldm, double* loss_sum, float* dlogits,
                     int64_t lddl, float* dtheta, float* dx, int64_t lddx, void* stream);

/* Negative-binomial loss on given means, as the reference exports it (modules/losses.py:169-212
 * compute_omics_recon_nb_loss(x, mu, theta, eps)): loss_sum += scale * sum_{r,c} -ll (fp64); dmu / dtheta
 * (nullable; dtheta is accumulated) = scale * d(sum -ll)/d.  theta: positive inverse dispersion [F].   */
int nc_nb_loss_means(const float* x, int64_t ldx, const float* mu, int64_t ldmu, const float* theta,
                     int32_t n_rows, int32_t F, float eps, float scale, double* loss_sum, float* dmu,
                     int64_t lddmu, float* dtheta, void* stream);

/* Cosine-similarity graph decoder (nn/decoders.py:56-62, nn/utils.py:8-36): rows [0,n_pairs)
 * of z against rows [n_pairs, 2*n_pairs).  norms: fp32[2*n_pairs] saved for the backward.   */
int nc_cosine_fwd(const float* z, int64_t ldz, int32_t P, int32_t n_pairs, float* logits,
                  float* norms, void* stream);
int nc_cosine_bwd(const float* z, int64_t ldz, int32_t P, int32_t n_pairs, const float* logits,
                  const float* norms, const float* dlogits, float* dz, int64_t lddz,
                  void* stream);
/* BCE-with-logits over the included edges (incl NULL = all): use_pos_weight != 0 weighs the positives
 * by #neg/#pos (modules/losses.py:123-166), 0 is the plain mean BCE of the covariate contrastive loss
 * (modules/losses.py:103-118); loss_out = mean loss, dlogits = scale * dloss/dlogits.       */
int nc_bce_fwd_bwd(const float* logits, const float* labels, const uint8_t* incl, int32_t n,
                   float scale, float* loss_out, float* dlogits, int use_pos_weight, void* stream);
/* KL regulariser (modules/losses.py:255-284): sum_out += kl (fp64); dmu/dlogstd (nullable)
 * = scale * dkl/d.                                                                          */
int nc_kl_fwd_bwd(const float* mu, int64_t ldmu, const float* logstd, int64_t ldls, int32_t R,
                  int32_t P, float scale, double* sum_out, float* dmu, int64_t lddmu,
                  float* dlogstd, int64_t lddls, void* stream);

/* ------------------------------------------------------------------------------------------
 * Latent plumbing (modules/vgpgae.py:626-651, modules/vgaemodulemixin.py:34-41)
 * ---------------------------------------------------------------------------------------- */
int nc_gather_rows_f32(const float* src, int64_t lds, const int32_t* idx, int32_t R, int32_t C,
                       float* dst, int64_t ldd, void* stream);
int nc_scatter_add_rows_f32(const float* src, int64_t lds, const int32_t* idx, int32_t R,
                            int32_t C, float* dst, int64_t ldd, void* stream);
/* z = (mu + eps * exp(logstd)) with inactive gene programs zeroed (eps NULL -> z = mu,
 * active NULL -> all active); optional bf16 copy for the decoder GEMM.                      */
int nc_reparam_fwd(const float* mu, int64_t ldmu, const float* logstd, int64_t ldls,
                   const float* eps, int64_t ldeps, const uint8_t* active, int32_t R, int32_t P,
                   float* z, int64_t ldz, void* z_bf16, int64_t ldzb, void* stream);
int nc_reparam_bwd(const float* logstd, int64_t ldls, const float* eps, int64_t ldeps,
                   const uint8_t* active, int32_t R, int32_t P, const float* dz, int64_t lddz,
                   float* dmu, int64_t lddmu, float* dlogstd, int64_t lddls, void* stream);
/* ema = momentum * mean_rows |mu| + (1 - momentum) * ema  (vgpgae.py:641-646)               */
int nc_abs_mean_ema(const float* mu, int64_t ldmu, int32_t R, int32_t P, float momentum,
                    float* ema, void* stream);

/* ------------------------------------------------------------------------------------------
 * Gene-program weight regularisers and active-GP statistics on a decoder weight W [F, P]
 * (modules/losses.py:215-252 group lasso, :287-339 L1; modules/vgpgae.py:1281-1312).
 * Column statistics are ACCUMULATED (call once per decoder): abs_sum[c] += sum_f |W| * l1mask,
 * sq_sum[c] += sum_f W^2, nnz[c] += #{W != 0}, mask_cnt[c] += #{mask != 0}.                 */
int nc_weight_col_stats(const float* W, int64_t ld, int32_t F, int32_t P, const uint8_t* l1mask,
                        const float* mask, int64_t ldmask, float* abs_sum, float* sq_sum,
                        float* nnz, float* mask_cnt, void* stream);
/* dW += gscale * (l1_scale * sign(W) * l1mask + gl_coef[c] * W); gscale: device scalar (the
 * upstream gradient of the loss term), NULL means 1.                                          */
int nc_weight_reg_grad(const float* W, int64_t ld, int32_t F, int32_t P, float l1_scale,
                       const uint8_t* l1mask, const float* gl_coef, const float* gscale, float* dW,
                       int64_t lddw, void* stream);

/* Fused Adam over a flat parameter segment with torch.optim.Adam semantics
 * (train/trainer.py:339-342,418-421): optional clip_grad_value_ (clip_value > 0), optional
 * refresh of the bf16 shadow copy used by the tensor-core GEMMs; zero_grad != 0 clears g after
 * it has been consumed (optimizer.zero_grad() of the next iteration, train/trainer.py:414).
 * With weight_decay == 0 a 16-byte group of parameters whose g, m and v are all zero keeps its
 * value bit for bit (torch's update is 0 / (0 + eps) there) and is not written back -- p_bf16
 * must therefore already hold the cast of p when the first step is taken.                     */
int nc_adam_step(float* p, float* g, float* m, float* v, int64_t n, float lr, float beta1,
                 float beta2, float eps, float weight_decay, int32_t step, float clip_value,
                 void* p_bf16, int zero_grad, void* stream);
/* The same update with {lr, step} read from device memory (dyn[0], dyn[1]; step >= 1 = the number of this
 * update) and the step advanced on the device afterwards: no launch argument depends on host state, so the
 * launch can be replayed from a captured CUDA graph.                                                      */
int nc_adam_step_dyn(float* p, float* g, float* m, float* v, int64_t n, float* dyn, float beta1,
                     float beta2, float eps, float weight_decay, float clip_value, void* p_bf16,
                     int zero_grad, void* stream);
/* fp32 [R, C] (lds) -> bf16 dst [R, fill_cols] (row pitch ldd; columns [C, fill_cols) zeroed;
 * fill_cols <= 0 means ldd), optionally multiplied by `mask` (nn/layercomponents.py:77-80:
 * weight * mask before the contraction).                                                    */
int nc_cast_bf16(const float* src, int64_t lds, const float* mask, int64_t ldmask, int32_t R,
                 int32_t C, void* dst, int64_t ldd, int32_t fill_cols, void* stream);

/* ------------------------------------------------------------------------------------------
 * GPU CSR neighbourhood sampler (PyG NeighborLoader / LinkNeighborLoader + pyg-lib
 * neighbor_sample + induced subgraph: data/dataloaders.py:64-100).
 * g_rowptr/g_col: CSC of the message-passing graph (incoming neighbours sorted by source id).
 * seeds: unique global node ids.  n_hops >= 1; k_neighbors < 0 keeps all neighbours, otherwise
 * the k slots with the smallest splitmix64(rng_seed, batch_id, node, slot) keys.
 * slot_ws: int32[n_graph_nodes] persistent workspace that must hold INT32_MAX everywhere on
 * entry and is restored on exit (nc_fill_i32 once).  Outputs: n_id[cap_nodes] (local -> global,
 * seeds first, then discovery order), local dst-CSR rowptr[cap_nodes+1] / col[cap_edges].
 * cand_ws: int32[2*cap_nodes + 3*cap_cand + 2] scratch.  counts: int32[8] on the device =
 * {n_sub, E_sub, overflow flag, longest row, ...}; nothing is read back by the library.      */
int nc_sample_subgraph(const int32_t* g_rowptr, const int32_t* g_col, const int32_t* seeds,
                       int32_t n_seed, int32_t n_hops, int32_t k_neighbors, int64_t rng_seed,
                       int64_t batch_id, int32_t* slot_ws, int32_t* n_id, int32_t cap_nodes,
                       int32_t* rowptr, int32_t* col, int32_t cap_edges, int32_t* cand_ws,
                       int32_t cap_cand, int32_t* counts, void* stream);

/* Two sampled sub-graphs (a: the node batch, b: the edge batch of one training iteration,
 * train/trainer.py:368-389 runs the encoder on each) merged into ONE block-diagonal encoder problem:
 * merged node order = [a seeds | b seeds | a others | b others]; m_nid[a_sub + b_sub] the global ids in
 * that order, m_rowptr[a_seed + b_seed + 1] / m_col the incoming-neighbour rows of the merged targets
 * (a's entries first; m_col needs room for a_rowptr[a_seed] + b_rowptr[b_seed] entries).           */
int nc_merge_subgraphs(const int32_t* a_rowptr, const int32_t* a_col, const int32_t* a_nid,
                       int32_t a_seed, int32_t a_sub, const int32_t* b_rowptr, const int32_t* b_col,
                       const int32_t* b_nid, int32_t b_seed, int32_t b_sub, int32_t* m_rowptr,
                       int32_t* m_col, int32_t* m_nid, void* stream);
/* The same merge into PADDED buffers (size buckets of the CUDA-graph training step): b_seed_pad >= b_seed merged
 * edge seeds, n_sub_pad nodes, e_pad entries; merged node order = [a seeds | b seeds | pad seeds | a others |
 * b others | pad others]; pad seeds have empty rows, pad nodes repeat a real node id and are referenced by no
 * edge.  m_rowptr[a_seed + b_seed_pad + 1], m_col[e_pad], m_nid[n_sub_pad].                                */
int nc_merge_subgraphs_padded(const int32_t* a_rowptr, const int32_t* a_col, const int32_t* a_nid,
                              int32_t a_seed, int32_t a_sub, const int32_t* b_rowptr, const int32_t* b_col,
                              const int32_t* b_nid, int32_t b_seed, int32_t b_sub, int32_t b_seed_pad,
                              int32_t n_sub_pad, int32_t e_pad, int32_t* m_rowptr, int32_t* m_col,
                              int32_t* m_nid, void* stream);
/* torch.unique(ids, return_inverse=True) for the link-batch seeds (sorted ascending), n <= 16384. */
int nc_unique_sorted_i32(const int32_t* ids, int32_t n, int32_t* uniq, int32_t* inverse,
                         int32_t* n_unique, void* stream);
int nc_fill_i32(int32_t* p, int64_t n, int32_t value, void* stream);

/* Bias gradients and ReLU backward of the encoder linears (autograd of nn/encoders.py:179):
 * out[c] += sum_r x[r, c];   out = dh * (h > 0) cast to the activation dtype.              */
int nc_col_sum(const void* x, int dtype, int64_t ld, int32_t R, int32_t C, float* out,
               void* stream);
int nc_relu_bwd(const float* dh, int64_t lddh, const void* h, int dtype, int64_t ldh, int32_t R,
                int32_t C, void* out, int64_t ldo, void* stream);
/* out [R, ldo] (dtype) = relu(x [R, ldx] fp32), pad columns zeroed: the activation between the two conv
 * layers of the 2-layer encoder (nn/encoders.py:191-194, ``hidden = activation(conv_l1(hidden))``).   */
int nc_relu_cast(const float* x, int64_t ldx, int32_t R, int32_t C, void* out, int dtype, int64_t ldo,
                 void* stream);

/* ------------------------------------------------------------------------------------------
 * Data-parallel training over NVLink peer memory (one process per GPU on one NVSwitch box).
 * The reference has no data parallelism (docs/installation.md:3-5); its optimiser step is
 * torch.optim.Adam.step() after loss.backward() (train/trainer.py:414-421).  nc_dp_adam_step replaces
 * ncclAllReduce(gradients) + nc_adam_step by ONE kernel: reduce-scatter of the gradient arenas through
 * peer loads, Adam on the owned shard, all-gather of the new parameters (+ bf16 shadow) and clearing of
 * the gradients through peer stores, cross-GPU barriers on flags in peer memory.
 * nc_peer_*: device memory that other processes on the box can map (cudaIpc); synchronous, no stream.
 * ---------------------------------------------------------------------------------------- */
int nc_peer_alloc(int64_t bytes, void** out_dev);
int nc_peer_free(void* dev);
int nc_peer_get_handle(void* dev, void* handle_host);          /* handle_host: 64 bytes */
int nc_peer_open(const void* handle_host, void** out_dev);
int nc_peer_close(void* dev);
/* *_ptrs_host: `world` device pointers each (entry `rank`: own buffer; others: nc_peer_open mappings);
 * shadow_ptrs_host NULL in fp32 mode; flag buffers: 2 * world uint32 each, zero-initialised; m, v: the
 * moments of this rank's shard only (ceil(numel / world / 4) * 4 floats); done_counter: one zeroed uint32.
 * Every rank calls with the same `step` sequence (1, 2, ...): it is the bias-correction step AND the
 * barrier epoch.  dyn != NULL: {lr, step} are read from device memory instead (and the step advanced by the
 * kernel), so that the launch can be replayed from a CUDA graph.  aux_ptrs_host (nullable): `world` peer pointers to
 * aux_n floats each (the |mu| running mean, modules/vgpgae.py:641-651) that are replaced by their mean over the ranks
 * once every peer has read them (aux_avg: aux_n floats of local scratch); done_counter: TWO zeroed uint32.
 * n_ctas <= 0: one CTA per SM.                                                                     */
int nc_dp_adam_step(void* const* p_ptrs_host, void* const* g_ptrs_host, void* const* shadow_ptrs_host,
                    void* const* flag_ptrs_host, float* m, float* v, int64_t numel, int32_t rank,
                    int32_t world, float lr, float beta1, float beta2, float eps, float weight_decay,
                    int32_t step, float clip_value, uint32_t* done_counter, int32_t n_ctas, float* dyn,
                    void* const* aux_ptrs_host, float* aux_avg, int32_t aux_n, void* stream);

/* Halo rows of a spatially partitioned full-graph pass read straight from their owners' HBM over NVLink
 * (models/tiled_inference.py; the reference runs latent inference on one device, models/nichecompass.py:1264-1381):
 * out[r, :row_bytes] = peer[owner[r]][pos[r], :row_bytes].  peer_ptrs_host: `world` device pointers (own buffer +
 * nc_peer_open mappings), row pitch src_pitch bytes; all byte counts multiples of 16.  The caller orders the call
 * after every owner has written its rows.                                                                    */
int nc_halo_gather_rows(void* const* peer_ptrs_host, int32_t world, const int32_t* owner, const int64_t* pos,
                        int64_t n_rows, int64_t src_pitch, int64_t row_bytes, void* out, int64_t out_pitch,
                        void* stream);

#ifdef __cplusplus
}
#endif
#endif /* NICHECOMPASS_B200_H */
