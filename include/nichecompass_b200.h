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

/* Reverse (by-source) view of the first n_rows rows of a dst-CSR, self entries dropped and
 * one self edge i->i appended for every i < n_rows when add_self != 0.
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
 * with col[e] == i.  x: [*, ldx] of x_dtype; out: [n_rows, ldo] of out_dtype.
 * Also used for GCNConv message passing (PyG GCNConv.propagate) and its backward.        */
int nc_spmm_csr(const void* x, int x_dtype, int64_t ldx, int32_t n_cols,
                const int32_t* rowptr, const int32_t* col, const float* w, const float* wself,
                int skip_self, int32_t n_rows, void* out, int out_dtype, int64_t ldo,
                void* stream);

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

#ifdef __cplusplus
}
#endif
#endif /* NICHECOMPASS_B200_H */
