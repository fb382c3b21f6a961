// Graph helpers, one-hop aggregator SpMM, and feature-store gathers.
// Reference call sites: nn/aggregators.py:205-213,260-264 (torch_sparse spmm_sum + gcn_norm),
// data/datasets.py:59-71 + PyG loader x[n_id] slicing, modules/vgpgae.py:598-601.
#include "common.cuh"

extern void nc_count_launch(int n);

// ============================================================================ gcn-norm weights
__global__ void k_zero_i32(int32_t* p, int32_t n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = 0;
}

__global__ void k_degree(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
                         int32_t n_nodes, int use_out_degree, int32_t* __restrict__ deg) {
    int v = blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n_nodes) return;
    int cnt = 0;
    for (int e = rowptr[v]; e < rowptr[v + 1]; ++e) {
        int j = col[e];
        if (j == v) continue;
        if (use_out_degree) atomicAdd(&deg[j], 1);
        else ++cnt;
    }
    if (!use_out_degree) deg[v] = cnt;
}

__global__ void k_norm_weights(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
                               int32_t n_rows, const int32_t* __restrict__ deg,
                               float* __restrict__ w, float* __restrict__ wself) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_rows) return;
    const float di = 1.0f / sqrtf((float)(deg[i] + 1));
    for (int e = rowptr[i]; e < rowptr[i + 1]; ++e) {
        int j = col[e];
        w[e] = (j == i) ? 0.f : (1.0f / sqrtf((float)(deg[j] + 1))) * di;
    }
    wself[i] = di * di;
}

extern "C" int nc_gcn_norm_weights(const int32_t* rowptr, const int32_t* col, int32_t n_nodes,
                                   int32_t n_rows, int use_out_degree, int32_t* deg_ws, float* w,
                                   float* wself, void* stream) {
    NC_REQUIRE(n_nodes >= 0 && n_rows >= 0 && n_rows <= n_nodes, NC_ESHAPE,
               "nc_gcn_norm_weights: need 0 <= n_rows <= n_nodes (got %d, %d)", n_rows, n_nodes);
    if (n_nodes == 0) return NC_OK;
    cudaStream_t s = (cudaStream_t)stream;
    const int T = 256;
    k_zero_i32<<<nc_cdiv(n_nodes, T), T, 0, s>>>(deg_ws, n_nodes);
    k_degree<<<nc_cdiv(n_nodes, T), T, 0, s>>>(rowptr, col, n_nodes, use_out_degree, deg_ws);
    if (n_rows > 0) k_norm_weights<<<nc_cdiv(n_rows, T), T, 0, s>>>(rowptr, col, n_rows, deg_ws, w, wself);
    nc_count_launch(n_rows > 0 ? 3 : 2);
    NC_CHECK_LAUNCH("nc_gcn_norm_weights");
    return NC_OK;
}

// ---------------------------------------------------------------------------- no-allocation variant
// Same three kernels, degree scratch in a file-scope device array instead of a caller buffer: for callers that cannot
// allocate (stream capture, serving loops).  No counterpart in the reference.
// 4 MiB of every GPU that loads the library, held until exit.  Every call on a device shares the array:
// it is cleared at the start of each call (k_zero_i32), and two calls that overlap on the GPU (different streams)
// race on it.
#define NC_NOALLOC_MAX_NODES (1 << 20)
static __device__ __align__(256) int32_t g_noalloc_deg[NC_NOALLOC_MAX_NODES];

// Device address of g_noalloc_deg on the CURRENT device (the name is only a host-side handle in host code),
// resolved once per device.
static int noalloc_deg_ptr(int32_t** out) {
    static int32_t* ptrs[64] = {nullptr};
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess || dev < 0 || dev >= 64) {
        nc_set_error("nc_gcn_norm_weights_noalloc: cudaGetDevice failed or device index %d out of range: %s", dev,
                     cudaGetErrorString(e));
        return NC_ECUDA;
    }
    int32_t* p = __atomic_load_n(&ptrs[dev], __ATOMIC_ACQUIRE);
    if (p == nullptr) {
        e = cudaGetSymbolAddress((void**)&p, g_noalloc_deg);
        if (e != cudaSuccess || p == nullptr) {
            cudaGetLastError();
            nc_set_error("nc_gcn_norm_weights_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev,
                         cudaGetErrorString(e));
            return NC_ECUDA;
        }
        __atomic_store_n(&ptrs[dev], p, __ATOMIC_RELEASE);       // every thread resolves the same address
    }
    *out = p;
    return NC_OK;
}

extern "C" int nc_gcn_norm_weights_noalloc(const int32_t* rowptr, const int32_t* col, int32_t n_nodes,
                                           int32_t n_rows, int use_out_degree, float* w, float* wself,
                                           void* stream) {
    NC_REQUIRE(n_nodes >= 0 && n_rows >= 0 && n_rows <= n_nodes, NC_ESHAPE,
               "nc_gcn_norm_weights_noalloc: need 0 <= n_rows <= n_nodes (got %d, %d)", n_rows, n_nodes);
    NC_REQUIRE(n_nodes <= NC_NOALLOC_MAX_NODES, NC_EUNSUPPORTED,
               "nc_gcn_norm_weights_noalloc: n_nodes = %d exceeds the static scratch of %d nodes; use "
               "nc_gcn_norm_weights with a caller-provided deg_ws", n_nodes, NC_NOALLOC_MAX_NODES);
    if (n_nodes == 0) return NC_OK;
    int32_t* deg = nullptr;
    int rc = noalloc_deg_ptr(&deg);
    if (rc != NC_OK) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    const int T = 256;
    k_zero_i32<<<nc_cdiv(n_nodes, T), T, 0, s>>>(deg, n_nodes);
    k_degree<<<nc_cdiv(n_nodes, T), T, 0, s>>>(rowptr, col, n_nodes, use_out_degree, deg);
    if (n_rows > 0) k_norm_weights<<<nc_cdiv(n_rows, T), T, 0, s>>>(rowptr, col, n_rows, deg, w, wself);
    nc_count_launch(n_rows > 0 ? 3 : 2);
    NC_CHECK_LAUNCH("nc_gcn_norm_weights_noalloc");
    return NC_OK;
}

// ============================================================================ reverse CSR
__global__ void k_rev_count(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
                            int32_t n_rows, int add_self, int32_t* __restrict__ cnt) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_rows) return;
    for (int e = rowptr[i]; e < rowptr[i + 1]; ++e) {
        int j = col[e];
        if (j != i || add_self == 2) atomicAdd(&cnt[j], 1);
    }
    if (add_self == 1) atomicAdd(&cnt[i], 1);
}

// single-CTA exclusive scan: out[0..n] (n+1 entries), in[0..n)
__global__ void k_exclusive_scan_i32(const int32_t* __restrict__ in, int32_t* __restrict__ out, int32_t n) {
    __shared__ int32_t warp_tot[32];
    __shared__ int32_t carry_s;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < n; base += blockDim.x) {
        int i = base + tid;
        int v = (i < n) ? in[i] : 0;
        int x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_tot[warp] = x;
        __syncthreads();
        if (warp == 0) {
            int t = (lane < (blockDim.x >> 5)) ? warp_tot[lane] : 0;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int y = __shfl_up_sync(0xffffffffu, t, o);
                if (lane >= o) t += y;
            }
            warp_tot[lane] = t;
        }
        __syncthreads();
        int carry = carry_s;
        int excl = carry + (warp > 0 ? warp_tot[warp - 1] : 0) + x - v;
        if (i < n) out[i] = excl;
        __syncthreads();
        if (tid == blockDim.x - 1) carry_s = carry + warp_tot[(blockDim.x >> 5) - 1];
        __syncthreads();
    }
    if (tid == 0) out[n] = carry_s;
}

__global__ void k_rev_fill(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
                           int32_t n_rows, int add_self, const int32_t* __restrict__ rptr,
                           int32_t* __restrict__ cursor, int32_t* __restrict__ rdst,
                           int32_t* __restrict__ rslot) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_rows) return;
    for (int e = rowptr[i]; e < rowptr[i + 1]; ++e) {
        int j = col[e];
        if (j == i && add_self != 2) continue;
        int pos = rptr[j] + atomicAdd(&cursor[j], 1);
        rdst[pos] = i;
        rslot[pos] = e + i;
    }
    if (add_self == 1) {
        int pos = rptr[i] + atomicAdd(&cursor[i], 1);
        rdst[pos] = i;
        rslot[pos] = rowptr[i + 1] + i;
    }
}

__global__ void k_rev_sort(const int32_t* __restrict__ rptr, int32_t n_nodes,
                           int32_t* __restrict__ rdst, int32_t* __restrict__ rslot) {
    int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n_nodes) return;
    const int lo = rptr[j], hi = rptr[j + 1];
    for (int a = lo + 1; a < hi; ++a) {           // insertion sort by slot (segments are tiny)
        int s = rslot[a], d = rdst[a], b = a - 1;
        while (b >= lo && rslot[b] > s) {
            rslot[b + 1] = rslot[b];
            rdst[b + 1] = rdst[b];
            --b;
        }
        rslot[b + 1] = s;
        rdst[b + 1] = d;
    }
}

extern "C" int nc_reverse_csr(const int32_t* rowptr, const int32_t* col, int32_t n_nodes,
                              int32_t n_rows, int add_self, int32_t* rptr, int32_t* rdst,
                              int32_t* rslot, int32_t* cursor_ws, void* stream) {
    NC_REQUIRE(n_nodes >= 0 && n_rows >= 0 && n_rows <= n_nodes, NC_ESHAPE,
               "nc_reverse_csr: need 0 <= n_rows <= n_nodes (got %d, %d)", n_rows, n_nodes);
    cudaStream_t s = (cudaStream_t)stream;
    const int T = 256;
    if (n_nodes > 0) k_zero_i32<<<nc_cdiv(n_nodes, T), T, 0, s>>>(cursor_ws, n_nodes);
    if (n_rows > 0) k_rev_count<<<nc_cdiv(n_rows, T), T, 0, s>>>(rowptr, col, n_rows, add_self, cursor_ws);
    k_exclusive_scan_i32<<<1, 1024, 0, s>>>(cursor_ws, rptr, n_nodes);
    if (n_nodes > 0) k_zero_i32<<<nc_cdiv(n_nodes, T), T, 0, s>>>(cursor_ws, n_nodes);
    if (n_rows > 0) {
        k_rev_fill<<<nc_cdiv(n_rows, T), T, 0, s>>>(rowptr, col, n_rows, add_self, rptr, cursor_ws, rdst, rslot);
        k_rev_sort<<<nc_cdiv(n_nodes, T), T, 0, s>>>(rptr, n_nodes, rdst, rslot);
    }
    nc_count_launch(6);
    NC_CHECK_LAUNCH("nc_reverse_csr");
    return NC_OK;
}

// ============================================================================ dense-x CSR SpMM
// grid = (n_rows, column chunks); block = 128 threads; each thread owns VPT float4 vectors.
constexpr int SPMM_T = 128;
constexpr int SPMM_VPT = 2;
constexpr int SPMM_CHUNK = SPMM_T * 4 * SPMM_VPT;   // 1024 columns per CTA
constexpr int SPMM_STAGE = 128;                      // edges staged in smem per pass

template <typename TX, typename TO, bool VEC>
__global__ void __launch_bounds__(SPMM_T)
k_spmm_csr(const TX* __restrict__ x, int64_t ldx, int32_t n_cols,
           const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
           const float* __restrict__ w, const float* __restrict__ wself, const float* __restrict__ bias,
           int skip_self, TO* __restrict__ out, int64_t ldo) {
    __shared__ int32_t s_col[SPMM_STAGE];
    __shared__ float s_w[SPMM_STAGE];
    const int i = blockIdx.x;
    const int c_base = blockIdx.y * SPMM_CHUNK;
    const int tid = threadIdx.x;
    float4 acc[SPMM_VPT];
#pragma unroll
    for (int v = 0; v < SPMM_VPT; ++v) acc[v] = make_float4(0.f, 0.f, 0.f, 0.f);
    const int e0 = rowptr[i], e1 = rowptr[i + 1];
    const bool has_self = (wself != nullptr);
    const int total = (e1 - e0) + (has_self ? 1 : 0);
    for (int base = 0; base < total; base += SPMM_STAGE) {
        const int n_here = min(SPMM_STAGE, total - base);
        __syncthreads();
        if (tid < n_here) {
            int t = base + tid;
            int j;
            float wt;
            if (t < e1 - e0) {
                j = col[e0 + t];
                wt = w ? w[e0 + t] : 1.f;
                if (skip_self && j == i) wt = 0.f;
            } else {
                j = i;
                wt = wself[i];
            }
            s_col[tid] = j;
            s_w[tid] = wt;
        }
        __syncthreads();
#pragma unroll 4
        for (int t = 0; t < n_here; ++t) {
            const float wt = s_w[t];
            if (wt == 0.f) continue;
            const TX* xr = x + (int64_t)s_col[t] * ldx;
#pragma unroll
            for (int v = 0; v < SPMM_VPT; ++v) {
                const int c = c_base + (v * SPMM_T + tid) * 4;
                if (VEC) {
                    if (c + 3 < n_cols) {
                        float4 q = Elem<TX>::ld4(xr + c);
                        acc[v].x += wt * q.x; acc[v].y += wt * q.y;
                        acc[v].z += wt * q.z; acc[v].w += wt * q.w;
                        continue;
                    }
                }
                if (c < n_cols) acc[v].x += wt * Elem<TX>::ld(xr + c);
                if (c + 1 < n_cols) acc[v].y += wt * Elem<TX>::ld(xr + c + 1);
                if (c + 2 < n_cols) acc[v].z += wt * Elem<TX>::ld(xr + c + 2);
                if (c + 3 < n_cols) acc[v].w += wt * Elem<TX>::ld(xr + c + 3);
            }
        }
    }
    TO* orow = out + (int64_t)i * ldo;
#pragma unroll
    for (int v = 0; v < SPMM_VPT; ++v) {
        const int c = c_base + (v * SPMM_T + tid) * 4;
        if (bias) {
            if (c < n_cols) acc[v].x += __ldg(bias + c);
            if (c + 1 < n_cols) acc[v].y += __ldg(bias + c + 1);
            if (c + 2 < n_cols) acc[v].z += __ldg(bias + c + 2);
            if (c + 3 < n_cols) acc[v].w += __ldg(bias + c + 3);
        }
        if (VEC && c + 3 < n_cols) {
            Elem<TO>::st4(orow + c, acc[v]);
        } else {
            if (c < n_cols) Elem<TO>::st(orow + c, acc[v].x);
            if (c + 1 < n_cols) Elem<TO>::st(orow + c + 1, acc[v].y);
            if (c + 2 < n_cols) Elem<TO>::st(orow + c + 2, acc[v].z);
            if (c + 3 < n_cols) Elem<TO>::st(orow + c + 3, acc[v].w);
        }
    }
}

template <typename TX, typename TO>
static int launch_spmm(const void* x, int64_t ldx, int32_t n_cols, const int32_t* rowptr,
                       const int32_t* col, const float* w, const float* wself, const float* bias, int skip_self,
                       int32_t n_rows, void* out, int64_t ldo, cudaStream_t s) {
    const size_t ax = sizeof(TX) * 4, ao = sizeof(TO) * 4;
    const bool vec = (ldx % 4 == 0) && (ldo % 4 == 0) && ((uintptr_t)x % ax == 0) && ((uintptr_t)out % ao == 0);
    dim3 grid(n_rows, (unsigned)nc_cdiv(n_cols, SPMM_CHUNK));
    if (vec)
        k_spmm_csr<TX, TO, true><<<grid, SPMM_T, 0, s>>>((const TX*)x, ldx, n_cols, rowptr, col, w, wself, bias, skip_self, (TO*)out, ldo);
    else
        k_spmm_csr<TX, TO, false><<<grid, SPMM_T, 0, s>>>((const TX*)x, ldx, n_cols, rowptr, col, w, wself, bias, skip_self, (TO*)out, ldo);
    nc_count_launch(1);
    return 0;
}

extern "C" int nc_spmm_csr(const void* x, int x_dtype, int64_t ldx, int32_t n_cols,
                           const int32_t* rowptr, const int32_t* col, const float* w,
                           const float* wself, const float* bias, int skip_self, int32_t n_rows,
                           void* out, int out_dtype, int64_t ldo, void* stream) {
    NC_REQUIRE(n_cols > 0 && n_rows >= 0 && ldx >= n_cols && ldo >= n_cols, NC_ESHAPE,
               "nc_spmm_csr: bad shape n_rows=%d n_cols=%d ldx=%lld ldo=%lld", n_rows, n_cols,
               (long long)ldx, (long long)ldo);
    if (n_rows == 0) return NC_OK;
    cudaStream_t s = (cudaStream_t)stream;
    if (x_dtype == NC_F32 && out_dtype == NC_F32)
        launch_spmm<float, float>(x, ldx, n_cols, rowptr, col, w, wself, bias, skip_self, n_rows, out, ldo, s);
    else if (x_dtype == NC_BF16 && out_dtype == NC_F32)
        launch_spmm<__nv_bfloat16, float>(x, ldx, n_cols, rowptr, col, w, wself, bias, skip_self, n_rows, out, ldo, s);
    else if (x_dtype == NC_BF16 && out_dtype == NC_BF16)
        launch_spmm<__nv_bfloat16, __nv_bfloat16>(x, ldx, n_cols, rowptr, col, w, wself, bias, skip_self, n_rows, out, ldo, s);
    else if (x_dtype == NC_F32 && out_dtype == NC_BF16)
        launch_spmm<float, __nv_bfloat16>(x, ldx, n_cols, rowptr, col, w, wself, bias, skip_self, n_rows, out, ldo, s);
    else
        NC_REQUIRE(false, NC_EINVAL, "nc_spmm_csr: unknown dtype %d/%d", x_dtype, out_dtype);
    NC_CHECK_LAUNCH("nc_spmm_csr");
    return NC_OK;
}

// ============================================================================ CSR-store SpMM
// One CTA per output row; the output row is staged in shared memory (n_cols floats) and the
// neighbours' sparse count rows are accumulated into it with shared-memory atomics.
constexpr int STORE_T = 256;

__global__ void __launch_bounds__(STORE_T)
k_spmm_csr_store(const int64_t* __restrict__ x_rowptr, const int32_t* __restrict__ x_col,
                 const float* __restrict__ x_val, int32_t n_cols, const int32_t* __restrict__ n_id,
                 const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
                 const float* __restrict__ w, const float* __restrict__ wself, int skip_self,
                 float* __restrict__ out, int64_t ldo, int chunk) {
    // blockIdx.y = column chunk [c0, c1) of the output row (rows wider than the shared-memory budget, e.g. the
    // 5 k genes + 50 k peaks of the RNA + ATAC layout): column ids are sorted inside a store row, so every
    // neighbour row is entered at lower_bound(c0) and left at the first id >= c1
    extern __shared__ float s_row[];
    const int i = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int c0 = blockIdx.y * chunk, c1 = min(n_cols, c0 + chunk);
    for (int c = tid; c < c1 - c0; c += STORE_T) s_row[c] = 0.f;
    __syncthreads();
    const int e0 = rowptr[i], e1 = rowptr[i + 1];
    const int total = (e1 - e0) + (wself ? 1 : 0);
    for (int t = warp; t < total; t += STORE_T / 32) {
        int j;
        float wt;
        if (t < e1 - e0) {
            j = col[e0 + t];
            wt = w ? w[e0 + t] : 1.f;
            if (skip_self && j == i) wt = 0.f;
        } else {
            j = i;
            wt = wself[i];
        }
        if (wt == 0.f) continue;
        const int64_t g = n_id ? (int64_t)n_id[j] : (int64_t)j;
        int64_t p0 = x_rowptr[g];
        const int64_t p1 = x_rowptr[g + 1];
        if (c0 > 0) {                                   // first entry with column id >= c0
            int64_t lo = p0, hi = p1;
            while (lo < hi) {
                const int64_t mid = (lo + hi) >> 1;
                if (__ldg(x_col + mid) < c0) lo = mid + 1; else hi = mid;
            }
            p0 = lo;
        }
        for (int64_t p = p0 + lane; p < p1; p += 32) {
            const int c = __ldg(x_col + p);
            if (c >= c1) break;
            atomicAdd(&s_row[c - c0], wt * __ldg(x_val + p));
        }
    }
    __syncthreads();
    float* orow = out + (int64_t)i * ldo + c0;
    for (int c = tid; c < c1 - c0; c += STORE_T) orow[c] = s_row[c];
}

extern "C" int nc_spmm_csr_store(const int64_t* x_rowptr, const int32_t* x_col, const float* x_val,
                                 int32_t n_cols, const int32_t* n_id, const int32_t* rowptr,
                                 const int32_t* col, const float* w, const float* wself,
                                 int skip_self, int32_t n_rows, float* out, int64_t ldo,
                                 void* stream) {
    NC_REQUIRE(n_cols > 0 && n_rows >= 0 && ldo >= n_cols, NC_ESHAPE,
               "nc_spmm_csr_store: bad shape n_rows=%d n_cols=%d ldo=%lld", n_rows, n_cols, (long long)ldo);
    constexpr int MAX_CHUNK = 200 * 1024 / (int)sizeof(float);
    const int n_chunks = (int)nc_cdiv(n_cols, MAX_CHUNK);
    const int chunk = (int)nc_cdiv(nc_cdiv(n_cols, n_chunks), 4) * 4;          // equal chunks, 16-byte aligned starts
    const size_t smem = (size_t)chunk * sizeof(float);
    if (n_rows == 0) return NC_OK;
    static bool attr_set[64] = {false};
    if (int rc = nc_set_max_smem((const void*)k_spmm_csr_store, 200 * 1024, attr_set, "nc_spmm_csr_store")) return rc;
    k_spmm_csr_store<<<dim3((unsigned)n_rows, (unsigned)n_chunks), STORE_T, smem, (cudaStream_t)stream>>>(
        x_rowptr, x_col, x_val, n_cols, n_id, rowptr, col, w, wself, skip_self, out, ldo, chunk);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_spmm_csr_store");
    return NC_OK;
}

// ============================================================================ row gathers
template <typename TO>
__global__ void __launch_bounds__(256)
k_csr_gather_rows(const int64_t* __restrict__ x_rowptr, const int32_t* __restrict__ x_col,
                  const float* __restrict__ x_val, int32_t n_cols, const int32_t* __restrict__ n_id,
                  int apply_log1p, TO* __restrict__ out, int64_t ldo) {
    extern __shared__ float s_row[];
    const int r = blockIdx.x, tid = threadIdx.x;
    const int width = (int)ldo;
    for (int c = tid; c < width; c += 256) s_row[c] = 0.f;
    __syncthreads();
    const int64_t g = n_id ? (int64_t)n_id[r] : (int64_t)r;
    const int64_t p0 = x_rowptr[g], p1 = x_rowptr[g + 1];
    for (int64_t p = p0 + tid; p < p1; p += 256) {
        float v = __ldg(x_val + p);
        s_row[__ldg(x_col + p)] = apply_log1p ? logf(1.0f + v) : v;
    }
    __syncthreads();
    TO* orow = out + (int64_t)r * ldo;
    if ((width & 3) == 0 && ((uintptr_t)orow % (sizeof(TO) * 4) == 0)) {
        for (int c = tid * 4; c < width; c += 256 * 4)
            Elem<TO>::st4(orow + c, make_float4(s_row[c], s_row[c + 1], s_row[c + 2], s_row[c + 3]));
    } else {
        for (int c = tid; c < width; c += 256) Elem<TO>::st(orow + c, s_row[c]);
    }
}

extern "C" int nc_csr_gather_rows(const int64_t* x_rowptr, const int32_t* x_col, const float* x_val,
                                  int32_t n_cols, const int32_t* n_id, int32_t n_rows,
                                  int apply_log1p, void* out, int out_dtype, int64_t ldo,
                                  void* stream) {
    NC_REQUIRE(n_cols > 0 && n_rows >= 0 && ldo >= n_cols, NC_ESHAPE,
               "nc_csr_gather_rows: bad shape n_rows=%d n_cols=%d ldo=%lld", n_rows, n_cols, (long long)ldo);
    const size_t smem = (size_t)ldo * sizeof(float);
    if (n_rows == 0) return NC_OK;
    if (smem > 200 * 1024)          // RNA + ATAC rows (55 k features): no staging, zero the row then scatter its non-zeros
        return nc_csr_gather_rows_bg(x_rowptr, x_col, x_val, n_cols, n_id, n_rows, apply_log1p, out, out_dtype, ldo,
                                     8 * nc_num_sms(), stream);
    static bool attr_f32[64] = {false}, attr_bf16[64] = {false};
    if (int rc = nc_set_max_smem((const void*)k_csr_gather_rows<float>, 200 * 1024, attr_f32, "nc_csr_gather_rows")) return rc;
    if (int rc = nc_set_max_smem((const void*)k_csr_gather_rows<__nv_bfloat16>, 200 * 1024, attr_bf16, "nc_csr_gather_rows")) return rc;
    cudaStream_t s = (cudaStream_t)stream;
    if (out_dtype == NC_F32)
        k_csr_gather_rows<float><<<n_rows, 256, smem, s>>>(x_rowptr, x_col, x_val, n_cols, n_id, apply_log1p, (float*)out, ldo);
    else if (out_dtype == NC_BF16)
        k_csr_gather_rows<__nv_bfloat16><<<n_rows, 256, smem, s>>>(x_rowptr, x_col, x_val, n_cols, n_id, apply_log1p, (__nv_bfloat16*)out, ldo);
    else
        NC_REQUIRE(false, NC_EINVAL, "nc_csr_gather_rows: unknown dtype %d", out_dtype);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_csr_gather_rows");
    return NC_OK;
}

// Background variant for a side stream: no shared memory and a grid of at most `max_ctas` CTAs that stride over
// the rows, so that the persistent GEMM CTAs of the compute stream (one per SM, ~all of its shared memory)
// can still be scheduled next to it.  With the store in pinned host memory the loads are bound by the host
// link; 74 CTAs x 256 threads x 4 loads keep ~600 KB in flight.  A row is zeroed, then its non-zeros are scattered.
template <typename TO>
__global__ void __launch_bounds__(256)
k_csr_gather_rows_bg(const int64_t* __restrict__ x_rowptr, const int32_t* __restrict__ x_col,
                     const float* __restrict__ x_val, int32_t n_rows, const int32_t* __restrict__ n_id,
                     int apply_log1p, TO* __restrict__ out, int64_t ldo) {
    const int tid = threadIdx.x;
    const int width = (int)ldo;
    for (int r = blockIdx.x; r < n_rows; r += gridDim.x) {
        const int64_t g = n_id ? (int64_t)n_id[r] : (int64_t)r;
        const int64_t p0 = x_rowptr[g], p1 = x_rowptr[g + 1];
        // issue the (slow) loads first, zero the row while they are in flight
        float v[4];
        int32_t c[4];
        int64_t p = p0 + tid;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int64_t q = p + (int64_t)u * 256;
            c[u] = q < p1 ? __ldg(x_col + q) : -1;
            v[u] = q < p1 ? __ldg(x_val + q) : 0.f;
        }
        TO* orow = out + (int64_t)r * ldo;
        if ((width & 3) == 0 && ((uintptr_t)orow % (sizeof(TO) * 4) == 0)) {
            for (int k = tid * 4; k < width; k += 256 * 4) Elem<TO>::st4(orow + k, make_float4(0.f, 0.f, 0.f, 0.f));
        } else {
            for (int k = tid; k < width; k += 256) Elem<TO>::st(orow + k, 0.f);
        }
        __syncthreads();                       // zeros of this CTA are ordered before its scattered values
#pragma unroll
        for (int u = 0; u < 4; ++u)
            if (c[u] >= 0) Elem<TO>::st(orow + c[u], apply_log1p ? logf(1.0f + v[u]) : v[u]);
        for (p += 4 * 256; p < p1; p += 256) {
            const float w = __ldg(x_val + p);
            Elem<TO>::st(orow + __ldg(x_col + p), apply_log1p ? logf(1.0f + w) : w);
        }
    }
}

extern "C" int nc_csr_gather_rows_bg(const int64_t* x_rowptr, const int32_t* x_col, const float* x_val,
                                     int32_t n_cols, const int32_t* n_id, int32_t n_rows, int apply_log1p,
                                     void* out, int out_dtype, int64_t ldo, int32_t max_ctas, void* stream) {
    NC_REQUIRE(n_cols > 0 && n_rows >= 0 && ldo >= n_cols, NC_ESHAPE,
               "nc_csr_gather_rows_bg: bad shape n_rows=%d n_cols=%d ldo=%lld", n_rows, n_cols, (long long)ldo);
    if (n_rows == 0) return NC_OK;
    if (max_ctas <= 0) max_ctas = nc_num_sms() / 2;      // measured: 74 or 148 CTAs leave the step undisturbed, 296 do not
    const int grid = n_rows < max_ctas ? n_rows : max_ctas;
    cudaStream_t s = (cudaStream_t)stream;
    if (out_dtype == NC_F32)
        k_csr_gather_rows_bg<float><<<grid, 256, 0, s>>>(x_rowptr, x_col, x_val, n_rows, n_id, apply_log1p, (float*)out, ldo);
    else if (out_dtype == NC_BF16)
        k_csr_gather_rows_bg<__nv_bfloat16><<<grid, 256, 0, s>>>(x_rowptr, x_col, x_val, n_rows, n_id, apply_log1p, (__nv_bfloat16*)out, ldo);
    else
        NC_REQUIRE(false, NC_EINVAL, "nc_csr_gather_rows_bg: unknown dtype %d", out_dtype);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_csr_gather_rows_bg");
    return NC_OK;
}

template <typename TO>
__global__ void __launch_bounds__(256)
k_dense_gather_rows(const float* __restrict__ x, int64_t ldx, int32_t n_cols,
                    const int32_t* __restrict__ n_id, int apply_log1p, TO* __restrict__ out, int64_t ldo) {
    const int r = blockIdx.x;
    const int64_t g = n_id ? (int64_t)n_id[r] : (int64_t)r;
    const float* xr = x + g * ldx;
    TO* orow = out + (int64_t)r * ldo;
    for (int c = threadIdx.x; c < (int)ldo; c += 256) {
        float v = 0.f;
        if (c < n_cols) {
            v = __ldg(xr + c);
            if (apply_log1p) v = logf(1.0f + v);
        }
        Elem<TO>::st(orow + c, v);
    }
}

extern "C" int nc_dense_gather_rows(const float* x, int64_t ldx, int32_t n_cols, const int32_t* n_id,
                                    int32_t n_rows, int apply_log1p, void* out, int out_dtype,
                                    int64_t ldo, void* stream) {
    NC_REQUIRE(n_cols > 0 && n_rows >= 0 && ldo >= n_cols && ldx >= n_cols, NC_ESHAPE,
               "nc_dense_gather_rows: bad shape n_rows=%d n_cols=%d", n_rows, n_cols);
    if (n_rows == 0) return NC_OK;
    cudaStream_t s = (cudaStream_t)stream;
    if (out_dtype == NC_F32)
        k_dense_gather_rows<float><<<n_rows, 256, 0, s>>>(x, ldx, n_cols, n_id, apply_log1p, (float*)out, ldo);
    else if (out_dtype == NC_BF16)
        k_dense_gather_rows<__nv_bfloat16><<<n_rows, 256, 0, s>>>(x, ldx, n_cols, n_id, apply_log1p, (__nv_bfloat16*)out, ldo);
    else
        NC_REQUIRE(false, NC_EINVAL, "nc_dense_gather_rows: unknown dtype %d", out_dtype);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_dense_gather_rows");
    return NC_OK;
}

// ------------------------------------------------------------------------- masked weight gradient (SDDMM)
// dW[f, p] += mval * sum_b dlT[f, b] * zT[p, b] for the stored (f, p) of a sparse gene-program mask
// (CSR by feature f).  Replaces the dense dlogits^T @ z followed by `* mask` of MaskedLinear's
// backward (nn/layercomponents.py:77-80): gene-program masks connect 5..50 of several thousand genes
// to a program, so only ~0.5 % of the [F, P] outer product is ever needed.  One warp per feature
// row: its column of dlogits stays in registers, every non-zero costs one coalesced row of zT.
namespace {
constexpr int SD_MAXB = 8;      // 32 * SD_MAXB rows of the batch held in registers per pass
__global__ void __launch_bounds__(128)
k_masked_sddmm(const float* __restrict__ dlT, int64_t lddl, const float* __restrict__ zT, int64_t ldz, int nb,
               const int32_t* __restrict__ mrowptr, const int32_t* __restrict__ mcol, const float* __restrict__ mval,
               int F, float* __restrict__ dW, int64_t lddw) {
    const int f = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (f >= F) return;
    const int p0 = mrowptr[f], p1 = mrowptr[f + 1];
    if (p0 == p1) return;
    const float* dr = dlT + (int64_t)f * lddl;
    for (int b0 = 0; b0 < nb; b0 += 32 * SD_MAXB) {
        float d[SD_MAXB];
#pragma unroll
        for (int q = 0; q < SD_MAXB; ++q) {
            const int b = b0 + q * 32 + lane;
            d[q] = b < nb ? __ldg(dr + b) : 0.f;
        }
        for (int e = p0; e < p1; ++e) {
            const float* zr = zT + (int64_t)mcol[e] * ldz;
            float acc = 0.f;
#pragma unroll
            for (int q = 0; q < SD_MAXB; ++q) {
                const int b = b0 + q * 32 + lane;
                if (b < nb) acc += d[q] * __ldg(zr + b);
            }
            acc = warp_sum(acc);
            if (lane == 0) dW[(int64_t)f * lddw + mcol[e]] += (mval ? mval[e] : 1.f) * acc;
        }
    }
}
}  // namespace

extern "C" int nc_masked_sddmm(const float* dlT, int64_t lddl, const float* zT, int64_t ldz, int32_t nb,
                               const int32_t* mrowptr, const int32_t* mcol, const float* mval, int32_t F,
                               float* dW, int64_t lddw, void* stream) {
    NC_REQUIRE(nb >= 0 && F >= 0 && lddl >= nb && ldz >= nb, NC_ESHAPE, "nc_masked_sddmm: bad shape nb=%d F=%d", nb, F);
    if (nb == 0 || F == 0) return NC_OK;
    k_masked_sddmm<<<(unsigned)nc_cdiv(F, 4), 128, 0, (cudaStream_t)stream>>>(dlT, lddl, zT, ldz, nb, mrowptr, mcol, mval,
                                                                            F, dW, lddw);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_masked_sddmm");
    return NC_OK;
}
