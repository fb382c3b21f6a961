// GPU CSR neighbourhood sampler: replaces PyG NeighborLoader / LinkNeighborLoader's CPU
// pyg-lib neighbor_sample + induced-subgraph extraction + torch.unique of the link seeds
// (data/dataloaders.py:64-100).  Bit-exact with oracle/sampler_ref.py:
//   * local node order: seeds in the given order, then newly discovered in-neighbours in
//     discovery order (per frontier node in order, over its CSC list in order);
//   * optional uniform sub-sampling of k neighbours by the smallest splitmix64 keys;
//   * induced subgraph = every edge of the full graph with both endpoints sampled, grouped by
//     local target id, CSC order inside a target.
// Everything runs on the stream without host synchronisation: sizes live in `counts` on the
// device and all grids are sized by the caller's capacities.
#include <limits.h>

#include <algorithm>

#include "common.cuh"

extern void nc_count_launch(int n);

namespace {
constexpr int ST = 256;
enum { C_NSUB = 0, C_ESUB = 1, C_OVERFLOW = 2, C_MAXROW = 3, C_FBEGIN = 4, C_EDST = 4 /* after the hops */, C_FEND = 5, C_NCAND = 6, C_FLEN = 7, C_NCOUNTS = 8 };

__device__ __forceinline__ uint64_t mix64(uint64_t x) {
    x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ULL;
    x ^= x >> 27; x *= 0x94D049BB133111EBULL;
    x ^= x >> 31;
    return x;
}

// exclusive scan of in[0..n) with n read from the device; out has n+1 entries. Single CTA.
__global__ void __launch_bounds__(1024)
k_scan_dev_n(const int32_t* __restrict__ in, int32_t* __restrict__ out, const int32_t* __restrict__ n_ptr,
             int32_t n_sub_off, int32_t cap) {
    __shared__ int32_t warp_tot[32];
    __shared__ int32_t carry_s;
    int n = *n_ptr - n_sub_off;
    n = max(0, min(n, cap));
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int i = base + tid;
        const int v = (i < n) ? in[i] : 0;
        int x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_tot[warp] = x;
        __syncthreads();
        if (warp == 0) {
            int t = warp_tot[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int y = __shfl_up_sync(0xffffffffu, t, o);
                if (lane >= o) t += y;
            }
            warp_tot[lane] = t;
        }
        __syncthreads();
        const int carry = carry_s;
        if (i < n) out[i] = carry + (warp > 0 ? warp_tot[warp - 1] : 0) + x - v;
        __syncthreads();
        if (tid == 1023) carry_s = carry + warp_tot[31];
        __syncthreads();
    }
    if (tid == 0) out[n] = carry_s;
}

__global__ void k_init(const int32_t* __restrict__ seeds, int n_seed, int32_t* __restrict__ slot,
                       int32_t* __restrict__ n_id, int32_t* __restrict__ counts) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n_seed) {
        const int v = seeds[i];
        slot[v] = -(i + 1);
        n_id[i] = v;
    }
    if (i == 0) {
        counts[C_NSUB] = n_seed; counts[C_ESUB] = 0; counts[C_OVERFLOW] = 0; counts[C_MAXROW] = 0;
        counts[C_FBEGIN] = 0; counts[C_FEND] = n_seed; counts[C_NCAND] = 0; counts[C_FLEN] = n_seed;
    }
}

__global__ void k_frontier_deg(const int32_t* __restrict__ g_rowptr, const int32_t* __restrict__ n_id,
                               const int32_t* __restrict__ counts, int k, int32_t* __restrict__ deg) {
    const int fb = counts[C_FBEGIN], fe = counts[C_FEND];
    for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < fe - fb; q += gridDim.x * blockDim.x) {
        const int v = n_id[fb + q];
        const int d = g_rowptr[v + 1] - g_rowptr[v];
        deg[q] = (k >= 0 && k < d) ? k : d;
    }
}

// one thread per frontier node: emit its (sub-sampled) neighbour list as candidates
__global__ void k_expand(const int32_t* __restrict__ g_rowptr, const int32_t* __restrict__ g_col,
                         const int32_t* __restrict__ n_id, int32_t* __restrict__ counts, int k,
                         uint64_t rng_base, const int32_t* __restrict__ cand_off, int32_t* __restrict__ cand,
                         int cap_cand, int32_t* __restrict__ slot) {
    const int fb = counts[C_FBEGIN], fe = counts[C_FEND];
    for (int q = blockIdx.x * blockDim.x + threadIdx.x; q < fe - fb; q += gridDim.x * blockDim.x) {
        const int v = n_id[fb + q];
        const int r0 = g_rowptr[v], d = g_rowptr[v + 1] - r0;
        int p = cand_off[q];
        if (k >= 0 && k < d) {
            const uint64_t nb = (uint64_t)v * 0xD1B54A32D192ED03ULL;
            for (int t = 0; t < d; ++t) {            // keep slot t iff fewer than k slots rank before it
                const uint64_t kt = mix64(rng_base ^ nb ^ (uint64_t)t);
                int rank = 0;
                for (int s = 0; s < d; ++s) {
                    const uint64_t ks = mix64(rng_base ^ nb ^ (uint64_t)s);
                    rank += (ks < kt || (ks == kt && s < t)) ? 1 : 0;
                }
                if (rank < k) {
                    if (p < cap_cand) {
                        const int u = g_col[r0 + t];
                        cand[p] = u;
                        atomicMin(&slot[u], p);
                    } else counts[C_OVERFLOW] = 1;
                    ++p;
                }
            }
        } else {
            for (int t = 0; t < d; ++t, ++p) {
                if (p < cap_cand) {
                    const int u = g_col[r0 + t];
                    cand[p] = u;
                    atomicMin(&slot[u], p);
                } else counts[C_OVERFLOW] = 1;
            }
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) counts[C_NCAND] = min(cand_off[fe - fb], cap_cand);
}

__global__ void k_flag(const int32_t* __restrict__ cand, const int32_t* __restrict__ counts,
                       const int32_t* __restrict__ slot, int32_t* __restrict__ flag) {
    const int n = counts[C_NCAND];
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x)
        flag[p] = (slot[cand[p]] == p) ? 1 : 0;
}

__global__ void k_assign(const int32_t* __restrict__ cand, const int32_t* __restrict__ flag,
                         const int32_t* __restrict__ newoff, const int32_t* __restrict__ counts,
                         int32_t* __restrict__ n_id, int cap_nodes, int32_t* __restrict__ slot,
                         int32_t* __restrict__ overflow) {
    const int n = counts[C_NCAND], base = counts[C_NSUB];
    for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < n; p += gridDim.x * blockDim.x) {
        if (!flag[p]) continue;
        const int local = base + newoff[p];
        const int v = cand[p];
        if (local < cap_nodes) {
            n_id[local] = v;
            slot[v] = -(local + 1);
        } else {
            slot[v] = INT_MAX;                       // dropped: restore and report
            *overflow = 1;
        }
    }
}

__global__ void k_advance(const int32_t* __restrict__ newoff, int32_t* __restrict__ counts, int cap_nodes) {
    const int n = counts[C_NCAND], base = counts[C_NSUB];
    const int n_new = min(base + newoff[n], cap_nodes);
    counts[C_FBEGIN] = base;
    counts[C_FEND] = n_new;
    counts[C_FLEN] = n_new - base;
    counts[C_NSUB] = n_new;
}

__global__ void k_row_count(const int32_t* __restrict__ g_rowptr, const int32_t* __restrict__ g_col,
                            const int32_t* __restrict__ n_id, int32_t* __restrict__ counts,
                            const int32_t* __restrict__ slot, int32_t* __restrict__ cnt) {
    const int n = counts[C_NSUB];
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
        const int v = n_id[t];
        int c = 0;
        for (int e = g_rowptr[v]; e < g_rowptr[v + 1]; ++e) c += (slot[g_col[e]] < 0) ? 1 : 0;
        cnt[t] = c;
        atomicMax(&counts[C_MAXROW], c);
    }
}

__global__ void k_row_fill(const int32_t* __restrict__ g_rowptr, const int32_t* __restrict__ g_col,
                           const int32_t* __restrict__ n_id, int32_t* __restrict__ counts,
                           const int32_t* __restrict__ slot, const int32_t* __restrict__ rowptr,
                           int32_t* __restrict__ col, int cap_edges, int n_seed) {
    const int n = counts[C_NSUB];
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) {
        const int v = n_id[t];
        int p = rowptr[t];
        for (int e = g_rowptr[v]; e < g_rowptr[v + 1]; ++e) {
            const int s = slot[g_col[e]];
            if (s < 0) {
                if (p < cap_edges) col[p] = -s - 1;
                else counts[C_OVERFLOW] = 1;
                ++p;
            }
        }
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        counts[C_ESUB] = rowptr[n];
        counts[C_EDST] = rowptr[min(n_seed, n)];         // entries of the seed rows: the only rows a 1-layer encoder reads
    }
}

__global__ void k_reset(const int32_t* __restrict__ n_id, const int32_t* __restrict__ counts, int32_t* __restrict__ slot) {
    const int n = counts[C_NSUB];
    for (int t = blockIdx.x * blockDim.x + threadIdx.x; t < n; t += gridDim.x * blockDim.x) slot[n_id[t]] = INT_MAX;
}

__global__ void k_fill_i32(int32_t* p, int64_t n, int32_t v) {
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (int64_t)gridDim.x * blockDim.x) p[i] = v;
}

// ------------------------------------------------------------------------- sorted unique + inverse
// Single CTA, n <= cap (power of two padded in shared memory).  torch.unique(return_inverse=True).
__global__ void __launch_bounds__(1024)
k_unique_sorted(const int32_t* __restrict__ ids, int n, int npad, int32_t* __restrict__ uniq,
                int32_t* __restrict__ inverse, int32_t* __restrict__ n_unique) {
    extern __shared__ int32_t s_keys[];           // npad keys + npad flags/offsets
    int32_t* s_off = s_keys + npad;
    __shared__ int32_t warp_tot[32];
    __shared__ int32_t carry_s;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    for (int i = tid; i < npad; i += 1024) s_keys[i] = (i < n) ? ids[i] : INT_MAX;
    __syncthreads();
    for (int k = 2; k <= npad; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < npad; i += 1024) {
                const int ixj = i ^ j;
                if (ixj > i) {
                    const int a = s_keys[i], b = s_keys[ixj];
                    const bool up = ((i & k) == 0);
                    if ((a > b) == up) { s_keys[i] = b; s_keys[ixj] = a; }
                }
            }
            __syncthreads();
        }
    }
    // flags of run starts, exclusive scan -> rank of each sorted position
    if (tid == 0) carry_s = 0;
    __syncthreads();
    for (int base = 0; base < n; base += 1024) {
        const int i = base + tid;
        const int v = (i < n && (i == 0 || s_keys[i] != s_keys[i - 1])) ? 1 : 0;
        int x = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            int y = __shfl_up_sync(0xffffffffu, x, o);
            if (lane >= o) x += y;
        }
        if (lane == 31) warp_tot[warp] = x;
        __syncthreads();
        if (warp == 0) {
            int t = warp_tot[lane];
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                int y = __shfl_up_sync(0xffffffffu, t, o);
                if (lane >= o) t += y;
            }
            warp_tot[lane] = t;
        }
        __syncthreads();
        const int carry = carry_s;
        const int incl = carry + (warp > 0 ? warp_tot[warp - 1] : 0) + x;
        if (i < n) {
            s_off[i] = incl - 1;                   // index of this key's run in the unique array
            if (v) uniq[incl - 1] = s_keys[i];
        }
        __syncthreads();
        if (tid == 1023) carry_s = carry + warp_tot[31];
        __syncthreads();
    }
    const int nu = carry_s;
    if (tid == 0) *n_unique = nu;
    // inverse: binary search of every original id in the sorted keys
    for (int i = tid; i < n; i += 1024) {
        const int key = ids[i];
        int lo = 0, hi = n - 1;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (s_keys[mid] < key) lo = mid + 1; else hi = mid;
        }
        inverse[i] = s_off[lo];
    }
}
}  // namespace

extern "C" int nc_fill_i32(int32_t* p, int64_t n, int32_t value, void* stream) {
    if (n <= 0) return NC_OK;
    k_fill_i32<<<(unsigned)std::min<int64_t>(nc_cdiv(n, 256), 4096), 256, 0, (cudaStream_t)stream>>>(p, n, value);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_fill_i32");
    return NC_OK;
}

extern "C" int nc_unique_sorted_i32(const int32_t* ids, int32_t n, int32_t* uniq, int32_t* inverse,
                                    int32_t* n_unique, void* stream) {
    NC_REQUIRE(n > 0 && n <= 16384, NC_EUNSUPPORTED, "nc_unique_sorted_i32: n=%d outside (0, 16384]", n);
    int npad = 1;
    while (npad < n) npad <<= 1;
    const size_t smem = (size_t)npad * 2 * sizeof(int32_t);
    static bool attr_set[64] = {false};
    if (int rc = nc_set_max_smem((const void*)k_unique_sorted, 16384 * 2 * 4, attr_set, "nc_unique_sorted_i32")) return rc;
    k_unique_sorted<<<1, 1024, smem, (cudaStream_t)stream>>>(ids, n, npad, uniq, inverse, n_unique);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_unique_sorted_i32");
    return NC_OK;
}

extern "C" int nc_sample_subgraph(const int32_t* g_rowptr, const int32_t* g_col, const int32_t* seeds,
                                  int32_t n_seed, int32_t n_hops, int32_t k_neighbors, int64_t rng_seed,
                                  int64_t batch_id, int32_t* slot_ws, int32_t* n_id, int32_t cap_nodes,
                                  int32_t* rowptr, int32_t* col, int32_t cap_edges, int32_t* cand_ws,
                                  int32_t cap_cand, int32_t* counts, void* stream) {
    NC_REQUIRE(n_seed > 0 && n_hops >= 1 && cap_nodes >= n_seed && cap_edges > 0 && cap_cand > 0, NC_ESHAPE,
               "nc_sample_subgraph: bad sizes (n_seed=%d hops=%d cap_nodes=%d cap_edges=%d cap_cand=%d)", n_seed,
               n_hops, cap_nodes, cap_edges, cap_cand);
    cudaStream_t s = (cudaStream_t)stream;
    // cand_ws layout: deg/cnt [cap_nodes] | off [cap_nodes+1] | cand [cap_cand] | flag [cap_cand] | newoff [cap_cand+1]
    int32_t* deg = cand_ws;
    int32_t* off = deg + cap_nodes;
    int32_t* cand = off + cap_nodes + 1;
    int32_t* flag = cand + cap_cand;
    int32_t* newoff = flag + cap_cand;
    const uint64_t rng_base = [&] {
        uint64_t x = (uint64_t)rng_seed ^ ((uint64_t)batch_id * 0x9E3779B97F4A7C15ULL);
        x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ULL; x ^= x >> 27; x *= 0x94D049BB133111EBULL; x ^= x >> 31;
        return x;
    }();
    const int gn = (int)std::min<int64_t>(nc_cdiv(cap_nodes, ST), 2 * nc_num_sms());
    const int gc = (int)std::min<int64_t>(nc_cdiv(cap_cand, ST), 2 * nc_num_sms());
    k_init<<<(unsigned)nc_cdiv(n_seed, ST), ST, 0, s>>>(seeds, n_seed, slot_ws, n_id, counts);
    int launches = 1;
    for (int h = 0; h < n_hops; ++h) {
        k_frontier_deg<<<gn, ST, 0, s>>>(g_rowptr, n_id, counts, k_neighbors, deg);
        k_scan_dev_n<<<1, 1024, 0, s>>>(deg, off, counts + C_FLEN, 0, cap_nodes);
        k_expand<<<gn, ST, 0, s>>>(g_rowptr, g_col, n_id, counts, k_neighbors, rng_base, off, cand, cap_cand, slot_ws);
        k_flag<<<gc, ST, 0, s>>>(cand, counts, slot_ws, flag);
        k_scan_dev_n<<<1, 1024, 0, s>>>(flag, newoff, counts + C_NCAND, 0, cap_cand);
        k_assign<<<gc, ST, 0, s>>>(cand, flag, newoff, counts, n_id, cap_nodes, slot_ws, counts + C_OVERFLOW);
        k_advance<<<1, 1, 0, s>>>(newoff, counts, cap_nodes);
        launches += 7;
    }
    k_row_count<<<gn, ST, 0, s>>>(g_rowptr, g_col, n_id, counts, slot_ws, deg);
    k_scan_dev_n<<<1, 1024, 0, s>>>(deg, rowptr, counts + C_NSUB, 0, cap_nodes);
    k_row_fill<<<gn, ST, 0, s>>>(g_rowptr, g_col, n_id, counts, slot_ws, rowptr, col, cap_edges, n_seed);
    k_reset<<<gn, ST, 0, s>>>(n_id, counts, slot_ws);
    nc_count_launch(launches + 4);
    NC_CHECK_LAUNCH("nc_sample_subgraph");
    return NC_OK;
}

// ------------------------------------------------------------------------- merged sub-graphs
// Two sampled sub-graphs (a: node batch, b: edge batch) as ONE block-diagonal problem for the encoder:
// merged node order = [a seeds | b seeds | a others | b others], so the targets are again the leading rows.
namespace {
__device__ __forceinline__ int merge_map(int j, int n_seed_self, int add_seed, int add_other) {
    return j < n_seed_self ? j + add_seed : j + add_other;
}
// b_seed_pad >= b_seed, n_sub_pad, e_pad: padded sizes of the merged problem (CUDA-graph size buckets of the
// training step): merged node order = [a seeds | b seeds | pad seeds | a others | b others | pad others]; pad seeds
// have empty rows, pad nodes carry the id of a real node (so every gather stays in bounds) and are referenced by no
// edge -- their rows cost work but change no result.
__global__ void k_merge_subgraphs(const int32_t* __restrict__ a_rowptr, const int32_t* __restrict__ a_col,
                                  const int32_t* __restrict__ a_nid, int a_seed, int a_sub,
                                  const int32_t* __restrict__ b_rowptr, const int32_t* __restrict__ b_col,
                                  const int32_t* __restrict__ b_nid, int b_seed, int b_sub,
                                  int32_t* __restrict__ m_rowptr, int32_t* __restrict__ m_col,
                                  int32_t* __restrict__ m_nid, int b_seed_pad, int n_sub_pad, int e_pad) {
    const int n_dst = a_seed + b_seed, n_dst_pad = a_seed + b_seed_pad;
    const int pad = b_seed_pad - b_seed;
    const int n_sub = a_sub + b_sub + pad;                    // real nodes + pad seeds
    const int Ea = a_rowptr[a_seed];
    const int Etot = Ea + b_rowptr[b_seed];
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x, gsz = gridDim.x * blockDim.x;
    const int filler = a_sub > 0 ? a_nid[0] : (b_sub > 0 ? b_nid[0] : 0);
    // node ids in merged order
    for (int i = gtid; i < n_sub_pad; i += gsz) {
        int v;
        if (i < a_seed) v = a_nid[i];
        else if (i < n_dst) v = b_nid[i - a_seed];
        else if (i < n_dst_pad) v = filler;
        else if (i < n_dst_pad + (a_sub - a_seed)) v = a_nid[i - b_seed_pad];
        else if (i < n_sub) v = b_nid[i - a_sub - pad];
        else v = filler;
        m_nid[i] = v;
    }
    // rows of the targets, column ids remapped: a: seeds stay, others + b_seed_pad; b: seeds + a_seed,
    // others + a_sub + pad
    for (int r = gtid; r <= n_dst_pad; r += gsz)
        m_rowptr[r] = r <= a_seed ? a_rowptr[r] : (r <= n_dst ? Ea + b_rowptr[r - a_seed] : Etot);
    for (int e = Etot + gtid; e < e_pad; e += gsz) m_col[e] = 0;
    const int warp = gtid >> 5, lane = gtid & 31, nwarp = gsz >> 5;
    for (int r = warp; r < n_dst; r += nwarp) {
        if (r < a_seed) {
            const int e0 = a_rowptr[r], e1 = a_rowptr[r + 1];
            for (int e = e0 + lane; e < e1; e += 32) m_col[e] = merge_map(a_col[e], a_seed, 0, b_seed_pad);
        } else {
            const int q = r - a_seed;
            const int e0 = b_rowptr[q], e1 = b_rowptr[q + 1];
            for (int e = e0 + lane; e < e1; e += 32) m_col[Ea + e] = merge_map(b_col[e], b_seed, a_seed, a_sub + pad);
        }
    }
}
}  // namespace

extern "C" int nc_merge_subgraphs(const int32_t* a_rowptr, const int32_t* a_col, const int32_t* a_nid,
                                  int32_t a_seed, int32_t a_sub, const int32_t* b_rowptr, const int32_t* b_col,
                                  const int32_t* b_nid, int32_t b_seed, int32_t b_sub, int32_t* m_rowptr,
                                  int32_t* m_col, int32_t* m_nid, void* stream) {
    NC_REQUIRE(a_seed >= 0 && b_seed >= 0 && a_sub >= a_seed && b_sub >= b_seed, NC_ESHAPE,
               "nc_merge_subgraphs: bad sizes");
    if (a_sub + b_sub == 0) return NC_OK;
    const int n = a_sub + b_sub;
    const int grid = (int)std::min<int64_t>(nc_cdiv(n, 256), 2 * nc_num_sms());
    k_merge_subgraphs<<<grid, 256, 0, (cudaStream_t)stream>>>(a_rowptr, a_col, a_nid, a_seed, a_sub, b_rowptr, b_col,
                                                             b_nid, b_seed, b_sub, m_rowptr, m_col, m_nid, b_seed,
                                                             a_sub + b_sub, 0);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_merge_subgraphs");
    return NC_OK;
}

extern "C" int nc_merge_subgraphs_padded(const int32_t* a_rowptr, const int32_t* a_col, const int32_t* a_nid,
                                         int32_t a_seed, int32_t a_sub, const int32_t* b_rowptr, const int32_t* b_col,
                                         const int32_t* b_nid, int32_t b_seed, int32_t b_sub, int32_t b_seed_pad,
                                         int32_t n_sub_pad, int32_t e_pad, int32_t* m_rowptr, int32_t* m_col,
                                         int32_t* m_nid, void* stream) {
    NC_REQUIRE(a_seed >= 0 && b_seed >= 0 && a_sub >= a_seed && b_sub >= b_seed && b_seed_pad >= b_seed &&
               n_sub_pad >= a_sub + b_sub + (b_seed_pad - b_seed) && e_pad >= 0, NC_ESHAPE,
               "nc_merge_subgraphs_padded: bad sizes");
    if (n_sub_pad == 0) return NC_OK;
    const int grid = (int)std::min<int64_t>(nc_cdiv(n_sub_pad, 256), 2 * nc_num_sms());
    k_merge_subgraphs<<<grid, 256, 0, (cudaStream_t)stream>>>(a_rowptr, a_col, a_nid, a_seed, a_sub, b_rowptr, b_col,
                                                             b_nid, b_seed, b_sub, m_rowptr, m_col, m_nid, b_seed_pad,
                                                             n_sub_pad, e_pad);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_merge_subgraphs_padded");
    return NC_OK;
}
