// Data-parallel gradient exchange fused with the optimiser step, over NVLink peer memory.
//
// The reference has no data parallelism (docs/installation.md:3-5); its optimiser step is
// torch.optim.Adam.step() after loss.backward() (train/trainer.py:414-421).  Here, with W ranks on one NVSwitch
// box, every rank keeps the parameter arena, the gradient arena and the bf16 shadow (train/optim.py FlatAdam) in
// memory that its peers have mapped (cudaIpc), and ONE kernel per step does what ncclAllReduce + k_adam did:
//
//   start barrier   every rank tells every peer "my gradients of step e are final" (a flag in the peer's memory)
//   reduce-scatter  rank r reads shard r of the gradient arena of all W ranks through NVLink (128-bit loads,
//                   coalesced, W independent streams in flight per thread) and averages them
//   Adam            on its shard of p / m / v (the moments exist only where they are used)
//   all-gather      the new fp32 parameters and their bf16 shadow are STORED into all W ranks' arenas
//   end barrier     a rank's kernel ends only when every peer has finished writing into its arenas
//   clear           every CTA clears its stripe of the LOCAL gradient arena (all peers have read it by now)
//
// No shared memory: the kernel could run beside persistent GEMM CTAs that own every SM's shared memory (the reason
// NCCL's copy kernels could not be overlapped); the grid is clamped to a co-resident size
// because every CTA waits for the last one.  Per rank and step, with N parameters and W ranks: (W-1)/W * 4 N bytes
// read and (W-1)/W * 6 N bytes written over the links (the same amounts arrive), plus one local pass over the
// shard's p / m / v -- against ~2 (W-1)/W * 4 N per direction for an all-reduce followed by a separate 34 N-byte
// Adam pass.  Measured: DESIGN.md section 6 (link-bound at W = 8).
#include <math.h>

#include "common.cuh"

extern void nc_count_launch(int n);

namespace {
constexpr int DP_MAX_WORLD = 16;

struct DpPeers {
    float* p[DP_MAX_WORLD];
    float* g[DP_MAX_WORLD];
    __nv_bfloat16* sh[DP_MAX_WORLD];     // may be all NULL (fp32 mode)
    uint32_t* flags[DP_MAX_WORLD];       // [2 * world] per rank: [0, W) start flags, [W, 2W) end flags
    float* aux[DP_MAX_WORLD];            // small per-rank statistic averaged with the step (may be all NULL)
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 ld_peer4(const float* p) {
    float4 v;
    asm volatile("ld.relaxed.sys.global.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float adam1(float pi, float gi, float& mi, float& vi, float lr_bc1, float b1, float b2,
                                       float eps, float wd, float bc2_sqrt, float clip) {
    if (clip > 0.f) gi = fminf(fmaxf(gi, -clip), clip);
    if (wd != 0.f) gi += wd * pi;
    mi = b1 * mi + (1.f - b1) * gi;
    vi = b2 * vi + (1.f - b2) * gi * gi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    return pi - lr_bc1 * (mi / denom);
}

// Waits until flags[lo + k] >= epoch for every k < W (epochs only grow; wrap-around after 4e9 steps).
__device__ __forceinline__ void wait_all(const uint32_t* flags, int lo, int W, uint32_t epoch) {
    for (int k = 0; k < W; ++k)
        while ((int32_t)(ld_acquire_sys(flags + lo + k) - epoch) < 0) __nanosleep(64);
}

template <int W>
__global__ void __launch_bounds__(512)
k_dp_adam(const DpPeers P, float* __restrict__ m, float* __restrict__ v, int64_t lo, int64_t hi, int64_t numel, int rank,
          uint32_t epoch,
          float lr, float b1, float b2, float eps, float wd, float bc1, float bc2_sqrt, float clip,
          unsigned int* __restrict__ done_counter, float* __restrict__ dyn, float* __restrict__ aux_avg, int aux_n) {
    if (dyn) {                      // {lr, step} on the device (CUDA-graph replay): step is also the barrier epoch
        lr = dyn[0];
        const double step = (double)dyn[1];
        epoch = (uint32_t)dyn[1];
        bc1 = (float)(1.0 - pow((double)b1, step));
        bc2_sqrt = (float)sqrt(1.0 - pow((double)b2, step));
    }
    // ---- start barrier
    if (threadIdx.x == 0) {
        if (blockIdx.x == 0) {
            __threadfence_system();
#pragma unroll
            for (int k = 0; k < W; ++k) st_release_sys(P.flags[k] + rank, epoch);
        }
        wait_all(P.flags[rank], 0, W, epoch);
    }
    __syncthreads();
    const float lr_bc1 = lr / bc1, invW = 1.0f / (float)W;
    // the |mu| running mean that drives gene-program pruning (modules/vgpgae.py:641-651 is per-process state in the
    // reference): averaged over the ranks by the last CTA of the grid into a local scratch, published below
    if (aux_n > 0 && blockIdx.x == gridDim.x - 1) {
        for (int i = threadIdx.x; i < aux_n; i += blockDim.x) {
            float a = 0.f;
#pragma unroll
            for (int k = 0; k < W; ++k) {
                float t;
                asm volatile("ld.relaxed.sys.global.f32 %0, [%1];" : "=f"(t) : "l"(P.aux[k] + i) : "memory");
                a += t;
            }
            aux_avg[i] = a * invW;
        }
    }
    const bool shadow = P.sh[0] != nullptr;
    const int64_t n4 = (hi - lo) >> 2;                        // lo and the shard length are multiples of 4 (host)
    const int64_t t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nt = (int64_t)gridDim.x * blockDim.x;
    // two 128-bit chunks per thread and iteration: 2 W peer loads in flight per thread (an NVLink round trip is ~2 us;
    // one chunk per thread left the links half empty)
    for (int64_t q0 = t0; q0 < n4; q0 += 2 * nt) {
        const int64_t q1 = q0 + nt;
        const bool two = q1 < n4;
        const int64_t i0 = lo + (q0 << 2), i1 = lo + ((two ? q1 : q0) << 2);
        float4 ga[W], gb[W];
#pragma unroll
        for (int k = 0; k < W; ++k) ga[k] = ld_peer4(P.g[(rank + k) % W] + i0);
#pragma unroll
        for (int k = 0; k < W; ++k) gb[k] = ld_peer4(P.g[(rank + k) % W] + i1);
#pragma unroll
        for (int half = 0; half < 2; ++half) {
            if (half == 1 && !two) break;
            const int64_t i = half ? i1 : i0;
            float4 pp = *reinterpret_cast<const float4*>(P.p[rank] + i);
            float4 mm = *reinterpret_cast<const float4*>(m + (i - lo));
            float4 vv = *reinterpret_cast<const float4*>(v + (i - lo));
            float4 g = half ? gb[0] : ga[0];
#pragma unroll
            for (int k = 1; k < W; ++k) {
                const float4 t = half ? gb[k] : ga[k];
                g.x += t.x; g.y += t.y; g.z += t.z; g.w += t.w;
            }
            pp.x = adam1(pp.x, g.x * invW, mm.x, vv.x, lr_bc1, b1, b2, eps, wd, bc2_sqrt, clip);
            pp.y = adam1(pp.y, g.y * invW, mm.y, vv.y, lr_bc1, b1, b2, eps, wd, bc2_sqrt, clip);
            pp.z = adam1(pp.z, g.z * invW, mm.z, vv.z, lr_bc1, b1, b2, eps, wd, bc2_sqrt, clip);
            pp.w = adam1(pp.w, g.w * invW, mm.w, vv.w, lr_bc1, b1, b2, eps, wd, bc2_sqrt, clip);
            *reinterpret_cast<float4*>(m + (i - lo)) = mm;
            *reinterpret_cast<float4*>(v + (i - lo)) = vv;
            __nv_bfloat162 b_lo = __floats2bfloat162_rn(pp.x, pp.y), b_hi = __floats2bfloat162_rn(pp.z, pp.w);
            uint2 sq;
            sq.x = *reinterpret_cast<uint32_t*>(&b_lo);
            sq.y = *reinterpret_cast<uint32_t*>(&b_hi);
#pragma unroll
            for (int k = 0; k < W; ++k) {
                const int peer = (rank + k) % W;
                *reinterpret_cast<float4*>(P.p[peer] + i) = pp;
                if (shadow) *reinterpret_cast<uint2*>(P.sh[peer] + i) = sq;
            }
        }
    }
    // ---- end barrier: the last CTA of this rank to finish publishes "my writes are done" and waits for the peers;
    // every CTA then clears its stripe of THIS rank's gradient arena (all peers have read it by now): a local HBM
    // write instead of 4 bytes per parameter and peer over the links
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned int ticket = atomicAdd(done_counter, 1u);
        if (ticket == gridDim.x - 1) {
            *done_counter = 0u;
            __threadfence_system();
#pragma unroll
            for (int k = 0; k < W; ++k) st_release_sys(P.flags[k] + W + rank, epoch);
            wait_all(P.flags[rank], W, W, epoch);
            if (dyn) dyn[1] += 1.0f;                     // every CTA has read the step: this is the last one running
            __threadfence();
            atomicExch(done_counter + 1, epoch);         // go: peers are done with this rank's gradients and statistic
        }
        while ((int32_t)(*reinterpret_cast<volatile unsigned int*>(done_counter + 1) - epoch) < 0) __nanosleep(64);
    }
    __syncthreads();
    {
        float4* gz = reinterpret_cast<float4*>(P.g[rank]);
        const float4 zero = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int64_t q = t0; q < (numel >> 2); q += nt) gz[q] = zero;
    }
    if (aux_n > 0 && blockIdx.x == gridDim.x - 1)        // (after the go signal above) publish the mean of the statistic
        for (int i = threadIdx.x; i < aux_n; i += blockDim.x) P.aux[rank][i] = aux_avg[i];
}
}  // namespace

// ---------------------------------------------------------------------------- peer memory (cudaIpc)
extern "C" int nc_peer_alloc(int64_t bytes, void** out_dev) {
    NC_REQUIRE(bytes > 0 && out_dev != nullptr, NC_EINVAL, "nc_peer_alloc: bad arguments");
    cudaError_t e = cudaMalloc(out_dev, (size_t)bytes);
    if (e == cudaSuccess) e = cudaMemset(*out_dev, 0, (size_t)bytes);
    if (e != cudaSuccess) {
        cudaGetLastError();
        nc_set_error("nc_peer_alloc: %lld B: %s", (long long)bytes, cudaGetErrorString(e));
        return NC_ECUDA;
    }
    return NC_OK;
}

extern "C" int nc_peer_free(void* dev) {
    if (dev == nullptr) return NC_OK;
    cudaError_t e = cudaFree(dev);
    if (e != cudaSuccess) { cudaGetLastError(); nc_set_error("nc_peer_free: %s", cudaGetErrorString(e)); return NC_ECUDA; }
    return NC_OK;
}

// handle_host: 64 bytes (cudaIpcMemHandle_t) for a pointer returned by nc_peer_alloc
extern "C" int nc_peer_get_handle(void* dev, void* handle_host) {
    NC_REQUIRE(dev != nullptr && handle_host != nullptr, NC_EINVAL, "nc_peer_get_handle: NULL argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "cudaIpcMemHandle_t is 64 bytes");
    cudaError_t e = cudaIpcGetMemHandle((cudaIpcMemHandle_t*)handle_host, dev);
    if (e != cudaSuccess) { cudaGetLastError(); nc_set_error("nc_peer_get_handle: %s", cudaGetErrorString(e)); return NC_ECUDA; }
    return NC_OK;
}

extern "C" int nc_peer_open(const void* handle_host, void** out_dev) {
    NC_REQUIRE(handle_host != nullptr && out_dev != nullptr, NC_EINVAL, "nc_peer_open: NULL argument");
    cudaIpcMemHandle_t h;
    memcpy(&h, handle_host, sizeof(h));
    cudaError_t e = cudaIpcOpenMemHandle(out_dev, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) { cudaGetLastError(); nc_set_error("nc_peer_open: %s", cudaGetErrorString(e)); return NC_ECUDA; }
    return NC_OK;
}

extern "C" int nc_peer_close(void* dev) {
    if (dev == nullptr) return NC_OK;
    cudaError_t e = cudaIpcCloseMemHandle(dev);
    if (e != cudaSuccess) { cudaGetLastError(); nc_set_error("nc_peer_close: %s", cudaGetErrorString(e)); return NC_ECUDA; }
    return NC_OK;
}

// p_ptrs_host / g_ptrs_host / shadow_ptrs_host / flag_ptrs_host: `world` device pointers each (entry `rank` = this
// rank's own buffer, the others = peer mappings); shadow_ptrs_host may be NULL.  m, v: this rank's moments for its
// shard only, [shard] floats with shard = ceil(numel / world / 4) * 4.  done_counter: one zero-initialised uint32
// in this rank's memory.  `step` >= 1 is also the barrier epoch: every rank must call with the same sequence.
extern "C" int nc_dp_adam_step(void* const* p_ptrs_host, void* const* g_ptrs_host, void* const* shadow_ptrs_host,
                               void* const* flag_ptrs_host, float* m, float* v, int64_t numel, int32_t rank,
                               int32_t world, float lr, float beta1, float beta2, float eps, float weight_decay,
                               int32_t step, float clip_value, uint32_t* done_counter, int32_t n_ctas, float* dyn,
                               void* const* aux_ptrs_host, float* aux_avg, int32_t aux_n, void* stream) {
    NC_REQUIRE(world >= 1 && world <= DP_MAX_WORLD && rank >= 0 && rank < world && numel > 0 && (step >= 1 || dyn), NC_EINVAL,
               "nc_dp_adam_step: bad arguments (world=%d rank=%d numel=%lld step=%d)", world, rank, (long long)numel, step);
    NC_REQUIRE(numel % 4 == 0, NC_ESHAPE, "nc_dp_adam_step: the arena length must be a multiple of 4");
    DpPeers P;
    for (int k = 0; k < DP_MAX_WORLD; ++k) { P.p[k] = nullptr; P.g[k] = nullptr; P.sh[k] = nullptr; P.flags[k] = nullptr; P.aux[k] = nullptr; }
    if (aux_ptrs_host == nullptr || aux_avg == nullptr) aux_n = 0;
    for (int k = 0; k < world && aux_n > 0; ++k) P.aux[k] = (float*)aux_ptrs_host[k];
    for (int k = 0; k < world; ++k) {
        P.p[k] = (float*)p_ptrs_host[k];
        P.g[k] = (float*)g_ptrs_host[k];
        P.sh[k] = shadow_ptrs_host ? (__nv_bfloat16*)shadow_ptrs_host[k] : nullptr;
        P.flags[k] = (uint32_t*)flag_ptrs_host[k];
        NC_REQUIRE(P.p[k] && P.g[k] && P.flags[k], NC_EINVAL, "nc_dp_adam_step: NULL peer pointer %d", k);
    }
    const int64_t shard = nc_cdiv(nc_cdiv(numel, world), 4) * 4;
    const int64_t lo = std::min<int64_t>((int64_t)rank * shard, numel), hi = std::min<int64_t>(lo + shard, numel);
    const float bc1 = dyn ? 1.f : (float)(1.0 - pow((double)beta1, (double)step));
    const float bc2_sqrt = dyn ? 1.f : (float)sqrt(1.0 - pow((double)beta2, (double)step));
    if (n_ctas <= 0) n_ctas = nc_num_sms();
    cudaStream_t s = (cudaStream_t)stream;
    // every CTA waits for the last one after the end barrier: the whole grid must be resident at once
#define DP_LAUNCH(WW)                                                                                              \
    {                                                                                                              \
        int occ_ = 1;                                                                                              \
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_, k_dp_adam<WW>, 512, 0) != cudaSuccess || occ_ < 1) { \
            cudaGetLastError();                                                                                    \
            occ_ = 1;                                                                                              \
        }                                                                                                          \
        if (n_ctas > occ_ * nc_num_sms()) n_ctas = occ_ * nc_num_sms();                                            \
    }                                                                                                              \
    k_dp_adam<WW><<<n_ctas, 512, 0, s>>>(P, m, v, lo, hi, numel, rank, (uint32_t)step, lr, beta1, beta2, eps, weight_decay, bc1, \
                                         bc2_sqrt, clip_value, done_counter, dyn, aux_avg, aux_n)
    switch (world) {
        case 1: DP_LAUNCH(1); break;
        case 2: DP_LAUNCH(2); break;
        case 4: DP_LAUNCH(4); break;
        case 8: DP_LAUNCH(8); break;
        default: NC_REQUIRE(false, NC_EUNSUPPORTED, "nc_dp_adam_step: world size %d (1, 2, 4 or 8)", world);
    }
#undef DP_LAUNCH
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_dp_adam_step");
    return NC_OK;
}

// ---------------------------------------------------------------------------- halo rows straight from the peers' HBM
// Full-atlas latent inference over spatial tiles (models/tiled_inference.py; the reference walks a NeighborLoader over
// the whole graph on one device, models/nichecompass.py:1264-1381): every rank needs the `hidden` rows of the nodes
// just outside its tile.  Instead of index exchange + pack + all_to_all + unpack, each rank reads those rows where
// they are: out[r, :] = peer[owner[r]][pos[r], :] over NVLink, one warp per row, 128-bit loads.
namespace {
struct HaloPeers { const uint8_t* base[DP_MAX_WORLD]; };

__global__ void __launch_bounds__(256)
k_halo_gather(const HaloPeers P, const int32_t* __restrict__ owner, const int64_t* __restrict__ pos, int64_t n_rows,
              int64_t src_pitch, int64_t row_bytes, uint8_t* __restrict__ out, int64_t out_pitch) {
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5, nwarp = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t r = warp; r < n_rows; r += nwarp) {
        const uint8_t* src = P.base[owner[r]] + pos[r] * src_pitch;
        uint8_t* dst = out + r * out_pitch;
        for (int64_t b = (int64_t)lane * 16; b < row_bytes; b += 32 * 16) {
            int4 v;
            asm volatile("ld.relaxed.sys.global.v4.s32 {%0, %1, %2, %3}, [%4];"
                         : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(src + b) : "memory");
            *reinterpret_cast<int4*>(dst + b) = v;
        }
    }
}
}  // namespace

// peer_ptrs_host: `world` device pointers (own buffer + nc_peer_open mappings) to row-major buffers of pitch
// src_pitch bytes; owner[r] / pos[r]: rank and row index (in the owner's buffer) of halo row r; row_bytes and both
// pitches multiples of 16.  The caller guarantees that every owner has finished writing its rows (a barrier).
extern "C" int nc_halo_gather_rows(void* const* peer_ptrs_host, int32_t world, const int32_t* owner, const int64_t* pos,
                                   int64_t n_rows, int64_t src_pitch, int64_t row_bytes, void* out, int64_t out_pitch,
                                   void* stream) {
    NC_REQUIRE(world >= 1 && world <= DP_MAX_WORLD && n_rows >= 0 && row_bytes > 0 && row_bytes % 16 == 0 &&
               src_pitch % 16 == 0 && out_pitch % 16 == 0 && src_pitch >= row_bytes && out_pitch >= row_bytes, NC_EINVAL,
               "nc_halo_gather_rows: bad arguments (world=%d rows=%lld row_bytes=%lld)", world, (long long)n_rows,
               (long long)row_bytes);
    if (n_rows == 0) return NC_OK;
    HaloPeers P;
    for (int k = 0; k < DP_MAX_WORLD; ++k) P.base[k] = k < world ? (const uint8_t*)peer_ptrs_host[k] : nullptr;
    const int grid = (int)std::min<int64_t>(nc_cdiv(n_rows, 8), 8 * (int64_t)nc_num_sms());
    k_halo_gather<<<grid, 256, 0, (cudaStream_t)stream>>>(P, owner, pos, n_rows, src_pitch, row_bytes, (uint8_t*)out,
                                                         out_pitch);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_halo_gather_rows");
    return NC_OK;
}
