// GATv2 edge kernels, second generation ("ring" kernels): bf16 rows, hot-path shapes.
// Same arithmetic as gatv2.cu (PyG GATv2Conv message passing, nn/encoders.py:124-140,197-207, SURVEY.md
// App. A.1), different machine mapping -- built for HBM throughput:
//
//   * persistent CTAs, one per SM: 8 consumer warps that never synchronise with each other, one producer warp and
//     three warps that only hand their registers over (setmaxnreg).  Consumer warp w owns one or two (conv, head)
//     segments of the x_l / x_r rows (a long one and a short one: the plan below pairs them so that every warp
//     touches the same number of bytes) for every target of its CTA;
//   * the producer warp copies whole rows ([x_r[i] | x_l[j] for every neighbour j | x_l[i]] per target) with one
//     cp.async.bulk each into a shared-memory ring of full / empty mbarriers: 6 stages x 25.6 KB in flight per SM at
//     the 1 M-cell configuration (round 1: 4 rows of 12 KB behind a __syncthreads per neighbour);
//   * scores, online softmax and the alpha-weighted sum of a head live inside one warp (shuffles only);
//   * forward writes the per-head outputs O (fp32) and the attention coefficients; the head mean
//     + bias is a second, tiny kernel;
//   * backward, per target: D_h = sum_t alpha_t dalpha_t = <g, O_h> / H is known BEFORE the
//     neighbours stream in (computed when the target's x_r row arrives), so d score, d x_r, d att and the LeakyReLU
//     sign bits of every (edge, element) are produced in ONE pass over the rows; the per-source pass then needs only
//     the sign bits (1 bit per element instead of re-reading x_l[j] + x_r[i]): d x_l[j] = sum over out
//     edges of alpha/H * g_i + d score * att * (bit ? slope : 1), and is fed by a producer ring of its own.
//
// Generic shapes (fp32 rows, widths that are not multiples of 4, more than 16 segments, very long
// segments) stay on the kernels of gatv2.cu: gat_ring_* return 1 ("not applicable") for them.
#include "common.cuh"

extern void nc_count_launch(int n);

namespace {
constexpr int RW_MAX = 8;            // consumer warps per CTA (two warpgroups)
// 8 consumer warps + one producer warpgroup (1 working warp, 3 that only hand their registers over):
// setmaxnreg moves the register file from the producer group (40 per thread) to the consumers (232 per thread)
constexpr int RING_THREADS = (RW_MAX + 4) * 32;
constexpr int RS_MAX = 16;           // ring stages per warp
constexpr int K_NBR = 0, K_XR = 1, K_LAST = 256;
constexpr int SMEM_MAX = 227 * 1024;

// element offset in the row, C / 4, output column, c * H + h; copy window of the segment inside a row
// (16-byte aligned superset, in bytes) and the byte offset of the segment's first element inside the stage
struct RingSeg { int off, C4, out_off, ch, st_off; };   // st_off: byte offset of the segment inside a row
struct RingPlan {
    RingSeg a[RW_MAX], b[RW_MAX];                  // b[w].C4 == 0: warp w has one segment only
    int n_warps, NL, n_stages, stage_bytes, row_bytes, sc_stride, NCH, out_width, max_deg1, meta_off, bar_off, sc_off;
    int lane_off;   // per-thread scratch [warp][chunk][lane] of 16 bytes: att (forward) / the target's dout (backward)
    int full;       // every warp's segments fill all chunks but (possibly) the last one: compile-time row offsets
    int e_mask_off, e_al_off;   // per-source kernel: offsets of the sign words / the alpha + d score rows in a stage
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    uint32_t ok;
    do {
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}"
            : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
    } while (!ok);
}
__device__ __forceinline__ void bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ float4 bf4(uint2 r) {
    float4 o;
    o.x = __uint_as_float(r.x << 16);
    o.y = __uint_as_float(r.x & 0xffff0000u);
    o.z = __uint_as_float(r.y << 16);
    o.w = __uint_as_float(r.y & 0xffff0000u);
    return o;
}
__device__ __forceinline__ uint2 pack_bf4(float4 v) {
    __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y);
    __nv_bfloat162 b = __floats2bfloat162_rn(v.z, v.w);
    uint2 r;
    r.x = *reinterpret_cast<uint32_t*>(&a);
    r.y = *reinterpret_cast<uint32_t*>(&b);
    return r;
}
__device__ __forceinline__ float lrelu(float v, float slope) { return fmaxf(v, v * slope); }   // 0 < slope < 1

// Packed fp32 pairs (FFMA2 / FADD2 on sm_100): half the issue slots of the scalar forms.  A 64-bit word of a row
// holds four bf16: (lo, hi) of .x -> elements 0, 1; of .y -> elements 2, 3.
__device__ __forceinline__ float2 bf2lo(uint2 r) { return make_float2(__uint_as_float(r.x << 16), __uint_as_float(r.x & 0xffff0000u)); }
__device__ __forceinline__ float2 bf2hi(uint2 r) { return make_float2(__uint_as_float(r.y << 16), __uint_as_float(r.y & 0xffff0000u)); }
__device__ __forceinline__ float2 f2(float a, float b) { return make_float2(a, b); }

// Sums of two values over the warp, both returned in every lane: lanes 0-15 reduce a, lanes 16-31 reduce b
// (7 shuffles with a dependent chain of 6, instead of 10 in a chain of 10 for two butterflies).
__device__ __forceinline__ void warp_sum2(float& a, float& b, int lane) {
    const bool up = (lane & 16) != 0;
    float v = up ? b : a;
    const float o = up ? a : b;
    v += __shfl_xor_sync(0xffffffffu, o, 16);
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    v += __shfl_xor_sync(0xffffffffu, v, 4);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    a = __shfl_sync(0xffffffffu, v, 0);
    b = __shfl_sync(0xffffffffu, v, 16);
}

// ---------------------------------------------------------------------------------- CTA pipeline
// One producer warp per CTA enumerates, for the targets ir = blockIdx.x, + gridDim.x, ..., the rows
// [x_r[ir] | x_l[j] for every CSR entry j != i | x_l[i]] and copies each WHOLE row (all segments, one bulk copy)
// into the next free stage of a shared-memory ring; the consumer warps wait on the stage's full barrier, pull
// their segments into registers and arrive on its empty barrier.  (First version: every consumer warp ran its own
// cursor and its own ring -- ~300 instructions of bookkeeping per row and warp, 40 % of all instructions issued.)
// The CSR range / first 32 column ids of the NEXT target are fetched one target ahead.
struct Cursor {
    int ir, e0, e1, t, cbase, jv;
    int ne0, ne1, njv;
};
struct Ring {
    uint8_t* base;       // n_stages * stage_bytes
    int4* meta;          // [n_stages]: {kind | flags, ir, slot, t}
    uint64_t* full;      // [n_stages]
    uint64_t* empty;     // [n_stages]
};
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ void cur_fetch(const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col, int ir,
                                          int n_dst, int dst_off, int lane, int& e0, int& e1, int& jv) {
    e0 = 0; e1 = 0; jv = -1;
    if (ir < n_dst) {
        e0 = __ldg(rowptr + ir + dst_off);
        e1 = __ldg(rowptr + ir + dst_off + 1);
        jv = (e0 + lane < e1) ? __ldg(col + e0 + lane) : -1;
    }
}
__device__ __forceinline__ void cur_init(Cursor& c, const int32_t* rowptr, const int32_t* col, int n_dst, int dst_off,
                                         int lane) {
    c.ir = blockIdx.x;
    c.t = -1;
    c.cbase = 0;
    cur_fetch(rowptr, col, c.ir, n_dst, dst_off, lane, c.e0, c.e1, c.jv);
    cur_fetch(rowptr, col, c.ir + (int)gridDim.x, n_dst, dst_off, lane, c.ne0, c.ne1, c.njv);
}

// Producer warp: starts the copy of the next row of the sequence into stage `ps` (all lanes call; lane 0 issues).
// Returns false when the sequence is exhausted.
// What the consumers of the per-target backward read from global memory when a target's x_r row arrives: its dout
// row, its per-head forward outputs and its alpha rows.  The producer asks for them in L2 (one bulk prefetch each)
// when it issues the x_r row, several rows before the consumers get there.
struct TargetPrefetch {
    const float* g; int64_t ldg_; uint32_t g_bytes;
    const float* o; int64_t ldo_; uint32_t o_bytes;
    const float* alpha; uint32_t a_row_bytes;
};
__device__ __forceinline__ void l2_prefetch(const void* p, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}

__device__ __forceinline__ bool issue_next(Cursor& c, const Ring& R, const RingPlan& P, uint32_t ps,
                                           const __nv_bfloat16* __restrict__ xl, const __nv_bfloat16* __restrict__ xr,
                                           int64_t ld, const int32_t* __restrict__ rowptr,
                                           const int32_t* __restrict__ col, int n_dst, int dst_off, int lane,
                                           const TargetPrefetch* pf) {
    const __nv_bfloat16* src;
    int4 meta;
    while (true) {
        if (c.ir >= n_dst) return false;
        const int i = c.ir + dst_off, deg = c.e1 - c.e0;
        if (c.t < 0) {                                               // the target's own x_r row opens its sequence
            src = xr + (int64_t)c.ir * ld;
            meta = make_int4(K_XR, c.ir, c.e0 + i, deg + 1);
            c.t = 0;
            if (pf != nullptr && lane == 0) {
                l2_prefetch(pf->g + (int64_t)c.ir * pf->ldg_, pf->g_bytes);
                l2_prefetch(pf->o + (int64_t)c.ir * pf->ldo_, pf->o_bytes);
                l2_prefetch(pf->alpha + (int64_t)(c.e0 + i) * (pf->a_row_bytes / 4u), (uint32_t)(deg + 1) * pf->a_row_bytes);
            }
            break;
        }
        if (c.t < deg) {
            if (c.t - c.cbase >= 32) {
                c.cbase += 32;
                c.jv = (c.e0 + c.cbase + lane < c.e1) ? __ldg(col + c.e0 + c.cbase + lane) : -1;
            }
            const int j = __shfl_sync(0xffffffffu, c.jv, c.t - c.cbase);
            const int t = c.t++;
            if (j == i) continue;                                    // PyG removes existing self loops ...
            src = xl + (int64_t)j * ld;
            meta = make_int4(K_NBR, c.ir, c.e0 + i + t, t);
            break;
        }
        src = xl + (int64_t)i * ld;                                  // ... and adds exactly one self edge
        meta = make_int4(K_NBR | K_LAST, c.ir, c.e1 + i, deg);
        c.ir += (int)gridDim.x;
        c.e0 = c.ne0; c.e1 = c.ne1; c.jv = c.njv;
        c.t = -1;
        c.cbase = 0;
        cur_fetch(rowptr, col, c.ir + (int)gridDim.x, n_dst, dst_off, lane, c.ne0, c.ne1, c.njv);
        break;
    }
    if (lane == 0) {
        R.meta[ps] = meta;
        // the consumers' generic-proxy reads of this stage (ordered before this point by their arrivals on the
        // empty barrier) must not be overtaken by the async-proxy write
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        mbar_expect_tx(&R.full[ps], (uint32_t)P.row_bytes);
        bulk_load(R.base + (size_t)ps * P.stage_bytes, src, (uint32_t)P.row_bytes, &R.full[ps]);
    }
    return true;
}

// All threads of the CTA call; barriers initialised by thread 0.
__device__ __forceinline__ Ring ring_setup(const RingPlan& P, uint8_t* smem) {
    Ring R;
    R.base = smem;
    R.meta = reinterpret_cast<int4*>(smem + P.meta_off);
    R.full = reinterpret_cast<uint64_t*>(smem + P.bar_off);
    R.empty = R.full + P.n_stages;
    if (threadIdx.x == 0) {
        for (int s = 0; s < P.n_stages; ++s) {
            mbar_init(&R.full[s], 1);
            mbar_init(&R.empty[s], (uint32_t)P.n_warps);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    return R;
}

// The producer warp's whole life: fill stages round-robin, waiting for a stage to be drained before reuse.
__device__ __forceinline__ void producer_loop(const Ring& R, const RingPlan& P, const __nv_bfloat16* __restrict__ xl,
                                              const __nv_bfloat16* __restrict__ xr, int64_t ld,
                                              const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
                                              int n_dst, int dst_off, int lane, const TargetPrefetch* pf = nullptr) {
    Cursor cur;
    cur_init(cur, rowptr, col, n_dst, dst_off, lane);
    uint32_t stage = 0, round = 0;
    while (true) {
        mbar_wait(&R.empty[stage], (round & 1) ^ 1);                 // passes at once in the first round
        if (!issue_next(cur, R, P, stage, xl, xr, ld, rowptr, col, n_dst, dst_off, lane, pf)) break;
        if (++stage == (uint32_t)P.n_stages) { stage = 0; ++round; }
    }
}

// ---------------------------------------------------------------------------------- forward
// Byte offset of chunk k of this lane inside a stage.  FULL plans (every warp's segments fill all chunks but
// possibly the last): compile-time offsets from one base register, only the last chunk of a segment selects
// between its own address and a valid dummy (offS) for the lanes past the segment's end.
#define ROFF(k)                                                                                                  \
    (FULL ? ((k) < NVA ? ((k) == NVA - 1 ? offAl : offA + 256u * (uint32_t)(k))                                  \
                       : ((k) == NV - 1 ? offBl : offB + 256u * (uint32_t)((k) - NVA)))                           \
          : ((((vmask) >> (k)) & 1u) ? ((k) < NVA ? offA + 256u * (uint32_t)(k)                                  \
                                                  : offB + 256u * (uint32_t)((k) - NVA)) : offS))

// Instruction diet of the inner loop (the first version of this kernel issued 15 instructions per row element and
// was bound by issue slots -- 2 warps per scheduler, IPC 0.5 -- not by memory; ncu: profiles/r2_gatv2_*):
//   * leaky_relu(v) = c1 v + c2 |v| with c1 = (1 + slope) / 2, c2 = (1 - slope) / 2, so the score is
//     c1 <att, x_l[j]> + c2 <att, |x_l[j] + x_r[i]|> (+ c1 <att, x_r[i]>, which is the same for every neighbour of i
//     and cancels in the softmax): half a packed add, half a packed multiply-add and one multiply-add with |.| as an
//     operand modifier per element, instead of add, multiply, max, multiply-add;
//   * a row is unpacked to fp32 ONCE and stays in registers from the score to the weighted accumulation (the stage
//     is released as soon as it is unpacked); att moved from 52 registers to a per-thread shared-memory slot (one
//     LDS.128 per chunk and row) to make room for it;
//   * fp32 arithmetic in pairs (FFMA2 / FADD2); the two head scores of a warp are reduced in one butterfly;
//   * lanes past the end of a segment read the first element of the slot (valid data) with att = 0 instead of
//     branching around every 128-bit chunk.
template <int NVA, int NVB, bool FULL>
__global__ void __launch_bounds__(RING_THREADS, 1)
k_gat_ring_fwd(const __grid_constant__ RingPlan P, const __nv_bfloat16* __restrict__ xl,
               const __nv_bfloat16* __restrict__ xr, int64_t ld, const float* __restrict__ att,
               const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col, int n_dst, int dst_off, float slope,
               float* __restrict__ head_out, float* __restrict__ alpha) {
    constexpr int NV = NVA + NVB;
    extern __shared__ __align__(128) uint8_t smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int n_left = (n_dst > (int)blockIdx.x) ? (n_dst - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
    if (n_left == 0) return;
    const Ring R = ring_setup(P, smem);
    if (warp >= RW_MAX) {                                     // producer warpgroup: warp 8 works, 9..11 only donate
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        if (warp == RW_MAX) producer_loop(R, P, xl, xr, ld, rowptr, col, n_dst, dst_off, lane);
        return;
    }
    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    const RingSeg sa = P.a[warp], sb = P.b[warp];
    const float c1 = 0.5f * (1.0f + slope), c2 = 0.5f * (1.0f - slope);
    uint32_t vmask = 0;
    // offS: a valid, 8-byte aligned element of the slot for the lanes past the end of a segment
    const uint32_t offA = (uint32_t)(sa.st_off + 8 * lane), offB = (uint32_t)(sb.st_off + 8 * lane), offS = (uint32_t)sa.st_off;
    // this thread's slot of the per-lane scratch: al_[k * 32] = att of chunk k (0 past the end of a segment)
    float4* al_ = reinterpret_cast<float4*>(smem + P.lane_off) + (size_t)warp * NV * 32 + lane;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        const bool isA = k < NVA;
        const int u = lane + 32 * (isA ? k : k - NVA);
        float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
        if (u < (isA ? sa.C4 : sb.C4)) {
            vmask |= 1u << k;
            a = __ldg(reinterpret_cast<const float4*>(att + (isA ? sa.off : sb.off) + 4 * u));
        }
        al_[k * 32] = a;
    }
    const uint32_t offAl = ((vmask >> (NVA - 1)) & 1u) ? offA + 256u * (uint32_t)(NVA - 1) : offS;
    const uint32_t offBl = (NVB > 0 && ((vmask >> (NV - 1)) & 1u)) ? offB + 256u * (uint32_t)(NVB > 0 ? NVB - 1 : 0) : offS;
    float* scA = reinterpret_cast<float*>(smem + P.sc_off + (size_t)warp * P.sc_stride);
    float* scB = scA + P.max_deg1;
    if (alpha) {
        for (int t = lane; t < P.max_deg1; t += 32) { scA[t] = -INFINITY; scB[t] = -INFINITY; }
        __syncwarp();
    }
    float2 xrv[NV][2], acc[NV][2];
#pragma unroll
    for (int k = 0; k < NV; ++k) xrv[k][0] = xrv[k][1] = acc[k][0] = acc[k][1] = f2(0.f, 0.f);
    float mA = -INFINITY, lA = 0.f, mB = -INFINITY, lB = 0.f;
    int slot0 = 0, nE = 0;
    uint32_t stage = 0, round = 0;
    while (true) {
        mbar_wait(&R.full[stage], round & 1);
        const uint8_t* row = R.base + (size_t)stage * P.stage_bytes;
        const int4 meta = R.meta[stage];
        float2 fv[NV][2];
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            const uint2 raw = *reinterpret_cast<const uint2*>(row + ROFF(k));
            fv[k][0] = bf2lo(raw);
            fv[k][1] = bf2hi(raw);
        }
        __syncwarp();                                        // the row now lives in registers: release the stage
        if (lane == 0) mbar_arrive(&R.empty[stage]);
        if ((meta.x & 255) == K_XR) {
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                xrv[k][0] = fv[k][0];
                xrv[k][1] = fv[k][1];
                acc[k][0] = acc[k][1] = f2(0.f, 0.f);
            }
            mA = mB = -INFINITY;
            lA = lB = 0.f;
            slot0 = meta.z;
            nE = min(meta.w, P.max_deg1);
        } else {
            float2 s1A0 = f2(0.f, 0.f), s1A1 = f2(0.f, 0.f), s1B = f2(0.f, 0.f);
            float s2A0 = 0.f, s2A1 = 0.f, s2A2 = 0.f, s2A3 = 0.f, s2B0 = 0.f, s2B1 = 0.f;
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                const float4 a = al_[k * 32];
                const float2 v0 = __fadd2_rn(fv[k][0], xrv[k][0]), v1 = __fadd2_rn(fv[k][1], xrv[k][1]);
                if (k < NVA) {
                    s1A0 = __ffma2_rn(f2(a.x, a.y), fv[k][0], s1A0);
                    s1A1 = __ffma2_rn(f2(a.z, a.w), fv[k][1], s1A1);
                    s2A0 = fmaf(a.x, fabsf(v0.x), s2A0); s2A1 = fmaf(a.y, fabsf(v0.y), s2A1);
                    s2A2 = fmaf(a.z, fabsf(v1.x), s2A2); s2A3 = fmaf(a.w, fabsf(v1.y), s2A3);
                } else {
                    s1B = __ffma2_rn(f2(a.x, a.y), fv[k][0], s1B);
                    s1B = __ffma2_rn(f2(a.z, a.w), fv[k][1], s1B);
                    s2B0 = fmaf(a.x, fabsf(v0.x), s2B0); s2B1 = fmaf(a.y, fabsf(v0.y), s2B1);
                    s2B0 = fmaf(a.z, fabsf(v1.x), s2B0); s2B1 = fmaf(a.w, fabsf(v1.y), s2B1);
                }
            }
            float pA = fmaf(c1, (s1A0.x + s1A0.y) + (s1A1.x + s1A1.y), c2 * ((s2A0 + s2A1) + (s2A2 + s2A3)));
            float pB = 0.f;
            if (NVB > 0) {
                pB = fmaf(c1, s1B.x + s1B.y, c2 * (s2B0 + s2B1));
                warp_sum2(pA, pB, lane);
            } else {
                pA = warp_sum(pA);
            }
            const int t = meta.w;
            if (alpha && lane == 0 && t < P.max_deg1) { scA[t] = pA; scB[t] = pB; }
            {
                const float mn = fmaxf(mA, pA), corr = __expf(mA - mn), w = __expf(pA - mn);
                lA = lA * corr + w;
                mA = mn;
                if (corr != 1.f) {
                    const float2 c = f2(corr, corr);
#pragma unroll
                    for (int k = 0; k < NVA; ++k) { acc[k][0] = __fmul2_rn(acc[k][0], c); acc[k][1] = __fmul2_rn(acc[k][1], c); }
                }
                const float2 w2 = f2(w, w);
#pragma unroll
                for (int k = 0; k < NVA; ++k) {
                    acc[k][0] = __ffma2_rn(w2, fv[k][0], acc[k][0]);
                    acc[k][1] = __ffma2_rn(w2, fv[k][1], acc[k][1]);
                }
            }
            if (NVB > 0) {
                const float mn = fmaxf(mB, pB), corr = __expf(mB - mn), w = __expf(pB - mn);
                lB = lB * corr + w;
                mB = mn;
                const float2 c = f2(corr, corr), w2 = f2(w, w);
#pragma unroll
                for (int k = NVA; k < NV; ++k) {
                    acc[k][0] = __ffma2_rn(w2, fv[k][0], __fmul2_rn(acc[k][0], c));
                    acc[k][1] = __ffma2_rn(w2, fv[k][1], __fmul2_rn(acc[k][1], c));
                }
            }
            if (meta.x & K_LAST) {
                const int ir = meta.y;
                const float iA = 1.0f / (lA + 1e-16f), iB = 1.0f / (lB + 1e-16f);
                float* orow = head_out + (int64_t)ir * ld;
#pragma unroll
                for (int k = 0; k < NV; ++k) {
                    if ((vmask >> k) & 1u) {
                        const bool isA = k < NVA;
                        const int u = lane + 32 * (isA ? k : k - NVA);
                        const float inv = isA ? iA : iB;
                        *reinterpret_cast<float4*>(orow + (isA ? sa.off : sb.off) + 4 * u) =
                            make_float4(acc[k][0].x * inv, acc[k][0].y * inv, acc[k][1].x * inv, acc[k][1].y * inv);
                    }
                }
                if (alpha) {
                    __syncwarp();
                    for (int tt = lane; tt < nE; tt += 32) {
                        float* al = alpha + (int64_t)(slot0 + tt) * P.NCH;
                        // a removed self entry (PyG drops it, cursor skipped it) never got a score: alpha = 0
                        const float sA = scA[tt], sB = scB[tt];
                        al[sa.ch] = sA == -INFINITY ? 0.f : __expf(sA - mA) * iA;
                        if (sb.C4 > 0) al[sb.ch] = sB == -INFINITY ? 0.f : __expf(sB - mB) * iB;
                    }
                    __syncwarp();
                    for (int tt = lane; tt < nE; tt += 32) { scA[tt] = -INFINITY; scB[tt] = -INFINITY; }
                    __syncwarp();
                }
                if (--n_left == 0) break;
            }
        }
        if (++stage == (uint32_t)P.n_stages) { stage = 0; ++round; }
    }
}

#undef ROFF

// out[ir, out_off_c + k] = mean_h O[ir, off_c + h * C_c + k] + bias   (one CTA per target)
__global__ void __launch_bounds__(256)
k_gat_head_mean(const float* __restrict__ head_out, int64_t ld, const int32_t* __restrict__ seg, int n_conv, int H,
                const float* __restrict__ bias, float* __restrict__ out, int64_t ldo) {
    const int ir = blockIdx.x;
    const float invH = 1.0f / (float)H;
    const float* o = head_out + (int64_t)ir * ld;
    for (int c = 0; c < n_conv; ++c) {
        const int off = seg[c * 4], C = seg[c * 4 + 1], oo = seg[c * 4 + 2];
        for (int k = threadIdx.x * 4; k < C; k += 256 * 4) {
            float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int h = 0; h < H; ++h) {
                const float4 v = *reinterpret_cast<const float4*>(o + off + h * C + k);
                s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
            }
            float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
            if (bias) b = __ldg(reinterpret_cast<const float4*>(bias + oo + k));
            *reinterpret_cast<float4*>(out + (int64_t)ir * ldo + oo + k) =
                make_float4(s.x * invH + b.x, s.y * invH + b.y, s.z * invH + b.z, s.w * invH + b.w);
        }
    }
}

// ---------------------------------------------------------------------------------- backward, per target
// Byte offset of chunk k of this lane inside a stage.  FULL plans (every warp's segments fill all chunks but
// possibly the last): compile-time offsets from one base register, only the last chunk of a segment selects
// between its own address and a valid dummy (offS) for the lanes past the segment's end.
#define ROFF(k)                                                                                                  \
    (FULL ? ((k) < NVA ? ((k) == NVA - 1 ? offAl : offA + 256u * (uint32_t)(k))                                  \
                       : ((k) == NV - 1 ? offBl : offB + 256u * (uint32_t)((k) - NVA)))                           \
          : ((((vmask) >> (k)) & 1u) ? ((k) < NVA ? offA + 256u * (uint32_t)(k)                                  \
                                                  : offB + 256u * (uint32_t)((k) - NVA)) : offS))

// Instruction diet (the first ring version of this kernel issued ~16 instructions per row element, kept 52 + 52 +
// 52 + 26 values per thread live and spilled in the inner loop -- 3.9 long-scoreboard stall cycles per issue):
//   * the target's dout segment lives in a per-thread shared-memory slot (written once per target, one LDS.128 per
//     chunk and row) instead of 52 registers: no spills;
//   * D_h = <g, O_h> / H is computed here when the target's x_r row arrives (was a kernel of its own + a round trip);
//   * v = x_l[j] + x_r[i] is added in round-down mode, where an exact zero comes out as -0 (x_r is stored with its
//     zeros as -0, so no pair of operands is (+0, +0)): the sign bit of v is exactly "v <= 0", what torch's
//     leaky_relu backward tests.  The sign bits are shifted into the mask word with one funnel shift per element,
//     and d lrelu = ds * (c1 + c2 * sgn v) is E + (q xor sign bit): one LOP3 and half a packed add per element;
//   * all fp32 arithmetic in pairs (FFMA2 / FADD2).
template <int NVA, int NVB, bool FULL>
__global__ void __launch_bounds__(RING_THREADS, 1)
k_gat_ring_bwd(const __grid_constant__ RingPlan P, const __nv_bfloat16* __restrict__ xl,
               const __nv_bfloat16* __restrict__ xr, int64_t ld, const float* __restrict__ att,
               const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col, int n_dst, float slope,
               float invH, const float* __restrict__ dout, int64_t ldo, const float* __restrict__ alpha,
               const float* __restrict__ head_out, float* __restrict__ ds, __nv_bfloat16* __restrict__ dxr,
               float* __restrict__ datt, uint32_t* __restrict__ mask) {
    constexpr int NV = NVA + NVB;
    extern __shared__ __align__(128) uint8_t smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int n_left = (n_dst > (int)blockIdx.x) ? (n_dst - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
    if (n_left == 0) return;
    const Ring R = ring_setup(P, smem);
    if (warp >= RW_MAX) {                                     // producer warpgroup: warp 8 works, 9..11 only donate
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        if (warp == RW_MAX) {
            TargetPrefetch pf;
            pf.g = dout; pf.ldg_ = ldo; pf.g_bytes = (uint32_t)P.out_width * 4u;
            pf.o = head_out; pf.ldo_ = ld; pf.o_bytes = (uint32_t)P.NL * 4u;
            pf.alpha = alpha; pf.a_row_bytes = (uint32_t)P.NCH * 4u;
            producer_loop(R, P, xl, xr, ld, rowptr, col, n_dst, 0, lane, (P.NCH % 4 == 0) ? &pf : nullptr);
        }
        return;
    }
    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    const RingSeg sa = P.a[warp], sb = P.b[warp];
    const float c1 = 0.5f * (1.0f + slope), c2 = 0.5f * (1.0f - slope);
    uint32_t vmask = 0;
    const uint32_t offA = (uint32_t)(sa.st_off + 8 * lane), offB = (uint32_t)(sb.st_off + 8 * lane), offS = (uint32_t)sa.st_off;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        const bool isA = k < NVA;
        const int u = lane + 32 * (isA ? k : k - NVA);
        if (u < (isA ? sa.C4 : sb.C4)) vmask |= 1u << k;
    }
    const uint32_t offAl = ((vmask >> (NVA - 1)) & 1u) ? offA + 256u * (uint32_t)(NVA - 1) : offS;
    const uint32_t offBl = (NVB > 0 && ((vmask >> (NV - 1)) & 1u)) ? offB + 256u * (uint32_t)(NVB > 0 ? NVB - 1 : 0) : offS;
    // this thread's slot of the per-lane scratch: gl[k * 32] = dout of chunk k of the current target
    float4* gl = reinterpret_cast<float4*>(smem + P.lane_off) + (size_t)warp * NV * 32 + lane;
    float2 nx[NV][2], dxa[NV][2], dat[NV][2];
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        nx[k][0] = nx[k][1] = dxa[k][0] = dxa[k][1] = dat[k][0] = dat[k][1] = f2(0.f, 0.f);
    }
    float DA = 0.f, DB = 0.f, alA = 0.f, alB = 0.f;       // alA / alB: alpha of entry `lane` of the current target
    int slot0 = 0;
    uint32_t stage = 0, round = 0;
    while (true) {
        mbar_wait(&R.full[stage], round & 1);
        const uint8_t* row = R.base + (size_t)stage * P.stage_bytes;
        const int4 meta = R.meta[stage];
        if ((meta.x & 255) == K_XR) {
            const int ir = meta.y;
            slot0 = meta.z;
            const float* g = dout + (int64_t)ir * ldo;
            const float* o = head_out + (int64_t)ir * ld;
            float2 dA2 = f2(0.f, 0.f), dB2 = f2(0.f, 0.f);
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                const bool isA = k < NVA;
                const int u4 = 4 * (lane + 32 * (isA ? k : k - NVA));
                const float4 f = bf4(*reinterpret_cast<const uint2*>(row + ROFF(k)));
                nx[k][0] = f2(f.x == 0.f ? -0.f : f.x, f.y == 0.f ? -0.f : f.y);
                nx[k][1] = f2(f.z == 0.f ? -0.f : f.z, f.w == 0.f ? -0.f : f.w);
                float4 gg = make_float4(0.f, 0.f, 0.f, 0.f), oo = make_float4(0.f, 0.f, 0.f, 0.f);
                if ((vmask >> k) & 1u) {
                    gg = __ldg(reinterpret_cast<const float4*>(g + (isA ? sa.out_off : sb.out_off) + u4));
                    oo = __ldg(reinterpret_cast<const float4*>(o + (isA ? sa.off : sb.off) + u4));
                }
                gl[k * 32] = gg;
                dxa[k][0] = dxa[k][1] = f2(0.f, 0.f);
                if (isA) {
                    dA2 = __ffma2_rn(f2(gg.x, gg.y), f2(oo.x, oo.y), dA2);
                    dA2 = __ffma2_rn(f2(gg.z, gg.w), f2(oo.z, oo.w), dA2);
                } else {
                    dB2 = __ffma2_rn(f2(gg.x, gg.y), f2(oo.x, oo.y), dB2);
                    dB2 = __ffma2_rn(f2(gg.z, gg.w), f2(oo.z, oo.w), dB2);
                }
            }
            DA = dA2.x + dA2.y;
            DB = dB2.x + dB2.y;
            if (NVB > 0) warp_sum2(DA, DB, lane); else DA = warp_sum(DA);
            DA *= invH;
            DB *= invH;
            alA = alB = 0.f;
            if (lane < meta.w) {
                alA = __ldg(alpha + (int64_t)(slot0 + lane) * P.NCH + sa.ch);
                if (sb.C4 > 0) alB = __ldg(alpha + (int64_t)(slot0 + lane) * P.NCH + sb.ch);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&R.empty[stage]);
        } else {
            // sweep 1: d alpha = <g, x_l[j, head]> / H
            float2 pa0 = f2(0.f, 0.f), pa1 = f2(0.f, 0.f), pb0 = f2(0.f, 0.f);
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                const uint2 raw = *reinterpret_cast<const uint2*>(row + ROFF(k));
                const float4 g4 = gl[k * 32];
                if (k < NVA) {
                    pa0 = __ffma2_rn(bf2lo(raw), f2(g4.x, g4.y), pa0);
                    pa1 = __ffma2_rn(bf2hi(raw), f2(g4.z, g4.w), pa1);
                } else {
                    pb0 = __ffma2_rn(bf2lo(raw), f2(g4.x, g4.y), pb0);
                    pb0 = __ffma2_rn(bf2hi(raw), f2(g4.z, g4.w), pb0);
                }
            }
            float pA = (pa0.x + pa0.y) + (pa1.x + pa1.y), pB = pb0.x + pb0.y;
            if (NVB > 0) warp_sum2(pA, pB, lane); else pA = warp_sum(pA);
            const int slot = meta.z, t = meta.w;
            const bool far = t >= 32;                       // entries past the 32 held in alA / alB: read them
            const float* al = alpha + (int64_t)slot * P.NCH;
            const float fA = far ? __ldg(al + sa.ch) : 0.f;
            const float fB = (far && sb.C4 > 0) ? __ldg(al + sb.ch) : 0.f;
            const float sA = __shfl_sync(0xffffffffu, alA, t & 31), sB = __shfl_sync(0xffffffffu, alB, t & 31);
            const float aA = far ? fA : sA, aB = far ? fB : sB;
            const float dsA = aA * fmaf(pA, invH, -DA);
            const float dsB = aB * fmaf(pB, invH, -DB);
            if (lane == 0) {
                ds[(int64_t)slot * P.NCH + sa.ch] = dsA;
                if (sb.C4 > 0) ds[(int64_t)slot * P.NCH + sb.ch] = dsB;
            }
            // sweep 2: d x_r, d att and the LeakyReLU sign bits of this edge
            const float2 EA = f2(dsA * c1, dsA * c1), EB = f2(dsB * c1, dsB * c1);
            const uint32_t qA = __float_as_uint(dsA * c2), qB = __float_as_uint(dsB * c2);
            uint32_t w0 = 0u, w1 = 0u;
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                const uint2 raw = *reinterpret_cast<const uint2*>(row + ROFF(k));
                // round-down add: an exact zero comes out as -0 (x_r is never +0), so the sign bit is "v <= 0"
                const float2 v0 = __fadd2_rd(bf2lo(raw), nx[k][0]), v1 = __fadd2_rd(bf2hi(raw), nx[k][1]);
                const uint32_t b0 = __float_as_uint(v0.x), b1 = __float_as_uint(v0.y);
                const uint32_t b2 = __float_as_uint(v1.x), b3 = __float_as_uint(v1.y);
                if (k < 8) {
                    w0 = __funnelshift_l(b0, w0, 1); w0 = __funnelshift_l(b1, w0, 1);
                    w0 = __funnelshift_l(b2, w0, 1); w0 = __funnelshift_l(b3, w0, 1);
                } else {
                    w1 = __funnelshift_l(b0, w1, 1); w1 = __funnelshift_l(b1, w1, 1);
                    w1 = __funnelshift_l(b2, w1, 1); w1 = __funnelshift_l(b3, w1, 1);
                }
                const uint32_t q = k < NVA ? qA : qB;
                const float2 E = k < NVA ? EA : EB;
                const float2 tt0 = __fadd2_rn(E, f2(__uint_as_float(q ^ (b0 & 0x80000000u)),
                                                     __uint_as_float(q ^ (b1 & 0x80000000u))));
                const float2 tt1 = __fadd2_rn(E, f2(__uint_as_float(q ^ (b2 & 0x80000000u)),
                                                     __uint_as_float(q ^ (b3 & 0x80000000u))));
                dxa[k][0] = __fadd2_rn(dxa[k][0], tt0);
                dxa[k][1] = __fadd2_rn(dxa[k][1], tt1);
                dat[k][0] = __ffma2_rn(tt0, v0, dat[k][0]);
                dat[k][1] = __ffma2_rn(tt1, v1, dat[k][1]);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&R.empty[stage]);
            uint32_t* mrow = mask + ((int64_t)slot * RW_MAX + warp) * 64;
            mrow[lane] = w0;
            if (NV > 8) mrow[32 + lane] = w1;
            if (meta.x & K_LAST) {
                // d x_r = att * sum tt: all att chunks are requested before the first one is used (one L2 latency per
                // target, not one per chunk); the two offsets are kept opaque so that they stay in registers
                const int ir = meta.y;
                int oA = sa.off + 4 * lane, oB = sb.off + 4 * lane;
                asm volatile("" : "+r"(oA), "+r"(oB));
                __nv_bfloat16* dA = dxr + (int64_t)ir * ld + oA;
                __nv_bfloat16* dB = dxr + (int64_t)ir * ld + oB;
                const float* tA = att + oA;
                const float* tB = att + oB;
                float4 av[NV];
#pragma unroll
                for (int k = 0; k < NV; ++k) {
                    const float* tp = ((vmask >> k) & 1u) ? (k < NVA ? tA + 128 * k : tB + 128 * (k - NVA)) : att;
                    av[k] = __ldg(reinterpret_cast<const float4*>(tp));
                }
#pragma unroll
                for (int k = 0; k < NV; ++k) {
                    const uint2 pk = pack_bf4(make_float4(av[k].x * dxa[k][0].x, av[k].y * dxa[k][0].y,
                                                          av[k].z * dxa[k][1].x, av[k].w * dxa[k][1].y));
                    if ((vmask >> k) & 1u) *reinterpret_cast<uint2*>(k < NVA ? dA + 128 * k : dB + 128 * (k - NVA)) = pk;
                }
                if (--n_left == 0) break;
            }
        }
        if (++stage == (uint32_t)P.n_stages) { stage = 0; ++round; }
    }
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        if ((vmask >> k) & 1u) {
            const bool isA = k < NVA;
            float* da = datt + (isA ? sa.off : sb.off) + 4 * (lane + 32 * (isA ? k : k - NVA));
            atomicAdd(da, dat[k][0].x); atomicAdd(da + 1, dat[k][0].y);
            atomicAdd(da + 2, dat[k][1].x); atomicAdd(da + 3, dat[k][1].y);
        }
    }
}
#undef ROFF

// ---------------------------------------------------------------------------------- backward, per source
// d x_l[j] = sum over the out edges (j -> i, slot) of  alpha / H * g_i  +  d score * att * (bit ? slope : 1), bit = "v <= 0".
// Same warp <-> segment plan as the per-target kernel (the sign bits are stored per warp and lane: element n of the
// first 32 of a lane sits at bit M0 - 1 - n of word 0, element 32 + n at bit M1 - 1 - n of word 1, M0 = 4 * min(NV, 8),
// M1 = 4 * (NV - 8): the order one funnel shift per element leaves them in).
//
// Third version.  The first two read (dst, slot) -> (alpha, d score, sign bits, dout row) with dependent global loads
// inside every warp: 6 long-scoreboard stall cycles per issued instruction, 140 us.  Now the same producer / ring
// scheme as the per-target kernels: a producer warp walks the reverse CSR of the sources b, b + grid, ... (index data
// of 32 sources fetched at a time, one lane each) and, per out edge, bulk-copies [dout row of the target | the sign
// words of the slot | alpha row | d score row] into the next free stage; the 8 consumer warps only touch shared
// memory and registers.  A source without out edges gets a stage that carries no data (its row is written as zeros).
// Arithmetic: with E = ds * c1, q = ds * c2 the coefficient of att is E + (bit ? -q : q); sum_e E_e is one scalar per
// head, sum_e (+-q_e) one packed add per element pair, both folded into the row when its last edge has arrived.
constexpr int E_DATA = 1, E_LAST = 2;

__device__ __forceinline__ void src_producer(const Ring& R, const RingPlan& P, const int32_t* __restrict__ rptr,
                                             const int32_t* __restrict__ rdst, const int32_t* __restrict__ rslot,
                                             const float* __restrict__ dout, int64_t ldo,
                                             const float* __restrict__ alpha, const float* __restrict__ ds,
                                             const uint32_t* __restrict__ mask, int n_src, int lane) {
    const int G = (int)gridDim.x;
    const uint32_t g_bytes = (uint32_t)P.out_width * 4u, a_bytes = (uint32_t)P.NCH * 4u;
    const uint32_t m_bytes = RW_MAX * 256u;
    uint32_t stage = 0, round = 0;
    for (int64_t jb = blockIdx.x; jb < n_src; jb += 32 * (int64_t)G) {
        const int64_t jl = jb + (int64_t)lane * G;
        int p0 = 0, p1 = 0;
        if (jl < n_src) { p0 = __ldg(rptr + jl); p1 = __ldg(rptr + jl + 1); }
        int d0 = 0, d1 = 0, d2 = 0, d3 = 0, s0 = 0, s1 = 0, s2 = 0, s3 = 0;
        if (p0 < p1) { d0 = __ldg(rdst + p0); s0 = __ldg(rslot + p0); }
        if (p0 + 1 < p1) { d1 = __ldg(rdst + p0 + 1); s1 = __ldg(rslot + p0 + 1); }
        if (p0 + 2 < p1) { d2 = __ldg(rdst + p0 + 2); s2 = __ldg(rslot + p0 + 2); }
        if (p0 + 3 < p1) { d3 = __ldg(rdst + p0 + 3); s3 = __ldg(rslot + p0 + 3); }
        for (int l = 0; l < 32; ++l) {
            const int64_t j = jb + (int64_t)l * G;
            if (j >= n_src) break;
            const int q0 = __shfl_sync(0xffffffffu, p0, l), q1 = __shfl_sync(0xffffffffu, p1, l);
            const int e0 = __shfl_sync(0xffffffffu, d0, l), e1 = __shfl_sync(0xffffffffu, d1, l);
            const int e2 = __shfl_sync(0xffffffffu, d2, l), e3 = __shfl_sync(0xffffffffu, d3, l);
            const int t0 = __shfl_sync(0xffffffffu, s0, l), t1 = __shfl_sync(0xffffffffu, s1, l);
            const int t2 = __shfl_sync(0xffffffffu, s2, l), t3 = __shfl_sync(0xffffffffu, s3, l);
            if (q0 == q1) {                                              // no out edge: a stage without data
                mbar_wait(&R.empty[stage], (round & 1) ^ 1);
                if (lane == 0) {
                    R.meta[stage] = make_int4(E_LAST, (int)j, 0, 0);
                    mbar_arrive(&R.full[stage]);
                }
                if (++stage == (uint32_t)P.n_stages) { stage = 0; ++round; }
                continue;
            }
            for (int p = q0; p < q1; ++p) {
                const int q = p - q0;
                int i, slot;
                if (q < 4) {
                    i = q == 0 ? e0 : (q == 1 ? e1 : (q == 2 ? e2 : e3));
                    slot = q == 0 ? t0 : (q == 1 ? t1 : (q == 2 ? t2 : t3));
                } else {
                    i = __ldg(rdst + p);
                    slot = __ldg(rslot + p);
                }
                mbar_wait(&R.empty[stage], (round & 1) ^ 1);
                if (lane == 0) {
                    R.meta[stage] = make_int4(E_DATA | (p + 1 == q1 ? E_LAST : 0), (int)j, 0, 0);
                    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                    uint8_t* dst = R.base + (size_t)stage * P.stage_bytes;
                    mbar_expect_tx(&R.full[stage], g_bytes + m_bytes + 2u * a_bytes);
                    bulk_load(dst, dout + (int64_t)i * ldo, g_bytes, &R.full[stage]);
                    bulk_load(dst + P.e_mask_off, mask + (int64_t)slot * RW_MAX * 64, m_bytes, &R.full[stage]);
                    bulk_load(dst + P.e_al_off, alpha + (int64_t)slot * P.NCH, a_bytes, &R.full[stage]);
                    bulk_load(dst + P.e_al_off + a_bytes, ds + (int64_t)slot * P.NCH, a_bytes, &R.full[stage]);
                }
                if (++stage == (uint32_t)P.n_stages) { stage = 0; ++round; }
            }
        }
    }
}

// One out edge of the current source: acc += alpha / H * g, T += (bit ? -q : q).  FIRST: the source's first edge
// initialises the two accumulators instead (no zeroing of 104 registers per source).
template <int NVA, int NVB, bool FIRST>
__device__ __forceinline__ void src_edge(const uint8_t* st, uint32_t gA, uint32_t gB, uint32_t w0, uint32_t w1,
                                         uint32_t qA, uint32_t qB, float aA, float aB,
                                         float2 (&acc)[NVA + NVB][2], float2 (&T)[NVA + NVB][2]) {
    constexpr int NV = NVA + NVB;
    constexpr int TOP0 = 4 * (NV < 8 ? NV : 8) - 1, TOP1 = NV > 8 ? 4 * (NV - 8) - 1 : 0;
    const float2 awA = f2(aA, aA), awB = f2(aB, aB);
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        const bool isA = k < NVA;
        const float4 gg = *reinterpret_cast<const float4*>(st + (isA ? gA + 512u * (uint32_t)k
                                                                    : gB + 512u * (uint32_t)(k - NVA)));
        const uint32_t wv = (k < 8) ? w0 : w1;
        const int top = ((k < 8) ? TOP0 : TOP1) - 4 * (k & 7);          // bit of element 0 of this chunk
        const uint32_t q = isA ? qA : qB;
        const float2 aw = isA ? awA : awB;
        const float2 sq0 = f2(__uint_as_float(q ^ ((wv << (31 - top)) & 0x80000000u)),
                              __uint_as_float(q ^ ((wv << (32 - top)) & 0x80000000u)));
        const float2 sq1 = f2(__uint_as_float(q ^ ((wv << (33 - top)) & 0x80000000u)),
                              __uint_as_float(q ^ ((wv << (34 - top)) & 0x80000000u)));
        if (FIRST) {
            T[k][0] = sq0;
            T[k][1] = sq1;
            acc[k][0] = __fmul2_rn(aw, f2(gg.x, gg.y));
            acc[k][1] = __fmul2_rn(aw, f2(gg.z, gg.w));
        } else {
            T[k][0] = __fadd2_rn(T[k][0], sq0);
            T[k][1] = __fadd2_rn(T[k][1], sq1);
            acc[k][0] = __ffma2_rn(aw, f2(gg.x, gg.y), acc[k][0]);
            acc[k][1] = __ffma2_rn(aw, f2(gg.z, gg.w), acc[k][1]);
        }
    }
}

template <int NVA, int NVB>
__global__ void __launch_bounds__(RING_THREADS, 1)
k_gat_ring_bwd_src(const __grid_constant__ RingPlan P, const float* __restrict__ att,
                   const int32_t* __restrict__ rptr, const int32_t* __restrict__ rdst,
                   const int32_t* __restrict__ rslot, float slope, float invH, const float* __restrict__ dout,
                   int64_t ldo, const float* __restrict__ alpha, const float* __restrict__ ds,
                   const uint32_t* __restrict__ mask, __nv_bfloat16* __restrict__ dxl, int64_t ld, int n_src) {
    constexpr int NV = NVA + NVB;
    extern __shared__ __align__(128) uint8_t smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int n_left = (n_src > (int)blockIdx.x) ? (n_src - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
    if (n_left == 0) return;
    const Ring R = ring_setup(P, smem);
    if (warp >= RW_MAX) {                                     // producer warpgroup: warp 8 works, 9..11 only donate
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        if (warp == RW_MAX) src_producer(R, P, rptr, rdst, rslot, dout, ldo, alpha, ds, mask, n_src, lane);
        return;
    }
    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    const RingSeg sa = P.a[warp], sb = P.b[warp];
    const float c1 = 0.5f * (1.0f + slope), c2 = 0.5f * (1.0f - slope);
    uint32_t vmask = 0;
    float4 attv[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        const bool isA = k < NVA;
        const int u = lane + 32 * (isA ? k : k - NVA);
        attv[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (u < (isA ? sa.C4 : sb.C4)) {
            vmask |= 1u << k;
            attv[k] = __ldg(reinterpret_cast<const float4*>(att + (isA ? sa.off : sb.off) + 4 * u));
        }
    }
    // lanes past the end of a segment read on into the stage (the dout row is followed by 2 KB of sign words): what
    // they accumulate is never written
    const uint32_t gA = (uint32_t)(sa.out_off + 4 * lane) * 4u, gB = (uint32_t)(sb.out_off + 4 * lane) * 4u;
    float2 acc[NV][2], T[NV][2];
#pragma unroll
    for (int k = 0; k < NV; ++k) acc[k][0] = acc[k][1] = T[k][0] = T[k][1] = f2(0.f, 0.f);
    float SEA = 0.f, SEB = 0.f;
    bool first = true;                                       // the next edge is the first one of its source
    uint32_t stage = 0, round = 0;
    while (true) {
        mbar_wait(&R.full[stage], round & 1);
        const uint8_t* st = R.base + (size_t)stage * P.stage_bytes;
        const int4 meta = R.meta[stage];
        if (meta.x & E_DATA) {
            const float* al = reinterpret_cast<const float*>(st + P.e_al_off);
            const float* dsv = al + P.NCH;
            const uint32_t* mw = reinterpret_cast<const uint32_t*>(st + P.e_mask_off) + warp * 64 + lane;
            const float aA = al[sa.ch] * invH, dA = dsv[sa.ch];
            const float aB = (NVB > 0) ? al[sb.ch] * invH : 0.f, dB = (NVB > 0) ? dsv[sb.ch] : 0.f;
            const uint32_t w0 = mw[0];
            const uint32_t w1 = (NV > 8) ? mw[32] : 0u;
            const uint32_t qA = __float_as_uint(dA * c2), qB = __float_as_uint(dB * c2);
            if (first) {
                SEA = dA * c1;
                SEB = dB * c1;
                src_edge<NVA, NVB, true>(st, gA, gB, w0, w1, qA, qB, aA, aB, acc, T);
            } else {
                SEA = fmaf(dA, c1, SEA);
                SEB = fmaf(dB, c1, SEB);
                src_edge<NVA, NVB, false>(st, gA, gB, w0, w1, qA, qB, aA, aB, acc, T);
            }
            first = false;
        }
        __syncwarp();
        if (lane == 0) mbar_arrive(&R.empty[stage]);
        if (meta.x & E_LAST) {
            if (first) {                                     // a source without out edges: its row is zero
#pragma unroll
                for (int k = 0; k < NV; ++k) acc[k][0] = acc[k][1] = T[k][0] = T[k][1] = f2(0.f, 0.f);
                SEA = SEB = 0.f;
            }
            __nv_bfloat16* drow = dxl + (int64_t)meta.y * ld;
            int oA = sa.off + 4 * lane, oB = sb.off + 4 * lane;
            asm volatile("" : "+r"(oA), "+r"(oB));           // keep the two offsets in registers
            const float2 seA = f2(SEA, SEA), seB = f2(SEB, SEB);
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                const bool isA = k < NVA;
                const float2 o0 = __ffma2_rn(f2(attv[k].x, attv[k].y), __fadd2_rn(T[k][0], isA ? seA : seB), acc[k][0]);
                const float2 o1 = __ffma2_rn(f2(attv[k].z, attv[k].w), __fadd2_rn(T[k][1], isA ? seA : seB), acc[k][1]);
                const uint2 pk = pack_bf4(make_float4(o0.x, o0.y, o1.x, o1.y));
                if ((vmask >> k) & 1u) *reinterpret_cast<uint2*>(drow + (isA ? oA + 128 * k : oB + 128 * (k - NVA))) = pk;
            }
            first = true;
            if (--n_left == 0) break;
        }
        if (++stage == (uint32_t)P.n_stages) { stage = 0; ++round; }
    }
}

// ---------------------------------------------------------------------------------- host side
struct RingCfg { RingPlan P; int nva, nvb; size_t smem; };

// Builds the warp <-> segment plan; false when the shape is not one the ring kernels take.
// fwd: the per-warp score arrays (attention logits of one target) live in shared memory as well.
bool ring_plan(const int32_t* seg_host, int n_conv, int H, int64_t ld, int max_deg, bool fwd, RingCfg& cfg) {
    const int S = n_conv * H;
    if (S < RW_MAX || S > 2 * RW_MAX || ld % 8 != 0 || max_deg < 0) return false;   // all 8 consumer warps busy
    RingSeg segs[2 * RW_MAX];
    int NL = 0, ow = 0;
    for (int c = 0; c < n_conv; ++c) {
        const int off = seg_host[c * 4], C = seg_host[c * 4 + 1], oo = seg_host[c * 4 + 2];
        if (C <= 0 || C % 4 != 0 || off % 4 != 0 || oo % 4 != 0 || off < 0 || oo < 0) return false;
        for (int h = 0; h < H; ++h) segs[c * H + h] = RingSeg{off + h * C, C / 4, oo, c * H + h, (off + h * C) * 2};
        NL = off + H * C > NL ? off + H * C : NL;
        ow = oo + C > ow ? oo + C : ow;
    }
    if (NL % 8 != 0 || NL > ld) return false;                // the row copy is a multiple of 16 bytes inside the row
    for (int a = 1; a < S; ++a) {                            // stable insertion sort, longest segment first
        RingSeg key = segs[a];
        int b = a - 1;
        while (b >= 0 && segs[b].C4 < key.C4) { segs[b + 1] = segs[b]; --b; }
        segs[b + 1] = key;
    }
    RingPlan& P = cfg.P;
    P.n_warps = S < RW_MAX ? S : RW_MAX;
    int maxA = 0, maxB = 0;
    for (int w = 0; w < RW_MAX; ++w) {
        P.a[w] = RingSeg{0, 0, 0, 0, 0};
        P.b[w] = RingSeg{0, 0, 0, 0, 0};
        if (w >= P.n_warps) continue;
        P.a[w] = segs[w];
        if (S - 1 - w >= RW_MAX) P.b[w] = segs[S - 1 - w];
        maxA = P.a[w].C4 > maxA ? P.a[w].C4 : maxA;
        maxB = P.b[w].C4 > maxB ? P.b[w].C4 : maxB;
    }
    const int nva = (maxA + 31) / 32, nvb = (maxB + 31) / 32;
    cfg.nva = nva <= 4 ? 4 : (nva <= 12 ? 12 : 0);
    cfg.nvb = nvb == 0 ? 0 : (nvb <= 1 ? 1 : (nvb <= 4 ? 4 : -1));
    if (cfg.nva == 0 || cfg.nvb < 0 || (cfg.nva == 12 && cfg.nvb == 4)) return false;
    P.NL = NL;
    P.NCH = S;
    P.out_width = ow;
    P.max_deg1 = max_deg + 1;
    P.row_bytes = NL * 2;
    P.stage_bytes = (P.row_bytes + 127) / 128 * 128;
    P.sc_stride = fwd ? ((2 * P.max_deg1 * 4 + 15) / 16 * 16) : 0;
    const int lane_bytes = RW_MAX * (cfg.nva + cfg.nvb) * 512;
    const int fixed = P.sc_stride * P.n_warps + lane_bytes + 512;
    const int per_stage = P.stage_bytes + 16 + 16;          // data + meta + full / empty barriers
    const int ns = (SMEM_MAX - fixed) / per_stage;
    if (ns < 3) return false;
    P.n_stages = ns > RS_MAX ? RS_MAX : ns;
    P.meta_off = P.n_stages * P.stage_bytes;
    P.bar_off = P.meta_off + P.n_stages * 16;
    P.sc_off = (P.bar_off + P.n_stages * 16 + 127) / 128 * 128;
    P.lane_off = (P.sc_off + P.sc_stride * P.n_warps + 127) / 128 * 128;
    cfg.smem = (size_t)P.lane_off + (size_t)lane_bytes;
    P.full = (P.n_warps == RW_MAX) ? 1 : 0;
    for (int w = 0; w < P.n_warps && P.full; ++w) {
        if ((P.a[w].C4 + 31) / 32 != cfg.nva) P.full = 0;
        if (cfg.nvb > 0 && (P.b[w].C4 + 31) / 32 != cfg.nvb) P.full = 0;
    }
    return cfg.smem <= (size_t)SMEM_MAX;
}

template <int TAG, typename K>
bool ring_set_smem(K kernel) {
    static bool done[64] = {false};
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return false;
    if (done[dev]) return true;
    if (cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_MAX) != cudaSuccess) {
        cudaGetLastError();
        return false;
    }
    done[dev] = true;
    return true;
}

#define RING_DISPATCH(cfg, CALL)                                   \
    do {                                                           \
        if (cfg.nva == 4 && cfg.nvb == 0) { CALL(4, 0); }          \
        else if (cfg.nva == 4 && cfg.nvb == 1) { CALL(4, 1); }     \
        else if (cfg.nva == 4 && cfg.nvb == 4) { CALL(4, 4); }     \
        else if (cfg.nva == 12 && cfg.nvb == 0) { CALL(12, 0); }   \
        else { CALL(12, 1); }                                      \
    } while (0)
}  // namespace

// Returns NC_OK when the ring kernels ran, 1 when the shape is not theirs (caller falls back), < 0 on error.
int gat_ring_fwd(const void* xl, const void* xr, int64_t ld, const float* att, const float* bias,
                 const int32_t* rowptr, const int32_t* col, int32_t dst_off, int32_t n_dst, const int32_t* seg,
                 const int32_t* seg_host, int32_t n_conv, int32_t H, int32_t max_deg, float slope, float* out,
                 int64_t ldo, float* alpha, float* head_ws, void* stream) {
    if (head_ws == nullptr || n_dst <= 0 || !(slope > 0.f && slope < 1.f)) return 1;
    if ((uintptr_t)xl % 16 || (uintptr_t)xr % 16 || (uintptr_t)att % 16 || (uintptr_t)head_ws % 16 ||
        (uintptr_t)out % 16 || ldo % 4 || (bias && (uintptr_t)bias % 16))
        return 1;
    RingCfg cfg;
    if (!ring_plan(seg_host, n_conv, H, ld, max_deg, true, cfg)) return 1;
    const RingPlan& P = cfg.P;
    cudaStream_t s = (cudaStream_t)stream;
    const int grid = n_dst < nc_num_sms() ? n_dst : nc_num_sms();
#define CALL_F(A, B, F)                                                                                              \
    do {                                                                                                             \
        if (!ring_set_smem<(A) * 32 + (B) * 2 + (F)>(k_gat_ring_fwd<A, B, F != 0>)) return 1;                        \
        k_gat_ring_fwd<A, B, F != 0><<<grid, RING_THREADS, cfg.smem, s>>>(P, (const __nv_bfloat16*)xl,             \
            (const __nv_bfloat16*)xr, ld, att, rowptr, col, n_dst, dst_off, slope, head_ws, alpha);                  \
    } while (0)
#define CALL(A, B)                                                                                                   \
    do {                                                                                                             \
        if (P.full) CALL_F(A, B, 1); else CALL_F(A, B, 0);                                                           \
    } while (0)
    RING_DISPATCH(cfg, CALL);
#undef CALL
#undef CALL_F
    k_gat_head_mean<<<n_dst, 256, 0, s>>>(head_ws, ld, seg, n_conv, H, bias, out, ldo);
    nc_count_launch(2);
    NC_CHECK_LAUNCH("nc_gatv2_fwd (ring)");
    return NC_OK;
}

// bwd_ws: n_slots * RW_MAX * 256 bytes of sign bits (+ n_dst * n_conv * H floats that the first ring version kept
// D in; the size stays so that callers' workspaces do not change).
int64_t gat_ring_bwd_ws_bytes(int64_t n_slots, int64_t n_dst, int n_conv, int H) {
    return n_slots * RW_MAX * 256 + (n_dst * n_conv * H * 4 + 255) / 256 * 256;
}

int gat_ring_bwd(const void* xl, const void* xr, int64_t ld, const float* att, const int32_t* rowptr,
                 const int32_t* col, int32_t n_dst, int32_t n_src, const int32_t* rptr, const int32_t* rdst,
                 const int32_t* rslot, const int32_t* seg, const int32_t* seg_host, int32_t n_conv, int32_t H,
                 int32_t max_deg, float slope, const float* dout, int64_t ldo, const float* alpha, float* ds_ws,
                 void* dxl, void* dxr, float* datt, float* dbias, const float* head_ws, void* bwd_ws,
                 int64_t n_slots, void* stream) {
    if (head_ws == nullptr || bwd_ws == nullptr || n_dst <= 0 || !(slope > 0.f && slope < 1.f)) return 1;
    if ((uintptr_t)xl % 16 || (uintptr_t)xr % 16 || (uintptr_t)att % 16 || (uintptr_t)head_ws % 16 ||
        (uintptr_t)dout % 16 || ldo % 4 || (uintptr_t)dxl % 16 || (uintptr_t)dxr % 16 || (uintptr_t)bwd_ws % 256)
        return 1;
    RingCfg cfg;
    if (!ring_plan(seg_host, n_conv, H, ld, max_deg, false, cfg)) return 1;
    const RingPlan& P = cfg.P;
    cudaStream_t s = (cudaStream_t)stream;
    uint32_t* mask = (uint32_t*)bwd_ws;
    const float invH = 1.0f / (float)H;
    const int grid = n_dst < nc_num_sms() ? n_dst : nc_num_sms();
    // per-source kernel: its own ring, one stage per out edge = [dout row | sign words | alpha row | d score row]
    RingPlan P2 = P;
    P2.e_mask_off = (P.out_width * 4 + 127) / 128 * 128;
    P2.e_al_off = P2.e_mask_off + RW_MAX * 256;
    P2.stage_bytes = (P2.e_al_off + 2 * P.NCH * 4 + 127) / 128 * 128;
    const int tail = 512 * (cfg.nva > cfg.nvb ? cfg.nva : cfg.nvb) + 512;   // lanes past a segment's end read on
    int ns2 = (SMEM_MAX - 1024 - tail) / (P2.stage_bytes + 32);
    ns2 = ns2 > RS_MAX ? RS_MAX : ns2;
    if (ns2 < 3 || P.NCH % 4 != 0 || (uintptr_t)alpha % 16 || (uintptr_t)ds_ws % 16) return 1;
    P2.n_stages = ns2;
    P2.meta_off = ns2 * P2.stage_bytes;
    P2.bar_off = P2.meta_off + ns2 * 16;
    const size_t src_smem = (size_t)P2.bar_off + (size_t)ns2 * 16 + (size_t)tail;
#define CALL_F(A, B, F)                                                                                              \
    do {                                                                                                             \
        if (!ring_set_smem<1024 + (A) * 32 + (B) * 2 + (F)>(k_gat_ring_bwd<A, B, F != 0>)) return 1;                 \
        k_gat_ring_bwd<A, B, F != 0><<<grid, RING_THREADS, cfg.smem, s>>>(P, (const __nv_bfloat16*)xl,             \
            (const __nv_bfloat16*)xr, ld, att, rowptr, col, n_dst, slope, invH, dout, ldo, alpha, head_ws, ds_ws,    \
            (__nv_bfloat16*)dxr, datt, mask);                                                                        \
    } while (0)
#define CALL(A, B)                                                                                                   \
    do {                                                                                                             \
        if (!ring_set_smem<2048 + (A) * 32 + (B) * 2>(k_gat_ring_bwd_src<A, B>)) return 1;                           \
        if (P.full) CALL_F(A, B, 1); else CALL_F(A, B, 0);                                                           \
        const int gsrc_ = n_src < nc_num_sms() ? n_src : nc_num_sms();                                               \
        k_gat_ring_bwd_src<A, B><<<gsrc_, RING_THREADS, src_smem, s>>>(                                              \
            P2, att, rptr, rdst, rslot, slope, invH, dout, ldo, alpha, ds_ws, mask, (__nv_bfloat16*)dxl, ld, n_src); \
    } while (0)
    RING_DISPATCH(cfg, CALL);
#undef CALL
#undef CALL_F
    nc_count_launch(2);
    NC_CHECK_LAUNCH("nc_gatv2_bwd (ring)");
    if (dbias) {
        int rc = nc_col_sum(dout, NC_F32, ldo, n_dst, P.out_width, dbias, stream);
        if (rc) return rc;
    }
    return NC_OK;
}
