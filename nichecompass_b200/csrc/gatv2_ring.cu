// GATv2 edge kernels, second generation ("ring" kernels): bf16 rows, hot-path shapes.
// Same arithmetic as gatv2.cu (PyG GATv2Conv message passing, nn/encoders.py:124-140,197-207, SURVEY.md
// App. A.1), different machine mapping -- built for HBM throughput:
//
//   * persistent CTAs, one per SM, 8 warps that never synchronise with each other.  Warp w owns one
//     or two (conv, head) segments of the x_l / x_r rows (a long one and a short one: the plan below
//     pairs them so that every warp touches the same number of bytes) for every target of its CTA;
//   * each warp runs its own software pipeline: a private shared-memory ring of n_stages slots that
//     holds ITS column slice of the next rows ([x_r slice | every neighbour's x_l slice | the self
//     row's slice] per target), filled by cp.async.bulk + mbarrier; the warp refills a slot right
//     after it has consumed it.  8 warps x 8 slots x 3.2 KB = 205 KB in flight per SM at C2 (before: 4
//     rows of 12 KB behind a __syncthreads per neighbour);
//   * scores, online softmax and the alpha-weighted sum of a head live inside one warp (shuffles only);
//   * forward writes the per-head outputs O (fp32) and the attention coefficients; the head mean
//     + bias is a second, tiny kernel;
//   * backward, per target: D_h = sum_t alpha_t dalpha_t = <g, O_h> / H is known BEFORE the
//     neighbours stream in (k_gat_D), so d score, d x_r, d att and the LeakyReLU sign bits of every
//     (edge, element) are produced in ONE pass over the rows; the per-source pass then needs only the
//     sign bits (1 bit per element instead of re-reading x_l[j] + x_r[i]): d x_l[j] = sum over out
//     edges of alpha/H * g_i + d score * att * (bit ? 1 : slope).
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
__device__ __forceinline__ bool issue_next(Cursor& c, const Ring& R, const RingPlan& P, uint32_t ps,
                                           const __nv_bfloat16* __restrict__ xl, const __nv_bfloat16* __restrict__ xr,
                                           int64_t ld, const int32_t* __restrict__ rowptr,
                                           const int32_t* __restrict__ col, int n_dst, int dst_off, int lane) {
    const __nv_bfloat16* src;
    int4 meta;
    while (true) {
        if (c.ir >= n_dst) return false;
        const int i = c.ir + dst_off, deg = c.e1 - c.e0;
        if (c.t < 0) {                                               // the target's own x_r row opens its sequence
            src = xr + (int64_t)c.ir * ld;
            meta = make_int4(K_XR, c.ir, c.e0 + i, deg + 1);
            c.t = 0;
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
                                              int n_dst, int dst_off, int lane) {
    Cursor cur;
    cur_init(cur, rowptr, col, n_dst, dst_off, lane);
    uint32_t stage = 0, round = 0;
    while (true) {
        mbar_wait(&R.empty[stage], (round & 1) ^ 1);                 // passes at once in the first round
        if (!issue_next(cur, R, P, stage, xl, xr, ld, rowptr, col, n_dst, dst_off, lane)) break;
        if (++stage == (uint32_t)P.n_stages) { stage = 0; ++round; }
    }
}

// ---------------------------------------------------------------------------------- forward
// Instruction diet of the inner loop (the first version of this kernel issued 15 instructions per row element and
// was bound by issue slots -- 2 warps per scheduler, IPC 0.5 -- not by memory; ncu: profiles/r2_gatv2_*):
//   * leaky_relu(v) = c1 v + c2 |v| with c1 = (1 + slope) / 2, c2 = (1 - slope) / 2, so the score is
//     c1 <att, x_l[j]> + c2 <att, |x_l[j] + x_r[i]|> (+ c1 <att, x_r[i]>, which is the same for every neighbour of i
//     and cancels in the softmax): one add and two fused multiply-adds per element (|.| is an operand modifier),
//     instead of add, multiply, max, multiply-add;
//   * x_r[i] and att live in registers as fp32 for the whole target (no re-unpacking per row);
//   * lanes past the end of a segment read the first element of the slot (valid data) with att = 0 instead of
//     branching around every 128-bit chunk.
template <int NVA, int NVB>
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
    float4 attv[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        const bool isA = k < NVA;
        const int kk = isA ? k : k - NVA;
        const int u = lane + 32 * kk;
        attv[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (u < (isA ? sa.C4 : sb.C4)) {
            vmask |= 1u << k;
            attv[k] = __ldg(reinterpret_cast<const float4*>(att + (isA ? sa.off : sb.off) + 4 * u));
        }
    }
#define BOFF(k) ((((vmask) >> (k)) & 1u) ? ((k) < NVA ? offA + 256u * (uint32_t)(k) : offB + 256u * (uint32_t)((k) - NVA)) : offS)
    float* scA = reinterpret_cast<float*>(smem + P.sc_off + (size_t)warp * P.sc_stride);
    float* scB = scA + P.max_deg1;
    if (alpha) {
        for (int t = lane; t < P.max_deg1; t += 32) { scA[t] = -INFINITY; scB[t] = -INFINITY; }
        __syncwarp();
    }
    float4 xrv[NV], acc[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) { xrv[k] = make_float4(0.f, 0.f, 0.f, 0.f); acc[k] = make_float4(0.f, 0.f, 0.f, 0.f); }
    float mA = -INFINITY, lA = 0.f, mB = -INFINITY, lB = 0.f;
    int slot0 = 0, nE = 0;
    uint32_t stage = 0, round = 0;
    while (true) {
        mbar_wait(&R.full[stage], round & 1);
        const uint8_t* row = R.base + (size_t)stage * P.stage_bytes;
        const int4 meta = R.meta[stage];
        uint2 raw[NV];
#pragma unroll
        for (int k = 0; k < NV; ++k) raw[k] = *reinterpret_cast<const uint2*>(row + BOFF(k));
        __syncwarp();                                        // the row now lives in registers: release the stage
        if (lane == 0) mbar_arrive(&R.empty[stage]);
        if ((meta.x & 255) == K_XR) {
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                xrv[k] = bf4(raw[k]);
                acc[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
            mA = mB = -INFINITY;
            lA = lB = 0.f;
            slot0 = meta.z;
            nE = min(meta.w, P.max_deg1);
        } else {
            // (the row stays packed in `raw` and is unpacked again for the accumulation: 26 registers instead of 52)
            float s1A = 0.f, s2A = 0.f, s1B = 0.f, s2B = 0.f;
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                const float4 f = bf4(raw[k]);
                const float4 a = attv[k], x = xrv[k];
                float t1 = a.x * f.x, t2 = a.x * fabsf(f.x + x.x);
                t1 = fmaf(a.y, f.y, t1); t2 = fmaf(a.y, fabsf(f.y + x.y), t2);
                t1 = fmaf(a.z, f.z, t1); t2 = fmaf(a.z, fabsf(f.z + x.z), t2);
                t1 = fmaf(a.w, f.w, t1); t2 = fmaf(a.w, fabsf(f.w + x.w), t2);
                if (k < NVA) { s1A += t1; s2A += t2; } else { s1B += t1; s2B += t2; }
            }
            float pA = warp_sum(fmaf(c1, s1A, c2 * s2A));
            float pB = 0.f;
            if (NVB > 0) pB = warp_sum(fmaf(c1, s1B, c2 * s2B));
            const int t = meta.w;
            if (alpha && lane == 0 && t < P.max_deg1) { scA[t] = pA; scB[t] = pB; }
            {
                const float mn = fmaxf(mA, pA), corr = __expf(mA - mn), w = __expf(pA - mn);
                lA = lA * corr + w;
                mA = mn;
                if (corr != 1.f) {
#pragma unroll
                    for (int k = 0; k < NVA; ++k) { acc[k].x *= corr; acc[k].y *= corr; acc[k].z *= corr; acc[k].w *= corr; }
                }
#pragma unroll
                for (int k = 0; k < NVA; ++k) {
                    const float4 f = bf4(raw[k]);
                    acc[k].x = fmaf(w, f.x, acc[k].x); acc[k].y = fmaf(w, f.y, acc[k].y);
                    acc[k].z = fmaf(w, f.z, acc[k].z); acc[k].w = fmaf(w, f.w, acc[k].w);
                }
            }
            if (NVB > 0) {
                const float mn = fmaxf(mB, pB), corr = __expf(mB - mn), w = __expf(pB - mn);
                lB = lB * corr + w;
                mB = mn;
#pragma unroll
                for (int k = NVA; k < NV; ++k) {
                    const float4 f = bf4(raw[k]);
                    acc[k].x = fmaf(w, f.x, acc[k].x * corr); acc[k].y = fmaf(w, f.y, acc[k].y * corr);
                    acc[k].z = fmaf(w, f.z, acc[k].z * corr); acc[k].w = fmaf(w, f.w, acc[k].w * corr);
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
                            make_float4(acc[k].x * inv, acc[k].y * inv, acc[k].z * inv, acc[k].w * inv);
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

#undef BOFF

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

// D[ir, c * H + h] = <dout[ir, out_off_c : + C_c], O[ir, off_c + h * C_c : + C_c]> / H  (= sum_t alpha_t dalpha_t)
__global__ void __launch_bounds__(256)
k_gat_D(const float* __restrict__ head_out, int64_t ld, const float* __restrict__ dout, int64_t ldo,
        const int32_t* __restrict__ seg, int n_conv, int H, float* __restrict__ Dbuf) {
    const int ir = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float invH = 1.0f / (float)H;
    for (int ch = warp; ch < n_conv * H; ch += 8) {
        const int c = ch / H, h = ch - c * H;
        const int off = seg[c * 4], C = seg[c * 4 + 1], oo = seg[c * 4 + 2];
        const float* o = head_out + (int64_t)ir * ld + off + h * C;
        const float* g = dout + (int64_t)ir * ldo + oo;
        float p = 0.f;
        for (int k = lane * 4; k < C; k += 128) {
            const float4 a = *reinterpret_cast<const float4*>(o + k);
            const float4 b = __ldg(reinterpret_cast<const float4*>(g + k));
            p += a.x * b.x + a.y * b.y + a.z * b.z + a.w * b.w;
        }
        p = warp_sum(p);
        if (lane == 0) Dbuf[(int64_t)ir * n_conv * H + ch] = p * invH;
    }
}

// ---------------------------------------------------------------------------------- backward, per target
template <int NVA, int NVB>
__global__ void __launch_bounds__(RING_THREADS, 1)
k_gat_ring_bwd(const __grid_constant__ RingPlan P, const __nv_bfloat16* __restrict__ xl,
               const __nv_bfloat16* __restrict__ xr, int64_t ld, const float* __restrict__ att,
               const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col, int n_dst, float slope,
               float invH, const float* __restrict__ dout, int64_t ldo, const float* __restrict__ alpha,
               const float* __restrict__ Dbuf, float* __restrict__ ds, __nv_bfloat16* __restrict__ dxr,
               float* __restrict__ datt, uint32_t* __restrict__ mask) {
    constexpr int NV = NVA + NVB;
    extern __shared__ __align__(128) uint8_t smem[];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int n_left = (n_dst > (int)blockIdx.x) ? (n_dst - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;
    if (n_left == 0) return;
    const Ring R = ring_setup(P, smem);
    if (warp >= RW_MAX) {                                     // producer warpgroup: warp 8 works, 9..11 only donate
        asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
        if (warp == RW_MAX) producer_loop(R, P, xl, xr, ld, rowptr, col, n_dst, 0, lane);
        return;
    }
    asm volatile("setmaxnreg.inc.sync.aligned.u32 232;");
    const RingSeg sa = P.a[warp], sb = P.b[warp];
    // lanes past the end of a segment read the first element of the slot (valid data) and carry g = 0: no branch
    // around the 128-bit chunks (same diet as the forward kernel)
    uint32_t vmask = 0;
    const uint32_t offA = (uint32_t)(sa.st_off + 8 * lane), offB = (uint32_t)(sb.st_off + 8 * lane), offS = (uint32_t)sa.st_off;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        const bool isA = k < NVA;
        const int u = lane + 32 * (isA ? k : k - NVA);
        if (u < (isA ? sa.C4 : sb.C4)) vmask |= 1u << k;
    }
#define BOFF(k) ((((vmask) >> (k)) & 1u) ? ((k) < NVA ? offA + 256u * (uint32_t)(k) : offB + 256u * (uint32_t)((k) - NVA)) : offS)
    float4 gv[NV], dxa[NV], dat[NV];
    uint2 xrv[NV];
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        gv[k] = dxa[k] = dat[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        xrv[k] = make_uint2(0u, 0u);
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
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                const bool isA = k < NVA;
                const int kk = isA ? k : k - NVA;
                xrv[k] = *reinterpret_cast<const uint2*>(row + BOFF(k));
                gv[k] = make_float4(0.f, 0.f, 0.f, 0.f);
                if ((vmask >> k) & 1u)
                    gv[k] = __ldg(reinterpret_cast<const float4*>(g + (isA ? sa.out_off : sb.out_off) + 4 * (lane + 32 * kk)));
                dxa[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
            DA = __ldg(Dbuf + (int64_t)ir * P.NCH + sa.ch);
            DB = (sb.C4 > 0) ? __ldg(Dbuf + (int64_t)ir * P.NCH + sb.ch) : 0.f;
            alA = alB = 0.f;
            if (lane < meta.w) {
                alA = __ldg(alpha + (int64_t)(slot0 + lane) * P.NCH + sa.ch);
                if (sb.C4 > 0) alB = __ldg(alpha + (int64_t)(slot0 + lane) * P.NCH + sb.ch);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&R.empty[stage]);
        } else {
            // sweep 1: d alpha = <g, x_l[j, head]> / H
            float pA = 0.f, pB = 0.f;
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                const float4 f = bf4(*reinterpret_cast<const uint2*>(row + BOFF(k)));
                const float s = fmaf(f.x, gv[k].x, fmaf(f.y, gv[k].y, fmaf(f.z, gv[k].z, f.w * gv[k].w)));
                if (k < NVA) pA += s; else pB += s;
            }
            pA = warp_sum(pA);
            if (NVB > 0) pB = warp_sum(pB);
            const int slot = meta.z, t = meta.w;
            float aA, aB = 0.f;
            if (t < 32) {
                aA = __shfl_sync(0xffffffffu, alA, t);
                aB = __shfl_sync(0xffffffffu, alB, t);
            } else {
                aA = __ldg(alpha + (int64_t)slot * P.NCH + sa.ch);
                if (sb.C4 > 0) aB = __ldg(alpha + (int64_t)slot * P.NCH + sb.ch);
            }
            const float dsA = aA * (pA * invH - DA);
            const float dsB = aB * (pB * invH - DB);
            if (lane == 0) {
                ds[(int64_t)slot * P.NCH + sa.ch] = dsA;
                if (sb.C4 > 0) ds[(int64_t)slot * P.NCH + sb.ch] = dsB;
            }
            // sweep 2: d x_r, d att and the LeakyReLU sign bits of this edge
            const float dsAs = dsA * slope, dsBs = dsB * slope;
            uint32_t w0 = 0u, w1 = 0u;
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                {
                    const float4 f = bf4(*reinterpret_cast<const uint2*>(row + BOFF(k))), x = bf4(xrv[k]);
                    const float d1 = k < NVA ? dsA : dsB, d0 = k < NVA ? dsAs : dsBs;
                    uint32_t bits = 0u;
                    float v, tt;
                    v = f.x + x.x; tt = v > 0.f ? d1 : d0; bits |= v > 0.f ? 1u : 0u; dxa[k].x += tt; dat[k].x += tt * v;
                    v = f.y + x.y; tt = v > 0.f ? d1 : d0; bits |= v > 0.f ? 2u : 0u; dxa[k].y += tt; dat[k].y += tt * v;
                    v = f.z + x.z; tt = v > 0.f ? d1 : d0; bits |= v > 0.f ? 4u : 0u; dxa[k].z += tt; dat[k].z += tt * v;
                    v = f.w + x.w; tt = v > 0.f ? d1 : d0; bits |= v > 0.f ? 8u : 0u; dxa[k].w += tt; dat[k].w += tt * v;
                    if (k < 8) w0 |= bits << (4 * (k & 7)); else w1 |= bits << (4 * (k & 7));
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&R.empty[stage]);
            uint32_t* mrow = mask + ((int64_t)slot * RW_MAX + warp) * 64;
            mrow[lane] = w0;
            if (NV > 8) mrow[32 + lane] = w1;
            if (meta.x & K_LAST) {
                const int ir = meta.y;
                __nv_bfloat16* drow = dxr + (int64_t)ir * ld;
#pragma unroll
                for (int k = 0; k < NV; ++k) {
                    if ((vmask >> k) & 1u) {
                        const bool isA = k < NVA;
                        const int e = (isA ? sa.off : sb.off) + 4 * (lane + 32 * (isA ? k : k - NVA));
                        const float4 a = __ldg(reinterpret_cast<const float4*>(att + e));
                        *reinterpret_cast<uint2*>(drow + e) =
                            pack_bf4(make_float4(a.x * dxa[k].x, a.y * dxa[k].y, a.z * dxa[k].z, a.w * dxa[k].w));
                    }
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
            atomicAdd(da, dat[k].x); atomicAdd(da + 1, dat[k].y); atomicAdd(da + 2, dat[k].z); atomicAdd(da + 3, dat[k].w);
        }
    }
}

#undef BOFF

// ---------------------------------------------------------------------------------- backward, per source
// d x_l[j] = sum over the out edges (j -> i, slot) of  alpha / H * g_i  +  d score * att * (bit ? 1 : slope).
// Same warp <-> segment plan as the per-target kernel (the sign bits are stored per warp and lane).
template <int NVA, int NVB>
__global__ void __launch_bounds__(RW_MAX * 32, 1)
k_gat_ring_bwd_src(const __grid_constant__ RingPlan P, const float* __restrict__ att,
                   const int32_t* __restrict__ rptr, const int32_t* __restrict__ rdst,
                   const int32_t* __restrict__ rslot, float slope, float invH, const float* __restrict__ dout,
                   int64_t ldo, const float* __restrict__ alpha, const float* __restrict__ ds,
                   const uint32_t* __restrict__ mask, __nv_bfloat16* __restrict__ dxl, int64_t ld, int n_src) {
    // persistent CTAs striding over the sources: att stays in registers (it was re-read from L2 for every edge),
    // the reverse-CSR range of the NEXT source is fetched while the current one is reduced
    constexpr int NV = NVA + NVB;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp >= P.n_warps) return;
    const RingSeg sa = P.a[warp], sb = P.b[warp];
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
    int j = blockIdx.x;
    int p0 = 0, p1 = 0;
    if (j < n_src) { p0 = __ldg(rptr + j); p1 = __ldg(rptr + j + 1); }
    for (; j < n_src; j += (int)gridDim.x) {
        const int jn = j + (int)gridDim.x;
        int np0 = 0, np1 = 0;
        if (jn < n_src) { np0 = __ldg(rptr + jn); np1 = __ldg(rptr + jn + 1); }
        float4 acc[NV];
#pragma unroll
        for (int k = 0; k < NV; ++k) acc[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int p = p0; p < p1; ++p) {
            const int i = __ldg(rdst + p), slot = __ldg(rslot + p);
            const float* al = alpha + (int64_t)slot * P.NCH;
            const float* dsl = ds + (int64_t)slot * P.NCH;
            const float aA = __ldg(al + sa.ch) * invH, dA = __ldg(dsl + sa.ch);
            const float aB = (sb.C4 > 0) ? __ldg(al + sb.ch) * invH : 0.f, dB = (sb.C4 > 0) ? __ldg(dsl + sb.ch) : 0.f;
            const uint32_t* mrow = mask + ((int64_t)slot * RW_MAX + warp) * 64;
            const uint32_t w0 = __ldg(mrow + lane);
            const uint32_t w1 = (NV > 8) ? __ldg(mrow + 32 + lane) : 0u;
            const float* g = dout + (int64_t)i * ldo;
            const float dAs = dA * slope, dBs = dB * slope;
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                const bool isA = k < NVA;
                const int u = lane + 32 * (isA ? k : k - NVA);
                float4 gg = make_float4(0.f, 0.f, 0.f, 0.f);
                if ((vmask >> k) & 1u)
                    gg = __ldg(reinterpret_cast<const float4*>(g + (isA ? sa.out_off : sb.out_off) + 4 * u));
                const float4 a = attv[k];
                const uint32_t bits = ((k < 8) ? w0 : w1) >> (4 * (k & 7));
                const float aw = isA ? aA : aB, d1 = isA ? dA : dB, d0 = isA ? dAs : dBs;
                acc[k].x = fmaf(aw, gg.x, fmaf((bits & 1u) ? d1 : d0, a.x, acc[k].x));
                acc[k].y = fmaf(aw, gg.y, fmaf((bits & 2u) ? d1 : d0, a.y, acc[k].y));
                acc[k].z = fmaf(aw, gg.z, fmaf((bits & 4u) ? d1 : d0, a.z, acc[k].z));
                acc[k].w = fmaf(aw, gg.w, fmaf((bits & 8u) ? d1 : d0, a.w, acc[k].w));
            }
        }
        __nv_bfloat16* drow = dxl + (int64_t)j * ld;
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            if ((vmask >> k) & 1u) {
                const bool isA = k < NVA;
                const int e = (isA ? sa.off : sb.off) + 4 * (lane + 32 * (isA ? k : k - NVA));
                *reinterpret_cast<uint2*>(drow + e) = pack_bf4(acc[k]);
            }
        }
        p0 = np0; p1 = np1;
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
    const int fixed = P.sc_stride * P.n_warps + 256;
    const int per_stage = P.stage_bytes + 16 + 16;          // data + meta + full / empty barriers
    const int ns = (SMEM_MAX - fixed) / per_stage;
    if (ns < 3) return false;
    P.n_stages = ns > RS_MAX ? RS_MAX : ns;
    P.meta_off = P.n_stages * P.stage_bytes;
    P.bar_off = P.meta_off + P.n_stages * 16;
    P.sc_off = (P.bar_off + P.n_stages * 16 + 127) / 128 * 128;
    cfg.smem = (size_t)P.sc_off + (size_t)P.sc_stride * P.n_warps;
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
#define CALL(A, B)                                                                                                   \
    do {                                                                                                             \
        if (!ring_set_smem<(A) * 16 + (B)>(k_gat_ring_fwd<A, B>)) return 1;                                                          \
        k_gat_ring_fwd<A, B><<<grid, RING_THREADS, cfg.smem, s>>>(P, (const __nv_bfloat16*)xl,                     \
            (const __nv_bfloat16*)xr, ld, att, rowptr, col, n_dst, dst_off, slope, head_ws, alpha);                  \
    } while (0)
    RING_DISPATCH(cfg, CALL);
#undef CALL
    k_gat_head_mean<<<n_dst, 256, 0, s>>>(head_ws, ld, seg, n_conv, H, bias, out, ldo);
    nc_count_launch(2);
    NC_CHECK_LAUNCH("nc_gatv2_fwd (ring)");
    return NC_OK;
}

// bwd_ws: n_slots * RW_MAX * 256 bytes of sign bits, then n_dst * n_conv * H floats (D).
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
    float* Dbuf = (float*)((uint8_t*)bwd_ws + n_slots * RW_MAX * 256);
    const float invH = 1.0f / (float)H;
    k_gat_D<<<n_dst, 256, 0, s>>>(head_ws, ld, dout, ldo, seg, n_conv, H, Dbuf);
    const int grid = n_dst < nc_num_sms() ? n_dst : nc_num_sms();
#define CALL(A, B)                                                                                                   \
    do {                                                                                                             \
        if (!ring_set_smem<1024 + (A) * 16 + (B)>(k_gat_ring_bwd<A, B>)) return 1;                                                          \
        k_gat_ring_bwd<A, B><<<grid, RING_THREADS, cfg.smem, s>>>(P, (const __nv_bfloat16*)xl,                     \
            (const __nv_bfloat16*)xr, ld, att, rowptr, col, n_dst, slope, invH, dout, ldo, alpha, Dbuf, ds_ws,       \
            (__nv_bfloat16*)dxr, datt, mask);                                                                        \
        int occ_ = 1;                                                                                                \
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ_, k_gat_ring_bwd_src<A, B>, P.n_warps * 32, 0) !=     \
            cudaSuccess || occ_ < 1) { cudaGetLastError(); occ_ = 1; }                                               \
        const int gsrc_ = n_src < occ_ * nc_num_sms() ? n_src : occ_ * nc_num_sms();   /* one resident wave */      \
        k_gat_ring_bwd_src<A, B><<<gsrc_, P.n_warps * 32, 0, s>>>(                                                   \
            P, att, rptr, rdst, rslot, slope, invH, dout, ldo, alpha, ds_ws, mask, (__nv_bfloat16*)dxl, ld, n_src);  \
    } while (0)
    RING_DISPATCH(cfg, CALL);
#undef CALL
    nc_count_launch(3);
    NC_CHECK_LAUNCH("nc_gatv2_bwd (ring)");
    if (dbias) {
        int rc = nc_col_sum(dout, NC_F32, ldo, n_dst, P.out_width, dbias, stream);
        if (rc) return rc;
    }
    return NC_OK;
}
