// bf16 GEMM on the 5th-gen tensor cores: TMA (cp.async.bulk.tensor) -> 128B-swizzled smem ring
// -> tcgen05.mma (accumulators in TMEM, double buffered) -> tcgen05.ld epilogue with fused
// bias / ReLU / gene-program mask -> 128B-swizzled smem staging -> TMA tensor store
// (cp.async.bulk.tensor) or, for accumulate / split-K, TMA tensor reduce-add
// (cp.reduce.async.bulk.tensor .add), so every global write is a full coalesced line and the
// bounds are clipped by the tensor map.  Persistent, one CTA per SM, warp-specialised:
//   warp 0 : TMA producer          warp 1 : TMEM allocator + MMA issuer (one elected lane)
//   warps 2-5 : epilogue (TMEM lane quarter = warp_id % 4), two 4 KB staging buffers each
// Work item = (output tile, K split); small-tile-count / deep-K problems are split along K and
// reduced in L2 by the TMA reduce-add.
//   C[M,N] (+)= act(opA(A) * opB(B) + bias) (* mask)
// A: [M,K] (K-major) or, trans_a, [K,M] (MN-major).  B: [N,K] (K-major) or, trans_b, [K,N].
// Replaces the cuBLAS call sites nn/encoders.py:179 (fc_l1), PyG GATv2Conv lin_l/lin_r,
// GCNConv lin, nn/layercomponents.py:77-80 (MaskedLinear) and all their backward GEMMs.
#include <cuda.h>
#include <cstdlib>

#include "common.cuh"

extern void nc_count_launch(int n);

namespace {

constexpr int BM = 128;
constexpr int BK = 64;                 // 64 bf16 = 128 B = one swizzle-128B atom
constexpr int UMMA_K = 16;
constexpr int NUM_THREADS = 192;
constexpr int A_STAGE_BYTES = BM * BK * 2;

// NCTA = 2: one 256 x BN output tile per CTA PAIR (thread-block cluster of two SMs of a TPC), computed by
// tcgen05.mma.cta_group::2 issued by the leader CTA: each CTA stages its own 128 rows of A and HALF of the
// B tile (BN/2 rows), so a k-block costs 32 KB of L2->smem traffic per SM instead of 48 KB and the ring is
// six stages deep; each CTA's TMEM receives its 128 accumulator rows and runs its own epilogue.
template <int BN, int NCTA = 1> struct Cfg {
    static constexpr int B_ROWS = BN / NCTA;       // rows of the B tile this CTA stages
    static constexpr int B_STAGE_BYTES = B_ROWS * BK * 2;
    static constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;
    static constexpr int STAGES = (B_ROWS == 256) ? 4 : 6;
    static constexpr int TMEM_COLS = 2 * BN;       // two accumulator stages (power of two)
    static constexpr int STAGING_BYTES = 4 /*warps*/ * 2 /*buffers*/ * 4096;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + STAGING_BYTES + 1024 /*align*/ + 256 /*barriers*/;
};

struct Epilogue {
    const float* bias;
    const float* mask;
    int64_t ldmask;
    int c_bf16;
    int relu;
    int reduce;        // TMA reduce-add instead of store (accumulate into C, or split-K partials)
    int bias_vec;      // bias is 16-byte aligned
    int mask_vec;      // mask rows are 16-byte aligned
    int splits;        // K splits per output tile
    int band;          // tile rows per raster band (see item_to_tile)
};

// ------------------------------------------------------------------------- PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra.uni WAIT_DONE;\n\t"
        "bra.uni WAIT_LOOP;\n\t"
        "WAIT_DONE:\n\t"
        "}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* tm, uint64_t* bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(dst)), "l"((uint64_t)tm), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
        : "memory");
}
// CTA-pair variants: the load lands in this CTA's smem, its bytes are counted on the LEADER's barrier
__device__ __forceinline__ void tma_load_2d_pair(void* dst, const CUtensorMap* tm, uint32_t leader_bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(dst)), "l"((uint64_t)tm), "r"(leader_bar), "r"(c0), "r"(c1)
        : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ uint32_t map_to_cta(uint32_t saddr, uint32_t rank) {      // shared::cta -> shared::cluster
    uint32_t r;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(saddr), "r"(rank));
    return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
    asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void umma_commit_pair(uint64_t* bar) {      // arrives on the same barrier of both CTAs
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void umma_bf16_pair(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum)
        : "memory");
}
// Raster order of the persistent schedule.  Concurrent work items (consecutive item ids) cover a BAND of
// `band` tile rows x all tile columns, so one wave touches band/num_m of A and all of B: an operand larger
// than L2 is streamed from HBM once instead of once per tile column.
__device__ __forceinline__ void item_to_tile(int item, int splits, int num_m, int num_n, int band, int& tm, int& tn,
                                             int& split) {
    const int tile = item / splits;
    split = item - tile * splits;
    const int per_band = band * num_n;
    const int b = tile / per_band, r = tile - b * per_band;
    const int rows = min(band, num_m - b * band);
    tn = r / rows;
    tm = b * band + (r - tn * rows);
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* tm, const void* src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"((uint64_t)tm), "r"(smem_u32(src)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_reduce_add_2d(const CUtensorMap* tm, const void* src, int c0, int c1) {
    asm volatile("cp.reduce.async.bulk.tensor.2d.global.shared::cta.add.tile.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"((uint64_t)tm), "r"(smem_u32(src)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_read_1() { asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void st_shared_v4(void* p, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(smem_u32(p)), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t"
        "}" ::"r"(tmem_d),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum)
        : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr)
        : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// smem matrix descriptor (sm_100 UMMA): start>>4 [0,14) | LBO>>4 [16,30) | SBO>>4 [32,46) |
// version=1 [46,48) | layout SWIZZLE_128B=2 [61,64)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)2 << 61;
    return d;
}

// instruction descriptor: D=f32 [4,6)=1, A=bf16 [7,10)=1, B=bf16 [10,13)=1, a_major bit15,
// b_major bit16 (1 = MN-major), N>>3 [17,23), M>>4 [24,29)
__host__ __device__ constexpr uint32_t make_idesc(int M, int N, bool a_mn, bool b_mn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((a_mn ? 1u : 0u) << 15) | ((b_mn ? 1u : 0u) << 16) |
           ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// ------------------------------------------------------------------------- epilogue helpers
// bias / ReLU / mask on the four accumulator values of columns [col, col+4) of one row.
__device__ __forceinline__ void epi_apply4(const Epilogue& e, float (&v)[4], const float* mrow, int col, int N,
                                           bool add_bias) {
    if (e.bias && add_bias) {
        if (e.bias_vec && col + 3 < N) {
            const float4 b = __ldg(reinterpret_cast<const float4*>(e.bias + col));
            v[0] += b.x; v[1] += b.y; v[2] += b.z; v[3] += b.w;
        } else {
#pragma unroll
            for (int t = 0; t < 4; ++t) if (col + t < N) v[t] += __ldg(e.bias + col + t);
        }
    }
    if (e.relu) {
#pragma unroll
        for (int t = 0; t < 4; ++t) v[t] = fmaxf(v[t], 0.f);
    }
    if (mrow) {
        if (e.mask_vec && col + 3 < N) {
            const float4 m = __ldg(reinterpret_cast<const float4*>(mrow + col));
            v[0] *= m.x; v[1] *= m.y; v[2] *= m.z; v[3] *= m.w;
        } else {
#pragma unroll
            for (int t = 0; t < 4; ++t) if (col + t < N) v[t] *= __ldg(mrow + col + t);
        }
    }
}

// One warp moves its 32 accumulator rows x (32 fp32 | 64 bf16) columns: registers -> 128B-swizzled
// staging buffer (row = lane; 16-byte unit j of a row lives at unit j ^ (row & 7), the layout a
// SWIZZLE_128B tensor map expects, and conflict-free for st.shared.v4) -> TMA store / reduce-add.
__device__ __forceinline__ void epi_flush(const Epilogue& e, const CUtensorMap* tmC, uint8_t* buf, int lane,
                                          const uint32_t (&packed)[32], int col0, int row0) {
    if (lane == 0) bulk_wait_read_1();           // the store issued from this buffer two chunks ago has read it
    __syncwarp();
    uint8_t* rowp = buf + lane * 128;
#pragma unroll
    for (int j = 0; j < 8; ++j)
        st_shared_v4(rowp + ((j ^ (lane & 7)) << 4), packed[4 * j], packed[4 * j + 1], packed[4 * j + 2],
                     packed[4 * j + 3]);
    fence_async_smem();
    __syncwarp();
    if (lane == 0) {
        if (e.reduce) tma_reduce_add_2d(tmC, buf, col0, row0);
        else tma_store_2d(tmC, buf, col0, row0);
        bulk_commit();
    }
}

// ------------------------------------------------------------------------- the kernel
template <int BN, bool TA, bool TB, int NCTA>
__global__ void __launch_bounds__(NUM_THREADS, 1)
k_gemm_bf16(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
            const __grid_constant__ CUtensorMap tmC, const Epilogue epi, int M, int N, int K) {
    using C = Cfg<BN, NCTA>;
    constexpr int TILE_M = BM * NCTA;                     // output rows per work item (CTA or CTA pair)
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + C::STAGES * A_STAGE_BYTES;
    uint8_t* smem_stage = smem + C::STAGES * C::STAGE_BYTES;               // 1024-aligned (stage sizes are)
    uint64_t* bars = (uint64_t*)(smem_stage + C::STAGING_BYTES);
    uint64_t* full_bar = bars;
    uint64_t* empty_bar = bars + C::STAGES;
    uint64_t* tmem_full = bars + 2 * C::STAGES;
    uint64_t* tmem_empty = tmem_full + 2;
    uint32_t* tmem_base_slot = (uint32_t*)(tmem_empty + 2);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t cta_rank = (NCTA == 2) ? cluster_ctarank() : 0u;
    const bool leader = cta_rank == 0;
    const int unit = blockIdx.x / NCTA, n_units = gridDim.x / NCTA;       // work is dealt to CTAs / CTA pairs
    const int num_m = (M + TILE_M - 1) / TILE_M, num_n = (N + BN - 1) / BN;
    const int num_kb = (K + BK - 1) / BK;
    const int splits = epi.splits;
    const int num_items = num_m * num_n * splits;
    const int kb_per = (num_kb + splits - 1) / splits;

    if (warp == 0 && lane == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)&tmA) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)&tmB) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)&tmC) : "memory");
        for (int s = 0; s < C::STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&tmem_full[s], 1);
            mbar_init(&tmem_empty[s], 128 * NCTA);        // the epilogue threads of every CTA of the pair
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        if (NCTA == 2) {
            asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_base_slot)),
                         "r"((uint32_t)C::TMEM_COLS) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
        } else {
            asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_base_slot)),
                         "r"((uint32_t)C::TMEM_COLS) : "memory");
            asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
        }
    }
    tcgen05_fence_before();
    if (NCTA == 2) cluster_sync_all(); else __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_base_slot;

    if (warp == 0) {
        // ===================================================== TMA producer (every CTA stages its own operands)
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int item = unit; item < num_items; item += n_units) {
                int tm, tn, split;
                item_to_tile(item, splits, num_m, num_n, epi.band, tm, tn, split);
                const int m0 = tm * TILE_M + (int)cta_rank * BM;
                const int n0 = tn * BN + (int)cta_rank * C::B_ROWS;
                const int kb0 = split * kb_per, kb1 = min(num_kb, kb0 + kb_per);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&empty_bar[stage], phase ^ 1);
                    uint8_t* sa = smem_a + stage * A_STAGE_BYTES;
                    uint8_t* sb = smem_b + stage * C::B_STAGE_BYTES;
                    const int k0 = kb * BK;
                    if (NCTA == 2) {
                        // the leader's barrier counts the bytes of both CTAs
                        if (leader) mbar_arrive_expect_tx(&full_bar[stage], 2 * C::STAGE_BYTES);
                        const uint32_t lbar = map_to_cta(smem_u32(&full_bar[stage]), 0);
                        if (TA) {
#pragma unroll
                            for (int s = 0; s < BM / 64; ++s)
                                tma_load_2d_pair(sa + s * (BK * 128), &tmA, lbar, m0 + s * 64, k0);
                        } else {
                            tma_load_2d_pair(sa, &tmA, lbar, k0, m0);
                        }
                        if (TB) {
#pragma unroll
                            for (int s = 0; s < C::B_ROWS / 64; ++s)
                                tma_load_2d_pair(sb + s * (BK * 128), &tmB, lbar, n0 + s * 64, k0);
                        } else {
                            tma_load_2d_pair(sb, &tmB, lbar, k0, n0);
                        }
                    } else {
                        mbar_arrive_expect_tx(&full_bar[stage], C::STAGE_BYTES);
                        if (TA) {
#pragma unroll
                            for (int s = 0; s < BM / 64; ++s)
                                tma_load_2d(sa + s * (BK * 128), &tmA, &full_bar[stage], m0 + s * 64, k0);
                        } else {
                            tma_load_2d(sa, &tmA, &full_bar[stage], k0, m0);
                        }
                        if (TB) {
#pragma unroll
                            for (int s = 0; s < BN / 64; ++s)
                                tma_load_2d(sb + s * (BK * 128), &tmB, &full_bar[stage], n0 + s * 64, k0);
                        } else {
                            tma_load_2d(sb, &tmB, &full_bar[stage], k0, n0);
                        }
                    }
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================================================== MMA issuer (one lane; the leader CTA of a pair)
        if (lane == 0 && leader) {
            constexpr uint32_t idesc = make_idesc(TILE_M, BN, TA, TB);
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int item = unit; item < num_items; item += n_units) {
                const int split = item % splits;
                const int kb0 = split * kb_per, kb1 = min(num_kb, kb0 + kb_per);
                mbar_wait(&tmem_empty[acc], acc_phase ^ 1);
                tcgen05_fence_after();
                const uint32_t d_tmem = tmem_base + (uint32_t)(acc * BN);
                for (int kb = kb0; kb < kb1; ++kb) {
                    mbar_wait(&full_bar[stage], phase);
                    tcgen05_fence_after();
                    const uint32_t sa = smem_u32(smem_a + stage * A_STAGE_BYTES);
                    const uint32_t sb = smem_u32(smem_b + stage * C::B_STAGE_BYTES);
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) {
                        // K-major: +32 B per UMMA_K inside the swizzle atom, SBO = 8 rows * 128 B.
                        // MN-major: 64-wide slabs LBO = BK*128 B apart, 8-row k-groups SBO = 1024 B apart,
                        //           +2 k-groups (2048 B) per UMMA_K.
                        const uint64_t adesc = TA ? make_desc(sa + k * 2048, BK * 128, 1024)
                                                  : make_desc(sa + k * 32, 16, 1024);
                        const uint64_t bdesc = TB ? make_desc(sb + k * 2048, BK * 128, 1024)
                                                  : make_desc(sb + k * 32, 16, 1024);
                        if (NCTA == 2) umma_bf16_pair(d_tmem, adesc, bdesc, idesc, ((kb - kb0) | k) != 0);
                        else umma_bf16(d_tmem, adesc, bdesc, idesc, ((kb - kb0) | k) != 0);
                    }
                    // frees the smem slot (of both CTAs) when the MMAs retire
                    if (NCTA == 2) umma_commit_pair(&empty_bar[stage]); else umma_commit(&empty_bar[stage]);
                    if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
                }
                if (NCTA == 2) umma_commit_pair(&tmem_full[acc]); else umma_commit(&tmem_full[acc]);   // accumulator complete
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else {
        // ===================================================== epilogue (warps 2..5), own 128 rows of the tile
        const int q = warp & 3;                              // TMEM lane quarter of this warp
        uint8_t* my_stage = smem_stage + q * 8192;
        int buf = 0;
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int item = unit; item < num_items; item += n_units) {
            int tm, tn, split;
            item_to_tile(item, splits, num_m, num_n, epi.band, tm, tn, split);
            const int m0 = tm * TILE_M + (int)cta_rank * BM, n0 = tn * BN;
            const bool has_k = split * kb_per < num_kb;      // trailing splits can be empty
            mbar_wait(&tmem_full[acc], acc_phase);
            tcgen05_fence_after();
            const int row0 = m0 + q * 32, row = row0 + lane;
            const uint32_t t_row = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN);
            const float* mrow = (epi.mask && row < M) ? epi.mask + (int64_t)row * epi.ldmask : nullptr;
            const bool add_bias = split == 0;
            if (row0 < M && has_k) {
                if (epi.c_bf16) {
#pragma unroll 1
                    for (int c = 0; c < BN; c += 64) {
                        if (n0 + c >= N) break;              // warp-uniform
                        uint32_t r0[32], r1[32], packed[32];
                        tmem_ld32(t_row + (uint32_t)c, r0);
                        tmem_ld32(t_row + (uint32_t)(c + 32), r1);
#pragma unroll
                        for (int g = 0; g < 16; ++g) {
                            float v[4];
#pragma unroll
                            for (int t = 0; t < 4; ++t)
                                v[t] = __uint_as_float(g < 8 ? r0[(g & 7) * 4 + t] : r1[(g & 7) * 4 + t]);
                            epi_apply4(epi, v, mrow, n0 + c + g * 4, N, add_bias);
                            __nv_bfloat162 lo = __floats2bfloat162_rn(v[0], v[1]), hi = __floats2bfloat162_rn(v[2], v[3]);
                            packed[g * 2] = *reinterpret_cast<uint32_t*>(&lo);
                            packed[g * 2 + 1] = *reinterpret_cast<uint32_t*>(&hi);
                        }
                        epi_flush(epi, &tmC, my_stage + buf * 4096, lane, packed, n0 + c, row0);
                        buf ^= 1;
                    }
                } else {
#pragma unroll 1
                    for (int c = 0; c < BN; c += 32) {
                        if (n0 + c >= N) break;              // warp-uniform
                        uint32_t r[32];
                        tmem_ld32(t_row + (uint32_t)c, r);
#pragma unroll
                        for (int g = 0; g < 8; ++g) {
                            float v[4];
#pragma unroll
                            for (int t = 0; t < 4; ++t) v[t] = __uint_as_float(r[g * 4 + t]);
                            epi_apply4(epi, v, mrow, n0 + c + g * 4, N, add_bias);
#pragma unroll
                            for (int t = 0; t < 4; ++t) r[g * 4 + t] = __float_as_uint(v[t]);
                        }
                        epi_flush(epi, &tmC, my_stage + buf * 4096, lane, r, n0 + c, row0);
                        buf ^= 1;
                    }
                }
            }
            tcgen05_fence_before();
            // the MMA issuer (leader CTA) may overwrite this accumulator stage once both CTAs have drained it
            if (NCTA == 2) mbar_arrive_cluster(map_to_cta(smem_u32(&tmem_empty[acc]), 0));
            else mbar_arrive(&tmem_empty[acc]);
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
        if (lane == 0) bulk_wait_all();                      // staging buffers must outlive the bulk reads
    }
    tcgen05_fence_before();
    // pair: neither CTA may exit (or free TMEM) while the other can still signal its barriers / read its smem
    if (NCTA == 2) cluster_sync_all(); else __syncthreads();
    if (warp == 1) {
        tcgen05_fence_after();
        if (NCTA == 2)
            asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)C::TMEM_COLS)
                         : "memory");
        else
            asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)C::TMEM_COLS)
                         : "memory");
    }
}

// ------------------------------------------------------------------------- host side
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                    const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

PFN_encodeTiled get_encode_fn() {
    static PFN_encodeTiled fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
            qres == cudaDriverEntryPointSuccess)
            fn = (PFN_encodeTiled)p;
    }
    return fn;
}

// 2-D bf16 tensor map: `inner` contiguous elements per row, `outer` rows, row pitch ld elements.
int make_tmap(CUtensorMap* tm, const void* ptr, int64_t inner, int64_t outer, int64_t ld, int box_inner, int box_outer,
              bool f32 = false) {
    PFN_encodeTiled fn = get_encode_fn();
    if (!fn) { nc_set_error("nc_gemm_bf16: cuTensorMapEncodeTiled unavailable (no CUDA driver?)"); return NC_ECUDA; }
    cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
    cuuint64_t strides[1] = {(cuuint64_t)ld * (f32 ? 4 : 2)};
    cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_outer};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = fn(tm, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), dims, strides, box, estr,
                    CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                    CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        nc_set_error("nc_gemm_bf16: cuTensorMapEncodeTiled failed (%d) inner=%lld outer=%lld ld=%lld", (int)r,
                     (long long)inner, (long long)outer, (long long)ld);
        return NC_ECUDA;
    }
    return NC_OK;
}

template <int BN, bool TA, bool TB, int NCTA>
int launch(const CUtensorMap& tmA, const CUtensorMap& tmB, const CUtensorMap& tmC, Epilogue epi, int M, int N,
           int K, cudaStream_t s) {
    using C = Cfg<BN, NCTA>;
    static bool attr_set = false;
    if (!attr_set) {
        cudaError_t e = cudaFuncSetAttribute(k_gemm_bf16<BN, TA, TB, NCTA>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             C::SMEM_BYTES);
        if (e != cudaSuccess) { nc_set_error("nc_gemm_bf16: cannot opt in to %d B smem: %s", C::SMEM_BYTES, cudaGetErrorString(e)); return NC_ECUDA; }
        attr_set = true;
    }
    const int num_m = (int)nc_cdiv(M, BM * NCTA), num_n = (int)nc_cdiv(N, BN);
    const int items = num_m * num_n * epi.splits;
    const int max_units = nc_num_sms() / NCTA;
    const int units = items < max_units ? items : max_units;
    int band = (int)nc_cdiv(units, num_n);                   // one wave ~ band tile rows x all tile columns
    epi.band = band < 1 ? 1 : (band > num_m ? num_m : band);
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)(units * NCTA));
    cfg.blockDim = dim3(NUM_THREADS);
    cfg.dynamicSmemBytes = C::SMEM_BYTES;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = NCTA;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = NCTA > 1 ? 1 : 0;
    cudaError_t e = cudaLaunchKernelEx(&cfg, k_gemm_bf16<BN, TA, TB, NCTA>, tmA, tmB, tmC, epi, M, N, K);
    if (e != cudaSuccess) { nc_set_error("nc_gemm_bf16: launch failed: %s", cudaGetErrorString(e)); return NC_ECUDA; }
    nc_count_launch(1);
    return NC_OK;
}

// NC_GEMM_PAIR=0 disables the CTA-pair (cta_group::2) configuration (A/B measurements)
bool pair_enabled() {
    static int v = -1;
    if (v < 0) {
        const char* e = getenv("NC_GEMM_PAIR");
        v = (e && e[0] == '0') ? 0 : 1;
    }
    return v == 1;
}

}  // namespace

extern "C" int nc_gemm_bf16(const void* A, int64_t lda, int trans_a, const void* B, int64_t ldb, int trans_b,
                            void* Cp, int c_dtype, int64_t ldc, int32_t M, int32_t N, int32_t K, const float* bias,
                            int relu, const float* mask, int64_t ldmask, int accumulate, void* stream) {
    NC_REQUIRE(M >= 0 && N >= 0 && K > 0, NC_ESHAPE, "nc_gemm_bf16: need M,N >= 0 and K > 0 (M=%d N=%d K=%d)", M, N, K);
    NC_REQUIRE(c_dtype == NC_F32 || c_dtype == NC_BF16, NC_EINVAL, "nc_gemm_bf16: bad c_dtype %d", c_dtype);
    NC_REQUIRE(((uintptr_t)A % 16 == 0) && ((uintptr_t)B % 16 == 0) && ((uintptr_t)Cp % 16 == 0), NC_EINVAL,
               "nc_gemm_bf16: A, B and C must be 16-byte aligned");
    NC_REQUIRE(lda % 8 == 0 && ldb % 8 == 0 && ldc % (c_dtype == NC_BF16 ? 8 : 4) == 0, NC_EINVAL,
               "nc_gemm_bf16: lda/ldb must be multiples of 8, ldc of %d (lda=%lld ldb=%lld ldc=%lld)",
               c_dtype == NC_BF16 ? 8 : 4, (long long)lda, (long long)ldb, (long long)ldc);
    NC_REQUIRE(ldc >= N && lda >= (trans_a ? M : K) && ldb >= (trans_b ? N : K), NC_ESHAPE,
               "nc_gemm_bf16: leading dimension too small (M=%d N=%d K=%d lda=%lld ldb=%lld ldc=%lld)", M, N, K,
               (long long)lda, (long long)ldb, (long long)ldc);
    if (M == 0 || N == 0) return NC_OK;
    // BN=256 tiles when they still give every SM a few tiles, else 128-wide tiles
    const int64_t tiles256 = nc_cdiv(M, BM) * nc_cdiv(N, 256);
    const bool wide = tiles256 >= 2 * (int64_t)nc_num_sms() || (N % 256 == 0 && tiles256 >= nc_num_sms());
    const int BNsel = wide ? 256 : 128;
    // CTA pairs (256 x 256 tiles, cta_group::2) as soon as they give every pair of SMs a tile
    const int64_t pair_tiles = nc_cdiv(M, 2 * BM) * nc_cdiv(N, 256);
    const bool pair = pair_enabled() && (wide || pair_tiles >= nc_num_sms() / 2);
    CUtensorMap tmA, tmB;
    int rc;
    if (trans_a) rc = make_tmap(&tmA, A, M, K, lda, 64, BK);
    else         rc = make_tmap(&tmA, A, K, M, lda, BK, BM);
    if (rc) return rc;
    if (trans_b) rc = make_tmap(&tmB, B, N, K, ldb, 64, BK);
    else         rc = make_tmap(&tmB, B, K, N, ldb, BK, pair ? 128 : BNsel);
    if (rc) return rc;
    // split K when the tiles alone cannot fill the machine and the reduction is deep (fp32 C, no ReLU):
    // the partial tiles are summed in L2 by the TMA reduce-add.
    const int num_kb = (int)nc_cdiv(K, BK);
    const int64_t tiles = nc_cdiv(M, BM) * nc_cdiv(N, BNsel);
    int splits = 1;
    if (c_dtype == NC_F32 && !relu && tiles * 2 <= nc_num_sms() && num_kb >= 8) {
        int want = (int)(nc_num_sms() / tiles);
        if (want > num_kb / 4) want = num_kb / 4;
        if (want > 16) want = 16;
        if (want > 1) {
            const int kb_per = (num_kb + want - 1) / want;
            splits = (num_kb + kb_per - 1) / kb_per;         // no empty split
        }
    }
    cudaStream_t s = (cudaStream_t)stream;
    if (splits > 1 && !accumulate) {
        cudaError_t e = cudaMemset2DAsync(Cp, (size_t)ldc * 4, 0, (size_t)N * 4, (size_t)M, s);
        if (e != cudaSuccess) { nc_set_error("nc_gemm_bf16: memset failed: %s", cudaGetErrorString(e)); return NC_ECUDA; }
    }
    CUtensorMap tmC;
    rc = make_tmap(&tmC, Cp, N, M, ldc, c_dtype == NC_BF16 ? 64 : 32, 32, c_dtype == NC_F32);
    if (rc) return rc;
    Epilogue epi{bias, mask, ldmask, c_dtype == NC_BF16, relu, (accumulate || splits > 1) ? 1 : 0,
                 ((uintptr_t)bias % 16 == 0) ? 1 : 0, (mask && (uintptr_t)mask % 16 == 0 && ldmask % 4 == 0) ? 1 : 0,
                 splits, 1};
#define DISPATCH(BN, NCTA)                                                                              \
    do {                                                                                                \
        if (trans_a && trans_b) rc = launch<BN, true, true, NCTA>(tmA, tmB, tmC, epi, M, N, K, s);      \
        else if (trans_a) rc = launch<BN, true, false, NCTA>(tmA, tmB, tmC, epi, M, N, K, s);           \
        else if (trans_b) rc = launch<BN, false, true, NCTA>(tmA, tmB, tmC, epi, M, N, K, s);           \
        else rc = launch<BN, false, false, NCTA>(tmA, tmB, tmC, epi, M, N, K, s);                       \
    } while (0)
    if (pair) DISPATCH(256, 2); else if (BNsel == 256) DISPATCH(256, 1); else DISPATCH(128, 1);
#undef DISPATCH
    if (rc) return rc;
    NC_CHECK_LAUNCH("nc_gemm_bf16");
    return NC_OK;
}
