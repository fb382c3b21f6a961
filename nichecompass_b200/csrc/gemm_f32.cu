// fp32 SIMT GEMM (parity mode; also the correctness witness for the tcgen05 kernel).
//   C[M,N] (+)= act(opA(A) * opB(B) + bias) (* mask)
// Replaces ATen addmm / cuBLAS SGEMM call sites (nn/encoders.py:179, PyG lin_l/lin_r/lin,
// nn/layercomponents.py:77-80) when the module runs in fp32.
#include "common.cuh"

extern void nc_count_launch(int n);

constexpr int GB_M = 64, GB_N = 64, GB_K = 16, G_T = 256;

template <bool TA, bool TB>
__global__ void __launch_bounds__(G_T)
k_gemm_f32(const float* __restrict__ A, int64_t lda, const float* __restrict__ B, int64_t ldb,
           float* __restrict__ C, int64_t ldc, int M, int N, int K, const float* __restrict__ bias,
           int relu, const float* __restrict__ mask, int64_t ldmask, int accumulate) {
    __shared__ float As[GB_K][GB_M + 4];
    __shared__ float Bs[GB_K][GB_N + 4];
    const int tid = threadIdx.x;
    const int m0 = blockIdx.y * GB_M, n0 = blockIdx.x * GB_N;
    const int tx = tid % 16, ty = tid / 16;       // 16 x 16 threads, 4 x 4 outputs each
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (int k0 = 0; k0 < K; k0 += GB_K) {
#pragma unroll
        for (int r = 0; r < (GB_M * GB_K) / G_T; ++r) {
            const int idx = tid + r * G_T;
            int m, k;
            if (TA) { m = idx % GB_M; k = idx / GB_M; } else { m = idx / GB_K; k = idx % GB_K; }
            const int gm = m0 + m, gk = k0 + k;
            float v = 0.f;
            if (gm < M && gk < K) v = TA ? __ldg(A + (int64_t)gk * lda + gm) : __ldg(A + (int64_t)gm * lda + gk);
            As[k][m] = v;
        }
#pragma unroll
        for (int r = 0; r < (GB_N * GB_K) / G_T; ++r) {
            const int idx = tid + r * G_T;
            int n, k;
            if (TB) { n = idx % GB_N; k = idx / GB_N; } else { n = idx / GB_K; k = idx % GB_K; }
            const int gn = n0 + n, gk = k0 + k;
            float v = 0.f;
            if (gn < N && gk < K) v = TB ? __ldg(B + (int64_t)gk * ldb + gn) : __ldg(B + (int64_t)gn * ldb + gk);
            Bs[k][n] = v;
        }
        __syncthreads();
        // two-level summation: the 16 products of a k-block are summed on their own and then added to the running
        // total, so the rounding error grows with K / 16 + 16 instead of K (a plain running sum over K = 5000 loses
        // ~1.5 digits against a blocked CPU SGEMM -- enough to flip the sign of near-zero GATv2 pre-activations
        // ten times as often, which the 1e-5 parity tests see as O(1e-3) gradient outliers)
        float part[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) part[i][j] = 0.f;
#pragma unroll
        for (int k = 0; k < GB_K; ++k) {
            float a[4], b[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) a[i] = As[k][ty * 4 + i];
#pragma unroll
            for (int j = 0; j < 4; ++j) b[j] = Bs[k][tx * 4 + j];
#pragma unroll
            for (int i = 0; i < 4; ++i)
#pragma unroll
                for (int j = 0; j < 4; ++j) part[i][j] = fmaf(a[i], b[j], part[i][j]);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) acc[i][j] += part[i][j];
        __syncthreads();
    }
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const int gm = m0 + ty * 4 + i;
        if (gm >= M) continue;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const int gn = n0 + tx * 4 + j;
            if (gn >= N) continue;
            float v = acc[i][j];
            if (bias) v += __ldg(bias + gn);
            if (relu) v = fmaxf(v, 0.f);
            if (mask) v *= __ldg(mask + (int64_t)gm * ldmask + gn);
            float* c = C + (int64_t)gm * ldc + gn;
            *c = accumulate ? (*c + v) : v;
        }
    }
}

extern "C" int nc_gemm_f32(const float* A, int64_t lda, int trans_a, const float* B, int64_t ldb,
                           int trans_b, float* C, int64_t ldc, int32_t M, int32_t N, int32_t K,
                           const float* bias, int relu, const float* mask, int64_t ldmask,
                           int accumulate, void* stream) {
    NC_REQUIRE(M >= 0 && N >= 0 && K >= 0, NC_ESHAPE, "nc_gemm_f32: negative dimension");
    NC_REQUIRE(ldc >= N && lda >= (trans_a ? M : K) && ldb >= (trans_b ? N : K), NC_ESHAPE,
               "nc_gemm_f32: leading dimension too small (M=%d N=%d K=%d lda=%lld ldb=%lld ldc=%lld)",
               M, N, K, (long long)lda, (long long)ldb, (long long)ldc);
    if (M == 0 || N == 0) return NC_OK;
    cudaStream_t s = (cudaStream_t)stream;
    dim3 grid((unsigned)nc_cdiv(N, GB_N), (unsigned)nc_cdiv(M, GB_M));
#define LAUNCH(TA, TB) k_gemm_f32<TA, TB><<<grid, G_T, 0, s>>>(A, lda, B, ldb, C, ldc, M, N, K, bias, relu, mask, ldmask, accumulate)
    if (trans_a && trans_b) LAUNCH(true, true);
    else if (trans_a) LAUNCH(true, false);
    else if (trans_b) LAUNCH(false, true);
    else LAUNCH(false, false);
#undef LAUNCH
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_gemm_f32");
    return NC_OK;
}
