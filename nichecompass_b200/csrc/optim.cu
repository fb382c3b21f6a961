// Latent sampling, row gather/scatter, gene-program weight regularisers, active-GP statistics,
// fused Adam and precision casts.  Reference: modules/vgaemodulemixin.py:34-41 (reparameterize),
// modules/vgpgae.py:626-651 (batch_idx gather, z masking, |mu| EMA), modules/losses.py:215-252
// (group lasso), :287-339 (L1), modules/vgpgae.py:1205-1334 (active-GP mask),
// train/trainer.py:339-342,421 (torch.optim.Adam).
#include <algorithm>

#include "common.cuh"

extern void nc_count_launch(int n);

namespace {
inline int grid_for(int64_t n, int t = 256) { return (int)std::min<int64_t>(std::max<int64_t>(nc_cdiv(n, t), 1), 8 * (int64_t)nc_num_sms()); }

// ------------------------------------------------------------------------- gather / scatter rows
__global__ void k_gather_rows_f32(const float* __restrict__ src, int64_t lds, const int32_t* __restrict__ idx,
                                  int R, int C, float* __restrict__ dst, int64_t ldd) {
    const int64_t total = (int64_t)R * C;
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (int64_t)gridDim.x * blockDim.x) {
        const int r = (int)(q / C), c = (int)(q % C);
        dst[(int64_t)r * ldd + c] = src[(int64_t)idx[r] * lds + c];
    }
}
__global__ void k_scatter_add_rows_f32(const float* __restrict__ src, int64_t lds, const int32_t* __restrict__ idx,
                                       int R, int C, float* __restrict__ dst, int64_t ldd) {
    const int64_t total = (int64_t)R * C;
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (int64_t)gridDim.x * blockDim.x) {
        const int r = (int)(q / C), c = (int)(q % C);
        atomicAdd(dst + (int64_t)idx[r] * ldd + c, src[(int64_t)r * lds + c]);
    }
}

// ------------------------------------------------------------------------- reparameterisation
// z = (mu + eps * exp(logstd)) * active;  optional bf16 copy for the decoder GEMM
__global__ void k_reparam_fwd(const float* __restrict__ mu, int64_t ldmu, const float* __restrict__ ls, int64_t ldls,
                              const float* __restrict__ eps, int64_t ldeps, const uint8_t* __restrict__ active,
                              int R, int P, float* __restrict__ z, int64_t ldz, __nv_bfloat16* __restrict__ zb,
                              int64_t ldzb) {
    const int64_t total = (int64_t)R * P;
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (int64_t)gridDim.x * blockDim.x) {
        const int r = (int)(q / P), c = (int)(q % P);
        float v = mu[(int64_t)r * ldmu + c];
        if (eps) v += eps[(int64_t)r * ldeps + c] * expf(ls[(int64_t)r * ldls + c]);
        if (active && !active[c]) v = 0.f;
        z[(int64_t)r * ldz + c] = v;
        if (zb) zb[(int64_t)r * ldzb + c] = __float2bfloat16_rn(v);
    }
}
__global__ void k_reparam_bwd(const float* __restrict__ ls, int64_t ldls, const float* __restrict__ eps, int64_t ldeps,
                              const uint8_t* __restrict__ active, int R, int P, const float* __restrict__ dz,
                              int64_t lddz, float* __restrict__ dmu, int64_t lddmu, float* __restrict__ dls, int64_t lddls) {
    const int64_t total = (int64_t)R * P;
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (int64_t)gridDim.x * blockDim.x) {
        const int r = (int)(q / P), c = (int)(q % P);
        float g = dz[(int64_t)r * lddz + c];
        if (active && !active[c]) g = 0.f;
        dmu[(int64_t)r * lddmu + c] = g;
        dls[(int64_t)r * lddls + c] = eps ? g * eps[(int64_t)r * ldeps + c] * expf(ls[(int64_t)r * ldls + c]) : 0.f;
    }
}

// ------------------------------------------------------------------------- |mu| EMA
// ema[c] = momentum * mean_r |mu[r,c]| + (1 - momentum) * ema[c]; one warp per column block
__global__ void k_abs_mean_ema(const float* __restrict__ mu, int64_t ldmu, int R, int P, float momentum,
                               float* __restrict__ ema) {
    const int c = blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= P) return;
    float s = 0.f;
    for (int r = 0; r < R; ++r) s += fabsf(mu[(int64_t)r * ldmu + c]);
    ema[c] = momentum * (s / (float)R) + (1.f - momentum) * ema[c];
}

// ------------------------------------------------------------------------- weight statistics
// Per gene-program (column) statistics of a decoder weight W [F, P] (row-major, ld):
//   abs_sum[c] += sum_f |W[f,c]| * (l1mask ? l1mask[f,c] : 1)
//   sq_sum[c]  += sum_f W[f,c]^2        nnz[c] += #{W[f,c] != 0}     mask_cnt[c] += #{mask[f,c] != 0}
// Any output may be NULL.  grid.x tiles columns (32 per CTA), grid.y tiles rows.
__global__ void __launch_bounds__(256)
k_col_stats(const float* __restrict__ W, int64_t ld, int F, int P, const uint8_t* __restrict__ l1mask,
            const float* __restrict__ mask, int64_t ldmask, float* __restrict__ abs_sum, float* __restrict__ sq_sum,
            float* __restrict__ nnz, float* __restrict__ mask_cnt) {
    __shared__ float s[4][8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + tx;
    float a = 0.f, q = 0.f, z = 0.f, mc = 0.f;
    if (c < P) {
        const int rows_per = (F + gridDim.y - 1) / gridDim.y;
        const int f0 = blockIdx.y * rows_per, f1 = min(F, f0 + rows_per);
        for (int f = f0 + ty; f < f1; f += 8) {
            const float w = W[(int64_t)f * ld + c];
            const float lm = l1mask ? (float)l1mask[(int64_t)f * P + c] : 1.f;
            a += fabsf(w) * lm;
            q += w * w;
            z += (w != 0.f) ? 1.f : 0.f;
            if (mask) mc += (mask[(int64_t)f * ldmask + c] != 0.f) ? 1.f : 0.f;
        }
    }
    s[0][ty][tx] = a; s[1][ty][tx] = q; s[2][ty][tx] = z; s[3][ty][tx] = mc;
    __syncthreads();
    if (ty == 0 && c < P) {
        float ra = 0.f, rq = 0.f, rz = 0.f, rm = 0.f;
        for (int k = 0; k < 8; ++k) { ra += s[0][k][tx]; rq += s[1][k][tx]; rz += s[2][k][tx]; rm += s[3][k][tx]; }
        if (abs_sum) atomicAdd(abs_sum + c, ra);
        if (sq_sum) atomicAdd(sq_sum + c, rq);
        if (nnz) atomicAdd(nnz + c, rz);
        if (mask_cnt) atomicAdd(mask_cnt + c, rm);
    }
}

// dW += l1_scale * sign(W) * l1mask + gl_coef[c] * W   (gl_coef[c] = lambda * sqrt(cnt_c) / ||W[:,c]||, 0 if norm 0)
__global__ void k_reg_grad(const float* __restrict__ W, int64_t ld, int F, int P, float l1_scale,
                           const uint8_t* __restrict__ l1mask, const float* __restrict__ gl_coef,
                           const float* __restrict__ gscale, float* __restrict__ dW, int64_t lddw) {
    const int64_t total = (int64_t)F * P;
    const float gs = gscale ? __ldg(gscale) : 1.f;
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (int64_t)gridDim.x * blockDim.x) {
        const int f = (int)(q / P), c = (int)(q % P);
        const float w = W[(int64_t)f * ld + c];
        float g = 0.f;
        if (l1_scale != 0.f && (!l1mask || l1mask[q])) g += l1_scale * (w > 0.f ? 1.f : (w < 0.f ? -1.f : 0.f));
        if (gl_coef) g += gl_coef[c] * w;
        dW[(int64_t)f * lddw + c] += gs * g;
    }
}

// ------------------------------------------------------------------------- Adam (torch semantics)
__device__ __forceinline__ float adam_elem(float pi, float gi, float& mi, float& vi, float lr_bc1, float b1, float b2,
                                           float eps, float wd, float bc2_sqrt, float clip) {
    if (clip > 0.f) gi = fminf(fmaxf(gi, -clip), clip);
    if (wd != 0.f) gi += wd * pi;
    mi = b1 * mi + (1.f - b1) * gi;
    vi = b2 * vi + (1.f - b2) * gi * gi;
    const float denom = sqrtf(vi) / bc2_sqrt + eps;
    return pi - lr_bc1 * (mi / denom);
}

// 128-bit accesses over the 4-aligned body, scalar tail.  zero_g: the gradient buffer is cleared
// after it has been consumed (saves the separate memset of optimizer.zero_grad()).
__global__ void __launch_bounds__(256)
k_adam(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
       int64_t n, float lr, float b1, float b2, float eps, float wd, float bc1, float bc2_sqrt,
       float clip, __nv_bfloat16* __restrict__ p_bf16, int zero_g, const float* __restrict__ dyn) {
    if (dyn) {                      // {lr, step} live on the device: the launch can be replayed from a CUDA graph
        lr = dyn[0];
        const double step = (double)dyn[1];
        bc1 = (float)(1.0 - pow((double)b1, step));
        bc2_sqrt = (float)sqrt(1.0 - pow((double)b2, step));
    }
    const float lr_bc1 = lr / bc1;
    const int64_t n4 = n >> 2;
    const int64_t t0 = (int64_t)blockIdx.x * blockDim.x + threadIdx.x, nt = (int64_t)gridDim.x * blockDim.x;
    for (int64_t i = t0; i < n4; i += nt) {
        float4 pp = reinterpret_cast<float4*>(p)[i];
        const float4 gg = reinterpret_cast<const float4*>(g)[i];
        float4 mm = reinterpret_cast<float4*>(m)[i];
        float4 vv = reinterpret_cast<float4*>(v)[i];
        // A chunk that has never seen a gradient (g = m = v = 0: the masked-out entries of the gene-program decoders,
        // about a quarter of the arena at the 5 k-gene configuration) keeps its value exactly: update = 0 / (0 + eps).
        // Nothing to write back, the bf16 shadow included.
        if (wd == 0.f && gg.x == 0.f && gg.y == 0.f && gg.z == 0.f && gg.w == 0.f && mm.x == 0.f && mm.y == 0.f &&
            mm.z == 0.f && mm.w == 0.f && vv.x == 0.f && vv.y == 0.f && vv.z == 0.f && vv.w == 0.f)
            continue;
        pp.x = adam_elem(pp.x, gg.x, mm.x, vv.x, lr_bc1, b1, b2, eps, wd, bc2_sqrt, clip);
        pp.y = adam_elem(pp.y, gg.y, mm.y, vv.y, lr_bc1, b1, b2, eps, wd, bc2_sqrt, clip);
        pp.z = adam_elem(pp.z, gg.z, mm.z, vv.z, lr_bc1, b1, b2, eps, wd, bc2_sqrt, clip);
        pp.w = adam_elem(pp.w, gg.w, mm.w, vv.w, lr_bc1, b1, b2, eps, wd, bc2_sqrt, clip);
        reinterpret_cast<float4*>(p)[i] = pp;
        reinterpret_cast<float4*>(m)[i] = mm;
        reinterpret_cast<float4*>(v)[i] = vv;
        if (zero_g) reinterpret_cast<float4*>(g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (p_bf16) {
            __nv_bfloat162 lo = __floats2bfloat162_rn(pp.x, pp.y), hi = __floats2bfloat162_rn(pp.z, pp.w);
            uint2 q;
            q.x = *reinterpret_cast<uint32_t*>(&lo);
            q.y = *reinterpret_cast<uint32_t*>(&hi);
            reinterpret_cast<uint2*>(p_bf16)[i] = q;
        }
    }
    for (int64_t i = (n4 << 2) + t0; i < n; i += nt) {
        float mi = m[i], vi = v[i];
        const float np = adam_elem(p[i], g[i], mi, vi, lr_bc1, b1, b2, eps, wd, bc2_sqrt, clip);
        p[i] = np; m[i] = mi; v[i] = vi;
        if (zero_g) g[i] = 0.f;
        if (p_bf16) p_bf16[i] = __float2bfloat16_rn(np);
    }
}

// fp32 -> bf16 cast (optionally multiplied by a mask of the same shape), 2-D with leading dims
__global__ void k_cast_bf16(const float* __restrict__ src, int64_t lds, const float* __restrict__ mask, int64_t ldm,
                            int R, int C, __nv_bfloat16* __restrict__ dst, int64_t ldd, int W) {
    const int64_t total = (int64_t)R * W;
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (int64_t)gridDim.x * blockDim.x) {
        const int r = (int)(q / W), c = (int)(q % W);
        float v = 0.f;
        if (c < C) {
            v = src[(int64_t)r * lds + c];
            if (mask) v *= mask[(int64_t)r * ldm + c];
        }
        dst[(int64_t)r * ldd + c] = __float2bfloat16_rn(v);
    }
}
// 4 columns per thread, grid.y tiles rows (all pitches multiples of 4, 16 / 8-byte aligned bases)
__global__ void __launch_bounds__(256)
k_cast_bf16_v4(const float* __restrict__ src, int64_t lds, const float* __restrict__ mask, int64_t ldm,
               int R, int C, __nv_bfloat16* __restrict__ dst, int64_t ldd, int W) {
    const int nv = W >> 2;
    for (int r = blockIdx.y; r < R; r += gridDim.y) {
        for (int v = blockIdx.x * blockDim.x + threadIdx.x; v < nv; v += gridDim.x * blockDim.x) {
            const int c = v * 4;
            float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
            if (c + 3 < C) {
                o = __ldg(reinterpret_cast<const float4*>(src + (int64_t)r * lds + c));
                if (mask) {
                    const float4 m = __ldg(reinterpret_cast<const float4*>(mask + (int64_t)r * ldm + c));
                    o.x *= m.x; o.y *= m.y; o.z *= m.z; o.w *= m.w;
                }
            } else {
                float t[4] = {0.f, 0.f, 0.f, 0.f};
                for (int k = 0; k < 4; ++k)
                    if (c + k < C) t[k] = src[(int64_t)r * lds + c + k] * (mask ? mask[(int64_t)r * ldm + c + k] : 1.f);
                o = make_float4(t[0], t[1], t[2], t[3]);
            }
            Elem<__nv_bfloat16>::st4(dst + (int64_t)r * ldd + c, o);
        }
    }
}
}  // namespace

extern "C" int nc_gather_rows_f32(const float* src, int64_t lds, const int32_t* idx, int32_t R, int32_t C, float* dst,
                                  int64_t ldd, void* stream) {
    NC_REQUIRE(R >= 0 && C > 0 && lds >= C && ldd >= C, NC_ESHAPE, "nc_gather_rows_f32: bad shape");
    if (R == 0) return NC_OK;
    k_gather_rows_f32<<<grid_for((int64_t)R * C), 256, 0, (cudaStream_t)stream>>>(src, lds, idx, R, C, dst, ldd);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_gather_rows_f32");
    return NC_OK;
}

extern "C" int nc_scatter_add_rows_f32(const float* src, int64_t lds, const int32_t* idx, int32_t R, int32_t C,
                                       float* dst, int64_t ldd, void* stream) {
    NC_REQUIRE(R >= 0 && C > 0 && lds >= C && ldd >= C, NC_ESHAPE, "nc_scatter_add_rows_f32: bad shape");
    if (R == 0) return NC_OK;
    k_scatter_add_rows_f32<<<grid_for((int64_t)R * C), 256, 0, (cudaStream_t)stream>>>(src, lds, idx, R, C, dst, ldd);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_scatter_add_rows_f32");
    return NC_OK;
}

extern "C" int nc_reparam_fwd(const float* mu, int64_t ldmu, const float* logstd, int64_t ldls, const float* eps,
                              int64_t ldeps, const uint8_t* active, int32_t R, int32_t P, float* z, int64_t ldz,
                              void* z_bf16, int64_t ldzb, void* stream) {
    NC_REQUIRE(R >= 0 && P > 0, NC_ESHAPE, "nc_reparam_fwd: bad shape");
    if (R == 0) return NC_OK;
    k_reparam_fwd<<<grid_for((int64_t)R * P), 256, 0, (cudaStream_t)stream>>>(mu, ldmu, logstd, ldls, eps, ldeps, active,
                                                                            R, P, z, ldz, (__nv_bfloat16*)z_bf16, ldzb);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_reparam_fwd");
    return NC_OK;
}

extern "C" int nc_reparam_bwd(const float* logstd, int64_t ldls, const float* eps, int64_t ldeps, const uint8_t* active,
                              int32_t R, int32_t P, const float* dz, int64_t lddz, float* dmu, int64_t lddmu,
                              float* dlogstd, int64_t lddls, void* stream) {
    NC_REQUIRE(R >= 0 && P > 0, NC_ESHAPE, "nc_reparam_bwd: bad shape");
    if (R == 0) return NC_OK;
    k_reparam_bwd<<<grid_for((int64_t)R * P), 256, 0, (cudaStream_t)stream>>>(logstd, ldls, eps, ldeps, active, R, P, dz,
                                                                            lddz, dmu, lddmu, dlogstd, lddls);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_reparam_bwd");
    return NC_OK;
}

extern "C" int nc_abs_mean_ema(const float* mu, int64_t ldmu, int32_t R, int32_t P, float momentum, float* ema,
                               void* stream) {
    NC_REQUIRE(R > 0 && P > 0, NC_ESHAPE, "nc_abs_mean_ema: bad shape");
    k_abs_mean_ema<<<(unsigned)nc_cdiv(P, 128), 128, 0, (cudaStream_t)stream>>>(mu, ldmu, R, P, momentum, ema);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_abs_mean_ema");
    return NC_OK;
}

extern "C" int nc_weight_col_stats(const float* W, int64_t ld, int32_t F, int32_t P, const uint8_t* l1mask,
                                   const float* mask, int64_t ldmask, float* abs_sum, float* sq_sum, float* nnz,
                                   float* mask_cnt, void* stream) {
    NC_REQUIRE(F > 0 && P > 0 && ld >= P, NC_ESHAPE, "nc_weight_col_stats: bad shape");
    dim3 grid((unsigned)nc_cdiv(P, 32), (unsigned)std::min<int64_t>(nc_cdiv(F, 64), 64));
    k_col_stats<<<grid, 256, 0, (cudaStream_t)stream>>>(W, ld, F, P, l1mask, mask, ldmask, abs_sum, sq_sum, nnz, mask_cnt);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_weight_col_stats");
    return NC_OK;
}

extern "C" int nc_weight_reg_grad(const float* W, int64_t ld, int32_t F, int32_t P, float l1_scale,
                                  const uint8_t* l1mask, const float* gl_coef, const float* gscale, float* dW,
                                  int64_t lddw, void* stream) {
    NC_REQUIRE(F > 0 && P > 0 && ld >= P && lddw >= P, NC_ESHAPE, "nc_weight_reg_grad: bad shape");
    k_reg_grad<<<grid_for((int64_t)F * P), 256, 0, (cudaStream_t)stream>>>(W, ld, F, P, l1_scale, l1mask, gl_coef, gscale, dW, lddw);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_weight_reg_grad");
    return NC_OK;
}

extern "C" int nc_adam_step(float* p, float* g, float* m, float* v, int64_t n, float lr, float beta1,
                            float beta2, float eps, float weight_decay, int32_t step, float clip_value, void* p_bf16,
                            int zero_grad, void* stream) {
    NC_REQUIRE(n >= 0 && step >= 1, NC_EINVAL, "nc_adam_step: need n >= 0 and step >= 1");
    if (n == 0) return NC_OK;
    const float bc1 = (float)(1.0 - pow((double)beta1, (double)step));
    const float bc2_sqrt = (float)sqrt(1.0 - pow((double)beta2, (double)step));
    NC_REQUIRE(((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) % 16 == 0 && (uintptr_t)p_bf16 % 8 == 0,
               NC_EINVAL, "nc_adam_step: buffers must be 16-byte aligned");
    k_adam<<<grid_for((n + 3) / 4), 256, 0, (cudaStream_t)stream>>>(p, g, m, v, n, lr, beta1, beta2, eps, weight_decay,
                                                                   bc1, bc2_sqrt, clip_value, (__nv_bfloat16*)p_bf16,
                                                                   zero_grad, nullptr);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_adam_step");
    return NC_OK;
}

__global__ void k_bump_step(float* dyn) { dyn[1] += 1.0f; }

// Same update with the learning rate and the step number read from device memory (dyn = {lr, step}, step >= 1 is the
// number of THIS update) and the step advanced by a second one-thread launch: nothing in the launch sequence
// depends on host state, so a captured training step can be replayed (train/trainer.py CUDA-graph buckets).
extern "C" int nc_adam_step_dyn(float* p, float* g, float* m, float* v, int64_t n, float* dyn, float beta1,
                                float beta2, float eps, float weight_decay, float clip_value, void* p_bf16,
                                int zero_grad, void* stream) {
    NC_REQUIRE(n >= 0 && dyn != nullptr, NC_EINVAL, "nc_adam_step_dyn: need n >= 0 and the device state");
    if (n == 0) return NC_OK;
    NC_REQUIRE(((uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v) % 16 == 0 && (uintptr_t)p_bf16 % 8 == 0,
               NC_EINVAL, "nc_adam_step_dyn: buffers must be 16-byte aligned");
    k_adam<<<grid_for((n + 3) / 4), 256, 0, (cudaStream_t)stream>>>(p, g, m, v, n, 0.f, beta1, beta2, eps, weight_decay,
                                                                   1.f, 1.f, clip_value, (__nv_bfloat16*)p_bf16,
                                                                   zero_grad, dyn);
    k_bump_step<<<1, 1, 0, (cudaStream_t)stream>>>(dyn);
    nc_count_launch(2);
    NC_CHECK_LAUNCH("nc_adam_step_dyn");
    return NC_OK;
}

extern "C" int nc_cast_bf16(const float* src, int64_t lds, const float* mask, int64_t ldmask, int32_t R, int32_t C,
                            void* dst, int64_t ldd, int32_t fill_cols, void* stream) {
    if (fill_cols <= 0) fill_cols = (int32_t)ldd;
    NC_REQUIRE(R >= 0 && C > 0 && lds >= C && fill_cols >= C && ldd >= fill_cols, NC_ESHAPE, "nc_cast_bf16: bad shape");
    if (R == 0) return NC_OK;
    if (R > 1 && lds % 4 == 0 && (!mask || (ldmask % 4 == 0 && (uintptr_t)mask % 16 == 0)) && ldd % 4 == 0 &&
        fill_cols % 4 == 0 && (uintptr_t)src % 16 == 0 && (uintptr_t)dst % 8 == 0) {
        dim3 grid((unsigned)std::max<int64_t>(1, std::min<int64_t>(nc_cdiv(fill_cols / 4, 256), 8)),
                  (unsigned)std::min<int64_t>(R, 2048));
        k_cast_bf16_v4<<<grid, 256, 0, (cudaStream_t)stream>>>(src, lds, mask, ldmask, R, C, (__nv_bfloat16*)dst, ldd,
                                                               fill_cols);
        nc_count_launch(1);
        NC_CHECK_LAUNCH("nc_cast_bf16");
        return NC_OK;
    }
    k_cast_bf16<<<grid_for((int64_t)R * fill_cols), 256, 0, (cudaStream_t)stream>>>(src, lds, mask, ldmask, R, C,
                                                                                   (__nv_bfloat16*)dst, ldd, fill_cols);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_cast_bf16");
    return NC_OK;
}

// ------------------------------------------------------------------------- bias gradients / ReLU backward
namespace {
// out[c] (+)= sum_r x[r, c]; grid.x tiles columns (32), grid.y tiles rows, atomics across row tiles
template <typename T>
__global__ void __launch_bounds__(256)
k_col_sum(const T* __restrict__ x, int64_t ld, int R, int C, float* __restrict__ out) {
    __shared__ float s[8][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int c = blockIdx.x * 32 + tx;
    float a = 0.f;
    if (c < C) {
        const int rows_per = (R + gridDim.y - 1) / gridDim.y;
        const int r0 = blockIdx.y * rows_per, r1 = min(R, r0 + rows_per);
        for (int r = r0 + ty; r < r1; r += 8) a += Elem<T>::ld(x + (int64_t)r * ld + c);
    }
    s[ty][tx] = a;
    __syncthreads();
    if (ty == 0 && c < C) {
        float t = 0.f;
        for (int k = 0; k < 8; ++k) t += s[k][tx];
        atomicAdd(out + c, t);
    }
}

// 4 columns per thread (8 / 16-byte loads), 8 row groups per CTA, one atomicAdd per column and CTA
template <typename T>
__global__ void __launch_bounds__(256)
k_col_sum_v4(const T* __restrict__ x, int64_t ld, int R, int C, float* __restrict__ out) {
    __shared__ float4 s[8][32];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int c = (blockIdx.x * 32 + tx) * 4;
    float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < C) {
        const int rows_per = (R + gridDim.y - 1) / gridDim.y;
        const int r0 = blockIdx.y * rows_per, r1 = min(R, r0 + rows_per);
#pragma unroll 4
        for (int r = r0 + ty; r < r1; r += 8) {
            const float4 v = Elem<T>::ld4(x + (int64_t)r * ld + c);
            a.x += v.x; a.y += v.y; a.z += v.z; a.w += v.w;
        }
    }
    s[ty][tx] = a;
    __syncthreads();
    if (ty == 0 && c < C) {
        float4 t = s[0][tx];
        for (int k = 1; k < 8; ++k) { t.x += s[k][tx].x; t.y += s[k][tx].y; t.z += s[k][tx].z; t.w += s[k][tx].w; }
        atomicAdd(out + c, t.x);
        if (c + 1 < C) atomicAdd(out + c + 1, t.y);
        if (c + 2 < C) atomicAdd(out + c + 2, t.z);
        if (c + 3 < C) atomicAdd(out + c + 3, t.w);
    }
}

// out = dh * (h > 0), cast to T; pad columns [C, ldo) are zeroed
template <typename T>
__global__ void k_relu_bwd(const float* __restrict__ dh, int64_t lddh, const T* __restrict__ h, int64_t ldh, int R,
                           int C, T* __restrict__ out, int64_t ldo) {
    const int64_t total = (int64_t)R * ldo;
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (int64_t)gridDim.x * blockDim.x) {
        const int r = (int)(q / ldo), c = (int)(q % ldo);
        float v = 0.f;
        if (c < C && Elem<T>::ld(h + (int64_t)r * ldh + c) > 0.f) v = dh[(int64_t)r * lddh + c];
        Elem<T>::st(out + q, v);
    }
}

// same, 4 columns per thread (all row pitches multiples of 4, 16-byte aligned bases, C % 4 == 0 or ld padding
// readable): grid.y tiles rows
template <typename T>
__global__ void __launch_bounds__(256)
k_relu_bwd_v4(const float* __restrict__ dh, int64_t lddh, const T* __restrict__ h, int64_t ldh, int R, int C,
              T* __restrict__ out, int64_t ldo) {
    const int nv = (int)(ldo >> 2);
    for (int r = blockIdx.y; r < R; r += gridDim.y) {
        for (int v = blockIdx.x * blockDim.x + threadIdx.x; v < nv; v += gridDim.x * blockDim.x) {
            const int c = v * 4;
            float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
            if (c + 3 < C) {
                const float4 hv = Elem<T>::ld4(h + (int64_t)r * ldh + c);
                const float4 d = __ldg(reinterpret_cast<const float4*>(dh + (int64_t)r * lddh + c));
                o.x = hv.x > 0.f ? d.x : 0.f; o.y = hv.y > 0.f ? d.y : 0.f;
                o.z = hv.z > 0.f ? d.z : 0.f; o.w = hv.w > 0.f ? d.w : 0.f;
            } else {
                float t[4] = {0.f, 0.f, 0.f, 0.f};
                for (int k = 0; k < 4; ++k)
                    if (c + k < C && Elem<T>::ld(h + (int64_t)r * ldh + c + k) > 0.f) t[k] = dh[(int64_t)r * lddh + c + k];
                o = make_float4(t[0], t[1], t[2], t[3]);
            }
            Elem<T>::st4(out + (int64_t)r * ldo + c, o);
        }
    }
}
// out = relu(x) cast to T; pad columns [C, ldo) are zeroed
template <typename T>
__global__ void k_relu_cast(const float* __restrict__ x, int64_t ldx, int R, int C, T* __restrict__ out, int64_t ldo) {
    const int64_t total = (int64_t)R * ldo;
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (int64_t)gridDim.x * blockDim.x) {
        const int r = (int)(q / ldo), c = (int)(q % ldo);
        Elem<T>::st(out + q, c < C ? fmaxf(x[(int64_t)r * ldx + c], 0.f) : 0.f);
    }
}
}  // namespace

extern "C" int nc_relu_cast(const float* x, int64_t ldx, int32_t R, int32_t C, void* out, int dtype, int64_t ldo,
                            void* stream) {
    NC_REQUIRE(R >= 0 && C > 0 && ldx >= C && ldo >= C, NC_ESHAPE, "nc_relu_cast: bad shape");
    NC_REQUIRE(dtype == NC_F32 || dtype == NC_BF16, NC_EINVAL, "nc_relu_cast: bad dtype %d", dtype);
    if (R == 0) return NC_OK;
    const int g = grid_for((int64_t)R * ldo);
    if (dtype == NC_F32) k_relu_cast<float><<<g, 256, 0, (cudaStream_t)stream>>>(x, ldx, R, C, (float*)out, ldo);
    else k_relu_cast<__nv_bfloat16><<<g, 256, 0, (cudaStream_t)stream>>>(x, ldx, R, C, (__nv_bfloat16*)out, ldo);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_relu_cast");
    return NC_OK;
}

extern "C" int nc_col_sum(const void* x, int dtype, int64_t ld, int32_t R, int32_t C, float* out, void* stream) {
    NC_REQUIRE(R >= 0 && C > 0 && ld >= C, NC_ESHAPE, "nc_col_sum: bad shape");
    if (R == 0) return NC_OK;
    NC_REQUIRE(dtype == NC_F32 || dtype == NC_BF16, NC_EINVAL, "nc_col_sum: bad dtype %d", dtype);
    const size_t esz = dtype == NC_F32 ? 4 : 2;
    if (ld % 4 == 0 && (uintptr_t)x % (4 * esz) == 0 && C >= 64) {
        const unsigned gx = (unsigned)nc_cdiv(C, 128);
        int64_t gy = nc_cdiv(4 * (int64_t)nc_num_sms(), gx);
        gy = std::max<int64_t>(1, std::min<int64_t>(gy, nc_cdiv(R, 32)));
        dim3 grid(gx, (unsigned)gy);
        if (dtype == NC_F32) k_col_sum_v4<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)x, ld, R, C, out);
        else k_col_sum_v4<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, ld, R, C, out);
        nc_count_launch(1);
        NC_CHECK_LAUNCH("nc_col_sum");
        return NC_OK;
    }
    dim3 grid((unsigned)nc_cdiv(C, 32), (unsigned)std::min<int64_t>(nc_cdiv(R, 64), 128));
    if (dtype == NC_F32) k_col_sum<float><<<grid, 256, 0, (cudaStream_t)stream>>>((const float*)x, ld, R, C, out);
    else if (dtype == NC_BF16) k_col_sum<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)x, ld, R, C, out);
    else NC_REQUIRE(false, NC_EINVAL, "nc_col_sum: bad dtype %d", dtype);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_col_sum");
    return NC_OK;
}

extern "C" int nc_relu_bwd(const float* dh, int64_t lddh, const void* h, int dtype, int64_t ldh, int32_t R, int32_t C,
                           void* out, int64_t ldo, void* stream) {
    NC_REQUIRE(R >= 0 && C > 0 && lddh >= C && ldh >= C && ldo >= C, NC_ESHAPE, "nc_relu_bwd: bad shape");
    if (R == 0) return NC_OK;
    NC_REQUIRE(dtype == NC_F32 || dtype == NC_BF16, NC_EINVAL, "nc_relu_bwd: bad dtype %d", dtype);
    const size_t esz = dtype == NC_F32 ? 4 : 2;
    if (lddh % 4 == 0 && ldh % 4 == 0 && ldo % 4 == 0 && (uintptr_t)dh % 16 == 0 && (uintptr_t)h % (4 * esz) == 0 &&
        (uintptr_t)out % (4 * esz) == 0) {
        dim3 grid((unsigned)std::max<int64_t>(1, std::min<int64_t>(nc_cdiv(ldo / 4, 256), 8)),
                  (unsigned)std::min<int64_t>(R, 2048));
        if (dtype == NC_F32) k_relu_bwd_v4<float><<<grid, 256, 0, (cudaStream_t)stream>>>(dh, lddh, (const float*)h, ldh, R, C, (float*)out, ldo);
        else k_relu_bwd_v4<__nv_bfloat16><<<grid, 256, 0, (cudaStream_t)stream>>>(dh, lddh, (const __nv_bfloat16*)h, ldh, R, C, (__nv_bfloat16*)out, ldo);
        nc_count_launch(1);
        NC_CHECK_LAUNCH("nc_relu_bwd");
        return NC_OK;
    }
    const int g = grid_for((int64_t)R * ldo);
    if (dtype == NC_F32) k_relu_bwd<float><<<g, 256, 0, (cudaStream_t)stream>>>(dh, lddh, (const float*)h, ldh, R, C, (float*)out, ldo);
    else if (dtype == NC_BF16) k_relu_bwd<__nv_bfloat16><<<g, 256, 0, (cudaStream_t)stream>>>(dh, lddh, (const __nv_bfloat16*)h, ldh, R, C, (__nv_bfloat16*)out, ldo);
    else NC_REQUIRE(false, NC_EINVAL, "nc_relu_bwd: bad dtype %d", dtype);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_relu_bwd");
    return NC_OK;
}
