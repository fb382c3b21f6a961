// Loss-side kernels: fused softmax -> NB means -> column select -> negative-binomial NLL
// (forward value + gradients in one pass), cosine-similarity graph decoder, weighted BCE,
// KL regulariser.  Reference: nn/layers.py:143, nn/decoders.py:168, modules/vgpgae.py:714-749,
// modules/losses.py:123-166 (BCE), :169-212 (NB), :255-284 (KL), nn/utils.py:8-36 (cosine).
#include <algorithm>
#include <cooperative_groups.h>

#include "common.cuh"

extern void nc_count_launch(int n);

namespace {

// digamma for x > 0 (fp64): recurrence up to x >= 6, then the asymptotic series
__device__ double digamma_d(double x) {
    double r = 0.0;
    while (x < 6.0) { r -= 1.0 / x; x += 1.0; }
    const double f = 1.0 / (x * x);
    return r + log(x) - 0.5 / x - f * (1.0 / 12.0 - f * (1.0 / 120.0 - f * (1.0 / 252.0 - f * (1.0 / 240.0 - f * (1.0 / 132.0)))));
}

constexpr int NB_T = 512;
constexpr int NB_CL = 4;      // CTAs (one thread-block cluster) per row

// One 4-CTA cluster per row: each CTA owns a quarter of the columns, the row-wide reductions
// (max, library size, softmax denominator, D) are exchanged through distributed shared memory.
// logits: decoder pre-activations over ALL F features; sel[c] != 0 marks the reconstructed
// columns (softmax is over all F, library size and loss over selected columns).
__global__ void __launch_bounds__(NB_T)
k_nb_fused(const float* __restrict__ logits, int64_t ldl, const float* __restrict__ x, int64_t ldx,
           const uint8_t* __restrict__ sel, const float* __restrict__ theta_param, int F, float scale,
           float* __restrict__ means, int64_t ldm, double* __restrict__ loss_sum,
           float* __restrict__ dlogits, int64_t lddl, float* __restrict__ dtheta, float* __restrict__ dx,
           int64_t lddx) {
    namespace cg = cooperative_groups;
    cg::cluster_group cluster = cg::this_cluster();
    __shared__ float red[32];
    __shared__ double redd[32];
    __shared__ float xf[2][2];     // [phase parity][value] partials read by the other CTAs of the cluster
    __shared__ double xd[1];
    const int rank = (int)cluster.block_rank();
    const int r = blockIdx.x / NB_CL, tid = threadIdx.x;
    const int per = (F + NB_CL - 1) / NB_CL;
    const int c0 = rank * per, c1 = min(F, c0 + per);
    const float* lr = logits + (int64_t)r * ldl;
    const float* xr = x + (int64_t)r * ldx;
    // row max, library size
    float m = -INFINITY, lib = 0.f;
    for (int c = c0 + tid; c < c1; c += NB_T) {
        m = fmaxf(m, lr[c]);
        if (sel[c]) lib += xr[c];
    }
    m = block_max(m, red);
    lib = block_sum(lib, red);
    if (tid == 0) { xf[0][0] = m; xf[0][1] = lib; }
    cluster.sync();
    m = -INFINITY; lib = 0.f;
#pragma unroll
    for (int k = 0; k < NB_CL; ++k) {
        const float* o = cluster.map_shared_rank(&xf[0][0], k);
        m = fmaxf(m, o[0]);
        lib += o[1];
    }
    float se = 0.f;
    for (int c = c0 + tid; c < c1; c += NB_T) se += expf(lr[c] - m);
    se = block_sum(se, red);
    if (tid == 0) xf[1][0] = se;
    cluster.sync();
    se = 0.f;
#pragma unroll
    for (int k = 0; k < NB_CL; ++k) se += *cluster.map_shared_rank(&xf[1][0], k);
    const double L = (double)lib;           // exp(log(library size))
    const double inv_se = 1.0 / (double)se;
    // loss and D = sum_c dmu_c * mu_c
    double loss = 0.0, D = 0.0;
    const double eps = 1e-8;
    for (int c = c0 + tid; c < c1; c += NB_T) {
        const double p = (double)expf(lr[c] - m) * inv_se;
        const double mu = L * p;
        if (means) means[(int64_t)r * ldm + c] = (float)mu;
        if (!sel[c]) continue;
        const double xv = (double)xr[c];
        const double th = exp((double)theta_param[c]);
        const double ltm = log(th + mu + eps);
        double ll = th * (log(th + eps) - ltm) + xv * (log(mu + eps) - ltm);
        if (xv != 0.0) ll += lgamma(xv + th) - lgamma(th) - lgamma(xv + 1.0);
        loss -= ll;
        const double dmu = -(xv / (mu + eps) - (xv + th) / (th + mu + eps));   // d(-ll)/dmu
        D += dmu * mu;
        if (dtheta) {
            double dth = log(th + eps) + th / (th + eps) - ltm - (th + xv) / (th + mu + eps);
            if (xv != 0.0) dth += digamma_d(xv + th) - digamma_d(th);
            atomicAdd(dtheta + c, (float)(-(double)scale * dth * th));
        }
    }
    loss = block_sum_d(loss, redd);
    D = block_sum_d(D, redd);
    if (tid == 0) { atomicAdd(loss_sum, loss * (double)scale); xd[0] = D; }
    cluster.sync();
    D = 0.0;
#pragma unroll
    for (int k = 0; k < NB_CL; ++k) D += *cluster.map_shared_rank(&xd[0], k);
    cluster.sync();                          // nobody leaves while its shared memory may still be read
    if (dlogits) {
        float* dr = dlogits + (int64_t)r * lddl;
        for (int c = c0 + tid; c < c1; c += NB_T) {
            const double p = (double)expf(lr[c] - m) * inv_se;
            const double mu = L * p;
            double dmu = 0.0;
            if (sel[c]) {
                const double xv = (double)xr[c];
                const double th = exp((double)theta_param[c]);
                dmu = -(xv / (mu + eps) - (xv + th) / (th + mu + eps));
            }
            dr[c] = (float)((double)scale * (mu * dmu - p * D));
        }
    }
    // gradient w.r.t. the labels (only the attention node-label aggregator produces labels that carry one):
    // d(-ll_k)/dx_k directly, plus D / L through the library size L = sum_sel x that scales every mean
    if (dx) {
        float* dxr = dx + (int64_t)r * lddx;
        const double DL = D / L;
        for (int c = c0 + tid; c < c1; c += NB_T) {
            double gx = 0.0;
            if (sel[c]) {
                const double p = (double)expf(lr[c] - m) * inv_se;
                const double mu = L * p;
                const double xv = (double)xr[c];
                const double th = exp((double)theta_param[c]);
                gx = -(log(mu + eps) - log(th + mu + eps) + digamma_d(xv + th) - digamma_d(xv + 1.0)) + DL;
            }
            dxr[c] = (float)((double)scale * gx);
        }
    }
}

// ------------------------------------------------------------------------- NB loss on given means
// modules/losses.py:169-212 as the reference exports it: x, mu [R, F], theta [F] (already positive).
// loss_sum += scale * sum -ll (fp64); dmu / dtheta (nullable) = scale * d(sum -ll)/d.
__global__ void __launch_bounds__(256)
k_nb_means(const float* __restrict__ x, int64_t ldx, const float* __restrict__ mu, int64_t ldmu,
           const float* __restrict__ theta, int R, int F, float eps_f, float scale, double* __restrict__ loss_sum,
           float* __restrict__ dmu, int64_t lddmu, float* __restrict__ dtheta) {
    __shared__ double redd[32];
    const double eps = (double)eps_f;
    double loss = 0.0;
    const int64_t total = (int64_t)R * F;
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (int64_t)gridDim.x * blockDim.x) {
        const int r = (int)(q / F), c = (int)(q - (int64_t)r * F);
        const double xv = (double)x[(int64_t)r * ldx + c], m = (double)mu[(int64_t)r * ldmu + c], th = (double)theta[c];
        const double ltm = log(th + m + eps);
        loss -= th * (log(th + eps) - ltm) + xv * (log(m + eps) - ltm) + lgamma(xv + th) - lgamma(th) - lgamma(xv + 1.0);
        if (dmu) dmu[(int64_t)r * lddmu + c] = (float)(-(double)scale * (xv / (m + eps) - (xv + th) / (th + m + eps)));
        if (dtheta) {
            const double d = log(th + eps) + th / (th + eps) - ltm - (th + xv) / (th + m + eps) + digamma_d(xv + th) - digamma_d(th);
            atomicAdd(dtheta + c, (float)(-(double)scale * d));
        }
    }
    loss = block_sum_d(loss, redd);
    if (threadIdx.x == 0) atomicAdd(loss_sum, loss * (double)scale);
}

// ------------------------------------------------------------------------- cosine decoder
// pair e: a = z[e], b = z[n_pairs + e]; logits[e] = <a/max(|a|,eps), b/max(|b|,eps)>
__global__ void __launch_bounds__(128)
k_cosine_fwd(const float* __restrict__ z, int64_t ldz, int P, int n_pairs, float* __restrict__ logits,
             float* __restrict__ norms) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= n_pairs) return;
    const float* a = z + (int64_t)warp * ldz;
    const float* b = z + (int64_t)(n_pairs + warp) * ldz;
    float aa = 0.f, bb = 0.f, ab = 0.f;
    for (int k = lane; k < P; k += 32) {
        const float u = a[k], v = b[k];
        aa += u * u; bb += v * v; ab += u * v;
    }
    aa = warp_sum(aa); bb = warp_sum(bb); ab = warp_sum(ab);
    if (lane == 0) {
        const float na = fmaxf(sqrtf(aa), 1e-8f), nb = fmaxf(sqrtf(bb), 1e-8f);
        logits[warp] = ab / (na * nb);
        norms[warp] = sqrtf(aa);
        norms[n_pairs + warp] = sqrtf(bb);
    }
}

__global__ void __launch_bounds__(128)
k_cosine_bwd(const float* __restrict__ z, int64_t ldz, int P, int n_pairs, const float* __restrict__ logits,
             const float* __restrict__ norms, const float* __restrict__ dlogits, float* __restrict__ dz, int64_t lddz) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= n_pairs) return;
    const float* a = z + (int64_t)warp * ldz;
    const float* b = z + (int64_t)(n_pairs + warp) * ldz;
    float* da = dz + (int64_t)warp * lddz;
    float* db = dz + (int64_t)(n_pairs + warp) * lddz;
    const float ra = norms[warp], rb = norms[n_pairs + warp];
    const float ina = 1.0f / fmaxf(ra, 1e-8f), inb = 1.0f / fmaxf(rb, 1e-8f);
    const float g = dlogits[warp], cs = logits[warp];
    // a_hat = a / na.  If |a| > eps: d cos/da = (b_hat - cos * a_hat) / na, else b_hat / na.
    const float ca = ra > 1e-8f ? cs : 0.f, cb = rb > 1e-8f ? cs : 0.f;
    const float ga = g * ina, gb = g * inb;
#pragma unroll 4
    for (int k = lane; k < P; k += 32) {
        const float ah = a[k] * ina, bh = b[k] * inb;
        da[k] = ga * (bh - ca * ah);
        db[k] = gb * (ah - cb * bh);
    }
}

// ------------------------------------------------------------------------- weighted BCE
// Single CTA.  incl == NULL -> all included.  loss = mean_incl( pw*y*softplus(-x) + (1-y)*softplus(x) ),
// pw = #neg/#pos.  dlogits = scale * d loss / d logits (0 for excluded).
__device__ __forceinline__ float softplusf(float v) { return fmaxf(v, 0.f) + log1pf(expf(-fabsf(v))); }

__global__ void __launch_bounds__(1024)
k_bce(const float* __restrict__ logits, const float* __restrict__ labels, const uint8_t* __restrict__ incl,
      int n, float scale, float* __restrict__ loss_out, float* __restrict__ dlogits, int use_pos_weight) {
    __shared__ float red[32];
    const int tid = threadIdx.x;
    float np = 0.f, nn = 0.f;
    for (int e = tid; e < n; e += 1024) {
        if (incl && !incl[e]) continue;
        if (labels[e] == 1.f) np += 1.f;
        else if (labels[e] == 0.f) nn += 1.f;
    }
    np = block_sum(np, red);
    nn = block_sum(nn, red);
    float cnt = 0.f;
    for (int e = tid; e < n; e += 1024) cnt += (incl && !incl[e]) ? 0.f : 1.f;
    cnt = block_sum(cnt, red);
    const float pw = use_pos_weight ? nn / np : 1.f;
    float acc = 0.f;
    for (int e = tid; e < n; e += 1024) {
        const bool in = !(incl && !incl[e]);
        const float xv = logits[e], y = labels[e];
        float d = 0.f;
        if (in) {
            acc += pw * y * softplusf(-xv) + (1.f - y) * softplusf(xv);
            const float sg = 1.f / (1.f + expf(-xv));
            d = (-(pw * y) * (1.f - sg) + (1.f - y) * sg) / cnt;
        }
        if (dlogits) dlogits[e] = scale * d;
    }
    acc = block_sum(acc, red);
    if (tid == 0) *loss_out = acc / cnt;
}

// ------------------------------------------------------------------------- KL
// kl = -0.5 * mean_rows sum_cols (1 + 2 ls - mu^2 - exp(ls)^2); grads scaled by `scale`.
__global__ void __launch_bounds__(256)
k_kl(const float* __restrict__ mu, int64_t ldmu, const float* __restrict__ ls, int64_t ldls, int R, int P,
     float scale, double* __restrict__ sum_out, float* __restrict__ dmu, int64_t lddmu, float* __restrict__ dls,
     int64_t lddls) {
    __shared__ double redd[32];
    const int64_t total = (int64_t)R * P;
    double acc = 0.0;
    for (int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; q < total; q += (int64_t)gridDim.x * blockDim.x) {
        const int r = (int)(q / P), c = (int)(q % P);
        const float m = mu[(int64_t)r * ldmu + c], l = ls[(int64_t)r * ldls + c];
        const float e2 = expf(l) * expf(l);
        acc += (double)(1.f + 2.f * l - m * m - e2);
        if (dmu) {
            dmu[(int64_t)r * lddmu + c] = scale * m / (float)R;
            dls[(int64_t)r * lddls + c] = scale * (e2 - 1.f) / (float)R;
        }
    }
    acc = block_sum_d(acc, redd);
    if (threadIdx.x == 0) atomicAdd(sum_out, -0.5 * acc / (double)R);
}

}  // namespace

extern "C" int nc_nb_loss_fused(const float* logits, int64_t ldl, const float* x, int64_t ldx, const uint8_t* sel,
                                const float* theta_param, int32_t n_rows, int32_t F, float scale, float* means,
                                int64_t ldm, double* loss_sum, float* dlogits, int64_t lddl, float* dtheta,
                                float* dx, int64_t lddx, void* stream) {
    NC_REQUIRE(n_rows >= 0 && F > 0 && ldl >= F && ldx >= F, NC_ESHAPE, "nc_nb_loss_fused: bad shape rows=%d F=%d", n_rows, F);
    if (n_rows == 0) return NC_OK;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)n_rows * NB_CL);
    cfg.blockDim = dim3(NB_T);
    cfg.dynamicSmemBytes = 0;
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = NB_CL;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t le = cudaLaunchKernelEx(&cfg, k_nb_fused, logits, ldl, x, ldx, sel, theta_param, (int)F, scale, means,
                                        ldm, loss_sum, dlogits, lddl, dtheta, dx, lddx);
    if (le != cudaSuccess) { nc_set_error("nc_nb_loss_fused: launch failed: %s", cudaGetErrorString(le)); return NC_ECUDA; }
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_nb_loss_fused");
    return NC_OK;
}

extern "C" int nc_cosine_fwd(const float* z, int64_t ldz, int32_t P, int32_t n_pairs, float* logits, float* norms,
                             void* stream) {
    NC_REQUIRE(n_pairs >= 0 && P > 0 && ldz >= P, NC_ESHAPE, "nc_cosine_fwd: bad shape");
    if (n_pairs == 0) return NC_OK;
    k_cosine_fwd<<<(unsigned)nc_cdiv(n_pairs, 4), 128, 0, (cudaStream_t)stream>>>(z, ldz, P, n_pairs, logits, norms);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_cosine_fwd");
    return NC_OK;
}

extern "C" int nc_cosine_bwd(const float* z, int64_t ldz, int32_t P, int32_t n_pairs, const float* logits,
                             const float* norms, const float* dlogits, float* dz, int64_t lddz, void* stream) {
    NC_REQUIRE(n_pairs >= 0 && P > 0 && ldz >= P && lddz >= P, NC_ESHAPE, "nc_cosine_bwd: bad shape");
    if (n_pairs == 0) return NC_OK;
    k_cosine_bwd<<<(unsigned)nc_cdiv(n_pairs, 4), 128, 0, (cudaStream_t)stream>>>(z, ldz, P, n_pairs, logits, norms,
                                                                                  dlogits, dz, lddz);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_cosine_bwd");
    return NC_OK;
}

extern "C" int nc_bce_fwd_bwd(const float* logits, const float* labels, const uint8_t* incl, int32_t n, float scale,
                              float* loss_out, float* dlogits, int use_pos_weight, void* stream) {
    NC_REQUIRE(n > 0, NC_ESHAPE, "nc_bce_fwd_bwd: empty input");
    k_bce<<<1, 1024, 0, (cudaStream_t)stream>>>(logits, labels, incl, n, scale, loss_out, dlogits, use_pos_weight);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_bce_fwd_bwd");
    return NC_OK;
}

extern "C" int nc_kl_fwd_bwd(const float* mu, int64_t ldmu, const float* logstd, int64_t ldls, int32_t R, int32_t P,
                             float scale, double* sum_out, float* dmu, int64_t lddmu, float* dlogstd, int64_t lddls,
                             void* stream) {
    NC_REQUIRE(R > 0 && P > 0, NC_ESHAPE, "nc_kl_fwd_bwd: empty input");
    const int64_t total = (int64_t)R * P;
    int blocks = (int)std::min<int64_t>(nc_cdiv(total, 256), 4 * nc_num_sms());
    k_kl<<<blocks, 256, 0, (cudaStream_t)stream>>>(mu, ldmu, logstd, ldls, R, P, scale, sum_out, dmu, lddmu, dlogstd, lddls);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_kl_fwd_bwd");
    return NC_OK;
}

extern "C" int nc_nb_loss_means(const float* x, int64_t ldx, const float* mu, int64_t ldmu, const float* theta,
                                int32_t n_rows, int32_t F, float eps, float scale, double* loss_sum, float* dmu,
                                int64_t lddmu, float* dtheta, void* stream) {
    NC_REQUIRE(n_rows > 0 && F > 0 && ldx >= F && ldmu >= F && (dmu == nullptr || lddmu >= F), NC_ESHAPE,
               "nc_nb_loss_means: bad shape n_rows=%d F=%d", n_rows, F);
    const int64_t total = (int64_t)n_rows * F;
    int blocks = (int)std::min<int64_t>(nc_cdiv(total, 256), 4 * nc_num_sms());
    k_nb_means<<<blocks, 256, 0, (cudaStream_t)stream>>>(x, ldx, mu, ldmu, theta, n_rows, F, eps, scale, loss_sum, dmu,
                                                         lddmu, dtheta);
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_nb_loss_means");
    return NC_OK;
}
