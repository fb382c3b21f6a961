// Fused GATv2 edge kernels (forward + backward) for the encoder's four attention convs.
// Replaces, per conv, PyG GATv2Conv's index_select x2 -> leaky_relu -> mul+sum -> segment
// softmax (scatter max / exp / scatter sum / div) -> mul -> scatter-add -> head mean -> + bias
// (nn/encoders.py:124-140,197-207) without materialising any [E,H,C] tensor.
//
// Layout: xl [n_src, ld], xr [n_dst, ld]: the outputs of lin_l / lin_r of ALL convs side by
// side; conv c, head h occupies columns [off_c + h*C_c, off_c + (h+1)*C_c).  `att` uses the
// same column indexing.  Self loops: CSR entries with col == i are ignored and exactly one
// self edge is appended per target (PyG remove_self_loops + add_self_loops).
// Slot of CSR entry e of row i: e + i; slot of the self edge: rowptr[i+1] + i.
#include "common.cuh"

extern void nc_count_launch(int n);

namespace {
constexpr int GT = 128;

struct Seg { int off, C, out_off; };
__device__ __forceinline__ Seg load_seg(const int32_t* seg, int c) {
    Seg s;
    s.off = seg[c * 4 + 0];
    s.C = seg[c * 4 + 1];
    s.out_off = seg[c * 4 + 2];
    return s;
}

template <typename T, bool VEC>
__device__ __forceinline__ float4 ldv(const T* p, int k, int C) {
    if (VEC) return Elem<T>::ld4(p + k);
    float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
    if (k < C) o.x = Elem<T>::ld(p + k);
    if (k + 1 < C) o.y = Elem<T>::ld(p + k + 1);
    if (k + 2 < C) o.z = Elem<T>::ld(p + k + 2);
    if (k + 3 < C) o.w = Elem<T>::ld(p + k + 3);
    return o;
}
template <bool VEC>
__device__ __forceinline__ float4 ldf(const float* p, int k, int C) {
    if (VEC) return __ldg(reinterpret_cast<const float4*>(p + k));
    float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
    if (k < C) o.x = __ldg(p + k);
    if (k + 1 < C) o.y = __ldg(p + k + 1);
    if (k + 2 < C) o.z = __ldg(p + k + 2);
    if (k + 3 < C) o.w = __ldg(p + k + 3);
    return o;
}
template <typename T, bool VEC>
__device__ __forceinline__ void stv(T* p, int k, int C, float4 v) {
    if (VEC) { Elem<T>::st4(p + k, v); return; }
    if (k < C) Elem<T>::st(p + k, v.x);
    if (k + 1 < C) Elem<T>::st(p + k + 1, v.y);
    if (k + 2 < C) Elem<T>::st(p + k + 2, v.z);
    if (k + 3 < C) Elem<T>::st(p + k + 3, v.w);
}

// ------------------------------------------------------------------------- forward
template <typename T, bool VEC>
__global__ void __launch_bounds__(GT)
k_gatv2_fwd(const T* __restrict__ xl, const T* __restrict__ xr, int64_t ld, const float* __restrict__ att,
            const float* __restrict__ bias, const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
            const int32_t* __restrict__ seg, int n_conv, int H, int max_deg, float slope,
            float* __restrict__ out, int64_t ldo, float* __restrict__ alpha) {
    extern __shared__ float s_mem[];
    const int stride = max_deg + 1;
    float* s_sc = s_mem;                               // [H][stride]
    int* s_j = (int*)(s_mem + H * stride);             // [stride]
    const int i = blockIdx.x, c = blockIdx.y;
    const Seg sg = load_seg(seg, c);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int e0 = rowptr[i], e1 = rowptr[i + 1];
    const int nE = e1 - e0 + 1;
    const int NCH = n_conv * H;
    for (int t = tid; t < nE; t += GT) s_j[t] = (t < nE - 1) ? col[e0 + t] : i;
    __syncthreads();
    const T* xri = xr + (int64_t)i * ld + sg.off;
    // pass 1: raw attention scores, one warp per incoming edge
    for (int t = warp; t < nE; t += GT / 32) {
        const int j = s_j[t];
        const bool skip = (t < nE - 1) && (j == i);
        const T* xlj = xl + (int64_t)j * ld + sg.off;
        for (int h = 0; h < H; ++h) {
            float p = 0.f;
            if (!skip) {
                const int base = h * sg.C;
                for (int k = lane * 4; k < sg.C; k += 128) {
                    float4 a = ldv<T, VEC>(xlj + base, k, sg.C);
                    float4 b = ldv<T, VEC>(xri + base, k, sg.C);
                    float4 w = ldf<VEC>(att + sg.off + base, k, sg.C);
                    p += w.x * leaky_relu(a.x + b.x, slope) + w.y * leaky_relu(a.y + b.y, slope) +
                         w.z * leaky_relu(a.z + b.z, slope) + w.w * leaky_relu(a.w + b.w, slope);
                }
                p = warp_sum(p);
            }
            if (lane == 0) s_sc[h * stride + t] = skip ? -INFINITY : p;
        }
    }
    __syncthreads();
    // segment softmax over the incoming edges, one warp per head
    for (int h = warp; h < H; h += GT / 32) {
        float m = -INFINITY;
        for (int t = lane; t < nE; t += 32) m = fmaxf(m, s_sc[h * stride + t]);
        m = warp_max(m);
        float s = 0.f;
        for (int t = lane; t < nE; t += 32) s += expf(s_sc[h * stride + t] - m);
        s = warp_sum(s) + 1e-16f;
        for (int t = lane; t < nE; t += 32) {
            const float a = expf(s_sc[h * stride + t] - m) / s;
            s_sc[h * stride + t] = a;
            alpha[(int64_t)(e0 + i + t) * NCH + c * H + h] = a;
        }
    }
    __syncthreads();
    // pass 2: weighted sum of x_l rows, mean over heads, + bias
    const float invH = 1.0f / (float)H;
    for (int k = tid * 4; k < sg.C; k += GT * 4) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int h = 0; h < H; ++h) {
            const int base = sg.off + h * sg.C;
            for (int t = 0; t < nE; ++t) {
                const float a = s_sc[h * stride + t];
                if (a == 0.f) continue;
                float4 v = ldv<T, VEC>(xl + (int64_t)s_j[t] * ld + base, k, sg.C);
                acc.x += a * v.x; acc.y += a * v.y; acc.z += a * v.z; acc.w += a * v.w;
            }
        }
        float4 b = make_float4(0.f, 0.f, 0.f, 0.f);
        if (bias) b = ldf<false>(bias + sg.out_off, k, sg.C);
        acc.x = acc.x * invH + b.x; acc.y = acc.y * invH + b.y;
        acc.z = acc.z * invH + b.z; acc.w = acc.w * invH + b.w;
        stv<float, false>(out + (int64_t)i * ldo + sg.out_off, k, sg.C, acc);
    }
}

// ------------------------------------------------------------------------- backward, per target
template <typename T, bool VEC>
__global__ void __launch_bounds__(GT)
k_gatv2_bwd_dst(const T* __restrict__ xl, const T* __restrict__ xr, int64_t ld, const float* __restrict__ att,
                const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
                const int32_t* __restrict__ seg, int n_conv, int H, int max_deg, float slope,
                const float* __restrict__ dout, int64_t ldo, const float* __restrict__ alpha,
                float* __restrict__ ds, T* __restrict__ dxr, float* __restrict__ datt, float* __restrict__ dbias) {
    extern __shared__ float s_mem[];
    const int stride = max_deg + 1;
    float* s_al = s_mem;                               // [H][stride] alpha
    float* s_ds = s_mem + H * stride;                  // [H][stride] d alpha -> d score
    int* s_j = (int*)(s_mem + 2 * H * stride);
    const int i = blockIdx.x, c = blockIdx.y;
    const Seg sg = load_seg(seg, c);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int e0 = rowptr[i], e1 = rowptr[i + 1];
    const int nE = e1 - e0 + 1;
    const int NCH = n_conv * H;
    const float invH = 1.0f / (float)H;
    for (int t = tid; t < nE; t += GT) s_j[t] = (t < nE - 1) ? col[e0 + t] : i;
    for (int q = tid; q < nE * H; q += GT) {
        const int t = q / H, h = q % H;
        s_al[h * stride + t] = alpha[(int64_t)(e0 + i + t) * NCH + c * H + h];
    }
    __syncthreads();
    const float* g = dout + (int64_t)i * ldo + sg.out_off;
    // d alpha[t,h] = (1/H) <g, x_l[j_t, h, :]>
    for (int t = warp; t < nE; t += GT / 32) {
        const T* xlj = xl + (int64_t)s_j[t] * ld + sg.off;
        for (int h = 0; h < H; ++h) {
            float p = 0.f;
            if (s_al[h * stride + t] != 0.f) {
                for (int k = lane * 4; k < sg.C; k += 128) {
                    float4 a = ldv<T, VEC>(xlj + h * sg.C, k, sg.C);
                    float4 gg = ldf<false>(g, k, sg.C);
                    p += a.x * gg.x + a.y * gg.y + a.z * gg.z + a.w * gg.w;
                }
            }
            p = warp_sum(p);
            if (lane == 0) s_ds[h * stride + t] = p * invH;
        }
    }
    __syncthreads();
    // softmax backward: ds = alpha * (dalpha - sum_t alpha * dalpha)
    for (int h = warp; h < H; h += GT / 32) {
        float dot = 0.f;
        for (int t = lane; t < nE; t += 32) dot += s_al[h * stride + t] * s_ds[h * stride + t];
        dot = warp_sum(dot);
        for (int t = lane; t < nE; t += 32) {
            const float d = s_al[h * stride + t] * (s_ds[h * stride + t] - dot);
            s_ds[h * stride + t] = d;
            ds[(int64_t)(e0 + i + t) * NCH + c * H + h] = d;
        }
    }
    __syncthreads();
    // d x_r[i], d att, d bias
    const T* xri = xr + (int64_t)i * ld + sg.off;
    for (int k = tid * 4; k < sg.C; k += GT * 4) {
        for (int h = 0; h < H; ++h) {
            const int base = h * sg.C;
            const float4 b = ldv<T, VEC>(xri + base, k, sg.C);
            const float4 w = ldf<VEC>(att + sg.off + base, k, sg.C);
            float4 ar = make_float4(0.f, 0.f, 0.f, 0.f), aa = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int t = 0; t < nE; ++t) {
                if (s_al[h * stride + t] == 0.f) continue;
                const float d = s_ds[h * stride + t];
                float4 a = ldv<T, VEC>(xl + (int64_t)s_j[t] * ld + sg.off + base, k, sg.C);
                float v;
                v = a.x + b.x; ar.x += d * (v > 0.f ? 1.f : slope); aa.x += d * leaky_relu(v, slope);
                v = a.y + b.y; ar.y += d * (v > 0.f ? 1.f : slope); aa.y += d * leaky_relu(v, slope);
                v = a.z + b.z; ar.z += d * (v > 0.f ? 1.f : slope); aa.z += d * leaky_relu(v, slope);
                v = a.w + b.w; ar.w += d * (v > 0.f ? 1.f : slope); aa.w += d * leaky_relu(v, slope);
            }
            ar.x *= w.x; ar.y *= w.y; ar.z *= w.z; ar.w *= w.w;
            stv<T, VEC>(dxr + (int64_t)i * ld + sg.off + base, k, sg.C, ar);
            float* da = datt + sg.off + base;
            if (k < sg.C) atomicAdd(da + k, aa.x);
            if (k + 1 < sg.C) atomicAdd(da + k + 1, aa.y);
            if (k + 2 < sg.C) atomicAdd(da + k + 2, aa.z);
            if (k + 3 < sg.C) atomicAdd(da + k + 3, aa.w);
        }
        if (dbias) {
            float4 gg = ldf<false>(g, k, sg.C);
            float* db = dbias + sg.out_off;
            if (k < sg.C) atomicAdd(db + k, gg.x);
            if (k + 1 < sg.C) atomicAdd(db + k + 1, gg.y);
            if (k + 2 < sg.C) atomicAdd(db + k + 2, gg.z);
            if (k + 3 < sg.C) atomicAdd(db + k + 3, gg.w);
        }
    }
}

// ------------------------------------------------------------------------- backward, per source
constexpr int SRC_STAGE = 64;
template <typename T, bool VEC>
__global__ void __launch_bounds__(GT)
k_gatv2_bwd_src(const T* __restrict__ xl, const T* __restrict__ xr, int64_t ld, const float* __restrict__ att,
                const int32_t* __restrict__ rptr, const int32_t* __restrict__ rdst, const int32_t* __restrict__ rslot,
                const int32_t* __restrict__ seg, int n_conv, int H, float slope,
                const float* __restrict__ dout, int64_t ldo, const float* __restrict__ alpha,
                const float* __restrict__ ds, T* __restrict__ dxl) {
    __shared__ int s_i[SRC_STAGE];
    __shared__ int s_slot[SRC_STAGE];
    const int j = blockIdx.x, c = blockIdx.y;
    const Seg sg = load_seg(seg, c);
    const int tid = threadIdx.x;
    const int p0 = rptr[j], p1 = rptr[j + 1];
    const int NCH = n_conv * H;
    const float invH = 1.0f / (float)H;
    const T* xlj = xl + (int64_t)j * ld + sg.off;
    T* dj = dxl + (int64_t)j * ld + sg.off;
    if (p1 - p0 <= SRC_STAGE) {
        for (int t = tid; t < p1 - p0; t += GT) { s_i[t] = rdst[p0 + t]; s_slot[t] = rslot[p0 + t]; }
        __syncthreads();
    }
    const bool staged = (p1 - p0 <= SRC_STAGE);
    for (int k = tid * 4; k < sg.C; k += GT * 4) {
        for (int h = 0; h < H; ++h) {
            const int base = h * sg.C;
            const float4 a = ldv<T, VEC>(xlj + base, k, sg.C);
            const float4 w = ldf<VEC>(att + sg.off + base, k, sg.C);
            float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int p = p0; p < p1; ++p) {
                const int i = staged ? s_i[p - p0] : rdst[p];
                const int slot = staged ? s_slot[p - p0] : rslot[p];
                const float al = __ldg(alpha + (int64_t)slot * NCH + c * H + h) * invH;
                const float d = __ldg(ds + (int64_t)slot * NCH + c * H + h);
                const float4 gg = ldf<false>(dout + (int64_t)i * ldo + sg.out_off, k, sg.C);
                const float4 b = ldv<T, VEC>(xr + (int64_t)i * ld + sg.off + base, k, sg.C);
                acc.x += al * gg.x + d * w.x * ((a.x + b.x) > 0.f ? 1.f : slope);
                acc.y += al * gg.y + d * w.y * ((a.y + b.y) > 0.f ? 1.f : slope);
                acc.z += al * gg.z + d * w.z * ((a.z + b.z) > 0.f ? 1.f : slope);
                acc.w += al * gg.w + d * w.w * ((a.w + b.w) > 0.f ? 1.f : slope);
            }
            stv<T, VEC>(dj + base, k, sg.C, acc);
        }
    }
}

bool vec_ok(const void* a, const void* b, int64_t ld, const int32_t* seg_host, int n_conv, size_t esz) {
    if (ld % 4 != 0) return false;
    if ((uintptr_t)a % (esz * 4) != 0 || (uintptr_t)b % (esz * 4) != 0) return false;
    for (int c = 0; c < n_conv; ++c)
        if (seg_host[c * 4] % 4 != 0 || seg_host[c * 4 + 1] % 4 != 0) return false;
    return true;
}
}  // namespace

extern "C" int nc_gatv2_fwd(const void* xl, const void* xr, int dtype, int64_t ld, const float* att,
                            const float* bias, const int32_t* rowptr, const int32_t* col, int32_t n_dst,
                            const int32_t* seg, const int32_t* seg_host, int32_t n_conv, int32_t H,
                            int32_t max_deg, float neg_slope, float* out, int64_t ldo, float* alpha,
                            void* stream) {
    NC_REQUIRE(n_dst >= 0 && n_conv > 0 && H > 0 && max_deg >= 0, NC_ESHAPE, "nc_gatv2_fwd: bad sizes");
    NC_REQUIRE(dtype == NC_F32 || dtype == NC_BF16, NC_EINVAL, "nc_gatv2_fwd: bad dtype %d", dtype);
    if (n_dst == 0) return NC_OK;
    const size_t smem = ((size_t)H * (max_deg + 1) + (max_deg + 1)) * 4;
    NC_REQUIRE(smem <= 48 * 1024, NC_EUNSUPPORTED, "nc_gatv2_fwd: max_deg=%d too large", max_deg);
    cudaStream_t s = (cudaStream_t)stream;
    dim3 grid(n_dst, n_conv);
    const bool v = vec_ok(xl, xr, ld, seg_host, n_conv, dtype == NC_F32 ? 4 : 2) && ((uintptr_t)att % 16 == 0);
#define L(T, V) k_gatv2_fwd<T, V><<<grid, GT, smem, s>>>((const T*)xl, (const T*)xr, ld, att, bias, rowptr, col, seg, n_conv, H, max_deg, neg_slope, out, ldo, alpha)
    if (dtype == NC_F32) { if (v) L(float, true); else L(float, false); }
    else { if (v) L(__nv_bfloat16, true); else L(__nv_bfloat16, false); }
#undef L
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_gatv2_fwd");
    return NC_OK;
}

extern "C" int nc_gatv2_bwd(const void* xl, const void* xr, int dtype, int64_t ld, const float* att,
                            const int32_t* rowptr, const int32_t* col, int32_t n_dst, int32_t n_src,
                            const int32_t* rptr, const int32_t* rdst, const int32_t* rslot,
                            const int32_t* seg, const int32_t* seg_host, int32_t n_conv, int32_t H,
                            int32_t max_deg, float neg_slope, const float* dout, int64_t ldo,
                            const float* alpha, float* ds_ws, void* dxl, void* dxr, float* datt,
                            float* dbias, void* stream) {
    NC_REQUIRE(n_dst >= 0 && n_src >= n_dst && n_conv > 0 && H > 0, NC_ESHAPE, "nc_gatv2_bwd: bad sizes");
    NC_REQUIRE(dtype == NC_F32 || dtype == NC_BF16, NC_EINVAL, "nc_gatv2_bwd: bad dtype %d", dtype);
    if (n_src == 0) return NC_OK;
    const size_t smem = ((size_t)2 * H * (max_deg + 1) + (max_deg + 1)) * 4;
    NC_REQUIRE(smem <= 48 * 1024, NC_EUNSUPPORTED, "nc_gatv2_bwd: max_deg=%d too large", max_deg);
    cudaStream_t s = (cudaStream_t)stream;
    const bool v = vec_ok(xl, xr, ld, seg_host, n_conv, dtype == NC_F32 ? 4 : 2) && ((uintptr_t)att % 16 == 0) &&
                   ((uintptr_t)dxl % 16 == 0) && ((uintptr_t)dxr % 16 == 0);
    dim3 gd(n_dst, n_conv), gs(n_src, n_conv);
#define LD(T, V) if (n_dst > 0) k_gatv2_bwd_dst<T, V><<<gd, GT, smem, s>>>((const T*)xl, (const T*)xr, ld, att, rowptr, col, seg, n_conv, H, max_deg, neg_slope, dout, ldo, alpha, ds_ws, (T*)dxr, datt, dbias)
#define LS(T, V) k_gatv2_bwd_src<T, V><<<gs, GT, 0, s>>>((const T*)xl, (const T*)xr, ld, att, rptr, rdst, rslot, seg, n_conv, H, neg_slope, dout, ldo, alpha, ds_ws, (T*)dxl)
    if (dtype == NC_F32) { if (v) { LD(float, true); LS(float, true); } else { LD(float, false); LS(float, false); } }
    else { if (v) { LD(__nv_bfloat16, true); LS(__nv_bfloat16, true); } else { LD(__nv_bfloat16, false); LS(__nv_bfloat16, false); } }
#undef LD
#undef LS
    nc_count_launch(2);
    NC_CHECK_LAUNCH("nc_gatv2_bwd");
    return NC_OK;
}
