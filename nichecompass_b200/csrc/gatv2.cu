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
// gatv2_ring.cu: the hot-path (bf16, vectorisable) kernels; return 1 when the shape is not theirs
int gat_ring_fwd(const void* xl, const void* xr, int64_t ld, const float* att, const float* bias,
                 const int32_t* rowptr, const int32_t* col, int32_t dst_off, int32_t n_dst, const int32_t* seg,
                 const int32_t* seg_host, int32_t n_conv, int32_t H, int32_t max_deg, float slope, float* out,
                 int64_t ldo, float* alpha, float* head_ws, void* stream);
int64_t gat_ring_bwd_ws_bytes(int64_t n_slots, int64_t n_dst, int n_conv, int H);
int gat_ring_bwd(const void* xl, const void* xr, int64_t ld, const float* att, const int32_t* rowptr,
                 const int32_t* col, int32_t n_dst, int32_t n_src, const int32_t* rptr, const int32_t* rdst,
                 const int32_t* rslot, const int32_t* seg, const int32_t* seg_host, int32_t n_conv, int32_t H,
                 int32_t max_deg, float slope, const float* dout, int64_t ldo, const float* alpha, float* ds_ws,
                 void* dxl, void* dxr, float* datt, float* dbias, const float* head_ws, void* bwd_ws,
                 int64_t n_slots, void* stream);

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
            float* __restrict__ out, int64_t ldo, float* __restrict__ alpha, int self_mode, int dst_off) {
    // self_mode 1: PyG GATv2Conv (existing self entries dropped, one self edge appended per target);
    // self_mode 0: edges exactly as given (attention node-label aggregator); out == NULL: alpha only.
    // dst_off: targets are rows [dst_off, dst_off + gridDim.x) of rowptr / xl; xr and out hold the tile
    // only (row blockIdx.x).  alpha == NULL (inference tiles): coefficients are not stored.
    extern __shared__ float s_mem[];
    const int stride = max_deg + 1;
    float* s_sc = s_mem;                               // [H][stride]
    int* s_j = (int*)(s_mem + H * stride);             // [stride]
    const int ir = blockIdx.x, i = ir + dst_off, c = blockIdx.y;
    const Seg sg = load_seg(seg, c);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int e0 = rowptr[i], e1 = rowptr[i + 1];
    const int nE = e1 - e0 + self_mode;
    const int NCH = n_conv * H;
    for (int t = tid; t < nE; t += GT) s_j[t] = (t < e1 - e0) ? col[e0 + t] : i;
    if (!self_mode && alpha && tid < H) alpha[(int64_t)(e1 + i) * NCH + c * H + tid] = 0.f;   // unused self slot
    __syncthreads();
    const T* xri = xr + (int64_t)ir * ld + sg.off;
    // pass 1: raw attention scores, one warp per incoming edge
    for (int t = warp; t < nE; t += GT / 32) {
        const int j = s_j[t];
        const bool skip = self_mode && (t < e1 - e0) && (j == i);
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
            if (alpha) alpha[(int64_t)(e0 + i + t) * NCH + c * H + h] = a;
        }
    }
    __syncthreads();
    if (out == nullptr) return;
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
        stv<float, false>(out + (int64_t)ir * ldo + sg.out_off, k, sg.C, acc);
    }
}

// ------------------------------------------------------------------------- forward, streaming (bf16)
// One CTA per (target, conv).  Every neighbour's x_l segment of this conv (H*C contiguous bf16) is
// pulled ONCE from HBM by a bulk async copy (cp.async.bulk -> mbarrier) into a 4-deep shared-memory
// ring and consumed from there: score (LeakyReLU + att dot, reduced over the NT/H threads of a head),
// online softmax (running max / sum with rescaling of the accumulators), alpha-weighted sum.
// Thread t works for head t / (NT/H) and owns NV 4-element vectors of that head's C columns.
constexpr int FS_NT = 256;
constexpr int FS_STAGES = 4;

__device__ __forceinline__ uint32_t fs_smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void fs_mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(fs_smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fs_mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(fs_smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void fs_mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "FS_WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra.uni FS_WAIT_DONE;\n\t"
        "bra.uni FS_WAIT_LOOP;\n\t"
        "FS_WAIT_DONE:\n\t"
        "}" ::"r"(fs_smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void fs_bulk_load(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(fs_smem_u32(dst)), "l"(src), "r"(bytes), "r"(fs_smem_u32(bar)) : "memory");
}
__device__ __forceinline__ float4 bf4_to_f4(uint2 r) {
    float4 o;
    o.x = __uint_as_float(r.x << 16);
    o.y = __uint_as_float(r.x & 0xffff0000u);
    o.z = __uint_as_float(r.y << 16);
    o.w = __uint_as_float(r.y & 0xffff0000u);
    return o;
}

template <int NV>
__global__ void __launch_bounds__(FS_NT, 2)
k_gatv2_fwd_stream(const __nv_bfloat16* __restrict__ xl, const __nv_bfloat16* __restrict__ xr, int64_t ld,
                   const float* __restrict__ att, const float* __restrict__ bias, const int32_t* __restrict__ rowptr,
                   const int32_t* __restrict__ col, const int32_t* __restrict__ seg, int n_conv, int H, int max_deg,
                   int seg_stride, float slope, float* __restrict__ out, int64_t ldo, float* __restrict__ alpha,
                   int dst_off) {
    extern __shared__ __align__(128) uint8_t fs_smem[];
    const int stride = max_deg + 1;
    uint8_t* ring = fs_smem;                                              // FS_STAGES * seg_stride bytes
    float* s_e = (float*)(ring + (size_t)FS_STAGES * seg_stride);         // [H][stride] raw scores
    int* s_j = (int*)(s_e + H * stride);                                  // [stride] compacted source ids
    int* s_t = s_j + stride;                                              // [stride] their entry index
    float* s_red = (float*)(s_t + stride);                                // [2][FS_NT/32]
    uint64_t* full = (uint64_t*)(((uintptr_t)(s_red + 2 * (FS_NT / 32)) + 7) & ~(uintptr_t)7);
    __shared__ int s_nv;

    const int ir = blockIdx.x, i = ir + dst_off, c = blockIdx.y;   // see k_gatv2_fwd for dst_off / NULL alpha
    const Seg sg = load_seg(seg, c);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int TPH = FS_NT / H, h = tid / TPH, lt = tid - h * TPH;         // TPH is a multiple of 32
    const int WPH = TPH / 32;                                             // warps per head
    const int e0 = rowptr[i], e1 = rowptr[i + 1];
    const int nE = e1 - e0 + 1;
    const int NCH = n_conv * H;
    const int C4 = sg.C >> 2;
    const uint32_t seg_bytes = (uint32_t)(H * sg.C * 2);

    if (tid == 0) {
        int n = 0;
        for (int t = 0; t < nE; ++t) {
            const int j = (t < nE - 1) ? col[e0 + t] : i;
            if (t < nE - 1 && j == i) continue;                           // PyG removes existing self loops
            s_j[n] = j; s_t[n] = t; ++n;
        }
        s_nv = n;
        for (int s = 0; s < FS_STAGES; ++s) fs_mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    for (int q = tid; q < H * stride; q += FS_NT) s_e[q] = -INFINITY;
    __syncthreads();
    const int nV = s_nv;
    if (tid == 0) {
        const int pre = nV < FS_STAGES ? nV : FS_STAGES;
        for (int n = 0; n < pre; ++n) {
            fs_mbar_expect_tx(&full[n], seg_bytes);
            fs_bulk_load(ring + (size_t)n * seg_stride, xl + (int64_t)s_j[n] * ld + sg.off, seg_bytes, &full[n]);
        }
    }
    // per-thread constants: x_r[i] and att of the owned vectors
    float4 xr4[NV], at4[NV], acc[NV];
    const __nv_bfloat16* xri = xr + (int64_t)ir * ld + sg.off + h * sg.C;
    const float* atth = att + sg.off + h * sg.C;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        const int u = lt + k * TPH;
        acc[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (u < C4) {
            xr4[k] = bf4_to_f4(__ldg(reinterpret_cast<const uint2*>(xri + 4 * u)));
            at4[k] = __ldg(reinterpret_cast<const float4*>(atth + 4 * u));
        } else {
            xr4[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            at4[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
    }
    float m_run = -INFINITY, l_run = 0.f;
    for (int it = 0; it < nV; ++it) {
        const int s = it % FS_STAGES;
        fs_mbar_wait(&full[s], (uint32_t)((it / FS_STAGES) & 1));
        const uint8_t* row = ring + (size_t)s * seg_stride + (size_t)h * sg.C * 2;
        float4 a[NV];
        float p = 0.f;
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            const int u = lt + k * TPH;
            if (u < C4) {
                a[k] = bf4_to_f4(*reinterpret_cast<const uint2*>(row + 8 * u));
                p += at4[k].x * leaky_relu(a[k].x + xr4[k].x, slope) + at4[k].y * leaky_relu(a[k].y + xr4[k].y, slope) +
                     at4[k].z * leaky_relu(a[k].z + xr4[k].z, slope) + at4[k].w * leaky_relu(a[k].w + xr4[k].w, slope);
            } else {
                a[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
        p = warp_sum(p);
        float* red = s_red + (it & 1) * (FS_NT / 32);
        if (lane == 0) red[warp] = p;
        __syncthreads();                       // scores visible; every thread has finished reading stage s
        if (tid == 0 && it + FS_STAGES < nV) {
            const int n = it + FS_STAGES;
            fs_mbar_expect_tx(&full[s], seg_bytes);
            fs_bulk_load(ring + (size_t)s * seg_stride, xl + (int64_t)s_j[n] * ld + sg.off, seg_bytes, &full[s]);
        }
        float e = 0.f;
        for (int w = 0; w < WPH; ++w) e += red[h * WPH + w];
        if (lt == 0) s_e[h * stride + s_t[it]] = e;
        const float m_new = fmaxf(m_run, e);
        const float corr = expf(m_run - m_new);          // 0 on the first neighbour (m_run = -inf)
        const float wgt = expf(e - m_new);
        l_run = l_run * corr + wgt;
        m_run = m_new;
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            acc[k].x = acc[k].x * corr + wgt * a[k].x;
            acc[k].y = acc[k].y * corr + wgt * a[k].y;
            acc[k].z = acc[k].z * corr + wgt * a[k].z;
            acc[k].w = acc[k].w * corr + wgt * a[k].w;
        }
    }
    __syncthreads();                           // s_e complete, ring free for the head-mean staging
    const float inv_l = 1.0f / (l_run + 1e-16f);
    // attention coefficients of every slot of this target (0 for the removed self entries)
    if (alpha)
        for (int t = lt; t < nE; t += TPH)
            alpha[(int64_t)(e0 + i + t) * NCH + c * H + h] = expf(s_e[h * stride + t] - m_run) * inv_l;
    float* s_o = (float*)ring;                 // [H][C]
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        const int u = lt + k * TPH;
        if (u < C4)
            *reinterpret_cast<float4*>(s_o + h * sg.C + 4 * u) =
                make_float4(acc[k].x * inv_l, acc[k].y * inv_l, acc[k].z * inv_l, acc[k].w * inv_l);
    }
    __syncthreads();
    const float invH = 1.0f / (float)H;
    for (int k = tid; k < sg.C; k += FS_NT) {
        float v = 0.f;
        for (int hh = 0; hh < H; ++hh) v += s_o[hh * sg.C + k];
        out[(int64_t)ir * ldo + sg.out_off + k] = v * invH + (bias ? __ldg(bias + sg.out_off + k) : 0.f);
    }
}

// ------------------------------------------------------------------------- backward, per target
template <typename T, bool VEC>
__global__ void __launch_bounds__(GT)
k_gatv2_bwd_dst(const T* __restrict__ xl, const T* __restrict__ xr, int64_t ld, const float* __restrict__ att,
                const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col,
                const int32_t* __restrict__ seg, int n_conv, int H, int max_deg, float slope,
                const float* __restrict__ dout, int64_t ldo, const float* __restrict__ alpha,
                float* __restrict__ ds, T* __restrict__ dxr, float* __restrict__ datt, float* __restrict__ dbias,
                int self_mode, const float* __restrict__ dalpha_ext) {
    // dalpha_ext != NULL: d alpha is supplied per slot (attention aggregator: the messages are not x_l)
    extern __shared__ float s_mem[];
    const int stride = max_deg + 1;
    float* s_al = s_mem;                               // [H][stride] alpha
    float* s_ds = s_mem + H * stride;                  // [H][stride] d alpha -> d score
    int* s_j = (int*)(s_mem + 2 * H * stride);
    const int i = blockIdx.x, c = blockIdx.y;
    const Seg sg = load_seg(seg, c);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int e0 = rowptr[i], e1 = rowptr[i + 1];
    const int nE = e1 - e0 + self_mode;
    const int NCH = n_conv * H;
    const float invH = 1.0f / (float)H;
    for (int t = tid; t < nE; t += GT) s_j[t] = (t < e1 - e0) ? col[e0 + t] : i;
    for (int q = tid; q < nE * H; q += GT) {
        const int t = q / H, h = q % H;
        s_al[h * stride + t] = alpha[(int64_t)(e0 + i + t) * NCH + c * H + h];
        if (dalpha_ext) s_ds[h * stride + t] = dalpha_ext[(int64_t)(e0 + i + t) * NCH + c * H + h];
    }
    if (!self_mode && tid < H) ds[(int64_t)(e1 + i) * NCH + c * H + tid] = 0.f;       // unused self slot
    __syncthreads();
    const float* g = dout ? dout + (int64_t)i * ldo + sg.out_off : nullptr;
    // d alpha[t,h] = (1/H) <g, x_l[j_t, h, :]>
    for (int t = warp; t < nE && !dalpha_ext; t += GT / 32) {
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
        if (dbias && g) {
            float4 gg = ldf<false>(g, k, sg.C);
            float* db = dbias + sg.out_off;
            if (k < sg.C) atomicAdd(db + k, gg.x);
            if (k + 1 < sg.C) atomicAdd(db + k + 1, gg.y);
            if (k + 2 < sg.C) atomicAdd(db + k + 2, gg.z);
            if (k + 3 < sg.C) atomicAdd(db + k + 3, gg.w);
        }
    }
}

// ------------------------------------------------------------------------- backward per target, streaming (bf16)
// One CTA per (chunk of TCH consecutive targets, conv, head group of HG heads); 256 threads, two CTAs
// per SM.  Every neighbour's segment of the head group is streamed ONCE through the shared-memory
// ring; the softmax backward
//   ds_t = alpha_t (dalpha_t - D),  D = sum_t alpha_t dalpha_t
// is folded into running sums so no second pass over the rows is needed:
//   d x_r = att * (A1 - D A2),  d att += B1 - D B2,
//   A1 = sum_t alpha_t dalpha_t s'_t, A2 = sum_t alpha_t s'_t, B1/B2 the same with s'_t (x_l[t] + x_r),
// s'_t = LeakyReLU'(x_l[t] + x_r).  d att is kept in registers over the chunk: one atomicAdd per
// element and chunk.
constexpr int BS_NT = 256;

template <int NV>
__global__ void __launch_bounds__(BS_NT, 2)
k_gatv2_bwd_dst_stream(const __nv_bfloat16* __restrict__ xl, const __nv_bfloat16* __restrict__ xr, int64_t ld,
                       const float* __restrict__ att, const int32_t* __restrict__ rowptr,
                       const int32_t* __restrict__ col, const int32_t* __restrict__ seg, int n_conv, int H, int HG,
                       int TCH, int n_dst, int max_deg, int seg_stride, float slope,
                       const float* __restrict__ dout, int64_t ldo, const float* __restrict__ alpha,
                       float* __restrict__ ds, __nv_bfloat16* __restrict__ dxr, float* __restrict__ datt) {
    extern __shared__ __align__(128) uint8_t fs_smem[];
    const int stride = max_deg + 1;
    uint8_t* ring = fs_smem;
    float* s_al = (float*)(ring + (size_t)FS_STAGES * seg_stride);        // [HG][stride] alpha
    float* s_da = s_al + HG * stride;                                     // [HG][stride] d alpha
    int* s_j = (int*)(s_da + HG * stride);                                // compacted source ids
    int* s_t = s_j + stride;                                              // their entry index
    int* s_raw = s_t + stride;                                            // raw source ids of the row
    float* s_red = (float*)(s_raw + stride);                              // [2][BS_NT/32]
    uint64_t* full = (uint64_t*)(((uintptr_t)(s_red + 2 * (BS_NT / 32)) + 7) & ~(uintptr_t)7);
    __shared__ int s_nv;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n_hg = H / HG;
    const int c = blockIdx.y / n_hg, hg = blockIdx.y - c * n_hg;
    const Seg sg = load_seg(seg, c);
    const int TPH = BS_NT / HG, hl = tid / TPH, lt = tid - hl * TPH;      // hl: head inside the group
    const int h = hg * HG + hl;
    const int WPH = TPH / 32;
    const int NCH = n_conv * H;
    const float invH = 1.0f / (float)H;
    const int C4 = sg.C >> 2;
    const uint32_t seg_bytes = (uint32_t)(HG * sg.C * 2);
    const int64_t seg_off = (int64_t)sg.off + (int64_t)hg * HG * sg.C;    // first element of the group's segment
    if (tid == 0) {
        for (int s = 0; s < FS_STAGES; ++s) fs_mbar_init(&full[s], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    float4 at4[NV], dacc[NV];
    const float* atth = att + sg.off + h * sg.C;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        const int u = lt + k * TPH;
        at4[k] = (u < C4) ? __ldg(reinterpret_cast<const float4*>(atth + 4 * u)) : make_float4(0.f, 0.f, 0.f, 0.f);
        dacc[k] = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    uint32_t it_glob = 0;                      // ring position runs on across the targets of the chunk
    const int i0 = blockIdx.x * TCH, i1 = min(n_dst, i0 + TCH);
    for (int i = i0; i < i1; ++i) {
        const int e0 = rowptr[i], e1 = rowptr[i + 1];
        const int nE = e1 - e0 + 1;
        __syncthreads();                       // previous target fully consumed; barrier init visible
        for (int t = tid; t < nE; t += BS_NT) s_raw[t] = (t < nE - 1) ? col[e0 + t] : i;
        for (int q = tid; q < nE * HG; q += BS_NT) {
            const int t = q / HG, hh = q - t * HG;
            s_al[hh * stride + t] = alpha[(int64_t)(e0 + i + t) * NCH + c * H + hg * HG + hh];
            s_da[hh * stride + t] = 0.f;
        }
        float4 xr4[NV], g4[NV], A1[NV], A2[NV], B1[NV], B2[NV];
        const __nv_bfloat16* xri = xr + (int64_t)i * ld + sg.off + h * sg.C;
        const float* gi = dout + (int64_t)i * ldo + sg.out_off;
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            const int u = lt + k * TPH;
            A1[k] = A2[k] = B1[k] = B2[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (u < C4) {
                xr4[k] = bf4_to_f4(__ldg(reinterpret_cast<const uint2*>(xri + 4 * u)));
                g4[k] = __ldg(reinterpret_cast<const float4*>(gi + 4 * u));
            } else {
                xr4[k] = g4[k] = make_float4(0.f, 0.f, 0.f, 0.f);
            }
        }
        __syncthreads();
        if (tid == 0) {
            int n = 0;
            for (int t = 0; t < nE; ++t) {
                const int j = s_raw[t];
                if (t < nE - 1 && j == i) continue;          // PyG removes existing self loops
                s_j[n] = j; s_t[n] = t; ++n;
            }
            s_nv = n;
            const int pre = n < FS_STAGES ? n : FS_STAGES;
            for (int q = 0; q < pre; ++q) {
                const int s = (int)((it_glob + q) % FS_STAGES);
                fs_mbar_expect_tx(&full[s], seg_bytes);
                fs_bulk_load(ring + (size_t)s * seg_stride, xl + (int64_t)s_j[q] * ld + seg_off, seg_bytes, &full[s]);
            }
        }
        __syncthreads();
        const int nV = s_nv;
        float D = 0.f;
        for (int it = 0; it < nV; ++it, ++it_glob) {
            const int s = (int)(it_glob % FS_STAGES);
            fs_mbar_wait(&full[s], (it_glob / FS_STAGES) & 1);
            const uint8_t* row = ring + (size_t)s * seg_stride + (size_t)hl * sg.C * 2;
            float4 a[NV];
            float p = 0.f;
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                const int u = lt + k * TPH;
                if (u < C4) {
                    a[k] = bf4_to_f4(*reinterpret_cast<const uint2*>(row + 8 * u));
                    p += a[k].x * g4[k].x + a[k].y * g4[k].y + a[k].z * g4[k].z + a[k].w * g4[k].w;
                } else {
                    a[k] = make_float4(0.f, 0.f, 0.f, 0.f);
                }
            }
            p = warp_sum(p);
            float* red = s_red + (it_glob & 1) * (BS_NT / 32);
            if (lane == 0) red[warp] = p;
            __syncthreads();
            if (tid == 0 && it + FS_STAGES < nV) {
                const int n = it + FS_STAGES;
                fs_mbar_expect_tx(&full[s], seg_bytes);
                fs_bulk_load(ring + (size_t)s * seg_stride, xl + (int64_t)s_j[n] * ld + seg_off, seg_bytes, &full[s]);
            }
            float da = 0.f;
            for (int w = 0; w < WPH; ++w) da += red[hl * WPH + w];
            da *= invH;
            const int t = s_t[it];
            const float al = s_al[hl * stride + t];
            if (lt == 0) s_da[hl * stride + t] = da;
            const float w1a = al * da;
            D += w1a;
#pragma unroll
            for (int k = 0; k < NV; ++k) {
                float v, sp;
                v = a[k].x + xr4[k].x; sp = v > 0.f ? 1.f : slope; A1[k].x += w1a * sp; A2[k].x += al * sp; B1[k].x += w1a * sp * v; B2[k].x += al * sp * v;
                v = a[k].y + xr4[k].y; sp = v > 0.f ? 1.f : slope; A1[k].y += w1a * sp; A2[k].y += al * sp; B1[k].y += w1a * sp * v; B2[k].y += al * sp * v;
                v = a[k].z + xr4[k].z; sp = v > 0.f ? 1.f : slope; A1[k].z += w1a * sp; A2[k].z += al * sp; B1[k].z += w1a * sp * v; B2[k].z += al * sp * v;
                v = a[k].w + xr4[k].w; sp = v > 0.f ? 1.f : slope; A1[k].w += w1a * sp; A2[k].w += al * sp; B1[k].w += w1a * sp * v; B2[k].w += al * sp * v;
            }
        }
        __nv_bfloat16* dxi = dxr + (int64_t)i * ld + sg.off + h * sg.C;
#pragma unroll
        for (int k = 0; k < NV; ++k) {
            const int u = lt + k * TPH;
            if (u < C4) {
                float4 r;
                r.x = at4[k].x * (A1[k].x - D * A2[k].x); r.y = at4[k].y * (A1[k].y - D * A2[k].y);
                r.z = at4[k].z * (A1[k].z - D * A2[k].z); r.w = at4[k].w * (A1[k].w - D * A2[k].w);
                Elem<__nv_bfloat16>::st4(dxi + 4 * u, r);
                dacc[k].x += B1[k].x - D * B2[k].x; dacc[k].y += B1[k].y - D * B2[k].y;
                dacc[k].z += B1[k].z - D * B2[k].z; dacc[k].w += B1[k].w - D * B2[k].w;
            }
        }
        __syncthreads();                       // s_da complete
        for (int t = lt; t < nE; t += TPH) {
            const float al = s_al[hl * stride + t];
            ds[(int64_t)(e0 + i + t) * NCH + c * H + h] = al * (s_da[hl * stride + t] - D);
        }
    }
    float* da = datt + sg.off + h * sg.C;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
        const int u = lt + k * TPH;
        if (u < C4) {
            atomicAdd(da + 4 * u, dacc[k].x); atomicAdd(da + 4 * u + 1, dacc[k].y);
            atomicAdd(da + 4 * u + 2, dacc[k].z); atomicAdd(da + 4 * u + 3, dacc[k].w);
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
                const float4 gg = dout ? ldf<false>(dout + (int64_t)i * ldo + sg.out_off, k, sg.C)
                                       : make_float4(0.f, 0.f, 0.f, 0.f);
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
// every conv's x_l segment can be fetched by cp.async.bulk: 16-byte aligned start, 16-byte multiple length,
// and the head <-> thread mapping of the streaming kernels needs 32*H | 256
bool stream_ok(const void* xl, int64_t ld, const int32_t* seg_host, int n_conv, int H) {
    if ((uintptr_t)xl % 16 != 0 || ld % 8 != 0) return false;
    if (!(H == 1 || H == 2 || H == 4 || H == 8)) return false;
    for (int c = 0; c < n_conv; ++c)
        if (seg_host[c * 4] % 8 != 0 || (H * seg_host[c * 4 + 1]) % 8 != 0) return false;
    return true;
}
}  // namespace

static int gatv2_fwd_impl(const void* xl, const void* xr, int dtype, int64_t ld, const float* att,
                          const float* bias, const int32_t* rowptr, const int32_t* col, int32_t dst_off, int32_t n_dst,
                          const int32_t* seg, const int32_t* seg_host, int32_t n_conv, int32_t H,
                          int32_t max_deg, float neg_slope, float* out, int64_t ldo, float* alpha,
                          float* head_ws, void* stream) {
    NC_REQUIRE(n_dst >= 0 && n_conv > 0 && H > 0 && max_deg >= 0, NC_ESHAPE, "nc_gatv2_fwd: bad sizes");
    NC_REQUIRE(dtype == NC_F32 || dtype == NC_BF16, NC_EINVAL, "nc_gatv2_fwd: bad dtype %d", dtype);
    if (n_dst == 0) return NC_OK;
    if (dtype == NC_BF16 && head_ws != nullptr && out != nullptr) {
        const int rc = gat_ring_fwd(xl, xr, ld, att, bias, rowptr, col, dst_off, n_dst, seg, seg_host, n_conv, H,
                                    max_deg, neg_slope, out, ldo, alpha, head_ws, stream);
        if (rc <= 0) return rc;
    }
    const size_t smem = ((size_t)H * (max_deg + 1) + (max_deg + 1)) * 4;
    NC_REQUIRE(smem <= 48 * 1024, NC_EUNSUPPORTED, "nc_gatv2_fwd: max_deg=%d too large", max_deg);
    cudaStream_t s = (cudaStream_t)stream;
    dim3 grid(n_dst, n_conv);
    const bool v = vec_ok(xl, xr, ld, seg_host, n_conv, dtype == NC_F32 ? 4 : 2) && ((uintptr_t)att % 16 == 0);
    // streaming kernel: bf16 rows, every conv segment a legal bulk-copy source (16-byte aligned, 16-byte multiple)
    if (dtype == NC_BF16 && v && stream_ok(xl, ld, seg_host, n_conv, H)) {
        int maxC = 0;
        for (int c = 0; c < n_conv; ++c) maxC = seg_host[c * 4 + 1] > maxC ? seg_host[c * 4 + 1] : maxC;
        const int TPH = FS_NT / H;
        const int nv = (maxC / 4 + TPH - 1) / TPH;
        const int seg_stride = (H * maxC * 2 + 127) / 128 * 128;
        const size_t sm = (size_t)FS_STAGES * seg_stride + ((size_t)H * (max_deg + 1) + 2 * (max_deg + 1) + 2 * (FS_NT / 32)) * 4 + 8 + FS_STAGES * 8;
        if (nv <= 8 && sm <= 100 * 1024) {
#define LS(NV)                                                                                                       \
    do {                                                                                                             \
        static bool attr_set[64] = {false};   /* one flag array per NV: the macro body is expanded per case */ \
        if (int rc_ = nc_set_max_smem((const void*)k_gatv2_fwd_stream<NV>, 100 * 1024, attr_set, "k_gatv2_fwd_stream")) return rc_; \
        k_gatv2_fwd_stream<NV><<<grid, FS_NT, sm, s>>>((const __nv_bfloat16*)xl, (const __nv_bfloat16*)xr, ld, att, \
                                                       bias, rowptr, col, seg, n_conv, H, max_deg, seg_stride,       \
                                                       neg_slope, out, ldo, alpha, dst_off);                         \
    } while (0)
            if (nv <= 2) LS(2); else if (nv <= 4) LS(4); else if (nv <= 6) LS(6); else LS(8);
#undef LS
            nc_count_launch(1);
            NC_CHECK_LAUNCH("nc_gatv2_fwd");
            return NC_OK;
        }
    }
#define L(T, V) k_gatv2_fwd<T, V><<<grid, GT, smem, s>>>((const T*)xl, (const T*)xr, ld, att, bias, rowptr, col, seg, n_conv, H, max_deg, neg_slope, out, ldo, alpha, 1, dst_off)
    if (dtype == NC_F32) { if (v) L(float, true); else L(float, false); }
    else { if (v) L(__nv_bfloat16, true); else L(__nv_bfloat16, false); }
#undef L
    nc_count_launch(1);
    NC_CHECK_LAUNCH("nc_gatv2_fwd");
    return NC_OK;
}

extern "C" int nc_gatv2_fwd(const void* xl, const void* xr, int dtype, int64_t ld, const float* att,
                            const float* bias, const int32_t* rowptr, const int32_t* col, int32_t n_dst,
                            const int32_t* seg, const int32_t* seg_host, int32_t n_conv, int32_t H,
                            int32_t max_deg, float neg_slope, float* out, int64_t ldo, float* alpha,
                            float* head_ws, void* stream) {
    NC_REQUIRE(alpha != nullptr, NC_EINVAL, "nc_gatv2_fwd: alpha must be given (the backward consumes it)");
    return gatv2_fwd_impl(xl, xr, dtype, ld, att, bias, rowptr, col, 0, n_dst, seg, seg_host, n_conv, H, max_deg,
                          neg_slope, out, ldo, alpha, head_ws, stream);
}

// Inference over a tile of targets of a large resident graph: targets are rows
// [dst_off, dst_off + n_dst) of rowptr (and of xl, for the self edge); xr and out hold the tile only.
// No attention coefficients are stored.
extern "C" int nc_gatv2_fwd_tile(const void* xl, const void* xr, int dtype, int64_t ld, const float* att,
                                 const float* bias, const int32_t* rowptr, const int32_t* col, int32_t dst_off,
                                 int32_t n_dst, const int32_t* seg, const int32_t* seg_host, int32_t n_conv,
                                 int32_t H, int32_t max_deg, float neg_slope, float* out, int64_t ldo,
                                 float* head_ws, void* stream) {
    NC_REQUIRE(dst_off >= 0 && out != nullptr, NC_EINVAL, "nc_gatv2_fwd_tile: bad dst_off / out");
    return gatv2_fwd_impl(xl, xr, dtype, ld, att, bias, rowptr, col, dst_off, n_dst, seg, seg_host, n_conv, H,
                          max_deg, neg_slope, out, ldo, nullptr, head_ws, stream);
}

extern "C" int nc_gatv2_bwd(const void* xl, const void* xr, int dtype, int64_t ld, const float* att,
                            const int32_t* rowptr, const int32_t* col, int32_t n_dst, int32_t n_src,
                            const int32_t* rptr, const int32_t* rdst, const int32_t* rslot,
                            const int32_t* seg, const int32_t* seg_host, int32_t n_conv, int32_t H,
                            int32_t max_deg, float neg_slope, const float* dout, int64_t ldo,
                            const float* alpha, float* ds_ws, void* dxl, void* dxr, float* datt,
                            float* dbias, const float* head_ws, void* bwd_ws, int64_t bwd_ws_bytes,
                            int64_t n_slots, void* stream) {
    NC_REQUIRE(n_dst >= 0 && n_src >= n_dst && n_conv > 0 && H > 0, NC_ESHAPE, "nc_gatv2_bwd: bad sizes");
    NC_REQUIRE(dtype == NC_F32 || dtype == NC_BF16, NC_EINVAL, "nc_gatv2_bwd: bad dtype %d", dtype);
    if (n_src == 0) return NC_OK;
    if (dtype == NC_BF16 && head_ws != nullptr && bwd_ws != nullptr && n_dst > 0 && dout != nullptr &&
        bwd_ws_bytes >= gat_ring_bwd_ws_bytes(n_slots, n_dst, n_conv, H)) {
        const int rc = gat_ring_bwd(xl, xr, ld, att, rowptr, col, n_dst, n_src, rptr, rdst, rslot, seg, seg_host,
                                    n_conv, H, max_deg, neg_slope, dout, ldo, alpha, ds_ws, dxl, dxr, datt, dbias,
                                    head_ws, bwd_ws, n_slots, stream);
        if (rc <= 0) return rc;
    }
    const size_t smem = ((size_t)2 * H * (max_deg + 1) + (max_deg + 1)) * 4;
    NC_REQUIRE(smem <= 48 * 1024, NC_EUNSUPPORTED, "nc_gatv2_bwd: max_deg=%d too large", max_deg);
    cudaStream_t s = (cudaStream_t)stream;
    const bool v = vec_ok(xl, xr, ld, seg_host, n_conv, dtype == NC_F32 ? 4 : 2) && ((uintptr_t)att % 16 == 0) &&
                   ((uintptr_t)dxl % 16 == 0) && ((uintptr_t)dxr % 16 == 0);
    dim3 gd(n_dst, n_conv), gs(n_src, n_conv);
    // streaming per-target kernel (bf16 rows, bulk-copyable segments); d bias = column sums of dout
    const int HG = H >= 2 ? 2 : 1;
    bool bs_ok = dtype == NC_BF16 && v && n_dst > 0 && stream_ok(xl, ld, seg_host, n_conv, H) && H % HG == 0 &&
                 ldo % 4 == 0 && (uintptr_t)dout % 16 == 0;
    for (int c = 0; c < n_conv && bs_ok; ++c)
        bs_ok = (HG * seg_host[c * 4 + 1]) % 8 == 0 && seg_host[c * 4 + 2] % 4 == 0;
    if (bs_ok) {
        int maxC = 0;
        for (int c = 0; c < n_conv; ++c) maxC = seg_host[c * 4 + 1] > maxC ? seg_host[c * 4 + 1] : maxC;
        const int TPH = BS_NT / HG;
        const int nv = (maxC / 4 + TPH - 1) / TPH;
        const int seg_stride = (HG * maxC * 2 + 127) / 128 * 128;
        const size_t sm = (size_t)FS_STAGES * seg_stride + ((size_t)2 * HG * (max_deg + 1) + 3 * (max_deg + 1) + 2 * (BS_NT / 32)) * 4 + 8 + FS_STAGES * 8;
        if (nv <= 4 && sm <= 100 * 1024) {
            const int TCH = n_dst >= 512 ? 4 : 1;
            dim3 grid((unsigned)nc_cdiv(n_dst, TCH), (unsigned)(n_conv * (H / HG)));
#define LBS(NV)                                                                                                      \
    do {                                                                                                             \
        static bool attr_set[64] = {false};   /* one flag array per NV: the macro body is expanded per case */ \
        if (int rc_ = nc_set_max_smem((const void*)k_gatv2_bwd_dst_stream<NV>, 100 * 1024, attr_set, "k_gatv2_bwd_dst_stream")) return rc_; \
        k_gatv2_bwd_dst_stream<NV><<<grid, BS_NT, sm, s>>>((const __nv_bfloat16*)xl, (const __nv_bfloat16*)xr, ld,  \
            att, rowptr, col, seg, n_conv, H, HG, TCH, n_dst, max_deg, seg_stride, neg_slope, dout, ldo, alpha,      \
            ds_ws, (__nv_bfloat16*)dxr, datt);                                                                       \
    } while (0)
            if (nv <= 1) LBS(1); else if (nv <= 2) LBS(2); else if (nv <= 3) LBS(3); else LBS(4);
#undef LBS
            nc_count_launch(1);
            NC_CHECK_LAUNCH("nc_gatv2_bwd (dst)");
            if (dbias) {
                int ow = 0;
                for (int c = 0; c < n_conv; ++c) {
                    const int e = seg_host[c * 4 + 2] + seg_host[c * 4 + 1];
                    ow = e > ow ? e : ow;
                }
                int rc = nc_col_sum(dout, NC_F32, ldo, n_dst, ow, dbias, stream);
                if (rc) return rc;
            }
            k_gatv2_bwd_src<__nv_bfloat16, true><<<gs, GT, 0, s>>>((const __nv_bfloat16*)xl, (const __nv_bfloat16*)xr, ld,
                att, rptr, rdst, rslot, seg, n_conv, H, neg_slope, dout, ldo, alpha, ds_ws, (__nv_bfloat16*)dxl);
            nc_count_launch(1);
            NC_CHECK_LAUNCH("nc_gatv2_bwd (src)");
            return NC_OK;
        }
    }
#define LD(T, V) if (n_dst > 0) k_gatv2_bwd_dst<T, V><<<gd, GT, smem, s>>>((const T*)xl, (const T*)xr, ld, att, rowptr, col, seg, n_conv, H, max_deg, neg_slope, dout, ldo, alpha, ds_ws, (T*)dxr, datt, dbias, 1, nullptr)
#define LS(T, V) k_gatv2_bwd_src<T, V><<<gs, GT, 0, s>>>((const T*)xl, (const T*)xr, ld, att, rptr, rdst, rslot, seg, n_conv, H, neg_slope, dout, ldo, alpha, ds_ws, (T*)dxl)
    if (dtype == NC_F32) { if (v) { LD(float, true); LS(float, true); } else { LD(float, false); LS(float, false); } }
    else { if (v) { LD(__nv_bfloat16, true); LS(__nv_bfloat16, true); } else { LD(__nv_bfloat16, false); LS(__nv_bfloat16, false); } }
#undef LD
#undef LS
    nc_count_launch(2);
    NC_CHECK_LAUNCH("nc_gatv2_bwd");
    return NC_OK;
}

// ============================================================================ attention node-label aggregator
// nn/aggregators.py:113-156: scores from g_l = linear_l_l(x), g_r = linear_r_l(x),
// messages are the RAW feature rows, identical for every head, so
//   out[i] = mean_h sum_j alpha_ijh x[j] = sum_j abar_ij x[j],  abar = mean_h alpha
// is the CSR SpMM of spmm.cu with abar as edge weights; the score / softmax machinery is the GATv2 one
// with the edges exactly as sampled (no self-loop re-normalisation).
namespace {
// w[e] = mean_h alpha[slot(e), h] for the CSR entries of the first n_dst rows (slot = e + row)
__global__ void k_agg_mean_alpha(const float* __restrict__ alpha, const int32_t* __restrict__ rowptr, int n_dst, int H,
                                 float* __restrict__ w) {
    const int i = blockIdx.x;
    const float invH = 1.0f / (float)H;
    for (int e = rowptr[i] + threadIdx.x; e < rowptr[i + 1]; e += blockDim.x) {
        float a = 0.f;
        for (int h = 0; h < H; ++h) a += alpha[(int64_t)(e + i) * H + h];
        w[e] = a * invH;
    }
}
// d alpha[slot(e), h] = (1/H) <dout[i], x[col[e]]>, one warp per entry
__global__ void __launch_bounds__(128)
k_agg_dalpha(const float* __restrict__ dout, int64_t ldo, const float* __restrict__ x, int64_t ldx, int F,
             const int32_t* __restrict__ rowptr, const int32_t* __restrict__ col, int n_dst, int H,
             float* __restrict__ dalpha) {
    const int i = blockIdx.x, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float* g = dout + (int64_t)i * ldo;
    const float invH = 1.0f / (float)H;
    for (int e = rowptr[i] + warp; e < rowptr[i + 1]; e += 4) {
        const float* xr = x + (int64_t)col[e] * ldx;
        float p = 0.f;
        for (int k = lane; k < F; k += 32) p += __ldg(g + k) * __ldg(xr + k);
        p = warp_sum(p) * invH;
        if (lane < H) dalpha[(int64_t)(e + i) * H + lane] = p;
    }
    if (threadIdx.x < H) dalpha[(int64_t)(rowptr[i + 1] + i) * H + threadIdx.x] = 0.f;   // unused self slot
}
}  // namespace

extern "C" int64_t nc_gatv2_bwd_ws_bytes(int64_t n_slots, int64_t n_dst, int32_t n_conv, int32_t H) {
    return gat_ring_bwd_ws_bytes(n_slots, n_dst, n_conv, H);
}

extern "C" int nc_attn_agg_fwd(const void* gl, const void* gr, int dtype, int64_t ld, const float* attn,
                               const int32_t* rowptr, const int32_t* col, int32_t n_dst, const int32_t* seg,
                               int32_t H, int32_t F, int32_t max_deg, float neg_slope, float* alpha, float* w,
                               void* stream) {
    NC_REQUIRE(n_dst >= 0 && H > 0 && H <= 32 && F > 0 && max_deg >= 0, NC_ESHAPE, "nc_attn_agg_fwd: bad sizes");
    NC_REQUIRE(dtype == NC_F32 || dtype == NC_BF16, NC_EINVAL, "nc_attn_agg_fwd: bad dtype %d", dtype);
    if (n_dst == 0) return NC_OK;
    const size_t smem = ((size_t)H * (max_deg + 1) + (max_deg + 1)) * 4;
    NC_REQUIRE(smem <= 48 * 1024, NC_EUNSUPPORTED, "nc_attn_agg_fwd: max_deg=%d too large", max_deg);
    cudaStream_t s = (cudaStream_t)stream;
    const int32_t seg_host[4] = {0, F, 0, 0};
    dim3 grid(n_dst, 1);
    const bool v = vec_ok(gl, gr, ld, seg_host, 1, dtype == NC_F32 ? 4 : 2) && ((uintptr_t)attn % 16 == 0);
#define L(T, V) k_gatv2_fwd<T, V><<<grid, GT, smem, s>>>((const T*)gl, (const T*)gr, ld, attn, nullptr, rowptr, col, seg, 1, H, max_deg, neg_slope, nullptr, 0, alpha, 0, 0)
    if (dtype == NC_F32) { if (v) L(float, true); else L(float, false); }
    else { if (v) L(__nv_bfloat16, true); else L(__nv_bfloat16, false); }
#undef L
    k_agg_mean_alpha<<<n_dst, 32, 0, s>>>(alpha, rowptr, n_dst, H, w);
    nc_count_launch(2);
    NC_CHECK_LAUNCH("nc_attn_agg_fwd");
    return NC_OK;
}

extern "C" int nc_attn_agg_bwd(const void* gl, const void* gr, int dtype, int64_t ld, const float* attn,
                               const int32_t* rowptr, const int32_t* col, int32_t n_dst, int32_t n_src,
                               const int32_t* rptr, const int32_t* rdst, const int32_t* rslot, const int32_t* seg,
                               int32_t H, int32_t F, int32_t max_deg, float neg_slope, const float* dout,
                               int64_t ldo, const float* x, int64_t ldx, const float* alpha, float* dalpha_ws,
                               float* ds_ws, void* dgl, void* dgr, float* dattn, void* stream) {
    NC_REQUIRE(n_dst >= 0 && n_src >= n_dst && H > 0 && H <= 32 && F > 0, NC_ESHAPE, "nc_attn_agg_bwd: bad sizes");
    NC_REQUIRE(dtype == NC_F32 || dtype == NC_BF16, NC_EINVAL, "nc_attn_agg_bwd: bad dtype %d", dtype);
    if (n_src == 0) return NC_OK;
    const size_t smem = ((size_t)2 * H * (max_deg + 1) + (max_deg + 1)) * 4;
    NC_REQUIRE(smem <= 48 * 1024, NC_EUNSUPPORTED, "nc_attn_agg_bwd: max_deg=%d too large", max_deg);
    cudaStream_t s = (cudaStream_t)stream;
    const int32_t seg_host[4] = {0, F, 0, 0};
    const bool v = vec_ok(gl, gr, ld, seg_host, 1, dtype == NC_F32 ? 4 : 2) && ((uintptr_t)attn % 16 == 0) &&
                   ((uintptr_t)dgl % 16 == 0) && ((uintptr_t)dgr % 16 == 0);
    dim3 gd(n_dst, 1), gs(n_src, 1);
    if (n_dst > 0) k_agg_dalpha<<<n_dst, 128, 0, s>>>(dout, ldo, x, ldx, F, rowptr, col, n_dst, H, dalpha_ws);
#define LD(T, V) if (n_dst > 0) k_gatv2_bwd_dst<T, V><<<gd, GT, smem, s>>>((const T*)gl, (const T*)gr, ld, attn, rowptr, col, seg, 1, H, max_deg, neg_slope, nullptr, 0, alpha, ds_ws, (T*)dgr, dattn, nullptr, 0, dalpha_ws)
#define LS(T, V) k_gatv2_bwd_src<T, V><<<gs, GT, 0, s>>>((const T*)gl, (const T*)gr, ld, attn, rptr, rdst, rslot, seg, 1, H, neg_slope, nullptr, 0, alpha, ds_ws, (T*)dgl)
    if (dtype == NC_F32) { if (v) { LD(float, true); LS(float, true); } else { LD(float, false); LS(float, false); } }
    else { if (v) { LD(__nv_bfloat16, true); LS(__nv_bfloat16, true); } else { LD(__nv_bfloat16, false); LS(__nv_bfloat16, false); } }
#undef LD
#undef LS
    nc_count_launch(3);
    NC_CHECK_LAUNCH("nc_attn_agg_bwd");
    return NC_OK;
}
