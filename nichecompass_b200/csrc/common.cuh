// Shared device/host helpers for the nichecompass_b200 kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

#include "../../include/nichecompass_b200.h"

// ---------------------------------------------------------------------------- errors
void nc_set_error(const char* fmt, ...);   // capi.cu
// once per (kernel, device), result checked; `done_per_device`: bool[64] private to ONE kernel (capi.cu)
int nc_set_max_smem(const void* func, int bytes, bool* done_per_device, const char* name);

#define NC_REQUIRE(cond, code, ...)                \
    do {                                           \
        if (!(cond)) {                             \
            nc_set_error(__VA_ARGS__);             \
            return (code);                         \
        }                                          \
    } while (0)

#define NC_CHECK_LAUNCH(name)                                                        \
    do {                                                                             \
        cudaError_t e__ = cudaGetLastError();                                        \
        if (e__ != cudaSuccess) {                                                    \
            nc_set_error("%s: CUDA error %d (%s)", name, (int)e__,                   \
                         cudaGetErrorString(e__));                                   \
            return NC_ECUDA;                                                         \
        }                                                                            \
    } while (0)

static inline int nc_num_sms() {
    static int sms = 0;
    if (sms == 0) {
        int dev = 0;
        if (cudaGetDevice(&dev) != cudaSuccess ||
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0)
            sms = 148;
    }
    return sms;
}

static inline int64_t nc_cdiv(int64_t a, int64_t b) { return (a + b - 1) / b; }

// ---------------------------------------------------------------------------- element access
// Rows are stored either as fp32 or bf16; kernels compute in fp32.
template <typename T> struct Elem;
template <> struct Elem<float> {
    static __device__ __forceinline__ float ld(const float* p) { return __ldg(p); }
    static __device__ __forceinline__ void st(float* p, float v) { *p = v; }
    // 4 consecutive elements, pointer must be 16 B aligned
    static __device__ __forceinline__ float4 ld4(const float* p) {
        return __ldg(reinterpret_cast<const float4*>(p));
    }
    static __device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
};
template <> struct Elem<__nv_bfloat16> {
    static __device__ __forceinline__ float ld(const __nv_bfloat16* p) {
        return __bfloat162float(__ldg(p));
    }
    static __device__ __forceinline__ void st(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
    // 4 consecutive elements, pointer must be 8 B aligned
    static __device__ __forceinline__ float4 ld4(const __nv_bfloat16* p) {
        uint2 r = __ldg(reinterpret_cast<const uint2*>(p));
        float4 o;
        o.x = __uint_as_float(r.x << 16);
        o.y = __uint_as_float(r.x & 0xffff0000u);
        o.z = __uint_as_float(r.y << 16);
        o.w = __uint_as_float(r.y & 0xffff0000u);
        return o;
    }
    static __device__ __forceinline__ void st4(__nv_bfloat16* p, float4 v) {
        __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y);
        __nv_bfloat162 b = __floats2bfloat162_rn(v.z, v.w);
        uint2 r;
        r.x = *reinterpret_cast<uint32_t*>(&a);
        r.y = *reinterpret_cast<uint32_t*>(&b);
        *reinterpret_cast<uint2*>(p) = r;
    }
};

// ---------------------------------------------------------------------------- reductions
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Block-wide sum (all threads get the result). `red` : >= 32 floats of shared memory.
__device__ __forceinline__ float block_sum(float v, float* red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float r = (lane < nw) ? red[lane] : 0.f;
    r = warp_sum(r);
    return r;
}
__device__ __forceinline__ float block_max(float v, float* red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    float r = (lane < nw) ? red[lane] : -INFINITY;
    r = warp_max(r);
    return r;
}
__device__ __forceinline__ double block_sum_d(double v, double* red) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum_d(v);
    __syncthreads();
    if (lane == 0) red[warp] = v;
    __syncthreads();
    double r = (lane < nw) ? red[lane] : 0.0;
    r = warp_sum_d(r);
    return r;
}

__device__ __forceinline__ float leaky_relu(float v, float slope) { return v > 0.f ? v : v * slope; }
