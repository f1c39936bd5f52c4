// Shared device helpers for the layerwise_b200 native library (sm_100a).
#pragma once
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "lwc_b200.h"

#define LWC_CUDA_CHECK(expr)                      \
    do {                                          \
        cudaError_t _e = (expr);                  \
        if (_e != cudaSuccess) return (int)_e;    \
    } while (0)

namespace lwc {

constexpr int kChunk = LWC_CHUNK;   // elements per CTA work item
constexpr int kThreads = 256;       // threads per CTA for the streaming kernels
constexpr int kQuadsPerThread = kChunk / (kThreads * 4);   // 8

// ---- dtype traits -------------------------------------------------------------------------------
template <typename T> struct DT;
template <> struct DT<float> {
    static __device__ __forceinline__ float ld(const float* p) { return *p; }
    static __device__ __forceinline__ void st(float* p, float v) { *p = v; }
    static __device__ __forceinline__ void ld4(const float* p, float v[4]) {
        float4 t = *reinterpret_cast<const float4*>(p);
        v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
    }
    static __device__ __forceinline__ void st4(float* p, const float v[4]) {
        *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    }
};
template <> struct DT<__nv_bfloat16> {
    static __device__ __forceinline__ float ld(const __nv_bfloat16* p) { return __bfloat162float(*p); }
    static __device__ __forceinline__ void st(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }
    static __device__ __forceinline__ void ld4(const __nv_bfloat16* p, float v[4]) {
        uint2 t = *reinterpret_cast<const uint2*>(p);
        __nv_bfloat162 a = *reinterpret_cast<__nv_bfloat162*>(&t.x);
        __nv_bfloat162 b = *reinterpret_cast<__nv_bfloat162*>(&t.y);
        v[0] = __low2float(a); v[1] = __high2float(a); v[2] = __low2float(b); v[3] = __high2float(b);
    }
    static __device__ __forceinline__ void st4(__nv_bfloat16* p, const float v[4]) {
        __nv_bfloat162 a = __floats2bfloat162_rn(v[0], v[1]);
        __nv_bfloat162 b = __floats2bfloat162_rn(v[2], v[3]);
        uint2 t;
        t.x = *reinterpret_cast<uint32_t*>(&a);
        t.y = *reinterpret_cast<uint32_t*>(&b);
        *reinterpret_cast<uint2*>(p) = t;
    }
};
template <> struct DT<__half> {
    static __device__ __forceinline__ float ld(const __half* p) { return __half2float(*p); }
    static __device__ __forceinline__ void st(__half* p, float v) { *p = __float2half_rn(v); }
    static __device__ __forceinline__ void ld4(const __half* p, float v[4]) {
        uint2 t = *reinterpret_cast<const uint2*>(p);
        __half2 a = *reinterpret_cast<__half2*>(&t.x);
        __half2 b = *reinterpret_cast<__half2*>(&t.y);
        v[0] = __low2float(a); v[1] = __high2float(a); v[2] = __low2float(b); v[3] = __high2float(b);
    }
    static __device__ __forceinline__ void st4(__half* p, const float v[4]) {
        __half2 a = __floats2half2_rn(v[0], v[1]);
        __half2 b = __floats2half2_rn(v[2], v[3]);
        uint2 t;
        t.x = *reinterpret_cast<uint32_t*>(&a);
        t.y = *reinterpret_cast<uint32_t*>(&b);
        *reinterpret_cast<uint2*>(p) = t;
    }
};

// Load up to four consecutive elements [j, j+4) of a chunk of length len (elements beyond len
// read as 0).  `vec` says whether base+j is suitably aligned for one vector access.
template <typename T>
__device__ __forceinline__ void load_quad(const T* base, int j, int len, bool vec, float v[4]) {
    if (vec && j + 4 <= len) {
        DT<T>::ld4(base + j, v);
    } else {
#pragma unroll
        for (int e = 0; e < 4; ++e) v[e] = (j + e < len) ? DT<T>::ld(base + j + e) : 0.0f;
    }
}
template <typename T>
__device__ __forceinline__ void store_quad(T* base, int j, int len, bool vec, const float v[4]) {
    if (vec && j + 4 <= len) {
        DT<T>::st4(base + j, v);
    } else {
#pragma unroll
        for (int e = 0; e < 4; ++e)
            if (j + e < len) DT<T>::st(base + j + e, v[e]);
    }
}
template <typename T>
__device__ __forceinline__ bool vec_ok(const T* p) {
    return (reinterpret_cast<uintptr_t>(p) & (4 * sizeof(T) - 1)) == 0;
}

// ---- warp / block reductions -----------------------------------------------------------------------
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
__device__ __forceinline__ unsigned warp_sum_u(unsigned v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
// Inclusive warp scan.
__device__ __forceinline__ unsigned warp_incl_scan(unsigned v) {
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned t = __shfl_up_sync(0xffffffffu, v, o);
        if (lane >= o) v += t;
    }
    return v;
}

// ---- cross-GPU ordering --------------------------------------------------------------------------
__device__ __forceinline__ void st_release_sys(uint64_t* p, uint64_t v) {
    asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ uint64_t ld_acquire_sys(const uint64_t* p) {
    uint64_t v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ uint64_t global_timer_ns() {
    uint64_t t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// ---- hashing / RNG -----------------------------------------------------------------------------------
__device__ __host__ __forceinline__ uint32_t mix32(uint32_t x) {   // lowbias32-style finaliser
    x ^= x >> 16; x *= 0x7feb352dU;
    x ^= x >> 15; x *= 0x846ca68bU;
    x ^= x >> 16;
    return x;
}
__device__ __host__ __forceinline__ uint64_t mix64(uint64_t x) {   // splitmix64 finaliser
    x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ULL;
    x ^= x >> 27; x *= 0x94D049BB133111EBULL;
    x ^= x >> 31;
    return x;
}
// Philox4x32-10: counter-based generator, one call yields four 32-bit words.
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
    constexpr uint32_t M0 = 0xD2511F53U, M1 = 0xCD9E8D57U, W0 = 0x9E3779B9U, W1 = 0xBB67AE85U;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
        uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += W0; key.y += W1;
    }
    return ctr;
}
__device__ __forceinline__ float u32_to_unit(uint32_t x) {   // [0,1) with 24 random bits
    return (x >> 8) * (1.0f / 16777216.0f);
}

}  // namespace lwc
