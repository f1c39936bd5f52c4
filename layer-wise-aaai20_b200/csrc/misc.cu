// Small model ops that used to run on stock ATen kernels: stand-alone ReLU, logit scale, dropout (Philox, 1-bit
// mask), adaptive / global average pooling (channels-last), argmax-correct and top-1 / top-5 counters.
// All are bandwidth-bound streaming kernels with 16-byte accesses.
//
// Reference ops: nn.ReLU after biased layers (CIFAR10/alexnet.py:31-36, CIFAR10/vgg16.py:24-30), Mul(0.125)
// (CIFAR10/torch_backend.py:72-77, dawn.py:54), nn.Dropout (alexnet.py:30,33; vgg16.py:26,29), AdaptiveAvgPool2d
// (IMAGENET/training/resnet.py:109, vgg16.py:22), Correct (torch_backend.py:88-90), top-k accuracy
// (IMAGENET/training/train_imagenet_nv.py:679-689).
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <string.h>

#include "common.cuh"

namespace lwc {
namespace misc {

template <typename T> struct Vec8;             // 8 elements = 16 bytes (bf16 / fp16) or 2 x 16 bytes (fp32)
template <> struct Vec8<float> {
    static __device__ __forceinline__ void ld(const float* p, float v[8]) {
        const float4 a = reinterpret_cast<const float4*>(p)[0], b = reinterpret_cast<const float4*>(p)[1];
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    }
    static __device__ __forceinline__ void st(float* p, const float v[8]) {
        reinterpret_cast<float4*>(p)[0] = make_float4(v[0], v[1], v[2], v[3]);
        reinterpret_cast<float4*>(p)[1] = make_float4(v[4], v[5], v[6], v[7]);
    }
};
template <> struct Vec8<__nv_bfloat16> {
    static __device__ __forceinline__ void ld(const __nv_bfloat16* p, float v[8]) {
        const uint4 u = *reinterpret_cast<const uint4*>(p);
        const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) { v[2 * e] = __uint_as_float(w[e] << 16); v[2 * e + 1] = __uint_as_float(w[e] & 0xffff0000u); }
    }
    static __device__ __forceinline__ void st(__nv_bfloat16* p, const float v[8]) {
        uint32_t w[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * e], v[2 * e + 1]);
            w[e] = *reinterpret_cast<uint32_t*>(&h);
        }
        *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
    }
};
template <> struct Vec8<__half> {
    static __device__ __forceinline__ void ld(const __half* p, float v[8]) {
        const uint4 u = *reinterpret_cast<const uint4*>(p);
        const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            const __half2 h = *reinterpret_cast<const __half2*>(&w[e]);
            v[2 * e] = __low2float(h); v[2 * e + 1] = __high2float(h);
        }
    }
    static __device__ __forceinline__ void st(__half* p, const float v[8]) {
        uint32_t w[4];
#pragma unroll
        for (int e = 0; e < 4; ++e) {
            __half2 h = __floats2half2_rn(v[2 * e], v[2 * e + 1]);
            w[e] = *reinterpret_cast<uint32_t*>(&h);
        }
        *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
    }
};

// ---- elementwise: y = f(x [, aux]) over n elements, 8 per thread, scalar tail ------------------------------------
// OP 0: relu(x)   1: alpha * x   2: dy * (y > 0)  (aux = y)
template <typename T, int OP>
__global__ void __launch_bounds__(256) k_eltwise(const T* __restrict__ x, const T* __restrict__ aux, T* __restrict__ y, long long n,
                                                 float alpha) {
    const long long nv = n >> 3;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += (long long)gridDim.x * blockDim.x) {
        float v[8], a[8];
        Vec8<T>::ld(x + i * 8, v);
        if (OP == 2) Vec8<T>::ld(aux + i * 8, a);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            if (OP == 0) v[k] = fmaxf(v[k], 0.0f);
            if (OP == 1) v[k] *= alpha;
            if (OP == 2) v[k] = a[k] > 0.0f ? v[k] : 0.0f;
        }
        Vec8<T>::st(y + i * 8, v);
    }
    if (blockIdx.x == 0) {
        for (long long i = (nv << 3) + threadIdx.x; i < n; i += blockDim.x) {
            float v = DT<T>::ld(x + i);
            if (OP == 0) v = fmaxf(v, 0.0f);
            if (OP == 1) v *= alpha;
            if (OP == 2) v = DT<T>::ld(aux + i) > 0.0f ? v : 0.0f;
            DT<T>::st(y + i, v);
        }
    }
}

// ---- dropout: Philox4x32-10 keyed by (seed, offset + group index); one mask byte per 8 elements --------------------
__device__ __forceinline__ uint4 philox4x32_10(uint4 ctr, uint2 key) {
    constexpr uint32_t M0 = 0xD2511F53u, M1 = 0xCD9E8D57u, W0 = 0x9E3779B9u, W1 = 0xBB67AE85u;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint32_t hi0 = __umulhi(M0, ctr.x), lo0 = M0 * ctr.x;
        const uint32_t hi1 = __umulhi(M1, ctr.z), lo1 = M1 * ctr.z;
        ctr = make_uint4(hi1 ^ ctr.y ^ key.x, lo1, hi0 ^ ctr.w ^ key.y, lo0);
        key.x += W0; key.y += W1;
    }
    return ctr;
}

template <typename T>
__global__ void __launch_bounds__(256) k_dropout_fwd(const T* __restrict__ x, T* __restrict__ y, unsigned char* __restrict__ mask,
                                                     long long n, float p, float scale, unsigned long long seed,
                                                     unsigned long long offset) {
    const long long nv = (n + 7) >> 3;
    const uint32_t thr = (uint32_t)(p * 65536.0f);           // drop when the 16-bit draw is below p
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += (long long)gridDim.x * blockDim.x) {
        const unsigned long long c = offset + (unsigned long long)i;
        const uint4 r = philox4x32_10(make_uint4((uint32_t)c, (uint32_t)(c >> 32), 0u, 0u), make_uint2((uint32_t)seed, (uint32_t)(seed >> 32)));
        const uint32_t rw[4] = {r.x, r.y, r.z, r.w};
        unsigned bits = 0;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const uint32_t d = (rw[k >> 1] >> ((k & 1) * 16)) & 0xffffu;
            bits |= (d >= thr ? 1u : 0u) << k;
        }
        if (i * 8 + 8 <= n) {
            float v[8];
            Vec8<T>::ld(x + i * 8, v);
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = ((bits >> k) & 1u) ? v[k] * scale : 0.0f;
            Vec8<T>::st(y + i * 8, v);
        } else {
            for (long long j = i * 8; j < n; ++j) DT<T>::st(y + j, ((bits >> (j - i * 8)) & 1u) ? DT<T>::ld(x + j) * scale : 0.0f);
        }
        mask[i] = (unsigned char)bits;
    }
}

template <typename T>
__global__ void __launch_bounds__(256) k_dropout_bwd(const T* __restrict__ dy, const unsigned char* __restrict__ mask, T* __restrict__ dx,
                                                     long long n, float scale) {
    const long long nv = (n + 7) >> 3;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < nv; i += (long long)gridDim.x * blockDim.x) {
        const unsigned bits = mask[i];
        if (i * 8 + 8 <= n) {
            float v[8];
            Vec8<T>::ld(dy + i * 8, v);
#pragma unroll
            for (int k = 0; k < 8; ++k) v[k] = ((bits >> k) & 1u) ? v[k] * scale : 0.0f;
            Vec8<T>::st(dx + i * 8, v);
        } else {
            for (long long j = i * 8; j < n; ++j) DT<T>::st(dx + j, ((bits >> (j - i * 8)) & 1u) ? DT<T>::ld(dy + j) * scale : 0.0f);
        }
    }
}

// ---- adaptive average pooling, channels-last [N][IH][IW][C] -> [N][OH][OW][C]; C % 8 == 0 --------------------------
// window of output o: [floor(o * I / O), ceil((o + 1) * I / O))
template <typename T>
__global__ void __launch_bounds__(256) k_apool_fwd(const T* __restrict__ x, T* __restrict__ y, int N, int IH, int IW, int OH, int OW, int C) {
    const int cg = C >> 3;
    const long long total = (long long)N * OH * OW * cg;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % cg);
        long long r = i / cg;
        const int ow = (int)(r % OW); r /= OW;
        const int oh = (int)(r % OH);
        const int n = (int)(r / OH);
        const int h0 = oh * IH / OH, h1 = ((oh + 1) * IH + OH - 1) / OH, w0 = ow * IW / OW, w1 = ((ow + 1) * IW + OW - 1) / OW;
        float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int h = h0; h < h1; ++h)
            for (int w = w0; w < w1; ++w) {
                float v[8];
                Vec8<T>::ld(x + (((size_t)n * IH + h) * IW + w) * C + c * 8, v);
#pragma unroll
                for (int k = 0; k < 8; ++k) s[k] += v[k];
            }
        const float inv = 1.0f / (float)((h1 - h0) * (w1 - w0));
#pragma unroll
        for (int k = 0; k < 8; ++k) s[k] *= inv;
        Vec8<T>::st(y + i * 8, s);
    }
}

template <typename T>
__global__ void __launch_bounds__(256) k_apool_bwd(const T* __restrict__ dy, T* __restrict__ dx, int N, int IH, int IW, int OH, int OW, int C) {
    const int cg = C >> 3;
    const long long total = (long long)N * IH * IW * cg;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
        const int c = (int)(i % cg);
        long long r = i / cg;
        const int w = (int)(r % IW); r /= IW;
        const int h = (int)(r % IH);
        const int n = (int)(r / IH);
        float s[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        // outputs whose window holds (h, w): oh in [floor(h * OH / IH) - 1, ...] -- test the few candidates exactly
        const int oh_lo = max(0, (int)((long long)h * OH / IH) - 1), oh_hi = min(OH - 1, (int)(((long long)h + 1) * OH / IH) + 1);
        const int ow_lo = max(0, (int)((long long)w * OW / IW) - 1), ow_hi = min(OW - 1, (int)(((long long)w + 1) * OW / IW) + 1);
        for (int oh = oh_lo; oh <= oh_hi; ++oh) {
            const int h0 = oh * IH / OH, h1 = ((oh + 1) * IH + OH - 1) / OH;
            if (h < h0 || h >= h1) continue;
            for (int ow = ow_lo; ow <= ow_hi; ++ow) {
                const int w0 = ow * IW / OW, w1 = ((ow + 1) * IW + OW - 1) / OW;
                if (w < w0 || w >= w1) continue;
                float v[8];
                Vec8<T>::ld(dy + (((size_t)n * OH + oh) * OW + ow) * C + c * 8, v);
                const float inv = 1.0f / (float)((h1 - h0) * (w1 - w0));
#pragma unroll
                for (int k = 0; k < 8; ++k) s[k] = fmaf(v[k], inv, s[k]);
            }
        }
        Vec8<T>::st(dx + i * 8, s);
    }
}

// ---- accuracy: one warp per sample.  rank = #{j : x_j > x_t or (x_j == x_t and j < t)} (torch.max / torch.topk order) ----
// correct (nullable): [N] bytes, 1 when rank == 0.  counters (nullable): float [2] += (top-1 hits, top-5 hits).
template <typename T>
__global__ void __launch_bounds__(256) k_accuracy(const T* __restrict__ logits, const long long* __restrict__ target,
                                                  unsigned char* __restrict__ correct, float* __restrict__ counters, int N, int C) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= N) return;
    const long long t = target[warp];
    int rank = C;                                   // an out-of-range target is never correct
    if (t >= 0 && t < C) {
        const T* row = logits + (size_t)warp * C;
        const float xt = DT<T>::ld(row + t);
        int cnt = 0;
        for (int j = lane; j < C; j += 32) {
            const float v = DT<T>::ld(row + j);
            cnt += (v > xt || (v == xt && j < t)) ? 1 : 0;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        rank = cnt;
    }
    if (lane == 0) {
        if (correct) correct[warp] = rank == 0 ? 1 : 0;
        if (counters) {
            if (rank == 0) atomicAdd(counters, 1.0f);
            if (rank < 5) atomicAdd(counters + 1, 1.0f);
        }
    }
}

inline int grid_for(long long work_items) {
    long long g = (work_items + 255) / 256;
    if (g > 148 * 16) g = 148 * 16;
    if (g < 1) g = 1;
    return (int)g;
}

}  // namespace misc
}  // namespace lwc

using namespace lwc::misc;

#define LWC_MISC_DISPATCH(dt, ...)                                             \
    switch (dt) {                                                              \
        case LWC_F32: { using T = float; __VA_ARGS__; } break;                 \
        case LWC_BF16: { using T = __nv_bfloat16; __VA_ARGS__; } break;        \
        case LWC_F16: { using T = __half; __VA_ARGS__; } break;                \
        default: return LWC_EINVAL;                                            \
    }

static bool aligned16(const void* p) { return (reinterpret_cast<uintptr_t>(p) & 15) == 0; }

// op: 0 = relu(x), 1 = alpha * x, 2 = relu backward dy * (aux > 0).  Pointers 16-byte aligned.
extern "C" int lwc_eltwise(const void* x, const void* aux, void* y, long long n, int op, float alpha, int dtype, void* stream) {
    if (!x || !y || n <= 0 || op < 0 || op > 2 || (op == 2 && !aux) || !aligned16(x) || !aligned16(y) || (aux && !aligned16(aux)))
        return LWC_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    const int g = grid_for(n >> 3);
    LWC_MISC_DISPATCH(dtype,
        if (op == 0) k_eltwise<T, 0><<<g, 256, 0, st>>>((const T*)x, (const T*)aux, (T*)y, n, alpha);
        else if (op == 1) k_eltwise<T, 1><<<g, 256, 0, st>>>((const T*)x, (const T*)aux, (T*)y, n, alpha);
        else k_eltwise<T, 2><<<g, 256, 0, st>>>((const T*)x, (const T*)aux, (T*)y, n, alpha));
    return (int)cudaGetLastError();
}

// y = mask * x / (1 - p); mask: ceil(n / 8) bytes, bit k of byte i <-> element 8 i + k (1 = kept).
extern "C" int lwc_dropout_forward(const void* x, void* y, void* mask, long long n, float p, unsigned long long seed,
                                   unsigned long long offset, int dtype, void* stream) {
    if (!x || !y || !mask || n <= 0 || !(p >= 0.0f && p < 1.0f) || !aligned16(x) || !aligned16(y)) return LWC_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    const int g = grid_for((n + 7) >> 3);
    const float scale = 1.0f / (1.0f - p);
    LWC_MISC_DISPATCH(dtype, k_dropout_fwd<T><<<g, 256, 0, st>>>((const T*)x, (T*)y, (unsigned char*)mask, n, p, scale, seed, offset));
    return (int)cudaGetLastError();
}

extern "C" int lwc_dropout_backward(const void* dy, const void* mask, void* dx, long long n, float p, int dtype, void* stream) {
    if (!dy || !dx || !mask || n <= 0 || !(p >= 0.0f && p < 1.0f) || !aligned16(dy) || !aligned16(dx)) return LWC_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    const int g = grid_for((n + 7) >> 3);
    LWC_MISC_DISPATCH(dtype, k_dropout_bwd<T><<<g, 256, 0, st>>>((const T*)dy, (const unsigned char*)mask, (T*)dx, n, 1.0f / (1.0f - p)));
    return (int)cudaGetLastError();
}

// Adaptive average pooling on channels-last tensors (C % 8 == 0); backward = 1 distributes dy over the windows.
extern "C" int lwc_adaptive_avgpool(const void* in, void* out, int N, int IH, int IW, int OH, int OW, int C, int backward, int dtype,
                                    void* stream) {
    if (!in || !out || N <= 0 || IH <= 0 || IW <= 0 || OH <= 0 || OW <= 0 || C <= 0 || (C % 8)) return LWC_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    if (!backward) {
        const int g = grid_for((long long)N * OH * OW * (C / 8));
        LWC_MISC_DISPATCH(dtype, k_apool_fwd<T><<<g, 256, 0, st>>>((const T*)in, (T*)out, N, IH, IW, OH, OW, C));
    } else {
        const int g = grid_for((long long)N * IH * IW * (C / 8));
        LWC_MISC_DISPATCH(dtype, k_apool_bwd<T><<<g, 256, 0, st>>>((const T*)in, (T*)out, N, IH, IW, OH, OW, C));
    }
    return (int)cudaGetLastError();
}

// logits [N][C], target int64 [N]; correct: [N] bytes or NULL; counters: float[2] (top-1, top-5 hit counts, accumulated) or NULL.
extern "C" int lwc_accuracy(const void* logits, const void* target_i64, void* correct, float* counters, int N, int C, int dtype,
                            void* stream) {
    if (!logits || !target_i64 || N <= 0 || C <= 0 || (!correct && !counters)) return LWC_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    const int g = (N * 32 + 255) / 256;
    LWC_MISC_DISPATCH(dtype, k_accuracy<T><<<g, 256, 0, st>>>((const T*)logits, (const long long*)target_i64, (unsigned char*)correct,
                                                             counters, N, C));
    return (int)cudaGetLastError();
}
