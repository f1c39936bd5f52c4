// Max pooling for channels-last activations (NHWC), forward with a one-byte argmax per element and a
// gather-style backward (no atomics): each input pixel sums dy over the <= ceil(k/s)^2 output windows
// that cover it and whose argmax points at it.  A thread owns 8 consecutive channels (16-byte vectors),
// consecutive threads own consecutive channel groups, so every access is coalesced.
// (reference ops: nn.MaxPool2d 2x2/2, 4x4/4 in CIFAR10/dawn.py:51,70-71; 3x3/2 pad 1 in
//  IMAGENET/training/resnet.py:104 -- ATen max_pool2d forward/backward kernels)
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <type_traits>
#include <string.h>

#include "common.cuh"

namespace lwc {
namespace pool {

struct Params {
    const void* x;
    void* y;
    uint8_t* idx;
    const void* dy;
    void* dx;
    int N, H, W, C, OH, OW, k, s, pad;
};

template <typename T> struct V8;
template <> struct V8<__nv_bfloat16> {
    static __device__ __forceinline__ void ld(const __nv_bfloat16* p, float v[8]) {
        const uint4 t = *reinterpret_cast<const uint4*>(p);
        const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&w[i]);
            v[2 * i] = __low2float(h); v[2 * i + 1] = __high2float(h);
        }
    }
    static __device__ __forceinline__ void st(__nv_bfloat16* p, const float v[8]) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]); w[i] = *reinterpret_cast<uint32_t*>(&h); }
        *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
    }
};
template <> struct V8<__half> {
    static __device__ __forceinline__ void ld(const __half* p, float v[8]) {
        const uint4 t = *reinterpret_cast<const uint4*>(p);
        const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const __half2 h = *reinterpret_cast<const __half2*>(&w[i]);
            v[2 * i] = __low2float(h); v[2 * i + 1] = __high2float(h);
        }
    }
    static __device__ __forceinline__ void st(__half* p, const float v[8]) {
        uint32_t w[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { __half2 h = __floats2half2_rn(v[2 * i], v[2 * i + 1]); w[i] = *reinterpret_cast<uint32_t*>(&h); }
        *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
    }
};
template <> struct V8<float> {
    static __device__ __forceinline__ void ld(const float* p, float v[8]) {
        const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
        v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
    }
    static __device__ __forceinline__ void st(float* p, const float v[8]) {
        *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
        *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
    }
};

template <typename T>
__global__ void __launch_bounds__(256) k_maxpool_fwd(Params p) {
    const int ncg = p.C >> 3;
    const long long total = (long long)p.N * p.OH * p.OW * ncg;
    for (long long id = (long long)blockIdx.x * blockDim.x + threadIdx.x; id < total; id += (long long)gridDim.x * blockDim.x) {
        const int cg = (int)(id % ncg);
        long long pix = id / ncg;
        const int ow = (int)(pix % p.OW); pix /= p.OW;
        const int oh = (int)(pix % p.OH);
        const int n = (int)(pix / p.OH);
        float best[8];
        unsigned arg[8];
#pragma unroll
        for (int c = 0; c < 8; ++c) { best[c] = -INFINITY; arg[c] = 0; }
        const int h0 = oh * p.s - p.pad, w0 = ow * p.s - p.pad;
        const T* xb = reinterpret_cast<const T*>(p.x) + (size_t)n * p.H * p.W * p.C + cg * 8;
        for (int kh = 0; kh < p.k; ++kh) {
            const int h = h0 + kh;
            if (h < 0 || h >= p.H) continue;
            for (int kw = 0; kw < p.k; ++kw) {
                const int w = w0 + kw;
                if (w < 0 || w >= p.W) continue;
                float v[8];
                V8<T>::ld(xb + ((size_t)h * p.W + w) * p.C, v);
                const unsigned code = kh * p.k + kw;
#pragma unroll
                for (int c = 0; c < 8; ++c)
                    if (v[c] > best[c]) { best[c] = v[c]; arg[c] = code; }
            }
        }
        const size_t o = (((size_t)n * p.OH + oh) * p.OW + ow) * p.C + cg * 8;
        V8<T>::st(reinterpret_cast<T*>(p.y) + o, best);
        if (p.idx) {
            uint2 packed;
            packed.x = arg[0] | (arg[1] << 8) | (arg[2] << 16) | (arg[3] << 24);
            packed.y = arg[4] | (arg[5] << 8) | (arg[6] << 16) | (arg[7] << 24);
            *reinterpret_cast<uint2*>(p.idx + o) = packed;
        }
    }
}

template <typename T>
__global__ void __launch_bounds__(256) k_maxpool_bwd(Params p) {
    const int ncg = p.C >> 3;
    const long long total = (long long)p.N * p.H * p.W * ncg;
    for (long long id = (long long)blockIdx.x * blockDim.x + threadIdx.x; id < total; id += (long long)gridDim.x * blockDim.x) {
        const int cg = (int)(id % ncg);
        long long pix = id / ncg;
        const int w = (int)(pix % p.W); pix /= p.W;
        const int h = (int)(pix % p.H);
        const int n = (int)(pix / p.H);
        float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        // output rows/cols whose window [o*s - pad, o*s - pad + k) contains h / w
        int oh_lo = (h + p.pad - p.k + p.s) / p.s;          // ceil((h + pad - k + 1) / s)
        if (h + p.pad - p.k + 1 < 0) oh_lo = 0;
        int oh_hi = (h + p.pad) / p.s;
        if (oh_hi > p.OH - 1) oh_hi = p.OH - 1;
        int ow_lo = (w + p.pad - p.k + p.s) / p.s;
        if (w + p.pad - p.k + 1 < 0) ow_lo = 0;
        int ow_hi = (w + p.pad) / p.s;
        if (ow_hi > p.OW - 1) ow_hi = p.OW - 1;
        for (int oh = oh_lo; oh <= oh_hi; ++oh) {
            const int kh = h - (oh * p.s - p.pad);
            for (int ow = ow_lo; ow <= ow_hi; ++ow) {
                const int kw = w - (ow * p.s - p.pad);
                const unsigned code = kh * p.k + kw;
                const size_t o = (((size_t)n * p.OH + oh) * p.OW + ow) * p.C + cg * 8;
                const uint2 packed = *reinterpret_cast<const uint2*>(p.idx + o);
                float d[8];
                V8<T>::ld(reinterpret_cast<const T*>(p.dy) + o, d);
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const unsigned a = ((c < 4 ? packed.x : packed.y) >> (8 * (c & 3))) & 0xffu;
                    if (a == code) acc[c] += d[c];
                }
            }
        }
        V8<T>::st(reinterpret_cast<T*>(p.dx) + (((size_t)n * p.H + h) * p.W + w) * p.C + cg * 8, acc);
    }
}

// Raw 8-element vectors: loaded without conversion so that several loads can be in flight before the first use.
template <typename T> struct Raw8 { uint4 v; };
template <> struct Raw8<float> { float4 a, b; };
template <typename T>
__device__ __forceinline__ void ldraw(const T* p, Raw8<T>& r) { r.v = *reinterpret_cast<const uint4*>(p); }
template <>
__device__ __forceinline__ void ldraw<float>(const float* p, Raw8<float>& r) {
    r.a = *reinterpret_cast<const float4*>(p);
    r.b = *reinterpret_cast<const float4*>(p + 4);
}
__device__ __forceinline__ void unpack(const Raw8<__nv_bfloat16>& r, float v[8]) {
    const uint32_t w[4] = {r.v.x, r.v.y, r.v.z, r.v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) { v[2 * i] = __uint_as_float(w[i] << 16); v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u); }
}
__device__ __forceinline__ void unpack(const Raw8<__half>& r, float v[8]) {
    const uint32_t w[4] = {r.v.x, r.v.y, r.v.z, r.v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const __half2 h = *reinterpret_cast<const __half2*>(&w[i]);
        v[2 * i] = __low2float(h); v[2 * i + 1] = __high2float(h);
    }
}
__device__ __forceinline__ void unpack(const Raw8<float>& r, float v[8]) {
    v[0] = r.a.x; v[1] = r.a.y; v[2] = r.a.z; v[3] = r.a.w; v[4] = r.b.x; v[5] = r.b.y; v[6] = r.b.z; v[7] = r.b.w;
}

// ---- 3x3 / stride 2 / pad 1 (the ResNet stem pool) specialisations ------------------------------------------
// Forward: the nine 16-byte loads of a window are issued back to back (compile-time window, 32-bit indexing).
template <typename T>
__global__ void __launch_bounds__(256) k_maxpool_fwd_k3s2(Params p) {
    const int ncg = p.C >> 3;
    const int total = p.N * p.OH * p.OW * ncg;
    for (int id = blockIdx.x * blockDim.x + threadIdx.x; id < total; id += gridDim.x * blockDim.x) {
        const int cg = id % ncg;
        int pix = id / ncg;
        const int ow = pix % p.OW; pix /= p.OW;
        const int oh = pix % p.OH;
        const int n = pix / p.OH;
        const int h0 = 2 * oh - 1, w0 = 2 * ow - 1;
        const T* xb = reinterpret_cast<const T*>(p.x) + (size_t)n * p.H * p.W * p.C + cg * 8;
        Raw8<T> raw[9];
        bool ok[9];
        // out-of-range rows / columns re-read the (always valid) centre row / column: no branch, and the tap is masked
        // out of the maximum below.  Row and column are clamped separately: three row bases, three column offsets.
        bool okh[3], okw[3];
        int roff[3], coff[3];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
            okh[r] = h0 + r >= 0 && h0 + r < p.H;
            okw[r] = w0 + r >= 0 && w0 + r < p.W;
            roff[r] = (okh[r] ? h0 + r : h0 + 1) * p.W * p.C;
            coff[r] = (okw[r] ? w0 + r : w0 + 1) * p.C;
        }
#pragma unroll
        for (int t = 0; t < 9; ++t) {
            ok[t] = okh[t / 3] && okw[t % 3];
            ldraw<T>(xb + roff[t / 3] + coff[t % 3], raw[t]);
        }
        const size_t o = (size_t)id * 8;                  // == ((n*OH + oh)*OW + ow)*C + cg*8
        if constexpr (std::is_same<T, __nv_bfloat16>::value) {
            // bf16: the running maximum and its tap index stay packed, two channels per 32-bit word (compare -> 16-bit
            // lane masks -> bitwise selects); no unpacking, no conversion on the way out.  The kernel was issue-bound.
            uint32_t best[4], arg[4];
#pragma unroll
            for (int w = 0; w < 4; ++w) { best[w] = 0xff80ff80u; arg[w] = 0u; }      // -inf, -inf
#pragma unroll
            for (int t = 0; t < 9; ++t) {
                const uint32_t v[4] = {raw[t].v.x, raw[t].v.y, raw[t].v.z, raw[t].v.w};
                const uint32_t okm = ok[t] ? 0xffffffffu : 0u;
                const uint32_t tp = (uint32_t)t * 0x00010001u;
#pragma unroll
                for (int w = 0; w < 4; ++w) {
                    const uint32_t m = __hgt2_mask(*reinterpret_cast<const __nv_bfloat162*>(&v[w]),
                                                   *reinterpret_cast<const __nv_bfloat162*>(&best[w])) & okm;
                    best[w] = (best[w] & ~m) | (v[w] & m);
                    arg[w] = (arg[w] & ~m) | (tp & m);
                }
            }
            *reinterpret_cast<uint4*>(reinterpret_cast<T*>(p.y) + o) = make_uint4(best[0], best[1], best[2], best[3]);
            if (p.idx) {
                uint2 packed;
                packed.x = __byte_perm(arg[0], arg[1], 0x6420);
                packed.y = __byte_perm(arg[2], arg[3], 0x6420);
                *reinterpret_cast<uint2*>(p.idx + o) = packed;
            }
        } else {
            float best[8];
            unsigned arg[8];
#pragma unroll
            for (int c = 0; c < 8; ++c) { best[c] = -INFINITY; arg[c] = 0; }
#pragma unroll
            for (int t = 0; t < 9; ++t) {
                float v[8];
                unpack(raw[t], v);
#pragma unroll
                for (int c = 0; c < 8; ++c)
                    if (ok[t] && v[c] > best[c]) { best[c] = v[c]; arg[c] = t; }
            }
            V8<T>::st(reinterpret_cast<T*>(p.y) + o, best);
            if (p.idx) {
                uint2 packed;
                packed.x = arg[0] | (arg[1] << 8) | (arg[2] << 16) | (arg[3] << 24);
                packed.y = arg[4] | (arg[5] << 8) | (arg[6] << 16) | (arg[7] << 24);
                *reinterpret_cast<uint2*>(p.idx + o) = packed;
            }
        }
    }
}

// Backward: a thread owns a 2x2 block of input pixels (rows 2i, 2i+1; columns 2j, 2j+1).  Only the four windows
// (i..i+1, j..j+1) can point into it, so their dy / argmax vectors are loaded once (4 x 24 bytes) for four outputs;
// the per-pixel gather needs nine window loads for the same block.  H and W even.
template <typename T>
__global__ void __launch_bounds__(256) k_maxpool_bwd_k3s2(Params p) {
    const int ncg = p.C >> 3;
    const int HB = p.H >> 1, WB = p.W >> 1;
    const int total = p.N * HB * WB * ncg;
    for (int id = blockIdx.x * blockDim.x + threadIdx.x; id < total; id += gridDim.x * blockDim.x) {
        const int cg = id % ncg;
        int blk = id / ncg;
        const int j = blk % WB; blk /= WB;
        const int i = blk % HB;
        const int n = blk / HB;
        Raw8<T> draw[4];
        uint2 a[4];
        bool ok[4];
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            const int oh = i + (t >> 1), ow = j + (t & 1);
            ok[t] = oh < p.OH && ow < p.OW;
            const size_t o = (((size_t)n * p.OH + (ok[t] ? oh : i)) * p.OW + (ok[t] ? ow : j)) * p.C + cg * 8;
            a[t] = *reinterpret_cast<const uint2*>(p.idx + o);
            ldraw<T>(reinterpret_cast<const T*>(p.dy) + o, draw[t]);
        }
        // window t = (dh, dw) relative to (i, j); the input pixel (2i + a_, 2j + b_) sits at kernel offset
        // kh = a_ + 1 - 2 dh, kw = b_ + 1 - 2 dw of that window (valid when 0 <= kh, kw <= 2)
        float acc[4][8];
#pragma unroll
        for (int q = 0; q < 4; ++q)
#pragma unroll
            for (int c = 0; c < 8; ++c) acc[q][c] = 0.0f;
#pragma unroll
        for (int t = 0; t < 4; ++t) {
            if (!ok[t]) continue;
            const int dh = t >> 1, dw = t & 1;
            float dv[8];
            unpack(draw[t], dv);
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int kh = (q >> 1) + 1 - 2 * dh, kw = (q & 1) + 1 - 2 * dw;
                if (kh < 0 || kh > 2 || kw < 0 || kw > 2) continue;      // compile-time after unrolling
                const unsigned code = kh * 3 + kw;
#pragma unroll
                for (int c = 0; c < 8; ++c) {
                    const unsigned ac = ((c < 4 ? a[t].x : a[t].y) >> (8 * (c & 3))) & 0xffu;
                    if (ac == code) acc[q][c] += dv[c];
                }
            }
        }
        T* dx = reinterpret_cast<T*>(p.dx) + (((size_t)n * p.H + 2 * i) * p.W + 2 * j) * p.C + cg * 8;
        V8<T>::st(dx, acc[0]);
        V8<T>::st(dx + p.C, acc[1]);
        V8<T>::st(dx + (size_t)p.W * p.C, acc[2]);
        V8<T>::st(dx + (size_t)p.W * p.C + p.C, acc[3]);
    }
}

}  // namespace pool
}  // namespace lwc

using namespace lwc::pool;

static int grid_for(long long total) {
    long long b = (total + 255) / 256;
    const long long cap = 148LL * 16;
    return (int)(b < cap ? (b < 1 ? 1 : b) : cap);
}

#define LWC_POOL_DISPATCH(dt, ...)                                             \
    switch (dt) {                                                              \
        case LWC_F32: { using T = float; __VA_ARGS__; } break;                 \
        case LWC_BF16: { using T = __nv_bfloat16; __VA_ARGS__; } break;        \
        case LWC_F16: { using T = __half; __VA_ARGS__; } break;                \
        default: return LWC_EINVAL;                                            \
    }

// idx (N*OH*OW*C bytes) may be NULL for inference.
extern "C" int lwc_maxpool_forward(const void* x, void* y, void* idx, int N, int H, int W, int C, int k, int stride,
                                   int pad, int dtype, void* stream) {
    if (!x || !y || N <= 0 || H <= 0 || W <= 0 || C <= 0 || (C & 7) || k < 1 || k > 15 || stride < 1 || pad < 0 || 2 * pad > k)
        return LWC_EINVAL;
    Params p;
    memset(&p, 0, sizeof(p));
    p.x = x; p.y = y; p.idx = (uint8_t*)idx;
    p.N = N; p.H = H; p.W = W; p.C = C; p.k = k; p.s = stride; p.pad = pad;
    p.OH = (H + 2 * pad - k) / stride + 1;
    p.OW = (W + 2 * pad - k) / stride + 1;
    if (p.OH < 1 || p.OW < 1) return LWC_EINVAL;
    const long long total = (long long)N * p.OH * p.OW * (C >> 3);
    if (k == 3 && stride == 2 && pad == 1 && (long long)N * H * W * C < (1LL << 31)) {
        LWC_POOL_DISPATCH(dtype, k_maxpool_fwd_k3s2<T><<<grid_for(total), 256, 0, (cudaStream_t)stream>>>(p));
        return (int)cudaGetLastError();
    }
    LWC_POOL_DISPATCH(dtype, k_maxpool_fwd<T><<<grid_for(total), 256, 0, (cudaStream_t)stream>>>(p));
    return (int)cudaGetLastError();
}

extern "C" int lwc_maxpool_backward(const void* dy, const void* idx, void* dx, int N, int H, int W, int C, int k,
                                    int stride, int pad, int dtype, void* stream) {
    if (!dy || !idx || !dx || N <= 0 || (C & 7) || k < 1 || k > 15 || stride < 1) return LWC_EINVAL;
    Params p;
    memset(&p, 0, sizeof(p));
    p.dy = dy; p.idx = (uint8_t*)const_cast<void*>(idx); p.dx = dx;
    p.N = N; p.H = H; p.W = W; p.C = C; p.k = k; p.s = stride; p.pad = pad;
    p.OH = (H + 2 * pad - k) / stride + 1;
    p.OW = (W + 2 * pad - k) / stride + 1;
    const long long total = (long long)N * H * W * (C >> 3);
    if (k == 3 && stride == 2 && pad == 1 && !(H & 1) && !(W & 1) && (long long)N * H * W * C < (1LL << 31)) {
        LWC_POOL_DISPATCH(dtype, k_maxpool_bwd_k3s2<T><<<grid_for(total / 4), 256, 0, (cudaStream_t)stream>>>(p));
        return (int)cudaGetLastError();
    }
    LWC_POOL_DISPATCH(dtype, k_maxpool_bwd<T><<<grid_for(total), 256, 0, (cudaStream_t)stream>>>(p));
    return (int)cudaGetLastError();
}
