// Input pipeline kernels: uint8 NCHW image batch -> normalised compute dtype, channels-last, one pass.
//
//   k_normalize_nhwc : y[n,h,w,c] = (x[n,c,h,w] - mean[c]) / std[c]                       (3 -> C_out channels, zero padded)
//   k_normalize_s2d  : the same, written in 2x2 space-to-depth form with a 2/1 pixel zero border, so that
//                      ResNet's 7x7 stride-2 pad-3 stem convolution becomes a dense 4x4 stride-1 convolution
//                      over 16 channels (12 used): Xs[n, bh+2, bw+2, (dy*2+dx)*3 + c] = norm(x[n, c, 2bh+dy, 2bw+dx]).
// Reference op: BatchTransformDataLoader.process_tensors -- .half()/.float(), sub_(mean), div_(std) as three
// elementwise kernels plus the layout change (IMAGENET/training/dataloader.py:88-93).
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#include <string.h>

#include "common.cuh"

namespace lwc {
namespace input {

struct Params {
    const uint8_t* x;
    void* y;
    int N, H, W, Cout;
    float mean[3], inv_std[3];
};

// one thread per output pixel: reads 3 (or 12) bytes, writes Cout (8 or 16) contiguous elements
template <typename T>
__global__ void __launch_bounds__(256) k_normalize_nhwc(Params p) {
    const long long total = (long long)p.N * p.H * p.W;
    const size_t plane = (size_t)p.H * p.W;
    for (long long id = (long long)blockIdx.x * blockDim.x + threadIdx.x; id < total; id += (long long)gridDim.x * blockDim.x) {
        const long long n = id / plane;
        const size_t hw = id % plane;
        const uint8_t* src = p.x + (size_t)n * 3 * plane + hw;
        float v[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
        for (int c = 0; c < 3; ++c) v[c] = ((float)src[c * plane] - p.mean[c]) * p.inv_std[c];
        T* dst = reinterpret_cast<T*>(p.y) + (size_t)id * p.Cout;
        if (p.Cout == 8) {
            store_quad<T>(dst, 0, 8, true, v);
            store_quad<T>(dst, 4, 8, true, v + 4);
        } else {
            for (int c = 0; c < p.Cout; ++c) DT<T>::st(dst + c, c < 3 ? v[c] : 0.0f);
        }
    }
}

// output: [N, H/2 + 3, W/2 + 3, 16]; one thread per output pixel (border pixels write zeros)
template <typename T>
__global__ void __launch_bounds__(256) k_normalize_s2d(Params p) {
    const int H2 = p.H / 2, W2 = p.W / 2, HP = H2 + 3, WP = W2 + 3;
    const long long total = (long long)p.N * HP * WP;
    const size_t plane = (size_t)p.H * p.W;
    for (long long id = (long long)blockIdx.x * blockDim.x + threadIdx.x; id < total; id += (long long)gridDim.x * blockDim.x) {
        const int pw = (int)(id % WP);
        const int ph = (int)((id / WP) % HP);
        const long long n = id / ((long long)WP * HP);
        float v[16];
#pragma unroll
        for (int k = 0; k < 16; ++k) v[k] = 0.0f;
        const int bh = ph - 2, bw = pw - 2;
        if (bh >= 0 && bh < H2 && bw >= 0 && bw < W2) {
            const uint8_t* src = p.x + (size_t)n * 3 * plane;
#pragma unroll
            for (int dy = 0; dy < 2; ++dy)
#pragma unroll
                for (int dx = 0; dx < 2; ++dx)
#pragma unroll
                    for (int c = 0; c < 3; ++c)
                        v[(dy * 2 + dx) * 3 + c] =
                            ((float)src[c * plane + (size_t)(2 * bh + dy) * p.W + (2 * bw + dx)] - p.mean[c]) * p.inv_std[c];
        }
        T* dst = reinterpret_cast<T*>(p.y) + (size_t)id * 16;
#pragma unroll
        for (int q = 0; q < 4; ++q) store_quad<T>(dst, 4 * q, 16, true, v + 4 * q);
    }
}

}  // namespace input
}  // namespace lwc

using namespace lwc::input;

#define LWC_IN_DISPATCH(dt, ...)                                               \
    switch (dt) {                                                              \
        case LWC_F32: { using T = float; __VA_ARGS__; } break;                 \
        case LWC_BF16: { using T = __nv_bfloat16; __VA_ARGS__; } break;        \
        case LWC_F16: { using T = __half; __VA_ARGS__; } break;                \
        default: return LWC_EINVAL;                                            \
    }

// mode 0: NHWC with c_out channels (3 real + zero padding, c_out in {3,4,8});  mode 1: space-to-depth (see above)
extern "C" int lwc_normalize_u8(const void* x_u8_nchw, void* y, int N, int H, int W, int c_out, int mode,
                                const float* mean3, const float* std3, int dtype, void* stream) {
    if (!x_u8_nchw || !y || !mean3 || !std3 || N <= 0 || H <= 0 || W <= 0) return LWC_EINVAL;
    if (mode == 1 && ((H | W) & 1)) return LWC_EINVAL;
    if (mode == 0 && (c_out < 3 || c_out > 8)) return LWC_EINVAL;
    Params p;
    memset(&p, 0, sizeof(p));
    p.x = (const uint8_t*)x_u8_nchw; p.y = y; p.N = N; p.H = H; p.W = W; p.Cout = c_out;
    for (int c = 0; c < 3; ++c) { p.mean[c] = mean3[c]; p.inv_std[c] = 1.0f / std3[c]; }
    const long long total = mode == 1 ? (long long)N * (H / 2 + 3) * (W / 2 + 3) : (long long)N * H * W;
    long long blocks = (total + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    cudaStream_t st = (cudaStream_t)stream;
    if (mode == 1) { LWC_IN_DISPATCH(dtype, k_normalize_s2d<T><<<(int)blocks, 256, 0, st>>>(p)); }
    else { LWC_IN_DISPATCH(dtype, k_normalize_nhwc<T><<<(int)blocks, 256, 0, st>>>(p)); }
    return (int)cudaGetLastError();
}
