// Cross-entropy over logits [N, C]: per-sample loss + log-sum-exp forward, softmax-minus-onehot backward.
// One warp per sample; works on bf16 / fp16 / fp32 logits with fp32 math.
// (reference ops: nn.CrossEntropyLoss(reduction='none') + .sum() in CIFAR10/dawn.py:85, core.py:310; mean-reduced
//  nn.CrossEntropyLoss in IMAGENET/training/train_imagenet_nv.py:187 -- ATen log_softmax + nll_loss kernels)
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include "common.cuh"

namespace lwc {
namespace loss {

template <typename T>
__global__ void __launch_bounds__(256) k_ce_fwd(const T* __restrict__ logits, const long long* __restrict__ target,
                                                float* __restrict__ loss, float* __restrict__ lse, int N, int C) {
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= N) return;
    const T* x = logits + (size_t)row * C;
    float m = -INFINITY;
    for (int c = lane; c < C; c += 32) m = fmaxf(m, DT<T>::ld(x + c));
    m = warp_max(m);
    float s = 0.0f;
    for (int c = lane; c < C; c += 32) s += __expf(DT<T>::ld(x + c) - m);
    s = warp_sum(s);
    if (lane == 0) {
        const float l = m + __logf(s);
        const long long t = target[row];
        lse[row] = l;
        loss[row] = (t >= 0 && t < C) ? l - DT<T>::ld(x + t) : 0.0f;
    }
}

// dlogits[n, c] = (exp(x - lse[n]) - [c == target[n]]) * gscale * (g ? g[n] : 1)
template <typename T>
__global__ void __launch_bounds__(256) k_ce_bwd(const T* __restrict__ logits, const long long* __restrict__ target,
                                                const float* __restrict__ lse, const float* __restrict__ g,
                                                T* __restrict__ dlogits, int N, int C, float gscale) {
    const int row = blockIdx.x * 8 + (threadIdx.x >> 5);
    const int lane = threadIdx.x & 31;
    if (row >= N) return;
    const T* x = logits + (size_t)row * C;
    T* d = dlogits + (size_t)row * C;
    const float l = lse[row];
    const long long t = target[row];
    const float w = gscale * (g ? g[row] : 1.0f);
    const bool valid = t >= 0 && t < C;
    for (int c = lane; c < C; c += 32) {
        float v = valid ? (__expf(DT<T>::ld(x + c) - l) - (c == t ? 1.0f : 0.0f)) * w : 0.0f;
        DT<T>::st(d + c, v);
    }
}

}  // namespace loss
}  // namespace lwc

using namespace lwc::loss;

#define LWC_CE_DISPATCH(dt, ...)                                               \
    switch (dt) {                                                              \
        case LWC_F32: { using T = float; __VA_ARGS__; } break;                 \
        case LWC_BF16: { using T = __nv_bfloat16; __VA_ARGS__; } break;        \
        case LWC_F16: { using T = __half; __VA_ARGS__; } break;                \
        default: return LWC_EINVAL;                                            \
    }

extern "C" int lwc_cross_entropy_forward(const void* logits, const void* target_i64, float* loss, float* lse, int N, int C,
                                         int dtype, void* stream) {
    if (!logits || !target_i64 || !loss || !lse || N <= 0 || C <= 0) return LWC_EINVAL;
    LWC_CE_DISPATCH(dtype, k_ce_fwd<T><<<(N + 7) / 8, 256, 0, (cudaStream_t)stream>>>(
        (const T*)logits, (const long long*)target_i64, loss, lse, N, C));
    return (int)cudaGetLastError();
}

extern "C" int lwc_cross_entropy_backward(const void* logits, const void* target_i64, const float* lse, const float* grad_per_sample,
                                          void* dlogits, int N, int C, int dtype, float gscale, void* stream) {
    if (!logits || !target_i64 || !lse || !dlogits || N <= 0 || C <= 0) return LWC_EINVAL;
    LWC_CE_DISPATCH(dtype, k_ce_bwd<T><<<(N + 7) / 8, 256, 0, (cudaStream_t)stream>>>(
        (const T*)logits, (const long long*)target_i64, lse, grad_per_sample, (T*)dlogits, N, C, gscale));
    return (int)cudaGetLastError();
}
