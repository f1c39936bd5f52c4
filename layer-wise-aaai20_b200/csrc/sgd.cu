// Fused multi-tensor SGD (momentum / Nesterov / weight decay / grad scale / low-precision copy).
//
// torch.optim.SGD semantics per element:
//   d = g * grad_scale + wd * w
//   v = first_step ? d : mu * v + (1 - dampening) * d          (only when mu != 0)
//   d = nesterov ? d + mu * v : v
//   w -= lr * d ;  model = cast(w)
// One launch covers every tensor of a parameter group through a chunk table; the reference's
// optimiser runs a Python loop with ~4 elementwise launches per tensor
// (CIFAR10/torch_backend.py:132-143; train_imagenet_nv.py:441) and a separate
// master->model copy pass under fp16 (fp16util.py:125-139).
#include <vector>

#include "common.cuh"

namespace lwc {
struct SgdChunk { int tensor; int len; long long start; };

template <typename GradT, typename ModelT>
__global__ void __launch_bounds__(kThreads) k_sgd(const LwcSgdTensor* __restrict__ tensors,
                                                  const SgdChunk* __restrict__ chunks, LwcSgdDesc d) {
    if (d.skip_flag && *d.skip_flag) return;           // the exchange timed out on a peer: leave the weights alone
    const SgdChunk c = chunks[blockIdx.x];
    const LwcSgdTensor t = tensors[c.tensor];
    float* w = reinterpret_cast<float*>(t.param) + c.start;
    const GradT* g = reinterpret_cast<const GradT*>(t.grad) + c.start;
    float* m = t.mom ? reinterpret_cast<float*>(t.mom) + c.start : nullptr;
    ModelT* md = t.model ? reinterpret_cast<ModelT*>(t.model) + c.start : nullptr;
    const bool wv = vec_ok(w), gv = vec_ok(g), mv = m && vec_ok(m), dv = md && vec_ok(md);
    const int tid = threadIdx.x;
    const float lr = d.lr_dev ? *d.lr_dev : d.lr;      // device-resident learning rate: replayable in a CUDA graph
#pragma unroll 2
    for (int it = 0; it < kQuadsPerThread; ++it) {
        const int j = (it * kThreads + tid) * 4;
        if (j >= c.len) break;
        float wq[4], gq[4], mq[4] = {0.f, 0.f, 0.f, 0.f};
        load_quad<float>(w, j, c.len, wv, wq);
        load_quad<GradT>(g, j, c.len, gv, gq);
        if (m && !d.first_step) load_quad<float>(m, j, c.len, mv, mq);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float dd = fmaf(d.weight_decay, wq[k], gq[k] * d.grad_scale);
            if (m) {
                mq[k] = d.first_step ? dd : fmaf(d.momentum, mq[k], (1.0f - d.dampening) * dd);
                dd = d.nesterov ? fmaf(d.momentum, mq[k], dd) : mq[k];
            }
            wq[k] = fmaf(-lr, dd, wq[k]);
        }
        store_quad<float>(w, j, c.len, wv, wq);
        if (m) store_quad<float>(m, j, c.len, mv, mq);
        if (md) store_quad<ModelT>(md, j, c.len, dv, wq);
    }
}
// Pure low-precision training (no fp32 masters: the CIFAR recipes with --dtype bf16): weights, momentum and gradients
// all in T; the arithmetic is fp32, the results are rounded once.
template <typename T>
__global__ void __launch_bounds__(kThreads) k_sgd_lowp(const LwcSgdTensor* __restrict__ tensors,
                                                       const SgdChunk* __restrict__ chunks, LwcSgdDesc d) {
    if (d.skip_flag && *d.skip_flag) return;
    const SgdChunk c = chunks[blockIdx.x];
    const LwcSgdTensor t = tensors[c.tensor];
    T* w = reinterpret_cast<T*>(t.param) + c.start;
    const T* g = reinterpret_cast<const T*>(t.grad) + c.start;
    T* m = t.mom ? reinterpret_cast<T*>(t.mom) + c.start : nullptr;
    const bool wv = vec_ok(w), gv = vec_ok(g), mv = m && vec_ok(m);
    const int tid = threadIdx.x;
    const float lr = d.lr_dev ? *d.lr_dev : d.lr;
#pragma unroll 2
    for (int it = 0; it < kQuadsPerThread; ++it) {
        const int j = (it * kThreads + tid) * 4;
        if (j >= c.len) break;
        float wq[4], gq[4], mq[4] = {0.f, 0.f, 0.f, 0.f};
        load_quad<T>(w, j, c.len, wv, wq);
        load_quad<T>(g, j, c.len, gv, gq);
        if (m && !d.first_step) load_quad<T>(m, j, c.len, mv, mq);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            float dd = fmaf(d.weight_decay, wq[k], gq[k] * d.grad_scale);
            if (m) {
                mq[k] = d.first_step ? dd : fmaf(d.momentum, mq[k], (1.0f - d.dampening) * dd);
                dd = d.nesterov ? fmaf(d.momentum, mq[k], dd) : mq[k];
            }
            wq[k] = fmaf(-lr, dd, wq[k]);
        }
        store_quad<T>(w, j, c.len, wv, wq);
        if (m) store_quad<T>(m, j, c.len, mv, mq);
    }
}
}  // namespace lwc

using namespace lwc;

struct LwcSgdPlan {
    void* d_tensors = nullptr;
    void* d_chunks = nullptr;
    int num_chunks = 0;
    int num_tensors = 0;
};

extern "C" int lwc_sgd_plan_create(LwcSgdPlan** out, const LwcSgdTensor* tensors, int32_t n) {
    if (!out || !tensors || n <= 0) return LWC_EINVAL;
    std::vector<SgdChunk> chunks;
    for (int t = 0; t < n; ++t) {
        if (!tensors[t].param || !tensors[t].grad || tensors[t].n < 0) return LWC_EINVAL;
        for (long long s = 0; s < tensors[t].n; s += kChunk) {
            SgdChunk c;
            c.tensor = t;
            c.start = s;
            c.len = (int)((tensors[t].n - s) < kChunk ? (tensors[t].n - s) : kChunk);
            chunks.push_back(c);
        }
    }
    LwcSgdPlan* P = new LwcSgdPlan();
    P->num_chunks = (int)chunks.size();
    P->num_tensors = n;
    cudaError_t e = cudaMalloc(&P->d_tensors, sizeof(LwcSgdTensor) * n);
    if (e == cudaSuccess) e = cudaMalloc(&P->d_chunks, sizeof(SgdChunk) * (chunks.size() ? chunks.size() : 1));
    if (e == cudaSuccess) e = cudaMemcpy(P->d_tensors, tensors, sizeof(LwcSgdTensor) * n, cudaMemcpyHostToDevice);
    if (e == cudaSuccess && !chunks.empty())
        e = cudaMemcpy(P->d_chunks, chunks.data(), sizeof(SgdChunk) * chunks.size(), cudaMemcpyHostToDevice);
    if (e != cudaSuccess) { lwc_sgd_plan_destroy(P); return (int)e; }
    *out = P;
    return 0;
}

extern "C" int lwc_sgd_plan_destroy(LwcSgdPlan* P) {
    if (!P) return 0;
    cudaFree(P->d_tensors);
    cudaFree(P->d_chunks);
    delete P;
    return 0;
}

template <typename GradT>
static int launch_sgd(LwcSgdPlan* P, const LwcSgdDesc& d, cudaStream_t st) {
    const LwcSgdTensor* t = (const LwcSgdTensor*)P->d_tensors;
    const SgdChunk* c = (const SgdChunk*)P->d_chunks;
    switch (d.model_dtype) {
        case LWC_F32: k_sgd<GradT, float><<<P->num_chunks, kThreads, 0, st>>>(t, c, d); break;
        case LWC_BF16: k_sgd<GradT, __nv_bfloat16><<<P->num_chunks, kThreads, 0, st>>>(t, c, d); break;
        case LWC_F16: k_sgd<GradT, __half><<<P->num_chunks, kThreads, 0, st>>>(t, c, d); break;
        default: return LWC_EINVAL;
    }
    return 0;
}

extern "C" int lwc_sgd_step(LwcSgdPlan* P, const LwcSgdDesc* d, void* stream) {
    if (!P || !d) return LWC_EINVAL;
    if (P->num_chunks == 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    int rc;
    if (d->param_dtype != LWC_F32) {
        // weights, momentum and gradients share one 16-bit type
        if (d->grad_dtype != d->param_dtype) return LWC_EINVAL;
        const LwcSgdTensor* t = (const LwcSgdTensor*)P->d_tensors;
        const SgdChunk* c = (const SgdChunk*)P->d_chunks;
        if (d->param_dtype == LWC_BF16) k_sgd_lowp<__nv_bfloat16><<<P->num_chunks, kThreads, 0, st>>>(t, c, *d);
        else if (d->param_dtype == LWC_F16) k_sgd_lowp<__half><<<P->num_chunks, kThreads, 0, st>>>(t, c, *d);
        else return LWC_EINVAL;
        return (int)cudaGetLastError();
    }
    switch (d->grad_dtype) {
        case LWC_F32: rc = launch_sgd<float>(P, *d, st); break;
        case LWC_BF16: rc = launch_sgd<__nv_bfloat16>(P, *d, st); break;
        case LWC_F16: rc = launch_sgd<__half>(P, *d, st); break;
        default: return LWC_EINVAL;
    }
    if (rc) return rc;
    return (int)cudaGetLastError();
}
