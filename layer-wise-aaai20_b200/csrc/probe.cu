// Diagnostics: issue rate of tcgen05.mma (M = 128, K = 16, kind::f16, operands in shared memory) for K-major and
// MN-major operand layouts.  No data movement at all -- one thread per CTA issues MMAs on a zeroed shared-memory tile in
// a loop and the cycle counter brackets the loop -- so the number is the tensor core's own operand-fetch + math rate.
// Used by tools/mma_rate_probe.py; DESIGN.md section 3 quotes the result.  No reference counterpart.
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "common.cuh"
#include "tc.cuh"

namespace lwc {
namespace probe {

using namespace lwc::tc;

template <int N>
__global__ void __launch_bounds__(128, 1) k_mma_rate(int a_mn, int b_mn, int iters, long long* cycles) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    __shared__ uint64_t bar;
    __shared__ uint32_t tmem_slot;
    // A: 128 x 64 bf16 = 16 KB, B: N x 64 bf16 <= 32 KB, both zero
    for (int i = threadIdx.x; i < (16384 + N * 128) / 16; i += blockDim.x) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0, 0, 0, 0);
    if (threadIdx.x == 0) { mbar_init(&bar, 1); fence_barrier_init(); }
    if (threadIdx.x < 32) tmem_alloc<512>(&tmem_slot);
    fence_proxy_async_smem();
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tmem = tmem_slot;
    if (threadIdx.x == 0) {
        const uint32_t a = smem_u32(smem), b = a + 16384;
        const uint64_t da0 = a_mn ? make_desc(a, 8192, 1024) : make_desc(a, 16, 1024);
        const uint64_t db0 = b_mn ? make_desc(b, 8192, 1024) : make_desc(b, 16, 1024);
        const uint64_t sa = a_mn ? 128 : 2, sb = b_mn ? 128 : 2;
        const uint32_t idesc = make_idesc(N, a_mn != 0, b_mn != 0);
        const long long t0 = clock64();
        for (int i = 0; i < iters; ++i) {
            // alternate between two accumulators so back-to-back MMAs are independent
            const uint32_t d = tmem + (uint32_t)((i & 1) * N);
#pragma unroll
            for (int k = 0; k < 4; ++k) umma(d, da0 + k * sa, db0 + k * sb, idesc, 1u);
        }
        umma_commit(&bar);
        mbar_wait(&bar, 0);
        const long long t1 = clock64();
        cycles[blockIdx.x] = t1 - t0;
    }
    fence_before();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc<512>(tmem);
}

}  // namespace probe
}  // namespace lwc

// Average cycles per MMA (M = 128, N = n in {64, 128, 256}, K = 16) over all SMs.  a_mn / b_mn: operand is MN-major.
extern "C" int lwc_probe_mma_rate(int a_mn, int b_mn, int n, int iters, double* cycles_per_mma) {
    using namespace lwc::probe;
    if (!cycles_per_mma || iters <= 0 || (n != 64 && n != 128 && n != 256)) return LWC_EINVAL;
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    long long* d = nullptr;
    LWC_CUDA_CHECK(cudaMalloc(&d, sizeof(long long) * sms));
    const int smem = 16384 + 256 * 128 + 1024;
    if (n == 64) {
        cudaFuncSetAttribute(k_mma_rate<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        k_mma_rate<64><<<sms, 128, smem>>>(a_mn, b_mn, iters, d);
    } else if (n == 128) {
        cudaFuncSetAttribute(k_mma_rate<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        k_mma_rate<128><<<sms, 128, smem>>>(a_mn, b_mn, iters, d);
    } else {
        cudaFuncSetAttribute(k_mma_rate<256>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        k_mma_rate<256><<<sms, 128, smem>>>(a_mn, b_mn, iters, d);
    }
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { cudaFree(d); return (int)e; }
    long long* h = new long long[sms];
    cudaMemcpy(h, d, sizeof(long long) * sms, cudaMemcpyDeviceToHost);
    double tot = 0;
    for (int i = 0; i < sms; ++i) tot += (double)h[i];
    delete[] h;
    cudaFree(d);
    *cycles_per_mma = tot / sms / ((double)iters * 4.0);
    return 0;
}
