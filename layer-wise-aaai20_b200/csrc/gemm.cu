// bf16 GEMM on 5th-generation tensor cores (tcgen05 + TMEM accumulators + TMA operand loads).
//
//   C[M,N] (+)= A[M,K] * B[N,K]^T  (+ bias[N]) (ReLU)      fp32 accumulation in tensor memory
//
// Each operand is either K-major (reduction dim contiguous) or MN-major (output dim contiguous),
// which covers the three GEMMs of a linear / 1x1-convolution layer without any transposition pass:
//   forward  Y  = X  W^T   : A = X  (K-major),  B = W  (K-major)
//   dgrad    dX = dY W     : A = dY (K-major),  B = W  (MN-major)
//   wgrad    dW = dY^T X   : A = dY (MN-major), B = X  (MN-major), split-K over the batch dim
// (reference ops: nn.Linear / 1x1 nn.Conv2d forward+backward through cuBLAS/cuDNN, e.g.
//  CIFAR10/dawn.py:53, CIFAR10/alexnet.py:31-36, CIFAR10/vgg16.py:24-30, IMAGENET/training/resnet.py:55-69,:110)
//
// Kernel structure (one 128 x BN output tile per CTA, 192 threads):
//   warp 0   : TMA producer   - cp.async.bulk.tensor tiles (128B swizzle) into a 4-stage smem ring
//   warp 1   : MMA issuer     - one thread issues tcgen05.mma (128 x BN x 16) with smem descriptors,
//              tcgen05.commit releases ring slots / signals the epilogue; also owns the TMEM allocation
//   warps 2-5: epilogue       - tcgen05.ld the fp32 accumulator (lane = row), bias/ReLU/accumulate,
//              convert, vector stores (or fp32 partials for split-K)
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace lwc {
namespace gemm {

constexpr int BM = 128;
constexpr int BK = 64;            // 64 bf16 = 128 bytes = one swizzle span
constexpr int STAGES = 4;
constexpr int THREADS = 192;
constexpr int UMMA_K = 16;

struct Params {
    void* C;
    float* ws;              // split-K partials [splits][M][N] fp32
    const float* bias;
    int M, N, K, ldc;
    int k_tiles, k_tiles_per_split, splits;
    int relu, accumulate, out_dtype;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t a = smem_u32(bar);
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE;\n\t"
        "bra WAIT_LOOP;\n\t"
        "DONE:\n\t}" ::"r"(a), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// 64-bit shared-memory matrix descriptor, 128B swizzle (cute::UMMA::SmemDescriptor layout):
//   [0,14) start>>4 | [16,30) LBO>>4 | [32,46) SBO>>4 | [46,48) version=1 | [61,64) layout=2 (SW128)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fffu);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32;
    d |= 1ULL << 46;
    d |= 2ULL << 61;
    return d;
}
// instruction descriptor, kind::f16: D=f32, A=B=bf16 (cute::UMMA::InstrDescriptor layout)
__host__ __device__ constexpr uint32_t make_idesc(int n, bool a_mn, bool b_mn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((a_mn ? 1u : 0u) << 15) | ((b_mn ? 1u : 0u) << 16) |
           ((uint32_t)(n >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
}
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
// The four K-steps of one 64-wide K tile in ONE asm block (descriptors advanced inside it), so the operands are moved
// into uniform registers (R2UR) once per tile instead of once per tcgen05.mma.  Measured: 1.4x on the stem kernel's
// small 128x64x16 MMAs (csrc/stem.cu), neutral here where an MMA occupies the tensor pipe for 64-128 cycles.
// GROUP = 1 or 2 (cta_group).
#define LWC_UMMA_X4_BODY(GROUP)                                                                      \
    asm volatile(                                                                                    \
        "{\n\t"                                                                                      \
        ".reg .pred p, pt;\n\t"                                                                      \
        ".reg .b64 xa, xb;\n\t"                                                                      \
        "setp.ne.b32 p, %4, 0;\n\t"                                                                  \
        "setp.eq.u32 pt, %3, %3;\n\t"                                                                \
        "mov.b64 xa, %1;\n\t"                                                                        \
        "mov.b64 xb, %2;\n\t"                                                                        \
        "tcgen05.mma.cta_group::" GROUP ".kind::f16 [%0], xa, xb, %3, p;\n\t"                        \
        "add.u64 xa, xa, %5;\n\t"                                                                    \
        "add.u64 xb, xb, %6;\n\t"                                                                    \
        "tcgen05.mma.cta_group::" GROUP ".kind::f16 [%0], xa, xb, %3, pt;\n\t"                       \
        "add.u64 xa, xa, %5;\n\t"                                                                    \
        "add.u64 xb, xb, %6;\n\t"                                                                    \
        "tcgen05.mma.cta_group::" GROUP ".kind::f16 [%0], xa, xb, %3, pt;\n\t"                       \
        "add.u64 xa, xa, %5;\n\t"                                                                    \
        "add.u64 xb, xb, %6;\n\t"                                                                    \
        "tcgen05.mma.cta_group::" GROUP ".kind::f16 [%0], xa, xb, %3, pt;\n\t"                       \
        "}"                                                                                          \
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate), "n"(A_STEP), "n"(B_STEP) : "memory")
template <uint32_t A_STEP, uint32_t B_STEP>
__device__ __forceinline__ void umma_x4(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    LWC_UMMA_X4_BODY("1");
}
template <uint32_t A_STEP, uint32_t B_STEP>
__device__ __forceinline__ void umma2_x4(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    LWC_UMMA_X4_BODY("2");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t r[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

template <int BN>
struct Smem {
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int BAR_OFF = STAGES * STAGE_BYTES;
    static constexpr int TOTAL = BAR_OFF + 256 + 1024;   // barriers + slack for 1024B alignment
};

template <int BN, bool A_MN, bool B_MN>
__global__ void __launch_bounds__(THREADS) k_gemm(const __grid_constant__ CUtensorMap tmA,
                                                  const __grid_constant__ CUtensorMap tmB, Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    using S = Smem<BN>;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + S::BAR_OFF);
    uint64_t* empty = full + STAGES;
    uint64_t* tmem_full = empty + STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
    const int kt0 = blockIdx.z * p.k_tiles_per_split;
    const int kt1 = min(kt0 + p.k_tiles_per_split, p.k_tiles);
    const int nk = kt1 - kt0;

    if (threadIdx.x == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        mbar_init(tmem_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(BN));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0 && lane == 0) {
        // ---------------- TMA producer ----------------
        for (int i = 0; i < nk; ++i) {
            const int s = i % STAGES;
            const uint32_t ph = (i / STAGES) & 1;
            mbar_wait(&empty[s], ph ^ 1);
            uint8_t* a = smem + s * S::STAGE_BYTES;
            uint8_t* b = a + S::A_BYTES;
            mbar_expect_tx(&full[s], S::STAGE_BYTES);
            const int k0 = (kt0 + i) * BK;
            if (A_MN) {
#pragma unroll
                for (int h = 0; h < BM / 64; ++h) tma_load_2d(a + h * 8192, &tmA, &full[s], m0 + h * 64, k0);
            } else {
                tma_load_2d(a, &tmA, &full[s], k0, m0);
            }
            if (B_MN) {
#pragma unroll
                for (int h = 0; h < BN / 64; ++h) tma_load_2d(b + h * 8192, &tmB, &full[s], n0 + h * 64, k0);
            } else {
                tma_load_2d(b, &tmB, &full[s], k0, n0);
            }
        }
    } else if (warp == 1 && lane == 0) {
        // ---------------- MMA issuer ----------------
        constexpr uint32_t idesc = make_idesc(BN, A_MN, B_MN);
        for (int i = 0; i < nk; ++i) {
            const int s = i % STAGES;
            const uint32_t ph = (i / STAGES) & 1;
            mbar_wait(&full[s], ph);
            tcgen05_fence_after();
            const uint32_t a = smem_u32(smem + s * S::STAGE_BYTES);
            const uint32_t b = a + S::A_BYTES;
            static_assert(BK / UMMA_K == 4, "umma_x4 issues four K-steps");
            // K-major: a K-step is 16 elements = 32 bytes inside the 128B swizzle row; 8-row groups 1024B apart.
            // MN-major: a K-step is 16 k-rows of 128 bytes = 2048 bytes; 64-wide MN blocks 8192B apart.
            // (steps below in descriptor start-address units of 16 bytes)
            const uint64_t da = A_MN ? make_desc(a, 8192, 1024) : make_desc(a, 16, 1024);
            const uint64_t db = B_MN ? make_desc(b, 8192, 1024) : make_desc(b, 16, 1024);
            umma_x4<(A_MN ? 128 : 2), (B_MN ? 128 : 2)>(tmem_base, da, db, idesc, i > 0 ? 1u : 0u);
            umma_commit(&empty[s]);           // frees the smem slot once these MMAs have read it
        }
        umma_commit(tmem_full);               // accumulator complete
    } else if (warp >= 2) {
        // ---------------- epilogue ----------------
        mbar_wait(tmem_full, 0);
        tcgen05_fence_after();
        const int q = warp & 3;                      // TMEM lane quarter this warp may access
        const int row = m0 + q * 32 + lane;
        const bool row_ok = row < p.M;
#pragma unroll 1
        for (int c = 0; c < BN; c += 32) {
            uint32_t r[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c, r);
            const int col0 = n0 + c;
            if (!row_ok || col0 >= p.N) continue;
            const int ncol = min(32, p.N - col0);
            if (p.splits > 1) {
                float* dst = p.ws + ((size_t)blockIdx.z * p.M + row) * p.N + col0;
                if (ncol == 32 && (p.N & 3) == 0) {
#pragma unroll
                    for (int j = 0; j < 32; j += 4)
                        *reinterpret_cast<float4*>(dst + j) = make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]),
                                                                          __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (j < ncol) dst[j] = __uint_as_float(r[j]);
                }
                continue;
            }
            float v[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) {
                float x = __uint_as_float(r[j]);
                if (p.bias && j < ncol) x += p.bias[col0 + j];
                v[j] = x;
            }
            if (p.out_dtype == LWC_F32) {
                float* dst = reinterpret_cast<float*>(p.C) + (size_t)row * p.ldc + col0;
#pragma unroll
                for (int j = 0; j < 32; ++j) {
                    if (j < ncol) {
                        float x = v[j] + (p.accumulate ? dst[j] : 0.0f);
                        dst[j] = p.relu ? fmaxf(x, 0.0f) : x;
                    }
                }
            } else {
                __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.C) + (size_t)row * p.ldc + col0;
                const bool vec = ncol == 32 && (p.ldc & 7) == 0 && (col0 & 7) == 0 &&
                                 (reinterpret_cast<uintptr_t>(p.C) & 15) == 0;
                if (vec) {
#pragma unroll
                    for (int j = 0; j < 32; j += 8) {
                        uint4 o;
                        uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
                        uint4 old = make_uint4(0, 0, 0, 0);
                        if (p.accumulate) old = *reinterpret_cast<const uint4*>(dst + j);
                        const uint32_t* oldw = reinterpret_cast<const uint32_t*>(&old);
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            float x0 = v[j + 2 * e], x1 = v[j + 2 * e + 1];
                            if (p.accumulate) {
                                __nv_bfloat162 ob = *reinterpret_cast<const __nv_bfloat162*>(&oldw[e]);
                                x0 += __low2float(ob);
                                x1 += __high2float(ob);
                            }
                            if (p.relu) { x0 = fmaxf(x0, 0.0f); x1 = fmaxf(x1, 0.0f); }
                            __nv_bfloat162 nb = __floats2bfloat162_rn(x0, x1);
                            ow[e] = *reinterpret_cast<uint32_t*>(&nb);
                        }
                        *reinterpret_cast<uint4*>(dst + j) = o;
                    }
                } else {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        if (j < ncol) {
                            float x = v[j] + (p.accumulate ? __bfloat162float(dst[j]) : 0.0f);
                            dst[j] = __float2bfloat16_rn(p.relu ? fmaxf(x, 0.0f) : x);
                        }
                    }
                }
            }
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(BN));
    }
}


// ======================================================================================================
// 2-CTA variant: a CTA pair (thread-block cluster of 2, cta_group::2) computes one 256 x 256 tile.
// Each CTA stages its own 128 rows of A and its own 128 of the 256 B rows (so operand traffic per MMA
// halves), the leader CTA issues tcgen05.mma.cta_group::2 (M = 256) which reads both CTAs' shared memory
// and writes 128 accumulator rows into each CTA's tensor memory; TMA loads of both CTAs complete on the
// leader's mbarrier, tcgen05.commit multicasts slot-free / accumulator-ready to both CTAs.
// ======================================================================================================
constexpr int BN2 = 256;
constexpr int STAGES2 = 6;
constexpr uint32_t kPeerBitMask = 0xFEFFFFFFu;     // clears the CTA-rank bit of a shared::cluster address

__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void tma_load_2d_2sm(void* dst, const CUtensorMap* map, uint64_t* leader_bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(leader_bar) & kPeerBitMask), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void umma2(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit_2sm(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;"
                 ::"r"(smem_u32(bar)), "h"((uint16_t)3) : "memory");
}
__host__ __device__ constexpr uint32_t make_idesc2(int m, int n, bool a_mn, bool b_mn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((a_mn ? 1u : 0u) << 15) | ((b_mn ? 1u : 0u) << 16) |
           ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

// One 32-column block of one accumulator row -> global memory (shared by both kernels).
__device__ __forceinline__ void epilogue_store(const Params& p, const uint32_t r[32], int row, int col0, int split) {
    if (row >= p.M || col0 >= p.N) return;
    const int ncol = min(32, p.N - col0);
    if (p.splits > 1) {
        float* dst = p.ws + ((size_t)split * p.M + row) * p.N + col0;
        if (ncol == 32 && (p.N & 3) == 0) {
#pragma unroll
            for (int j = 0; j < 32; j += 4)
                *reinterpret_cast<float4*>(dst + j) = make_float4(__uint_as_float(r[j]), __uint_as_float(r[j + 1]),
                                                                  __uint_as_float(r[j + 2]), __uint_as_float(r[j + 3]));
        } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
                if (j < ncol) dst[j] = __uint_as_float(r[j]);
        }
        return;
    }
    float v[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) {
        float x = __uint_as_float(r[j]);
        if (p.bias && j < ncol) x += p.bias[col0 + j];
        v[j] = x;
    }
    if (p.out_dtype == LWC_F32) {
        float* dst = reinterpret_cast<float*>(p.C) + (size_t)row * p.ldc + col0;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            if (j < ncol) {
                float x = v[j] + (p.accumulate ? dst[j] : 0.0f);
                dst[j] = p.relu ? fmaxf(x, 0.0f) : x;
            }
        }
        return;
    }
    __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.C) + (size_t)row * p.ldc + col0;
    const bool vec = ncol == 32 && (p.ldc & 7) == 0 && (col0 & 7) == 0 && (reinterpret_cast<uintptr_t>(p.C) & 15) == 0;
    if (vec) {
#pragma unroll
        for (int j = 0; j < 32; j += 8) {
            uint4 o;
            uint32_t* ow = reinterpret_cast<uint32_t*>(&o);
            uint4 old = make_uint4(0, 0, 0, 0);
            if (p.accumulate) old = *reinterpret_cast<const uint4*>(dst + j);
            const uint32_t* oldw = reinterpret_cast<const uint32_t*>(&old);
#pragma unroll
            for (int e = 0; e < 4; ++e) {
                float x0 = v[j + 2 * e], x1 = v[j + 2 * e + 1];
                if (p.accumulate) {
                    __nv_bfloat162 ob = *reinterpret_cast<const __nv_bfloat162*>(&oldw[e]);
                    x0 += __low2float(ob);
                    x1 += __high2float(ob);
                }
                if (p.relu) { x0 = fmaxf(x0, 0.0f); x1 = fmaxf(x1, 0.0f); }
                __nv_bfloat162 nb = __floats2bfloat162_rn(x0, x1);
                ow[e] = *reinterpret_cast<uint32_t*>(&nb);
            }
            *reinterpret_cast<uint4*>(dst + j) = o;
        }
    } else {
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            if (j < ncol) {
                float x = v[j] + (p.accumulate ? __bfloat162float(dst[j]) : 0.0f);
                dst[j] = __float2bfloat16_rn(p.relu ? fmaxf(x, 0.0f) : x);
            }
        }
    }
}

struct Smem2 {
    static constexpr int A_BYTES = BM * BK * 2;            // this CTA's 128 rows of A
    static constexpr int B_BYTES = (BN2 / 2) * BK * 2;     // this CTA's 128 of the 256 B rows
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int BAR_OFF = STAGES2 * STAGE_BYTES;
    static constexpr int TOTAL = BAR_OFF + 256 + 1024;
};

template <bool A_MN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(THREADS) k_gemm2(const __grid_constant__ CUtensorMap tmA,
                                                                              const __grid_constant__ CUtensorMap tmB, Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    using S = Smem2;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + S::BAR_OFF);
    uint64_t* empty = full + STAGES2;
    uint64_t* tmem_full = empty + STAGES2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t cta = cluster_ctarank();
    const bool leader = cta == 0;
    const int m0 = (blockIdx.x >> 1) * (2 * BM) + (int)cta * BM;      // my 128 rows of the 256-row tile
    const int n0 = blockIdx.y * BN2;
    const int nb0 = n0 + (int)cta * (BN2 / 2);                        // my half of the B rows
    const int kt0 = blockIdx.z * p.k_tiles_per_split;
    const int kt1 = min(kt0 + p.k_tiles_per_split, p.k_tiles);
    const int nk = kt1 - kt0;

    if (threadIdx.x == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
        for (int s = 0; s < STAGES2; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        mbar_init(tmem_full, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(BN2));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;");
    }
    tcgen05_fence_before();
    __syncthreads();
    cluster_sync_all();            // barriers initialised and TMEM allocated in BOTH CTAs before any cross-CTA traffic
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0 && lane == 0) {
        // ---------------- TMA producer (both CTAs; completion lands on the leader's barrier) ----------------
        for (int i = 0; i < nk; ++i) {
            const int s = i % STAGES2;
            const uint32_t ph = (i / STAGES2) & 1;
            mbar_wait(&empty[s], ph ^ 1);
            uint8_t* a = smem + s * S::STAGE_BYTES;
            uint8_t* b = a + S::A_BYTES;
            if (leader) mbar_expect_tx(&full[s], 2 * S::STAGE_BYTES);
            const int k0 = (kt0 + i) * BK;
            if (A_MN) {
#pragma unroll
                for (int h = 0; h < BM / 64; ++h) tma_load_2d_2sm(a + h * 8192, &tmA, &full[s], m0 + h * 64, k0);
            } else {
                tma_load_2d_2sm(a, &tmA, &full[s], k0, m0);
            }
            if (B_MN) {
#pragma unroll
                for (int h = 0; h < (BN2 / 2) / 64; ++h) tma_load_2d_2sm(b + h * 8192, &tmB, &full[s], nb0 + h * 64, k0);
            } else {
                tma_load_2d_2sm(b, &tmB, &full[s], k0, nb0);
            }
        }
    } else if (warp == 1 && lane == 0 && leader) {
        // ---------------- MMA issuer (leader CTA only) ----------------
        constexpr uint32_t idesc = make_idesc2(2 * BM, BN2, A_MN, B_MN);
        for (int i = 0; i < nk; ++i) {
            const int s = i % STAGES2;
            const uint32_t ph = (i / STAGES2) & 1;
            mbar_wait(&full[s], ph);
            tcgen05_fence_after();
            const uint32_t a = smem_u32(smem + s * S::STAGE_BYTES);
            const uint32_t b = a + S::A_BYTES;
            const uint64_t da = A_MN ? make_desc(a, 8192, 1024) : make_desc(a, 16, 1024);
            const uint64_t db = B_MN ? make_desc(b, 8192, 1024) : make_desc(b, 16, 1024);
            umma2_x4<(A_MN ? 128 : 2), (B_MN ? 128 : 2)>(tmem_base, da, db, idesc, i > 0 ? 1u : 0u);
            umma_commit_2sm(&empty[s]);
        }
        umma_commit_2sm(tmem_full);
    } else if (warp >= 2) {
        // ---------------- epilogue (both CTAs, own 128 accumulator rows) ----------------
        mbar_wait(tmem_full, 0);
        tcgen05_fence_after();
        const int q = warp & 3;
        const int row = m0 + q * 32 + lane;
#pragma unroll 1
        for (int c = 0; c < BN2; c += 32) {
            uint32_t r[32];
            tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)c, r);
            epilogue_store(p, r, row, n0 + c, blockIdx.z);
        }
    }
    tcgen05_fence_before();
    __syncthreads();
    cluster_sync_all();            // the peer may still be reading my shared memory / writing my TMEM until here
    if (warp == 1) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(BN2));
    }
}

// Sum split-K partials and apply the epilogue.
__global__ void k_splitk_reduce(Params p) {
    const size_t total = (size_t)p.M * p.N;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
        const int row = (int)(i / p.N), col = (int)(i % p.N);
        float x = 0.0f;
        for (int s = 0; s < p.splits; ++s) x += p.ws[(size_t)s * total + i];
        if (p.bias) x += p.bias[col];
        if (p.out_dtype == LWC_F32) {
            float* dst = reinterpret_cast<float*>(p.C) + (size_t)row * p.ldc + col;
            if (p.accumulate) x += *dst;
            *dst = p.relu ? fmaxf(x, 0.0f) : x;
        } else {
            __nv_bfloat16* dst = reinterpret_cast<__nv_bfloat16*>(p.C) + (size_t)row * p.ldc + col;
            if (p.accumulate) x += __bfloat162float(*dst);
            *dst = __float2bfloat16_rn(p.relu ? fmaxf(x, 0.0f) : x);
        }
    }
}

}  // namespace gemm
}  // namespace lwc

using namespace lwc::gemm;

// rows x cols (cols contiguous) bf16 matrix -> 2D tensor map with a {box_cols, box_rows} box, 128B swizzle
static int make_map(CUtensorMap* map, const void* ptr, uint64_t rows, uint64_t cols, uint64_t ld_elems,
                    uint32_t box_cols, uint32_t box_rows) {
    cuuint64_t gdim[2] = {cols, rows};
    cuuint64_t gstride[1] = {ld_elems * 2};
    cuuint32_t box[2] = {box_cols, box_rows};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = cuTensorMapEncodeTiled(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(ptr), gdim, gstride,
                                        box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS && getenv("LWC_DEBUG"))
        fprintf(stderr, "cuTensorMapEncodeTiled -> %d (ptr=%p rows=%llu cols=%llu ld=%llu box=%ux%u)\n", (int)r, ptr,
                (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld_elems, box_cols, box_rows);
    return r == CUDA_SUCCESS ? 0 : LWC_EINVAL;
}

template <int BN, bool A_MN, bool B_MN>
static int launch(const CUtensorMap& ta, const CUtensorMap& tb, const Params& p, cudaStream_t st) {
    using S = Smem<BN>;
    static bool attr_set = false;
    if (!attr_set) {
        LWC_CUDA_CHECK(cudaFuncSetAttribute(k_gemm<BN, A_MN, B_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL));
        attr_set = true;
    }
    dim3 grid((p.M + BM - 1) / BM, (p.N + BN - 1) / BN, p.splits);
    k_gemm<BN, A_MN, B_MN><<<grid, THREADS, S::TOTAL, st>>>(ta, tb, p);
    return 0;
}

template <bool A_MN, bool B_MN>
static int launch2(const CUtensorMap& ta, const CUtensorMap& tb, const Params& p, cudaStream_t st) {
    static bool attr_set = false;
    if (!attr_set) {
        LWC_CUDA_CHECK(cudaFuncSetAttribute(k_gemm2<A_MN, B_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem2::TOTAL));
        attr_set = true;
    }
    dim3 grid(2 * ((p.M + 2 * BM - 1) / (2 * BM)), (p.N + BN2 - 1) / BN2, p.splits);
    k_gemm2<A_MN, B_MN><<<grid, THREADS, Smem2::TOTAL, st>>>(ta, tb, p);
    return 0;
}

extern "C" size_t lwc_gemm_workspace_bytes(int M, int N, int splits) {
    return splits > 1 ? (size_t)splits * M * N * sizeof(float) : 0;
}

extern "C" int lwc_gemm_pick_splits(int M, int N, int K) {
    const int bn = N <= 64 ? 64 : 128;
    const long tiles = (long)((M + BM - 1) / BM) * ((N + bn - 1) / bn);
    const int k_tiles = (K + BK - 1) / BK;
    if (tiles >= 148 || k_tiles < 8) return 1;
    long s = (2 * 148 + tiles - 1) / tiles;
    if (s > k_tiles / 4) s = k_tiles / 4;
    if (s > 64) s = 64;
    return s < 1 ? 1 : (int)s;
}

// A: a_mn ? stored [K][M] (lda >= M) : stored [M][K] (lda >= K).  B likewise with N.
extern "C" int lwc_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                             int a_mn, int b_mn, int out_dtype, const float* bias, int relu, int accumulate,
                             int splits, float* ws, void* stream) {
#define LWC_GEMM_FAIL(why) do { if (getenv("LWC_DEBUG")) fprintf(stderr, "lwc_gemm_bf16: %s (M=%d N=%d K=%d lda=%d ldb=%d ldc=%d a_mn=%d b_mn=%d splits=%d)\n", why, M, N, K, lda, ldb, ldc, a_mn, b_mn, splits); return LWC_EINVAL; } while (0)
    if (!A || !B || !C || M <= 0 || N <= 0 || K <= 0) LWC_GEMM_FAIL("null pointer or empty shape");
    if ((lda & 7) || (ldb & 7)) LWC_GEMM_FAIL("leading dimensions must be multiples of 8 elements (TMA 16-byte strides)");
    if ((reinterpret_cast<uintptr_t>(A) & 15) || (reinterpret_cast<uintptr_t>(B) & 15)) LWC_GEMM_FAIL("operands must be 16-byte aligned");
    if (out_dtype != LWC_F32 && out_dtype != LWC_BF16) LWC_GEMM_FAIL("output dtype must be fp32 or bf16");
    if (splits < 1) splits = 1;
    if (splits > 1 && !ws) LWC_GEMM_FAIL("split-K needs a workspace");
    // cuTensorMapEncodeTiled is a driver-API call and needs a current context on THIS thread; autograd's
    // worker threads may not have touched the runtime yet (CUDA_ERROR_INVALID_CONTEXT otherwise).
    {
        CUcontext ctx = nullptr;
        if (cuCtxGetCurrent(&ctx) != CUDA_SUCCESS || !ctx) {      // no sync, no allocation: safe inside graph capture
            int dev = 0;
            cudaGetDevice(&dev);
            CUcontext prim = nullptr;
            if (cuDevicePrimaryCtxRetain(&prim, dev) == CUDA_SUCCESS) cuCtxSetCurrent(prim);
        }
    }
    cudaStream_t st = (cudaStream_t)stream;
    static int force_1cta = -1;
    if (force_1cta < 0) { const char* e = getenv("LWC_GEMM_1CTA"); force_1cta = (e && e[0] == '1') ? 1 : 0; }
    const bool two_cta = !force_1cta && M >= 256 && N >= 256;        // 256 x 256 tiles on CTA pairs
    const int BN = N <= 64 ? 64 : 128;
    Params p;
    memset(&p, 0, sizeof(p));
    p.C = C; p.ws = ws; p.bias = bias;
    p.M = M; p.N = N; p.K = K; p.ldc = ldc;
    p.k_tiles = (K + BK - 1) / BK;
    if (splits > p.k_tiles) splits = p.k_tiles;
    p.k_tiles_per_split = (p.k_tiles + splits - 1) / splits;
    splits = (p.k_tiles + p.k_tiles_per_split - 1) / p.k_tiles_per_split;    // no empty split
    p.splits = splits;
    p.relu = relu; p.accumulate = accumulate; p.out_dtype = out_dtype;
    CUtensorMap ta, tb;
    int rc;
    rc = a_mn ? make_map(&ta, A, K, M, lda, 64, BK) : make_map(&ta, A, M, K, lda, BK, BM);
    if (rc) LWC_GEMM_FAIL("cuTensorMapEncodeTiled(A) failed");
    rc = b_mn ? make_map(&tb, B, K, N, ldb, 64, BK) : make_map(&tb, B, N, K, ldb, BK, two_cta ? BN2 / 2 : BN);
    if (rc) LWC_GEMM_FAIL("cuTensorMapEncodeTiled(B) failed");
    if (two_cta) {
        if (!a_mn && !b_mn) rc = launch2<false, false>(ta, tb, p, st);
        else if (!a_mn && b_mn) rc = launch2<false, true>(ta, tb, p, st);
        else if (a_mn && b_mn) rc = launch2<true, true>(ta, tb, p, st);
        else rc = launch2<true, false>(ta, tb, p, st);
    } else {
#define LWC_GEMM_CASE(bn, amn, bmn) \
    if (BN == bn && (bool)a_mn == amn && (bool)b_mn == bmn) rc = launch<bn, amn, bmn>(ta, tb, p, st);
    LWC_GEMM_CASE(64, false, false) LWC_GEMM_CASE(64, false, true) LWC_GEMM_CASE(64, true, true) LWC_GEMM_CASE(64, true, false)
    LWC_GEMM_CASE(128, false, false) LWC_GEMM_CASE(128, false, true) LWC_GEMM_CASE(128, true, true) LWC_GEMM_CASE(128, true, false)
#undef LWC_GEMM_CASE
    }
    if (rc) return rc;
    if (splits > 1) {
        const size_t total = (size_t)M * N;
        int blocks = (int)((total + 255) / 256);
        if (blocks > 148 * 8) blocks = 148 * 8;
        k_splitk_reduce<<<blocks, 256, 0, st>>>(p);
    }
    return (int)cudaGetLastError();
}
