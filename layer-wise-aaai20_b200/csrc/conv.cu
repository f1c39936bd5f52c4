// Persistent implicit-GEMM convolution on tcgen05 / TMEM / TMA for channels-last bf16 activations.
//
//   Y[p, co] = sum_{tap, ci} X[p + tap, ci] * Wt[co, tap, ci]        p = output pixel (n, h, w)
//
// * 1x1 stride-1 convolutions are plain GEMMs over M = N*H*W pixels (2-D tensor maps);
// * RxS stride-1 convolutions walk the taps in the K loop: every (tap, 64-channel chunk) is one 4-D TMA box
//   load of a spatial patch of X shifted by the tap offset -- out-of-bounds rows/columns are zero-filled
//   by the TMA unit, which *is* the padding; no im2col buffer exists anywhere.
// The kernel is persistent (one CTA per SM loops over output tiles), the accumulator is double-buffered in
// tensor memory so the epilogue of tile i overlaps the TMA/MMA main loop of tile i+1, and the epilogue
// goes TMEM -> registers -> swizzled shared memory -> TMA store (coalesced, asynchronous).
// Optional epilogue fusion: per-channel sum / sum-of-squares of the (bf16-rounded) output tile, written as
// per-tile partials for the following batch-norm -- its separate statistics pass disappears.
//
// Warp roles (192 threads): warp 0 TMA producer, warp 1 MMA issuer + TMEM owner, warps 2-5 epilogue.
// Reference ops: nn.Conv2d forward / dgrad through cuDNN (IMAGENET/training/resnet.py:21-24,:55-69;
// CIFAR10/dawn.py:23-35) followed by nn.BatchNorm2d's statistics kernel.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace lwc {
namespace conv {

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int STAGES = 4;
constexpr int THREADS = 192;
constexpr int UMMA_K = 16;
constexpr int NCBUF = 3;             // output / addend staging buffers: the addend of tile t+2 is in flight while t is written

struct Params {
    int M, N, Kc, taps, kw, pad;     // rows (2-D mode), out channels, in channels per tap, taps, kernel width, padding
    int mode4d;
    int H, W, Nimg, Wb, Hb, Nb, tiles_w, tiles_h;
    int tiles_m, tiles_n, rows_per_tile;
    int relu;
    const float* bias;
    float* stats;                    // [tiles_m][2][N] per-tile column sums / sums of squares, or NULL
    const __nv_bfloat16* addend;     // optional [M][N] tensor added in the epilogue (2-D mode): Y = conv(X) + addend
    const unsigned char* mask;       // optional [M][N/8] bit mask (2-D mode, N % 128 == 0): outputs whose bit is 0 are zeroed
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t a = smem_u32(bar);
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "CONV_WAIT:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra CONV_DONE;\n\t"
        "bra CONV_WAIT;\n\t"
        "CONV_DONE:\n\t}" ::"r"(a), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_load_4d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2, int c3) {
    asm volatile(
        "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, const void* src, int c0, int c1) {
    asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                 ::"l"(map), "r"(smem_u32(src)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* map, const void* src, int c0, int c1, int c2, int c3) {
    asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
                 ::"l"(map), "r"(smem_u32(src)), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fffu);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32;
    d |= 1ULL << 46;
    d |= 2ULL << 61;
    return d;
}
__host__ __device__ constexpr uint32_t make_idesc(int n, bool b_mn) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((b_mn ? 1u : 0u) << 16) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
}
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
}
// The four K-steps of one 64-wide K tile in ONE asm block (descriptors advanced inside it): the operands are moved into
// uniform registers (R2UR) once per tile instead of once per tcgen05.mma (measured neutral here, 1.4x in csrc/stem.cu).
template <uint32_t A_STEP, uint32_t B_STEP>
__device__ __forceinline__ void umma_x4(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p, pt;\n\t"
        ".reg .b64 xa, xb;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "setp.eq.u32 pt, %3, %3;\n\t"
        "mov.b64 xa, %1;\n\t"
        "mov.b64 xb, %2;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], xa, xb, %3, p;\n\t"
        "add.u64 xa, xa, %5;\n\t"
        "add.u64 xb, xb, %6;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], xa, xb, %3, pt;\n\t"
        "add.u64 xa, xa, %5;\n\t"
        "add.u64 xb, xb, %6;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], xa, xb, %3, pt;\n\t"
        "add.u64 xa, xa, %5;\n\t"
        "add.u64 xb, xb, %6;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], xa, xb, %3, pt;\n\t"
        "}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate), "n"(A_STEP), "n"(B_STEP) : "memory");
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
// Issue without waiting (several loads in flight); tmem_ld_wait() + use_after_wait() before the registers are read.
__device__ __forceinline__ void tmem_ld32_issue(uint32_t taddr, uint32_t r[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
// Ties the loaded registers to a point after the wait: without a data dependency the compiler may schedule
// arithmetic on them above the wait.
__device__ __forceinline__ void use_after_wait(uint32_t r[32]) {
    asm volatile(""
                 : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]),
                   "+r"(r[8]), "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]),
                   "+r"(r[16]), "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]),
                   "+r"(r[24]), "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])
                 :: "memory");
}
__device__ __forceinline__ void epi_bar_sync() { asm volatile("bar.sync 1, 128;" ::: "memory"); }

template <int BN>
struct Smem {
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int C_BYTES = BM * BN * 2;                 // one staging buffer (BN/64 swizzled 128x64 sub-tiles)
    static constexpr int C_OFF = STAGES * STAGE_BYTES;
    static constexpr int BAR_OFF = C_OFF + NCBUF * C_BYTES;
    static constexpr int TOTAL = BAR_OFF + 256 + 1024;
};

// tile index -> coordinates
struct TileCoord { int mt, nt, m0, w0, h0, n0; };
__device__ __forceinline__ TileCoord tile_coord(const Params& p, int tile) {
    TileCoord t;
    t.mt = tile / p.tiles_n;
    t.nt = tile % p.tiles_n;
    t.m0 = t.mt * BM;
    t.w0 = t.h0 = t.n0 = 0;
    if (p.mode4d) {
        const int tw = t.mt % p.tiles_w;
        const int th = (t.mt / p.tiles_w) % p.tiles_h;
        const int tn = t.mt / (p.tiles_w * p.tiles_h);
        t.w0 = tw * p.Wb; t.h0 = th * p.Hb; t.n0 = tn * p.Nb;
    }
    return t;
}

template <int BN, bool B_MN>
__global__ void __launch_bounds__(THREADS, 1) k_conv(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                                                     const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmD,
                                                     Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    using S = Smem<BN>;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + S::BAR_OFF);
    uint64_t* empty = full + STAGES;
    uint64_t* tmem_full = empty + STAGES;      // [2]
    uint64_t* tmem_empty = tmem_full + 2;      // [2]
    uint64_t* add_full = tmem_empty + 2;       // [NCBUF] addend tile landed in the staging buffer
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(add_full + NCBUF);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int num_tiles = p.tiles_m * p.tiles_n;
    const int kchunks = p.Kc / BK;
    const int kiters = p.taps * kchunks;
    const uint32_t a_bytes = (uint32_t)p.rows_per_tile * (BK * 2);

    if (threadIdx.x == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmC) : "memory");
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(&tmem_full[a], 1); mbar_init(&tmem_empty[a], 4); }
        for (int a = 0; a < NCBUF; ++a) mbar_init(&add_full[a], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(2 * BN));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0 && lane == 0) {
        // ---------------- TMA producer ----------------
        uint32_t it = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            const TileCoord t = tile_coord(p, tile);
            const int co0 = t.nt * BN;
            for (int kk = 0; kk < kiters; ++kk, ++it) {
                const int s = it % STAGES;
                const uint32_t ph = (it / STAGES) & 1;
                mbar_wait(&empty[s], ph ^ 1);
                uint8_t* a = smem + s * S::STAGE_BYTES;
                uint8_t* b = a + S::A_BYTES;
                const int tap = kk / kchunks, c0 = (kk % kchunks) * BK;
                mbar_expect_tx(&full[s], a_bytes + S::B_BYTES);
                if (p.mode4d) {
                    const int dh = tap / p.kw - p.pad, dw = tap % p.kw - p.pad;
                    tma_load_4d(a, &tmA, &full[s], c0, t.w0 + dw, t.h0 + dh, t.n0);
                } else {
                    tma_load_2d(a, &tmA, &full[s], c0, t.m0);
                }
                const int kb = tap * p.Kc + c0;
                if (B_MN) {
#pragma unroll
                    for (int h = 0; h < BN / 64; ++h) tma_load_2d(b + h * 8192, &tmB, &full[s], co0 + h * 64, kb);
                } else {
                    tma_load_2d(b, &tmB, &full[s], kb, co0);
                }
            }
        }
    } else if (warp == 1 && lane == 0) {
        // ---------------- MMA issuer ----------------
        constexpr uint32_t idesc = make_idesc(BN, B_MN);
        uint32_t it = 0, tcount = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tcount) {
            const uint32_t as = tcount & 1, aph = (tcount >> 1) & 1;
            mbar_wait(&tmem_empty[as], aph ^ 1);              // epilogue has drained this accumulator
            tcgen05_fence_after();
            const uint32_t tacc = tmem_base + as * BN;
            for (int kk = 0; kk < kiters; ++kk, ++it) {
                const int s = it % STAGES;
                const uint32_t ph = (it / STAGES) & 1;
                mbar_wait(&full[s], ph);
                tcgen05_fence_after();
                const uint32_t a = smem_u32(smem + s * S::STAGE_BYTES);
                const uint32_t b = a + S::A_BYTES;
                static_assert(BK / UMMA_K == 4, "umma_x4 issues four K-steps");
                // K-major operands advance 32 bytes per K-step inside the 128-byte swizzle row, MN-major ones 16 k-rows
                // of 128 bytes (descriptor start-address units of 16 bytes)
                const uint64_t da = make_desc(a, 16, 1024);
                const uint64_t db = B_MN ? make_desc(b, 8192, 1024) : make_desc(b, 16, 1024);
                umma_x4<2, (B_MN ? 128 : 2)>(tacc, da, db, idesc, kk > 0 ? 1u : 0u);
                umma_commit(&empty[s]);
            }
            umma_commit(&tmem_full[as]);
        }
    } else if (warp >= 2) {
        // ---------------- epilogue: TMEM -> regs -> swizzled smem -> TMA store ----------------
        const int q = warp & 3;                        // TMEM lane quarter
        const int r = q * 32 + lane;                   // row inside the tile
        const bool leader = (warp == 2 && lane == 0);
        const bool has_add = p.addend != nullptr && !p.mode4d;
        // The addend tile (e.g. the skip connection's gradient) is TMA-prefetched into the staging buffer one tile
        // ahead; the epilogue adds the accumulator onto it in place and TMA-stores the sum.
        auto prefetch_addend = [&](int tile_, uint32_t buf) {
            const TileCoord tt = tile_coord(p, tile_);
            uint8_t* dst = smem + S::C_OFF + buf * S::C_BYTES;
            mbar_expect_tx(&add_full[buf], S::C_BYTES);
#pragma unroll
            for (int h = 0; h < BN / 64; ++h) tma_load_2d(dst + h * (BM * 128), &tmD, &add_full[buf], tt.nt * BN + h * 64, tt.m0);
        };
        if (has_add && leader) {               // two tiles ahead: the HBM latency of an addend tile exceeds one epilogue
            if ((int)blockIdx.x < num_tiles) prefetch_addend(blockIdx.x, 0);
            if ((int)(blockIdx.x + gridDim.x) < num_tiles) prefetch_addend(blockIdx.x + gridDim.x, 1);
        }
        uint32_t tcount = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tcount) {
            const TileCoord t = tile_coord(p, tile);
            const int co0 = t.nt * BN;
            const uint32_t as = tcount & 1, aph = (tcount >> 1) & 1;
            const uint32_t cbuf = tcount % NCBUF, cph = (tcount / NCBUF) & 1;
            uint8_t* cst = smem + S::C_OFF + cbuf * S::C_BYTES;
            mbar_wait(&tmem_full[as], aph);
            tcgen05_fence_after();
            if (has_add) {
                mbar_wait(&add_full[cbuf], cph);              // this tile's addend is in cst (its prefetch also guarded reuse)
            } else {
                if (leader) asm volatile("cp.async.bulk.wait_group.read 2;" ::: "memory");     // staging buffer free again
                epi_bar_sync();
            }
            // ReLU bit mask of this thread's output row (BN/8 bytes), requested before the accumulator wait
            uint32_t mbits[BN / 32];
            const bool has_mask = p.mask != nullptr && !p.mode4d;
            if (has_mask) {
                const bool in = t.m0 + r < p.M;
                const unsigned char* mp = p.mask + (size_t)(in ? t.m0 + r : 0) * (p.N >> 3) + (co0 >> 3);
                if constexpr (BN == 128) {
                    const uint4 mv = *reinterpret_cast<const uint4*>(mp);
                    mbits[0] = mv.x; mbits[1] = mv.y; mbits[2] = mv.z; mbits[3] = mv.w;
                } else {
                    const uint2 mv = *reinterpret_cast<const uint2*>(mp);
                    mbits[0] = mv.x; mbits[1] = mv.y;
                }
            }
            // all BN/32 accumulator chunks of this thread's row are requested back to back, one wait
            uint32_t vall[BN / 32][32];
#pragma unroll
            for (int cc = 0; cc < BN / 32; ++cc)
                tmem_ld32_issue(tmem_base + ((uint32_t)(q * 32) << 16) + as * BN + (uint32_t)(cc * 32), vall[cc]);
            tmem_ld_wait();
            tcgen05_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tmem_empty[as]);                   // the row is in registers: accumulator may be overwritten
#pragma unroll
            for (int cc = 0; cc < BN / 32; ++cc) {
                const int c = cc * 32;
                uint32_t* v = vall[cc];
                use_after_wait(v);
                if (has_add) {
                    const uint8_t* asub = cst + (c / 64) * (BM * 128) + r * 128;
                    const int achunk0 = (c % 64) / 8;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const uint4 a4 = *reinterpret_cast<const uint4*>(asub + (((achunk0 + j) ^ (r & 7)) * 16));
                        const uint32_t aw[4] = {a4.x, a4.y, a4.z, a4.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            v[j * 8 + 2 * e] = __float_as_uint(__uint_as_float(v[j * 8 + 2 * e]) + __uint_as_float(aw[e] << 16));
                            v[j * 8 + 2 * e + 1] = __float_as_uint(__uint_as_float(v[j * 8 + 2 * e + 1]) + __uint_as_float(aw[e] & 0xffff0000u));
                        }
                    }
                }
                if (has_mask) {                                           // bit j of mbits[cc] <-> column c + j
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (!((mbits[cc] >> j) & 1u)) v[j] = 0u;
                }
                uint8_t* sub = cst + (c / 64) * (BM * 128) + r * 128;     // 128x64 sub-tile, 128-byte rows
                const int chunk0 = (c % 64) / 8;
#pragma unroll
                for (int j = 0; j < 4; ++j) {                             // four 16-byte chunks (8 bf16 each)
                    uint32_t w[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        float x0 = __uint_as_float(v[j * 8 + 2 * e]), x1 = __uint_as_float(v[j * 8 + 2 * e + 1]);
                        if (p.bias) { x0 += p.bias[co0 + c + j * 8 + 2 * e]; x1 += p.bias[co0 + c + j * 8 + 2 * e + 1]; }
                        if (p.relu) { x0 = fmaxf(x0, 0.0f); x1 = fmaxf(x1, 0.0f); }
                        __nv_bfloat162 h = __floats2bfloat162_rn(x0, x1);
                        w[e] = *reinterpret_cast<uint32_t*>(&h);
                    }
                    const int chunk = (chunk0 + j) ^ (r & 7);             // 128B swizzle: chunk index XOR row-in-group
                    *reinterpret_cast<uint4*>(sub + chunk * 16) = make_uint4(w[0], w[1], w[2], w[3]);
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to TMA
            epi_bar_sync();
            if (p.stats) {
                // per-channel sum / sum of squares of the rounded outputs of this tile (valid rows only)
                const int col = threadIdx.x - 64;                          // 0..127
                if (col < BN && co0 + col < p.N) {
                    int valid = p.rows_per_tile;
                    if (!p.mode4d) valid = min(BM, p.M - t.m0);
                    const uint8_t* sub = cst + (col / 64) * (BM * 128);
                    const int cc = col % 64, chunk = cc / 8, within = (cc % 8) * 2;
                    float s1 = 0.0f, s2 = 0.0f;
                    if (!p.mode4d) {
                        // eight independent loads in flight per thread: the loop is latency-, not bandwidth-bound
                        float a1[8] = {0, 0, 0, 0, 0, 0, 0, 0}, a2[8] = {0, 0, 0, 0, 0, 0, 0, 0};
                        int rr = 0;
                        for (; rr + 8 <= valid; rr += 8) {
#pragma unroll
                            for (int u = 0; u < 8; ++u) {      // (rr + u) & 7 == u because rr is a multiple of 8
                                const __nv_bfloat16 hv = *reinterpret_cast<const __nv_bfloat16*>(sub + (rr + u) * 128 + ((chunk ^ u) * 16) + within);
                                const float x = __bfloat162float(hv);
                                a1[u] += x;
                                a2[u] = fmaf(x, x, a2[u]);
                            }
                        }
                        for (; rr < valid; ++rr) {
                            const __nv_bfloat16 hv = *reinterpret_cast<const __nv_bfloat16*>(sub + rr * 128 + ((chunk ^ (rr & 7)) * 16) + within);
                            const float x = __bfloat162float(hv);
                            a1[0] += x;
                            a2[0] = fmaf(x, x, a2[0]);
                        }
                        s1 = ((a1[0] + a1[1]) + (a1[2] + a1[3])) + ((a1[4] + a1[5]) + (a1[6] + a1[7]));
                        s2 = ((a2[0] + a2[1]) + (a2[2] + a2[3])) + ((a2[4] + a2[5]) + (a2[6] + a2[7]));
                    } else {
                        for (int rr = 0; rr < valid; ++rr) {
                            // rows of the spatial patch that fall outside the image are not outputs
                            const int w = t.w0 + rr % p.Wb, h = t.h0 + (rr / p.Wb) % p.Hb, n = t.n0 + rr / (p.Wb * p.Hb);
                            if (w < p.W && h < p.H && n < p.Nimg) {
                                const __nv_bfloat16 hv = *reinterpret_cast<const __nv_bfloat16*>(sub + rr * 128 + ((chunk ^ (rr & 7)) * 16) + within);
                                const float x = __bfloat162float(hv);
                                s1 += x;
                                s2 = fmaf(x, x, s2);
                            }
                        }
                    }
                    float* o = p.stats + (size_t)t.mt * 2 * p.N + co0 + col;
                    o[0] = s1;
                    o[p.N] = s2;
                }
            }
            if (leader) {
#pragma unroll
                for (int h = 0; h < BN / 64; ++h) {
                    if (co0 + h * 64 >= p.N) break;
                    if (p.mode4d) tma_store_4d(&tmC, cst + h * (BM * 128), co0 + h * 64, t.w0, t.h0, t.n0);
                    else tma_store_2d(&tmC, cst + h * (BM * 128), co0 + h * 64, t.m0);
                }
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                if (has_add && tile + 2 * (int)gridDim.x < num_tiles) {
                    // the other buffer was last read by the store issued one tile ago: wait for it, then refill it
                    asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
                    prefetch_addend(tile + 2 * gridDim.x, (tcount + 2) % NCBUF);
                }
            }
        }
        if (leader) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(2 * BN));
    }
}

}  // namespace conv
}  // namespace lwc

using namespace lwc::conv;

static void ensure_context() {
    CUcontext ctx = nullptr;
    if (cuCtxGetCurrent(&ctx) != CUDA_SUCCESS || !ctx) {
        int dev = 0;
        cudaGetDevice(&dev);
        CUcontext prim = nullptr;
        if (cuDevicePrimaryCtxRetain(&prim, dev) == CUDA_SUCCESS) cuCtxSetCurrent(prim);
    }
}

static int encode(CUtensorMap* map, const void* ptr, int rank, const cuuint64_t* gdim, const cuuint64_t* gstride_bytes,
                  const cuuint32_t* box) {
    cuuint32_t estr[4] = {1, 1, 1, 1};
    CUresult r = cuTensorMapEncodeTiled(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(ptr), gdim, gstride_bytes,
                                        box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS && getenv("LWC_DEBUG")) fprintf(stderr, "lwc_conv: cuTensorMapEncodeTiled rank %d -> %d\n", rank, (int)r);
    return r == CUDA_SUCCESS ? 0 : LWC_EINVAL;
}

template <int BN, bool B_MN>
static int launch(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc, const CUtensorMap& td, const Params& p,
                  cudaStream_t st) {
    using S = Smem<BN>;
    static bool attr_set = false;
    if (!attr_set) {
        LWC_CUDA_CHECK(cudaFuncSetAttribute(k_conv<BN, B_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize, S::TOTAL));
        attr_set = true;
    }
    int grid = p.tiles_m * p.tiles_n;
    if (grid > 148) grid = 148;
    k_conv<BN, B_MN><<<grid, THREADS, S::TOTAL, st>>>(ta, tb, tc, td, p);
    return 0;
}

// Number of M tiles of a problem (the caller sizes the statistics buffer [tiles_m][2][N] with it).
extern "C" int lwc_conv_tiles_m(int Nimg, int H, int W, int kh, int* rows_per_tile_out) {
    if (kh == 1) {
        const long long M = (long long)Nimg * H * W;
        if (rows_per_tile_out) *rows_per_tile_out = BM;
        return (int)((M + BM - 1) / BM);
    }
    // spatial patch Wb x Hb x Nb with Wb*Hb*Nb <= 128
    int Wb = W < 128 ? W : 128;
    if (W > 16 && W % 8 == 0 && W > 32) Wb = 8;            // wide images: 8 x 16 patches
    int Hb = 128 / Wb;
    if (Hb > H) Hb = H;
    int Nb = 128 / (Wb * Hb);
    if (Nb < 1) Nb = 1;
    if (Nb > Nimg) Nb = Nimg;
    if (rows_per_tile_out) *rows_per_tile_out = Wb * Hb * Nb;
    const int tw = (W + Wb - 1) / Wb, th = (H + Hb - 1) / Hb, tn = (Nimg + Nb - 1) / Nb;
    return tw * th * tn;
}

// Forward-style pass:  Y[N,H,W,Cout] = conv(X[N,H,W,Cin], Wt) (+bias)(ReLU), stride 1, "same" padding (kh = kw in {1,3}).
//   b_mn = 0 : Wt is [Cout][kh][kw][Cin] (channels-last weight), the forward pass
//   b_mn = 1 : 1x1 only; Wt is stored [Cin_of_this_pass][Cout_of_this_pass] (i.e. the forward weight [Cout_f][Cin_f]
//              used un-transposed for the data gradient dX = dY * W)
// stats (optional): [tiles_m][2][Cout] fp32 per-tile sums / sums of squares of the bf16 outputs.
static int conv_forward_impl(const void* x, const void* wt, void* y, int Nimg, int H, int W, int Cin, int Cout, int kh,
                             int b_mn, const float* bias, int relu, float* stats, const void* addend, const void* mask,
                             void* stream);

extern "C" int lwc_conv_forward(const void* x, const void* wt, void* y, int Nimg, int H, int W, int Cin, int Cout, int kh,
                                int b_mn, const float* bias, int relu, float* stats, void* stream) {
    return conv_forward_impl(x, wt, y, Nimg, H, W, Cin, Cout, kh, b_mn, bias, relu, stats, nullptr, nullptr, stream);
}

// 1x1 only: Y = conv(X) + addend, addend bf16 with the shape of Y (used for dX = dY * W + d_identity at residual forks).
extern "C" int lwc_conv1x1_add(const void* x, const void* wt, const void* addend, void* y, int Nimg, int H, int W, int Cin,
                               int Cout, int b_mn, void* stream) {
    if (!addend) return LWC_EINVAL;
    return conv_forward_impl(x, wt, y, Nimg, H, W, Cin, Cout, 1, b_mn, nullptr, 0, nullptr, addend, nullptr, stream);
}

// Same, and outputs whose bit in relu_mask ([M][Cout/8] bytes, bit j of byte b <-> channel 8 b + j; the layout written by
// lwc_bn_forward_ex) is 0 are zeroed: Y = (conv(X) + addend) * mask.  The data gradient at a residual fork arrives at the
// producing block already multiplied with that block's ReLU mask, so its batch-norm backward neither reads a mask nor
// writes a separate residual gradient.  addend may be NULL.  Cout % 128 == 0.
extern "C" int lwc_conv1x1_add_mask(const void* x, const void* wt, const void* addend, const void* relu_mask, void* y, int Nimg,
                                    int H, int W, int Cin, int Cout, int b_mn, void* stream) {
    if (!relu_mask || (Cout % 128)) return LWC_EINVAL;
    return conv_forward_impl(x, wt, y, Nimg, H, W, Cin, Cout, 1, b_mn, nullptr, 0, nullptr, addend, relu_mask, stream);
}

static int conv_forward_impl(const void* x, const void* wt, void* y, int Nimg, int H, int W, int Cin, int Cout, int kh,
                             int b_mn, const float* bias, int relu, float* stats, const void* addend, const void* mask,
                             void* stream) {
    if (!x || !wt || !y || Nimg <= 0 || H <= 0 || W <= 0) return LWC_EINVAL;
    if ((Cin % 64) || (Cout % 64) || (kh != 1 && kh != 3) || (b_mn && kh != 1)) return LWC_EINVAL;
    ensure_context();
    cudaStream_t st = (cudaStream_t)stream;
    Params p;
    memset(&p, 0, sizeof(p));
    p.N = Cout; p.Kc = Cin; p.taps = kh * kh; p.kw = kh; p.pad = kh / 2; p.relu = relu; p.bias = bias; p.stats = stats; p.addend = (const __nv_bfloat16*)addend; p.mask = (const unsigned char*)mask;
    p.H = H; p.W = W; p.Nimg = Nimg;
    const int BN = (Cout % 128 == 0) ? 128 : 64;
    p.tiles_n = Cout / BN;
    CUtensorMap ta, tb, tc;
    int rc;
    if (kh == 1) {
        const long long M = (long long)Nimg * H * W;
        if (M > 0x7fffffffLL) return LWC_ETOOBIG;
        p.M = (int)M; p.mode4d = 0; p.rows_per_tile = BM; p.tiles_m = (int)((M + BM - 1) / BM);
        cuuint64_t gd[2] = {(cuuint64_t)Cin, (cuuint64_t)M}, gs[1] = {(cuuint64_t)Cin * 2};
        cuuint32_t bx[2] = {BK, BM};
        rc = encode(&ta, x, 2, gd, gs, bx);
        if (rc) return rc;
        cuuint64_t cd[2] = {(cuuint64_t)Cout, (cuuint64_t)M}, cs[1] = {(cuuint64_t)Cout * 2};
        cuuint32_t cb[2] = {64, BM};
        rc = encode(&tc, y, 2, cd, cs, cb);
        if (rc) return rc;
    } else {
        int rows = 0;
        p.tiles_m = lwc_conv_tiles_m(Nimg, H, W, kh, &rows);
        p.rows_per_tile = rows;
        p.mode4d = 1;
        int Wb = W < 128 ? W : 128;
        if (W > 16 && W % 8 == 0 && W > 32) Wb = 8;
        int Hb = 128 / Wb; if (Hb > H) Hb = H;
        int Nb = 128 / (Wb * Hb); if (Nb < 1) Nb = 1; if (Nb > Nimg) Nb = Nimg;
        p.Wb = Wb; p.Hb = Hb; p.Nb = Nb;
        p.tiles_w = (W + Wb - 1) / Wb; p.tiles_h = (H + Hb - 1) / Hb;
        cuuint64_t gd[4] = {(cuuint64_t)Cin, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)Nimg};
        cuuint64_t gs[3] = {(cuuint64_t)Cin * 2, (cuuint64_t)W * Cin * 2, (cuuint64_t)H * W * Cin * 2};
        cuuint32_t bx[4] = {BK, (cuuint32_t)Wb, (cuuint32_t)Hb, (cuuint32_t)Nb};
        rc = encode(&ta, x, 4, gd, gs, bx);
        if (rc) return rc;
        cuuint64_t cd[4] = {(cuuint64_t)Cout, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)Nimg};
        cuuint64_t cs[3] = {(cuuint64_t)Cout * 2, (cuuint64_t)W * Cout * 2, (cuuint64_t)H * W * Cout * 2};
        cuuint32_t cb[4] = {64, (cuuint32_t)Wb, (cuuint32_t)Hb, (cuuint32_t)Nb};
        rc = encode(&tc, y, 4, cd, cs, cb);
        if (rc) return rc;
    }
    if (b_mn) {
        // stored [K = Cin][N = Cout]
        cuuint64_t bd[2] = {(cuuint64_t)Cout, (cuuint64_t)Cin}, bs[1] = {(cuuint64_t)Cout * 2};
        cuuint32_t bb[2] = {64, BK};
        rc = encode(&tb, wt, 2, bd, bs, bb);
    } else {
        const cuuint64_t Kt = (cuuint64_t)p.taps * Cin;
        cuuint64_t bd[2] = {Kt, (cuuint64_t)Cout}, bs[1] = {Kt * 2};
        cuuint32_t bb[2] = {BK, (cuuint32_t)BN};
        rc = encode(&tb, wt, 2, bd, bs, bb);
    }
    if (rc) return rc;
    CUtensorMap td = tc;                 // addend: same geometry as the output, different base pointer
    if (addend) {
        if (kh != 1) return LWC_EINVAL;
        cuuint64_t dd[2] = {(cuuint64_t)Cout, (cuuint64_t)p.M}, ds[1] = {(cuuint64_t)Cout * 2};
        cuuint32_t db[2] = {64, BM};
        rc = encode(&td, addend, 2, dd, ds, db);
        if (rc) return rc;
    }
    if (BN == 128) rc = b_mn ? launch<128, true>(ta, tb, tc, td, p, st) : launch<128, false>(ta, tb, tc, td, p, st);
    else rc = b_mn ? launch<64, true>(ta, tb, tc, td, p, st) : launch<64, false>(ta, tb, tc, td, p, st);
    if (rc) return rc;
    return (int)cudaGetLastError();
}
