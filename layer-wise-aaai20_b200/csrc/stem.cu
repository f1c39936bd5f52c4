// Weight gradient of the ResNet stem (7x7 / stride 2 / pad 3, 3 -> 64 channels) on tcgen05 / TMEM / TMA.
//
// The stem runs forward as a dense 4x4 / stride-1 convolution over the 2x2 space-to-depth input
//   xs[n][r][w][16]   (r, w in [0, OH+3) x [0, OW+3); 12 used channels (dh, dw, c); zero border)
// (ops/input.py).  Its weight gradient in that form is
//   dWs[bh][co][(bw, ch)] = sum_{n, oh, ow} dy[n][oh][ow][co] * xs[n][oh + bh][ow + bw][ch]
// i.e. for every INPUT row (n, r) one GEMM step   D_bh[64 co x 64 (bw,ch)] += dy[n][r - bh]^T [64 x OW] * X_r [OW x 64]
// where X_r[ow][(bw, ch)] = xs[n][r][ow + bw][ch] is a view of one xs row with OVERLAPPING windows: window ow
// starts 32 bytes after window ow - 1 and is 128 bytes long.  One 3-D TMA box load (inner 64 elements, 32-byte
// stride between windows) materialises it in shared memory as a standard 128B-swizzled MN-major operand tile --
// no im2col buffer, the input row is read from HBM once and feeds all four kernel rows bh.
//
// dy rows are staged in PAIRS (rows 2j, 2j+1 of an image -> one MN-major A tile with M = 128 = 2 x 64 channels),
// so each tcgen05.mma is the full 128 x 64 x 16 shape.  Which kernel row a half of an accumulator belongs to
// depends on the parity of r:
//     r = 2j+1:  pair j   -> acc A (lo: bh 1, hi: bh 0)     pair j-1 -> acc B (lo: bh 3, hi: bh 2)
//     r = 2j  :  pair j   -> acc D (lo: bh 0, hi: unused)   pair j-1 -> acc C (lo: bh 2, hi: bh 1)
//                pair j-2 -> acc E (lo: unused, hi: bh 3)
// Five fp32 accumulators of 64 columns live in tensor memory for the whole kernel; the CTA (one per SM, each
// owning a contiguous range of input rows) dumps them once at the end and a small second kernel sums the
// per-CTA partials in a fixed order and scatters them into the 7x7 layout.  dy pairs stay in a 5-slot ring
// (each is used by five consecutive input rows and requested two rows early), xs rows go through a ring of up
// to 8 stages: 100+ KB of loads in flight per SM.
//
// Replaces cuDNN's legacy (sm80 xmma) wgrad of the plain 3-channel formulation -- 0.65 ms + a 3-channel copy of
// the batch (0.11 ms) + cuDNN's channel padding kernels (0.10 ms) per ResNet-50 step -- by a kernel that streams
// dy (360 MB) and xs (95 MB) once.  Reference op: the `conv1` weight gradient of IMAGENET/training/resnet.py:100.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace lwc {
namespace stem {

constexpr int THREADS = 192;          // warp 0: TMA producer, warp 1: MMA issuer + TMEM owner, warps 2-5: epilogue
constexpr int PAIR_SLOTS = 5;         // three live pairs + two of slack: the producer may run four rows ahead of the
                                      // MMA completions that release slots (tcgen05.commit round trips are ~1-3 us)
constexpr int MAX_B_STAGES = 8;      // xs row tiles in flight (as many as fit next to the pair ring)
constexpr int PAIR_LOOKAHEAD = 2;     // a dy pair is requested this many input rows before its first use
constexpr int NACC = 5;
constexpr int TMEM_COLS = 512;        // 5 x 64 columns used

struct Params {
    int N, OH, OW, HB;                // images, dy rows / cols per image, xs rows per image (OH + 3)
    int rows_total;                   // N * HB
    int b_stages;
    int debug;                        // LWC_STEM_DEBUG: 1 = issue no MMAs, 2 = load nothing (timing experiments)
    float* partial;                   // [grid][NACC][128][64]
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
__device__ __forceinline__ void tma_load_3d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1, int c2) {
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
        ::"r"(smem_u32(dst)), "l"(map), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ void tcgen05_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tcgen05_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// shared-memory matrix descriptor, 128B swizzle: [0,14) start>>4 | [16,30) LBO>>4 | [32,46) SBO>>4 | version 1 | SW128
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3fffu);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16;
    d |= (uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32;
    d |= 1ULL << 46;
    d |= 2ULL << 61;
    return d;
}
// kind::f16 instruction descriptor: D = f32, A = B = bf16, both operands MN-major, N = 64, M = 128
constexpr uint32_t kIdesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) | ((64u >> 3) << 17) | ((128u >> 4) << 24);
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(da), "l"(db), "r"(kIdesc), "r"(accumulate) : "memory");
}
// All MMAs of one input row in ONE asm block: up to three (A tile, accumulator) uses x ksteps K-steps, descriptors
// advanced by 2048 bytes (16 k-rows) per step inside the block.  Issued one by one from C++, every tcgen05.mma costs
// ~120 cycles of operand shuffling into uniform registers (R2UR) against 32 cycles of tensor-pipe time; here the
// operands enter uniform registers once per row.
__device__ __forceinline__ void umma_row(uint32_t t0, uint32_t t1, uint32_t t2, uint64_t da0, uint64_t da1, uint64_t da2,
                                         uint64_t db, uint32_t use_mask, uint32_t acc_mask, uint32_t ksteps) {
    asm volatile(
        "{\n\t"
        ".reg .pred u0, u1, u2, a0, a1, a2, lp;\n\t"
        ".reg .b64 x0, x1, x2, xb;\n\t"
        ".reg .b32 i, t;\n\t"
        "and.b32 t, %7, 1;\n\t setp.ne.b32 u0, t, 0;\n\t"
        "and.b32 t, %7, 2;\n\t setp.ne.b32 u1, t, 0;\n\t"
        "and.b32 t, %7, 4;\n\t setp.ne.b32 u2, t, 0;\n\t"
        "and.b32 t, %8, 1;\n\t setp.ne.b32 a0, t, 0;\n\t"
        "and.b32 t, %8, 2;\n\t setp.ne.b32 a1, t, 0;\n\t"
        "and.b32 t, %8, 4;\n\t setp.ne.b32 a2, t, 0;\n\t"
        "mov.b64 x0, %3;\n\t mov.b64 x1, %4;\n\t mov.b64 x2, %5;\n\t mov.b64 xb, %6;\n\t mov.u32 i, 0;\n\t"
        "ROW_LOOP:\n\t"
        "@u0 tcgen05.mma.cta_group::1.kind::f16 [%0], x0, xb, %10, a0;\n\t"
        "@u1 tcgen05.mma.cta_group::1.kind::f16 [%1], x1, xb, %10, a1;\n\t"
        "@u2 tcgen05.mma.cta_group::1.kind::f16 [%2], x2, xb, %10, a2;\n\t"
        "add.u64 x0, x0, 128;\n\t add.u64 x1, x1, 128;\n\t add.u64 x2, x2, 128;\n\t add.u64 xb, xb, 128;\n\t"
        "setp.eq.u32 a0, i, i;\n\t setp.eq.u32 a1, i, i;\n\t setp.eq.u32 a2, i, i;\n\t"
        "add.u32 i, i, 1;\n\t"
        "setp.lt.u32 lp, i, %9;\n\t"
        "@lp bra ROW_LOOP;\n\t"
        "}"
        ::"r"(t0), "r"(t1), "r"(t2), "l"(da0), "l"(da1), "l"(da2), "l"(db), "r"(use_mask), "r"(acc_mask), "r"(ksteps), "r"(kIdesc)
        : "memory");
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

// accumulator ids
enum { ACC_A = 0, ACC_B = 1, ACC_C = 2, ACC_D = 3, ACC_E = 4 };

// The (pair, accumulator) uses of input row r: pair index j' in [0, J) with J = OH / 2.
struct RowUses {
    int n;            // number of uses (0..3)
    int pair[3];
    int acc[3];
};
__device__ __forceinline__ RowUses row_uses(int r, int J) {
    RowUses u;
    u.n = 0;
    const int j = r >> 1;
    if (r & 1) {
        if (j < J) { u.pair[u.n] = j; u.acc[u.n] = ACC_A; ++u.n; }
        if (j - 1 >= 0 && j - 1 < J) { u.pair[u.n] = j - 1; u.acc[u.n] = ACC_B; ++u.n; }
    } else {
        if (j < J) { u.pair[u.n] = j; u.acc[u.n] = ACC_D; ++u.n; }
        if (j - 1 >= 0 && j - 1 < J) { u.pair[u.n] = j - 1; u.acc[u.n] = ACC_C; ++u.n; }
        if (j - 2 >= 0 && j - 2 < J) { u.pair[u.n] = j - 2; u.acc[u.n] = ACC_E; ++u.n; }
    }
    return u;
}

// Pairs are loaded in increasing j within an image, consecutively; q = running index of a pair in this CTA's
// load sequence (ring slot q % 4, barrier phase (q / 4) & 1).  q0 = q of pair 0 of the current image (may be
// negative in the CTA's first, partially covered image).  Producer and MMA issuer run the same bookkeeping.
struct PairSeq {
    int q0, q_next;
    // first row of the CTA: returns the pairs to preload in [lo, hi) (possibly empty)
    __device__ __forceinline__ void start(int r0, int J, int& lo, int& hi) {
        const int j0 = r0 >> 1;
        lo = (r0 & 1) ? j0 - 1 : j0 - 2;
        hi = (r0 & 1) ? j0 + 1 : j0;              // odd first row also needs its own pair j0; an even one loads it itself
        if (lo < 0) lo = 0;
        if (hi > J) hi = J;
        if (hi < lo) hi = lo;
        q0 = -lo;
        q_next = hi - lo;
    }
    __device__ __forceinline__ void new_image() { q0 = q_next; }
    // row r introduces pair r/2 when r is even and the pair exists
    __device__ __forceinline__ bool introduces(int r, int J) const { return !(r & 1) && (r >> 1) < J; }
    __device__ __forceinline__ int q_of(int j) const { return q0 + j; }
};

__global__ void __launch_bounds__(THREADS, 1) k_stem_wgrad(const __grid_constant__ CUtensorMap tmA,
                                                           const __grid_constant__ CUtensorMap tmB, Params p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    const int slot_bytes = p.OW * 128;                       // one dy row [OW pixels][64 co] or one xs window tile
    const int pair_bytes = 2 * slot_bytes;
    uint8_t* pairs = smem;
    uint8_t* bst = smem + PAIR_SLOTS * pair_bytes;
    const int NB = p.b_stages;
    uint64_t* pair_full = reinterpret_cast<uint64_t*>(bst + NB * slot_bytes);
    uint64_t* pair_empty = pair_full + PAIR_SLOTS;
    uint64_t* b_full = pair_empty + PAIR_SLOTS;
    uint64_t* b_empty = b_full + MAX_B_STAGES;
    uint64_t* done = b_empty + MAX_B_STAGES;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int J = p.OH >> 1;
    const int R0 = (int)((long long)p.rows_total * blockIdx.x / gridDim.x);
    const int R1 = (int)((long long)p.rows_total * (blockIdx.x + 1) / gridDim.x);
    const int ksteps = p.OW >> 4;

    if (threadIdx.x == 0) {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmA) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&tmB) : "memory");
        for (int s = 0; s < PAIR_SLOTS; ++s) { mbar_init(&pair_full[s], 1); mbar_init(&pair_empty[s], 1); }
        for (int s = 0; s < NB; ++s) { mbar_init(&b_full[s], 1); mbar_init(&b_empty[s], 1); }
        mbar_init(done, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "n"(TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    tcgen05_fence_before();
    __syncthreads();
    tcgen05_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    unsigned long long* stamps = (p.debug & 16) && blockIdx.x == 0
        ? reinterpret_cast<unsigned long long*>(p.partial + (size_t)gridDim.x * NACC * 128 * 64) : nullptr;
    if (stamps && threadIdx.x == 0) stamps[0] = global_timer_ns();
    if (warp == 0 && lane == 0) {
        // ---------------- TMA producer ----------------
        PairSeq seq;
        int n = R0 / p.HB, r = R0 - n * p.HB;
        int q = 0;
        auto load_pair = [&](int img, int j) {
            const int s = q % PAIR_SLOTS;
            const uint32_t ph = (q / PAIR_SLOTS) & 1;
            mbar_wait(&pair_empty[s], ph ^ 1);
            uint8_t* dst = pairs + s * pair_bytes;
            const int row0 = (img * p.OH + 2 * j) * p.OW;             // pixel index of dy row 2j
            if (p.debug & 2) {
                asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&pair_full[s])) : "memory");
            } else {
                mbar_expect_tx(&pair_full[s], pair_bytes);
                tma_load_2d(dst, &tmA, &pair_full[s], 0, row0);
                tma_load_2d(dst + slot_bytes, &tmA, &pair_full[s], 0, row0 + p.OW);
            }
            ++q;
        };
        if (R0 < R1) {
            int lo, hi;
            seq.start(r, J, lo, hi);
            for (int j = lo; j < hi; ++j) load_pair(n, j);
        }
        // (cn, cr, cR): the next row whose newly introduced pair has not been requested yet.  Pairs are requested in
        // the order the issuer expects them, PAIR_LOOKAHEAD rows early: the slot a pair goes into was released
        // four rows before its first use, so waiting for it can never wait for an xs row not yet requested.
        int cn = n, cr = r, cR = R0;
        for (int R = R0, i = 0; R < R1; ++R, ++i) {
            const int ahead = R + PAIR_LOOKAHEAD < R1 ? R + PAIR_LOOKAHEAD : R1 - 1;
            for (; cR <= ahead; ++cR) {
                if (seq.introduces(cr, J)) load_pair(cn, cr >> 1);
                if (++cr == p.HB) { cr = 0; ++cn; }
            }
            const int s = i % NB;
            const uint32_t ph = (i / NB) & 1;
            mbar_wait(&b_empty[s], ph ^ 1);
            if (p.debug & 2) {
                asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(&b_full[s])) : "memory");
            } else {
                mbar_expect_tx(&b_full[s], slot_bytes);
                tma_load_3d(bst + s * slot_bytes, &tmB, &b_full[s], 0, 0, R);
            }
        }
        if (stamps) stamps[1] = global_timer_ns();
    } else if (warp == 1 && lane == 0) {
        // ---------------- MMA issuer ----------------
        // Measured (LWC_STEM_DEBUG=16 stamps): the loop is issue-bound, ~120 cycles per 32-cycle 128x64x16 MMA (every
        // operand moves through R2UR into uniform registers); running the whole warp with a lane-0 predicate on the
        // issue was slower still.  Larger MMAs (a dy-pair-major formulation with N = 128) are the next step.
        const bool leader = true;
        PairSeq seq;
        int n = R0 / p.HB, r = R0 - n * p.HB;
        if (R0 < R1) {
            int lo, hi;
            seq.start(r, J, lo, hi);
        }
        uint32_t started = 0;
        long long tw_b = 0, tw_p = 0, t_iss = 0, t_com = 0;
        for (int R = R0, i = 0; R < R1; ++R, ++i) {
            const long long c0 = stamps ? clock64() : 0;
            if (R != R0 && r == 0) seq.new_image();
            if (seq.introduces(r, J)) ++seq.q_next;
            // the (at most three) uses of this row, kept in scalar registers: use k reads pair j - k
            const int j = r >> 1;
            const bool odd = r & 1;
            const int s = i % NB;
            mbar_wait(&b_full[s], (i / NB) & 1);
            const long long c1 = stamps ? clock64() : 0;
            // descriptor of the A tile of pair pj (0 = not used by this row); everything stays in scalar registers
            auto prep = [&](int pj, bool wanted) -> uint64_t {
                if (!wanted || pj < 0 || pj >= J) return 0ull;
                const int q = seq.q_of(pj);
                const int ps = q % PAIR_SLOTS;
                mbar_wait(&pair_full[ps], (q / PAIR_SLOTS) & 1);
                // MN-major A: the two 64-channel halves (the two dy rows of the pair) are slot_bytes apart,
                // 8-row (pixel) groups 1024 bytes apart
                return make_desc(smem_u32(pairs + ps * pair_bytes), slot_bytes, 1024);
            };
            const uint64_t da0 = prep(j, true), da1 = prep(j - 1, true), da2 = prep(j - 2, !odd);
            const uint32_t acc0 = odd ? ACC_A : ACC_D, acc1 = odd ? ACC_B : ACC_C, acc2 = ACC_E;
            const long long c2 = stamps ? clock64() : 0;
            tcgen05_fence_after();
            const uint64_t db = make_desc(smem_u32(bst + s * slot_bytes), 8192, 1024);
            const uint32_t use_mask = (da0 ? 1u : 0u) | (da1 ? 2u : 0u) | (da2 ? 4u : 0u);
            const uint32_t acc_mask = ((started >> acc0) & 1u) | (((started >> acc1) & 1u) << 1) | (((started >> acc2) & 1u) << 2);
            if (leader && !(p.debug & 1))
                umma_row(tmem_base + acc0 * 64, tmem_base + acc1 * 64, tmem_base + acc2 * 64, da0, da1, da2, db, use_mask, acc_mask,
                         (uint32_t)ksteps);
            started |= (da0 ? 1u << acc0 : 0u) | (da1 ? 1u << acc1 : 0u) | (da2 ? 1u << acc2 : 0u);
            const long long c3 = stamps ? clock64() : 0;
            if (leader) umma_commit(&b_empty[s]);
            // pair j-2 of an even row is used for the last time here
            if (!(r & 1)) {
                const int jl = (r >> 1) - 2;
                if (leader && jl >= 0 && jl < J) umma_commit(&pair_empty[seq.q_of(jl) % PAIR_SLOTS]);
            }
            if (++r == p.HB) { r = 0; ++n; }
            const long long c4 = stamps ? clock64() : 0;
            tw_b += c1 - c0; tw_p += c2 - c1; t_iss += c3 - c2; t_com += c4 - c3;
        }
        if (leader) umma_commit(done);
        if (stamps && leader) {
            stamps[2] = global_timer_ns();
            stamps[5] = tw_b; stamps[6] = tw_p; stamps[7] = t_iss; stamps[8] = t_com;
        }
    } else if (warp >= 2) {
        // ---------------- epilogue: dump the five accumulators ----------------
        // which accumulators received anything (same bookkeeping as the issuer; an unused one holds garbage)
        uint32_t started = 0;
        {
            int n = R0 / p.HB, r = R0 - n * p.HB;
            for (int R = R0; R < R1 && started != 31u; ++R) {
                const RowUses u = row_uses(r, J);
                for (int k = 0; k < u.n; ++k) {
                    started |= 1u << u.acc[k];
                }
                if (++r == p.HB) r = 0;
            }
        }
        mbar_wait(done, 0);
        if (stamps && threadIdx.x == 64) stamps[3] = global_timer_ns();
        tcgen05_fence_after();
        const int qd = warp & 3;                      // TMEM lane quarter this warp may access
        const int row = qd * 32 + lane;
        float* out = p.partial + ((size_t)blockIdx.x * NACC * 128 + row) * 64;
#pragma unroll 1
        for (int a = 0; a < NACC; ++a) {
#pragma unroll 1
            for (int c = 0; c < 64; c += 32) {
                uint32_t v[32];
                tmem_ld32(tmem_base + ((uint32_t)(qd * 32) << 16) + (uint32_t)(a * 64 + c), v);
                if (a <= ACC_C && ((started >> (NACC + a)) & 1u)) {          // fold the duplicate accumulator in
                    uint32_t w[32];
                    tmem_ld32(tmem_base + ((uint32_t)(qd * 32) << 16) + (uint32_t)((NACC + a) * 64 + c), w);
#pragma unroll
                    for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(__uint_as_float(v[j]) + __uint_as_float(w[j]));
                }
                float* dst = out + (size_t)a * 128 * 64 + c;
                const bool live = (started >> a) & 1u;
#pragma unroll
                for (int j = 0; j < 32; j += 4)
                    *reinterpret_cast<float4*>(dst + j) =
                        live ? make_float4(__uint_as_float(v[j]), __uint_as_float(v[j + 1]), __uint_as_float(v[j + 2]),
                                           __uint_as_float(v[j + 3]))
                             : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            }
        }
    }
    if (stamps && threadIdx.x == 64) stamps[4] = global_timer_ns();
    tcgen05_fence_before();
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "n"(TMEM_COLS));
    }
}

// dW7[co][kh][kw][c] (the channels-last memory order of a [64, 3, 7, 7] weight) from the per-CTA partials:
// kernel row bh of the 4x4 form is spread over two accumulator halves (see the table at the top).
template <typename OutT>
__global__ void __launch_bounds__(256) k_stem_wgrad_reduce(const float* __restrict__ partial, OutT* __restrict__ out, int grid,
                                                           float scale) {
    const int id = blockIdx.x * blockDim.x + threadIdx.x;
    if (id >= 64 * 7 * 7 * 3) return;
    const int c = id % 3;
    const int kw = (id / 3) % 7;
    const int kh = (id / 21) % 7;
    const int co = id / 147;
    const int kh8 = kh + 1, kw8 = kw + 1;                       // the 7x7 kernel sits in an 8x8 one with a zero first row / column
    const int bh = kh8 >> 1, dh = kh8 & 1, bw = kw8 >> 1, dw = kw8 & 1;
    const int col = bw * 16 + dh * 6 + dw * 3 + c;
    // (accumulator, half) pairs holding kernel row bh
    const int acc0 = bh == 0 ? ACC_A : bh == 1 ? ACC_A : bh == 2 ? ACC_B : ACC_B;
    const int hi0 = bh == 0 ? 1 : bh == 1 ? 0 : bh == 2 ? 1 : 0;
    const int acc1 = bh == 0 ? ACC_D : bh == 1 ? ACC_C : bh == 2 ? ACC_C : ACC_E;
    const int hi1 = bh == 0 ? 0 : bh == 1 ? 1 : bh == 2 ? 0 : 1;
    const size_t o0 = ((size_t)acc0 * 128 + hi0 * 64 + co) * 64 + col;
    const size_t o1 = ((size_t)acc1 * 128 + hi1 * 64 + co) * 64 + col;
    const size_t cta_stride = (size_t)NACC * 128 * 64;
    float s0 = 0.0f, s1 = 0.0f, s2 = 0.0f, s3 = 0.0f;
    int g = 0;
    for (; g + 1 < grid; g += 2) {
        s0 += partial[g * cta_stride + o0];
        s1 += partial[g * cta_stride + o1];
        s2 += partial[(g + 1) * cta_stride + o0];
        s3 += partial[(g + 1) * cta_stride + o1];
    }
    if (g < grid) { s0 += partial[g * cta_stride + o0]; s1 += partial[g * cta_stride + o1]; }
    const float v = ((s0 + s1) + (s2 + s3)) * scale;
    if constexpr (sizeof(OutT) == 4) out[id] = v;
    else out[id] = __float2bfloat16_rn(v);
}

}  // namespace stem
}  // namespace lwc

using namespace lwc::stem;

static void ensure_context() {
    CUcontext ctx = nullptr;
    if (cuCtxGetCurrent(&ctx) != CUDA_SUCCESS || !ctx) {
        int dev = 0;
        cudaGetDevice(&dev);
        CUcontext prim = nullptr;
        if (cuDevicePrimaryCtxRetain(&prim, dev) == CUDA_SUCCESS) cuCtxSetCurrent(prim);
    }
}

static int stem_grid(int N, int OH) {
    const long long rows = (long long)N * (OH + 3);
    long long g = rows / 32;                        // >= 32 input rows per CTA: every accumulator gets used
    if (g > 148) g = 148;
    if (g < 1) g = 1;
    return (int)g;
}

// fp32 workspace the caller provides: grid * 5 * 128 * 64 floats.
extern "C" size_t lwc_stem_wgrad_workspace_bytes(int N, int OH) {
    return (size_t)stem_grid(N, OH) * NACC * 128 * 64 * sizeof(float) + 256;      // + timing stamps (LWC_STEM_DEBUG=16)
}

// dy  : [N][OH][OW][64] bf16 (channels-last gradient of the stem output)
// xs  : [N][OH+3][OW+3][16] bf16 (space-to-depth input produced by lwc_normalize_u8 mode 1)
// out : [64][7][7][3] (= channels-last [64,3,7,7]) in out_dtype (LWC_F32 / LWC_BF16), = scale * dW
// Requirements: OH even, OW a multiple of 16; shared memory (8 + stages) * OW * 128 bytes + barriers (OW <= 176).
extern "C" int lwc_stem_wgrad(const void* dy, const void* xs, void* out, float* workspace, int N, int OH, int OW, int out_dtype,
                              float scale, void* stream) {
    if (!dy || !xs || !out || !workspace || N <= 0 || OH <= 0 || (OH & 1) || OW < 16 || (OW & 15) || OW > 256) return LWC_EINVAL;
    if (out_dtype != LWC_F32 && out_dtype != LWC_BF16) return LWC_EINVAL;
    const int HB = OH + 3, WB = OW + 3;
    const size_t slot = (size_t)OW * 128;
    int nb = (int)((220 * 1024 - 2 * PAIR_SLOTS * slot) / slot);
    if (nb > MAX_B_STAGES) nb = MAX_B_STAGES;
    if (nb < 2) return LWC_ETOOBIG;
    const size_t smem_bytes = (2 * PAIR_SLOTS + nb) * slot + 512 + 1024;
    if ((long long)N * OH * OW >= (1LL << 31) || (long long)N * HB >= (1LL << 31)) return LWC_ETOOBIG;
    ensure_context();
    cudaStream_t st = (cudaStream_t)stream;
    CUtensorMap tmA, tmB;
    cuuint32_t estr[3] = {1, 1, 1};
    {
        cuuint64_t gdim[2] = {64, (cuuint64_t)N * OH * OW};
        cuuint64_t gstr[1] = {128};
        cuuint32_t box[2] = {64, (cuuint32_t)OW};
        CUresult r = cuTensorMapEncodeTiled(&tmA, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(dy), gdim, gstr, box, estr,
                                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                            CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) {
            if (getenv("LWC_DEBUG")) fprintf(stderr, "lwc_stem_wgrad: tensor map A -> %d\n", (int)r);
            return LWC_EINVAL;
        }
    }
    {
        // overlapping windows: 64 elements (4 pixels x 16 channels) every 32 bytes (1 pixel)
        cuuint64_t gdim[3] = {64, (cuuint64_t)OW, (cuuint64_t)N * HB};
        cuuint64_t gstr[2] = {32, (cuuint64_t)WB * 32};
        cuuint32_t box[3] = {64, (cuuint32_t)OW, 1};
        CUresult r = cuTensorMapEncodeTiled(&tmB, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(xs), gdim, gstr, box, estr,
                                            CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                            CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) {
            if (getenv("LWC_DEBUG")) fprintf(stderr, "lwc_stem_wgrad: tensor map B -> %d\n", (int)r);
            return LWC_EINVAL;
        }
    }
    static bool attr_set = false;
    if (!attr_set) {
        LWC_CUDA_CHECK(cudaFuncSetAttribute(k_stem_wgrad, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_set = true;
    }
    Params p;
    p.N = N; p.OH = OH; p.OW = OW; p.HB = HB; p.rows_total = N * HB; p.partial = workspace; p.b_stages = nb;
    p.debug = getenv("LWC_STEM_DEBUG") ? atoi(getenv("LWC_STEM_DEBUG")) : 0;
    const int grid = stem_grid(N, OH);
    k_stem_wgrad<<<grid, THREADS, smem_bytes, st>>>(tmA, tmB, p);
    const int total = 64 * 147;
    if (out_dtype == LWC_F32)
        k_stem_wgrad_reduce<float><<<(total + 255) / 256, 256, 0, st>>>(workspace, reinterpret_cast<float*>(out), grid, scale);
    else
        k_stem_wgrad_reduce<__nv_bfloat16><<<(total + 255) / 256, 256, 0, st>>>(workspace, reinterpret_cast<__nv_bfloat16*>(out), grid, scale);
    return (int)cudaGetLastError();
}
