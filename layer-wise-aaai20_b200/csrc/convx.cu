// Tap-table implicit-GEMM convolution on tcgen05 / TMEM / TMA: every convolution of the model zoo, forward,
// data gradient and weight gradient, channels-last bf16.
//
// Forward-like kernel k_fx (forward, data gradient, strided variants, small-Cin first layers, the 7x7/2 stem):
//
//   Y[p, co] = sum_{t < ntaps} sum_{c < Kc} A[(p * s + d_t), a_c0[t] + c] * B[co, b_k0[t] + c]
//
//   * the output tile is a Wb x Hb x Nb patch of pixels (<= 128 rows); each (tap, 64-channel chunk) of the K loop is
//     ONE 4-D TMA box load of that patch of the input shifted by the tap offset d_t -- out-of-bounds rows, columns
//     and channels are zero-filled by the TMA unit, which is the padding (also the channel padding of 3- / 8- /
//     16-channel inputs: the K chunk is padded in the descriptor, not in memory);
//   * stride 2 is the tensor map's traversal stride (elementStrides), the kernel only scales the start coordinate;
//   * 1x1 stride-1 layers are the same code with the pixels flattened into one axis;
//   * the data gradient of a stride-2 layer is four launches, one per output parity class, each with its own tap
//     subset and a strided output map; the 7x7/2 stem is four row-taps over overlapping 128-byte windows of the
//     space-to-depth input;
//   * BN in {64, 128, 256}: Cout <= 256 layers read the activation exactly once and write whole output rows;
//   * persistent CTAs (static tile schedule), accumulator double-buffered in tensor memory, epilogue
//     TMEM -> registers -> (bias, ReLU) -> swizzled shared memory -> TMA store through a ring of 64-column staging
//     buffers; optional per-channel sum / sum-of-squares of the rounded outputs accumulated per CTA in shared
//     memory in a fixed order (batch-norm statistics without a pass over the output).
//
// Weight-gradient kernel k_wx:   dW[co, t, ci] = sum_p dY[p, co] * X[p * s + d_t, ci]
//   D[(tap, ci) : 128][co : BN] accumulators in tensor memory (up to 512 columns = several (tap, ci-pair) slices
//   per CTA), both operands MN-major straight from the channels-last tensors (4-D TMA boxes of the same pixel patch,
//   the X box shifted by the tap), split over pixel patches across CTAs, fp32 partials reduced in a fixed order.
//
// All TMA loads here read this GPU's own activations and weights; none is a cluster-multicast load.
//
// Reference ops: nn.Conv2d forward / backward through cuDNN -- IMAGENET/training/resnet.py:18-21,59-66,100-101,122-126;
// CIFAR10/dawn.py:23-35; CIFAR10/alexnet.py:15-27; CIFAR10/vgg16.py:69-82.
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "tc.cuh"

namespace lwc {
namespace convx {

using namespace lwc::tc;

constexpr int BM = 128;
constexpr int BK = 64;
constexpr int THREADS = 320;          // k_fx: TMA warp, MMA warp, 8 epilogue warps
constexpr int WX_THREADS = 192;       // k_wx: TMA warp, MMA warp, 4 epilogue warps
constexpr int MAXT = 16;
constexpr int SUB_BYTES = BM * 128;          // one 128-row x 64-column bf16 staging sub-tile
constexpr int SMEM_BUDGET = 227 * 1024;

struct FxParams {
    int N;                                   // output channels
    int ntaps, kchunks;                      // K loop = ntaps * kchunks iterations of 64 channels
    int Wb, Hb, Nb, tiles_w, tiles_h;        // output patch and tile grid (tiles_m = tiles_w * tiles_h * tiles_nimg)
    int sw, sh;                              // input start coordinate = output coordinate * s + d
    int OW, OH, ON;                          // output extent (rows of a patch beyond it are neither stored nor counted)
    int tiles_m, tiles_n, rows_per_tile;
    int relu;
    int stages, ncs;                         // operand ring depth, staging ring depth
    int c_off, stat_off, bar_off;            // shared-memory offsets
    const float* bias;
    float* stats;                            // [gridDim.x][2][N] per-CTA sums / sums of squares, or NULL
    // batch-norm BACKWARD sums instead (data-gradient launches whose output dz is the gradient of a batch-norm + ReLU
    // output): the matching tile of the batch norm's input x is TMA-prefetched next to the staged dz tile and the
    // epilogue accumulates S1 = sum d, S2 = sum d * x with d = dz * [gamma * invstd * (x - mean) + beta > 0]
    int bnb, bnb_relu, x_off;
    const float *bn_gamma, *bn_beta, *bn_mean, *bn_invstd;
    // residual fork (data gradient of the block's first 1x1 convolution): out = mask * (acc + addend), where addend is the
    // skip connection's gradient (a tile of it is TMA-prefetched like the x tile, on the same barrier) and mask is the
    // one-bit-per-element ReLU mask of the batch norm that produced the block input ([pixel][N / 8] bytes, or NULL)
    int add, add_off;
    const uint8_t* mask;
    // halo mode (64 input channels, stride 1, several taps): the weights of all taps stay in shared memory for the whole
    // kernel and ONE TMA box per tile brings the input patch with its halo (pixel pitch 16); every tap is then the same
    // box read from a start address shifted by (dh * 16 + dw) rows -- the activation crosses L2 -> SM once, not per tap
    int halo, halo_dw0, halo_dh0;            // on/off, offset of the halo origin from the output origin
    int a_off, a_stage, halo_bytes;          // operand ring offset / stage size, bytes of one halo box
    int halo_bo;                             // debugging only: also put (start >> 7) & 7 into the descriptor's base-offset field
    short dw[MAXT], dh[MAXT];
    int a_c0[MAXT], b_k0[MAXT], b_n0[MAXT];   // b_n0: column offset of the tap in an MN-major weight matrix
};
constexpr int HALO_PW = 16;                  // pixel pitch of a halo box: accumulator row r = (r / 16, r % 16)

struct TileCoord { int mt, nt, w0, h0, n0; };
__device__ __forceinline__ TileCoord tile_coord(const FxParams& p, int tile) {
    TileCoord t;
    t.mt = tile / p.tiles_n;
    t.nt = tile - t.mt * p.tiles_n;
    const int tw = t.mt % p.tiles_w;
    const int r = t.mt / p.tiles_w;
    const int th = r % p.tiles_h;
    t.w0 = tw * p.Wb; t.h0 = th * p.Hb; t.n0 = (r / p.tiles_h) * p.Nb;
    return t;
}

// REGSUM: the variant that keeps per-channel sums in registers (batch statistics of the outputs, batch-norm backward
// sums); the plain variant carries neither the registers nor the code.  (Tried and dropped: the statistics as one more
// MMA per staged sub-tile, D += [Y^T ; 1^T] Y with both operands MN-major straight from the staging tile, issued by the
// epilogue's leader thread next to its TMA store.  Correct, but slower than the CUDA-core sums it saved: the eight
// descriptor-building MMA issues sit on the one thread all eight epilogue warps wait for at the next barrier.  The
// operand layout is not the reason -- tools/mma_rate_probe.py measures the same 48 / 64 / 128 cycles per MMA at
// N = 64 / 128 / 256 for K-major and MN-major operands alike.  Issued from an eleventh, otherwise idle warp instead
// (mbarrier hand-shake per staging slot) the statistics cost 6 % less than the register sums at 128-wide tiles, still
// more than the 64-wide two-CTAs-per-SM variant with register sums, and the extra warp slowed the plain path by 6-10 %:
// dropped as well.)
// TWO: compiled for two resident CTAs per SM (<= 96 registers; the host halves the shared-memory plan and the tensor
// memory of both fits for BN <= 128).  The epilogue of one CTA is a chain of dependent steps with two barriers per
// sub-tile, eight warps in step; a second CTA at another point of its chain fills the issue slots the first leaves idle.
template <int BN, bool B_MN, bool FORK, bool REGSUM, bool TWO>
__global__ void __launch_bounds__(THREADS, TWO ? 2 : 1) k_fx(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                                                   const __grid_constant__ CUtensorMap tmC, const __grid_constant__ CUtensorMap tmX,
                                                   const __grid_constant__ CUtensorMap tmD, const FxParams p) {
    // 1024-byte aligned by declaration (the 128-byte swizzle atoms need it): no alignment slack in the budget, which is
    // what lets the two-CTAs-per-SM plans with an auxiliary ring fit
    extern __shared__ __align__(1024) uint8_t smem_fx[];
    uint8_t* smem = smem_fx;
    constexpr int A_BYTES = BM * BK * 2;
    constexpr int B_BYTES = BN * BK * 2;
    constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    constexpr int MAXS = 8;
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + p.bar_off);
    uint64_t* empty = full + MAXS;
    uint64_t* tmem_full = empty + MAXS;        // [2]
    uint64_t* tmem_empty = tmem_full + 2;      // [2]
    uint64_t* xfull = tmem_empty + 2;          // [4] x tile of a chunk has landed (batch-norm backward sums)
    uint64_t* wfull = xfull + 4;               // [1] resident weights have landed (halo mode)
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(wfull + 1);
    constexpr int TCOLS = 2 * BN;

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int num_tiles = p.tiles_m * p.tiles_n;
    const int kiters = p.ntaps * p.kchunks;
    const int stages = p.stages;
    const uint32_t a_bytes = (uint32_t)p.rows_per_tile * (BK * 2);

    if (threadIdx.x == 0) {
        prefetch_tensormap(&tmA); prefetch_tensormap(&tmB); prefetch_tensormap(&tmC);
        for (int s = 0; s < stages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        for (int a = 0; a < 2; ++a) { mbar_init(&tmem_full[a], 1); mbar_init(&tmem_empty[a], 8); }
        for (int a = 0; a < 4; ++a) mbar_init(&xfull[a], 1);
        mbar_init(wfull, 1);
        if (p.bnb) prefetch_tensormap(&tmX);
        if (FORK) prefetch_tensormap(&tmD);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<TCOLS>(tmem_slot);
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0 && lane == 0) {
        // ---------------- TMA producer ----------------
        uint32_t s = 0, ph = 0;
        if (p.halo) {
            // the grid is a multiple of tiles_n: this CTA's output channels never change, its weights are loaded once
            const int co0 = ((int)blockIdx.x % p.tiles_n) * BN;
            mbar_expect_tx(wfull, (uint32_t)p.ntaps * B_BYTES);
            for (int tap = 0; tap < p.ntaps; ++tap) {
                uint8_t* b = smem + tap * B_BYTES;
                if (B_MN) {
#pragma unroll
                    for (int h = 0; h < BN / 64; ++h) tma_load_2d(b + h * 8192, &tmB, wfull, p.b_n0[tap] + co0 + h * 64, p.b_k0[tap]);
                } else {
                    tma_load_2d(b, &tmB, wfull, p.b_k0[tap], co0);
                }
            }
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
                const TileCoord t = tile_coord(p, tile);
                mbar_wait(&empty[s], ph ^ 1);
                mbar_expect_tx(&full[s], (uint32_t)p.halo_bytes);
                tma_load_4d(smem + p.a_off + s * p.a_stage, &tmA, &full[s], p.a_c0[0], t.w0 + p.halo_dw0, t.h0 + p.halo_dh0, t.n0);
                if (++s == (uint32_t)stages) { s = 0; ph ^= 1; }
            }
        } else
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
            const TileCoord t = tile_coord(p, tile);
            const int co0 = t.nt * BN;
            const int wi = t.w0 * p.sw, hi = t.h0 * p.sh;
            for (int tap = 0; tap < p.ntaps; ++tap) {
                const int aw = wi + p.dw[tap], ah = hi + p.dh[tap], ac = p.a_c0[tap], bk = p.b_k0[tap], bn0 = p.b_n0[tap];
                for (int ch = 0; ch < p.kchunks; ++ch) {
                    mbar_wait(&empty[s], ph ^ 1);
                    uint8_t* a = smem + s * STAGE_BYTES;
                    uint8_t* b = a + A_BYTES;
                    mbar_expect_tx(&full[s], a_bytes + B_BYTES);
                    tma_load_4d(a, &tmA, &full[s], ac + ch * BK, aw, ah, t.n0);
                    if (B_MN) {
#pragma unroll
                        for (int h = 0; h < BN / 64; ++h) tma_load_2d(b + h * 8192, &tmB, &full[s], bn0 + co0 + h * 64, bk + ch * BK);
                    } else {
                        tma_load_2d(b, &tmB, &full[s], bk + ch * BK, co0);
                    }
                    if (++s == (uint32_t)stages) { s = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 1 && lane == 0) {
        // ---------------- MMA issuer ----------------
        constexpr uint32_t idesc = make_idesc(BN, false, B_MN);
        uint32_t s = 0, ph = 0, tcount = 0;
        if (p.halo) {
            mbar_wait(wfull, 0);
            const uint32_t wsm = smem_u32(smem);
            for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tcount) {
                const uint32_t as = tcount & 1, aph = (tcount >> 1) & 1;
                mbar_wait(&tmem_empty[as], aph ^ 1);
                mbar_wait(&full[s], ph);
                fence_after();
                const uint32_t tacc = tmem_base + as * BN;
                const uint32_t a = smem_u32(smem + p.a_off + s * p.a_stage);
                for (int tap = 0; tap < p.ntaps; ++tap) {
                    // tap = the same box, `off` pixel rows (128 bytes each) further on.  The start is then no longer
                    // 1024-byte aligned, and that is fine as it stands: the 128-byte swizzle is applied to the shared-memory
                    // ADDRESS (bits 4-6 ^= bits 7-9), by the TMA unit that wrote the box and by the tensor core that reads
                    // it alike, so a row is found where it was put whatever row the descriptor starts at.  Measured on
                    // B200: correct with the base-offset field left 0, wrong with (start >> 7) & 7 in it.
                    const uint32_t off = (uint32_t)((p.dh[tap] - p.halo_dh0) * HALO_PW + (p.dw[tap] - p.halo_dw0));
                    uint64_t da = make_desc(a + off * 128u, 16, 1024);
                    if (p.halo_bo) da |= (uint64_t)(off & 7u) << 49;
                    const uint32_t b = wsm + tap * B_BYTES;
                    const uint64_t db = B_MN ? make_desc(b, 8192, 1024) : make_desc(b, 16, 1024);
                    umma_x4<2, (B_MN ? 128 : 2)>(tacc, da, db, idesc, tap > 0 ? 1u : 0u);
                }
                umma_commit(&empty[s]);
                umma_commit(&tmem_full[as]);
                if (++s == (uint32_t)stages) { s = 0; ph ^= 1; }
            }
        } else
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tcount) {
            const uint32_t as = tcount & 1, aph = (tcount >> 1) & 1;
            mbar_wait(&tmem_empty[as], aph ^ 1);
            fence_after();
            const uint32_t tacc = tmem_base + as * BN;
            for (int kk = 0; kk < kiters; ++kk) {
                mbar_wait(&full[s], ph);
                fence_after();
                const uint32_t a = smem_u32(smem + s * STAGE_BYTES);
                const uint32_t b = a + A_BYTES;
                const uint64_t da = make_desc(a, 16, 1024);
                const uint64_t db = B_MN ? make_desc(b, 8192, 1024) : make_desc(b, 16, 1024);
                umma_x4<2, (B_MN ? 128 : 2)>(tacc, da, db, idesc, kk > 0 ? 1u : 0u);
                umma_commit(&empty[s]);
                if (++s == (uint32_t)stages) { s = 0; ph ^= 1; }
            }
            umma_commit(&tmem_full[as]);
        }
    } else if (warp >= 2) {
        // ---------------- epilogue: 8 warps, two per TMEM lane quarter (each takes 32 of a chunk's 64 columns) --------
        const int ew = warp - 2;                       // 0..7
        const int q = warp & 3;                        // TMEM lane quarter this warp may read
        const int hsel = ew >> 2;                      // which 32-column half of every 64-column chunk
        const int r = q * 32 + lane;                   // accumulator row = pixel of the patch
        const int et = threadIdx.x - 64;               // 0..255
        const bool leader = et == 0;
        // pixel of the patch this accumulator row is, and the row of the (compact) staging tile it goes to: in halo mode
        // the accumulator has the box's 16-pixel pitch and the Wb-pixel-pitch staging tile drops the overhang columns
        // (they land behind the patch, zeroed, where the TMA store does not look)
        const int r_w = p.halo ? (r & (HALO_PW - 1)) : r % p.Wb;
        const int r_h = p.halo ? (r / HALO_PW) : (r / p.Wb) % p.Hb;
        const int r_n = p.halo ? 0 : r / (p.Wb * p.Hb);
        const bool r_in = p.halo ? (r_w < p.Wb && r_h < p.Hb) : (r < p.rows_per_tile);
        int dr = r;
        if (p.halo) dr = r_in ? r_h * p.Wb + r_w : p.Wb * p.Hb + r_h * (HALO_PW - p.Wb) + (r_w < p.Wb ? 0 : r_w - p.Wb);
        const bool slow = p.bias != nullptr || p.relu != 0;
        constexpr int NCH = BN / 64;
        // auxiliary-tile prefetch (x tile of the batch-norm backward sums, addend tile of the residual fork): the leader
        // keeps the loads two chunks ahead of the epilogue, one barrier per ring slot covers both tiles
        const bool aux = p.bnb || FORK;
        int pf_tile = blockIdx.x, pf_h = 0;
        uint32_t pf_q = 0, q_cnt = 0;
        auto prefetch_x = [&]() {
            if (pf_tile >= num_tiles) return;
            const TileCoord tt = tile_coord(p, pf_tile);
            const uint32_t slot = pf_q % (uint32_t)p.ncs;
            mbar_expect_tx(&xfull[slot], (p.bnb && FORK) ? 2 * a_bytes : a_bytes);
            if (p.bnb) tma_load_4d(smem + p.x_off + slot * SUB_BYTES, &tmX, &xfull[slot], tt.nt * BN + pf_h * 64, tt.w0, tt.h0, tt.n0);
            if (FORK) tma_load_4d(smem + p.add_off + slot * SUB_BYTES, &tmD, &xfull[slot], tt.nt * BN + pf_h * 64, tt.w0, tt.h0, tt.n0);
            ++pf_q;
            if (++pf_h == NCH) { pf_h = 0; pf_tile += gridDim.x; }
        };
        // prefetch distance: two sub-tiles, one when the ring has only two slots
        if (aux && leader) { prefetch_x(); if (p.ncs >= 3) prefetch_x(); }
        constexpr int NS = REGSUM ? NCH : 1;
        float sacc[NS][16];                            // [chunk][sum x 8 channels | sum of squares x 8 channels]
#pragma unroll
        for (int h = 0; h < NS; ++h)
#pragma unroll
            for (int k = 0; k < 16; ++k) sacc[h][k] = 0.0f;
        uint32_t tcount = 0, sub = 0;
        for (int tile = blockIdx.x; tile < num_tiles; tile += gridDim.x, ++tcount) {
            const TileCoord t = tile_coord(p, tile);
            const int co0 = t.nt * BN;
            const uint32_t as = tcount & 1, aph = (tcount >> 1) & 1;
            // rows of the patch that are not output pixels are zeroed in the staging tile: the TMA store clips them
            // anyway, and the statistics must not see them
            const bool row_live = r_in && t.w0 + r_w < p.OW && t.h0 + r_h < p.OH && t.n0 + r_n < p.ON;
            // ReLU-mask bits of this thread's 32 channels of every chunk: fetched before the accumulator is waited for
            uint32_t mbits[NCH];
#pragma unroll
            for (int h = 0; h < NCH; ++h) mbits[h] = 0xffffffffu;
            if (FORK && p.mask && row_live) {
                const size_t pix = ((size_t)(t.n0 + r_n) * p.OH + (t.h0 + r_h)) * p.OW + (t.w0 + r_w);
                const uint8_t* mrow = p.mask + pix * (size_t)(p.N >> 3) + ((co0 + hsel * 32) >> 3);
#pragma unroll
                for (int h = 0; h < NCH; ++h) mbits[h] = __ldg(reinterpret_cast<const uint32_t*>(mrow + h * 8));
            }
            mbar_wait(&tmem_full[as], aph);
            fence_after();
            const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + as * BN + (uint32_t)(hsel * 32);
            // accumulator loads run one 64-column chunk ahead of the conversion -- except in the two-CTAs-per-SM variant
            // with register sums at 128-wide tiles, where the second buffer does not fit in 96 registers (the other
            // CTA covers the latency there)
            constexpr int NVV = (TWO && REGSUM && BN >= 128) ? 1 : 2;
            uint32_t vv[NVV][32];
            tmem_ld32_issue(taddr, vv[0]);
#pragma unroll
            for (int h = 0; h < NCH; ++h) {
                uint8_t* cst = smem + p.c_off + sub * SUB_BYTES;
                uint32_t* v = vv[h % NVV];
                if (NVV == 1 && h > 0) tmem_ld32_issue(taddr + (uint32_t)(h * 64), vv[0]);
                // the store that read this staging buffer `ncs` chunks ago must have finished reading it
                if (leader) {
                    if (p.ncs >= 4) tma_wait_group_read<3>(); else if (p.ncs == 3) tma_wait_group_read<2>(); else tma_wait_group_read<1>();
                }
                tmem_ld_wait();
                if (h + 1 < NCH) {
                    if (NVV == 2) tmem_ld32_issue(taddr + (uint32_t)((h + 1) * 64), vv[(h + 1) % NVV]);     // next chunk in flight
                } else {                                // whole row is in registers: the accumulator may be overwritten
                    fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&tmem_empty[as]);
                }
                epi_bar_sync_n<256>();                         // staging buffer free (leader waited above)
                if (aux && leader) prefetch_x();        // chunk q + 2: its slot was last read two chunks ago
                use_after_wait(v);
                const uint32_t slot = q_cnt % (uint32_t)p.ncs;
                if (aux) mbar_wait(&xfull[slot], (q_cnt / (uint32_t)p.ncs) & 1u);
                if (FORK) {
                    const uint8_t* arow = smem + p.add_off + slot * SUB_BYTES + r * 128;
                    const uint32_t mb = mbits[h];
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const uint4 u = *reinterpret_cast<const uint4*>(arow + (((hsel * 4 + j) ^ (r & 7)) * 16));
                        const uint32_t uw[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float lo = __uint_as_float(v[j * 8 + 2 * e]) + __uint_as_float(uw[e] << 16);
                            const float hi = __uint_as_float(v[j * 8 + 2 * e + 1]) + __uint_as_float(uw[e] & 0xffff0000u);
                            v[j * 8 + 2 * e] = ((mb >> (j * 8 + 2 * e)) & 1u) ? __float_as_uint(lo) : 0u;
                            v[j * 8 + 2 * e + 1] = ((mb >> (j * 8 + 2 * e + 1)) & 1u) ? __float_as_uint(hi) : 0u;
                        }
                    }
                }
                if (slow) {
#pragma unroll
                    for (int e = 0; e < 32; ++e) {
                        float x = __uint_as_float(v[e]);
                        if (p.bias) x += __ldg(p.bias + co0 + h * 64 + hsel * 32 + e);
                        if (p.relu) x = fmaxf(x, 0.0f);
                        v[e] = __float_as_uint(x);
                    }
                }
                if (!row_live) {
#pragma unroll
                    for (int e = 0; e < 32; ++e) v[e] = 0u;
                }
                uint8_t* rowp = cst + dr * 128;
#pragma unroll
                for (int j = 0; j < 4; ++j) {           // four 16-byte chunks (8 bf16 each) of this thread's half row
                    uint32_t w[4];
#pragma unroll
                    for (int e = 0; e < 4; ++e) {
                        __nv_bfloat162 hb = __floats2bfloat162_rn(__uint_as_float(v[j * 8 + 2 * e]), __uint_as_float(v[j * 8 + 2 * e + 1]));
                        w[e] = *reinterpret_cast<uint32_t*>(&hb);
                    }
                    *reinterpret_cast<uint4*>(rowp + (((hsel * 4 + j) ^ (dr & 7)) * 16)) = make_uint4(w[0], w[1], w[2], w[3]);
                }
                fence_proxy_async_smem();
                epi_bar_sync_n<256>();
                if (leader) {
                    if (co0 + h * 64 < p.N) tma_store_4d(&tmC, cst, co0 + h * 64, t.w0, t.h0, t.n0);
                    tma_commit_group();                 // one group per chunk keeps the count in step with the ring
                }
                if constexpr (REGSUM) {
                if (p.bnb) {
                    // batch-norm backward sums over this chunk: dz from the staging tile, x from the prefetched tile
                    const uint8_t* xs = smem + p.x_off + slot * SUB_BYTES;
                    const int c = et & 7, rg = et >> 3;
                    float fa[8], fb[8];
                    {
                        const int ch0 = co0 + h * 64 + c * 8;
#pragma unroll
                        for (int k = 0; k < 8; ++k) {
                            const int ch = ch0 + k < p.N ? ch0 + k : p.N - 1;
                            fa[k] = (p.bn_gamma ? __ldg(p.bn_gamma + ch) : 1.0f) * __ldg(p.bn_invstd + ch);
                            fb[k] = (p.bn_beta ? __ldg(p.bn_beta + ch) : 0.0f) - __ldg(p.bn_mean + ch) * fa[k];
                        }
                    }
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int row = rg * 4 + i;
                        if (row >= p.rows_per_tile) continue;      // rows past the patch: the x tile holds nothing there
                        const int off = row * 128 + ((c ^ (row & 7)) * 16);
                        const uint4 ud = *reinterpret_cast<const uint4*>(cst + off);
                        const uint4 ux = *reinterpret_cast<const uint4*>(xs + off);
                        const uint32_t dw4[4] = {ud.x, ud.y, ud.z, ud.w}, xw4[4] = {ux.x, ux.y, ux.z, ux.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            float d0 = __uint_as_float(dw4[e] << 16), d1 = __uint_as_float(dw4[e] & 0xffff0000u);
                            const float x0 = __uint_as_float(xw4[e] << 16), x1 = __uint_as_float(xw4[e] & 0xffff0000u);
                            if (p.bnb_relu) {
                                if (!(fmaf(x0, fa[2 * e], fb[2 * e]) > 0.0f)) d0 = 0.0f;
                                if (!(fmaf(x1, fa[2 * e + 1], fb[2 * e + 1]) > 0.0f)) d1 = 0.0f;
                            }
                            sacc[h][2 * e] += d0; sacc[h][8 + 2 * e] = fmaf(d0, x0, sacc[h][8 + 2 * e]);
                            sacc[h][2 * e + 1] += d1; sacc[h][8 + 2 * e + 1] = fmaf(d1, x1, sacc[h][8 + 2 * e + 1]);
                        }
                    }
                } else if (p.stats) {
                    // column sums of the rounded 128 x 64 sub-tile, kept in registers for the whole kernel: thread =
                    // (group of 4 rows, 16-byte chunk of 8 channels); the host sizes the grid so that a CTA always
                    // works on the same BN output channels, i.e. chunk h means the same channels in every tile
                    const int c = et & 7, rg = et >> 3;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int row = rg * 4 + i;
                        const uint4 u = *reinterpret_cast<const uint4*>(cst + row * 128 + ((c ^ (row & 7)) * 16));
                        const uint32_t uw[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float lo = __uint_as_float(uw[e] << 16), hi = __uint_as_float(uw[e] & 0xffff0000u);
                            sacc[h][2 * e] += lo; sacc[h][8 + 2 * e] = fmaf(lo, lo, sacc[h][8 + 2 * e]);
                            sacc[h][2 * e + 1] += hi; sacc[h][8 + 2 * e + 1] = fmaf(hi, hi, sacc[h][8 + 2 * e + 1]);
                        }
                    }
                }
                }
                if (aux) ++q_cnt;
                if (++sub == (uint32_t)p.ncs) sub = 0;
            }
        }
        if (leader) tma_wait_group<0>();
        if constexpr (REGSUM) {
        if (p.stats) {
            // once per kernel: row groups of a warp by shuffles, the eight warps through shared memory (the staging
            // ring is free now) in a fixed order, then one partial row [2][N] per CTA (zero outside its BN channels)
            epi_bar_sync_n<256>();
            float* scratch = reinterpret_cast<float*>(smem + p.c_off);     // [8 warps][NCH][2][64]
#pragma unroll
            for (int h = 0; h < NCH; ++h) {
#pragma unroll
                for (int k = 0; k < 16; ++k) {
                    float v = sacc[h][k];
                    v += __shfl_xor_sync(0xffffffffu, v, 8);
                    v += __shfl_xor_sync(0xffffffffu, v, 16);
                    sacc[h][k] = v;
                }
                if (lane < 8) {
                    float* o = scratch + (ew * NCH + h) * 128 + lane * 8;
#pragma unroll
                    for (int k = 0; k < 8; ++k) { o[k] = sacc[h][k]; o[64 + k] = sacc[h][8 + k]; }
                }
            }
            epi_bar_sync_n<256>();
            float* o = p.stats + (size_t)blockIdx.x * 2 * p.N;
            const int nt_mine = ((int)blockIdx.x < num_tiles) ? (int)blockIdx.x % p.tiles_n : 0;
            for (int i = et; i < 2 * p.N; i += 256) {
                const int which = i / p.N, cg = i - which * p.N;
                const int local = cg - nt_mine * BN;
                float tot = 0.0f;
                if (local >= 0 && local < BN) {
                    const float* sp = scratch + (local >> 6) * 128 + which * 64 + (local & 63);
                    constexpr int WS = NCH * 128;
                    tot = ((sp[0] + sp[WS]) + (sp[2 * WS] + sp[3 * WS])) + ((sp[4 * WS] + sp[5 * WS]) + (sp[6 * WS] + sp[7 * WS]));
                }
                o[i] = tot;
            }
        }
        }
    }
    fence_before();
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        tmem_dealloc<TCOLS>(tmem_base);
    }
}

// ======================================================================================================
// Weight gradient.  One CTA = (slice group g, pixel split z):
//   * boxes are (tap, 64-input-channel chunk) pairs in tap-major order; a slice is two consecutive boxes (M = 128
//     accumulator rows), a group is `nacc` consecutive slices that share the same BN output channels;
//   * every pipeline stage holds one patch of PK pixels: the dY tile [PK][BN] (B operand, MN-major) and for each
//     slice two X boxes [PK][64] shifted by their tap (A operand, MN-major);  nacc * PK / 16 MMAs per stage;
//   * the CTA walks patches z, z + ksplit, ...; at the end the nacc accumulators go to the fp32 workspace
//     [ksplit][nboxes][64][Cout], which k_wx_reduce sums over the splits in a fixed order and transposes into the
//     channels-last weight layout [Cout][taps][Cin].
// ======================================================================================================
struct WxParams {
    int Cout, Cin_pad;                       // Cin_pad = 64 * cchunks (channels past the tensor's extent read as zero)
    int ntaps, cchunks, nboxes;              // nboxes = ntaps * cchunks
    int nacc, groups_per_co;                 // slices per CTA, groups per output-channel tile
    int PK;                                  // pixels per stage (multiple of 16)
    int Wb, Hb, Nb, tiles_w, tiles_h, npatches;
    int sw, sh;
    int stages, stage_bytes, bar_off;
    int ksplit;
    float* ws;                               // [ksplit][nboxes][64][Cout]
    short dw[MAXT], dh[MAXT];
};

template <int BN>
__global__ void __launch_bounds__(WX_THREADS, 1) k_wx(const __grid_constant__ CUtensorMap tmX, const __grid_constant__ CUtensorMap tmDY,
                                                   const WxParams p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    constexpr int MAXS = 8;
    constexpr int TCOLS = 512;               // nacc * BN <= 512
    uint64_t* full = reinterpret_cast<uint64_t*>(smem + p.bar_off);
    uint64_t* empty = full + MAXS;
    uint64_t* tmem_full = empty + MAXS;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tmem_full + 1);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int g = blockIdx.x, z = blockIdx.y;
    const int co_tile = g / p.groups_per_co;
    const int box0 = (g % p.groups_per_co) * p.nacc * 2;
    const int co0 = co_tile * BN;
    int nslices = (p.nboxes - box0 + 1) / 2;
    if (nslices > p.nacc) nslices = p.nacc;
    if (nslices < 0) nslices = 0;
    const int box_bytes = p.PK * 128;
    const int b_bytes = (BN / 64) * box_bytes;
    const int npat = z < p.npatches ? (p.npatches - z + p.ksplit - 1) / p.ksplit : 0;

    if (threadIdx.x == 0) {
        prefetch_tensormap(&tmX); prefetch_tensormap(&tmDY);
        for (int s = 0; s < p.stages; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
        mbar_init(tmem_full, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc<TCOLS>(tmem_slot);
    fence_before();
    __syncthreads();
    fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ---------------- TMA producer: one lane per box, so the boxes of a stage are issued side by side -------------
        // (a single thread walking the boxes -- two integer divisions and a 4-D TMA issue each -- was the bottleneck:
        //  ncu showed the MMA thread waiting on `full` with the tensor pipe 25 % busy)
        int nlive = p.nboxes - box0;                       // boxes of this group that exist
        if (nlive > 2 * nslices) nlive = 2 * nslices;
        if (nlive < 0) nlive = 0;
        constexpr int NDY = BN / 64;
        const bool is_dy = lane < NDY;
        const bool act = lane < NDY + nlive;
        int my_c = 0, my_dw = 0, my_dh = 0, my_sw = 1, my_sh = 1;
        uint32_t my_off = 0;
        if (is_dy) {
            my_c = co0 + lane * 64;
            my_off = (uint32_t)(lane * box_bytes);
        } else if (act) {
            const int bx = box0 + lane - NDY;
            const int tap = bx / p.cchunks, cc = bx - tap * p.cchunks;
            my_c = cc * 64; my_dw = p.dw[tap]; my_dh = p.dh[tap]; my_sw = p.sw; my_sh = p.sh;
            my_off = (uint32_t)(b_bytes + (lane - NDY) * box_bytes);
        }
        const CUtensorMap* my_map = is_dy ? &tmDY : &tmX;
        // patch index z, z + ksplit, ... in mixed radix (tiles_w, tiles_h, images): no division inside the loop
        int tw = z % p.tiles_w, th = (z / p.tiles_w) % p.tiles_h, tn = z / (p.tiles_w * p.tiles_h);
        const int step_w = p.ksplit % p.tiles_w, step_h = (p.ksplit / p.tiles_w) % p.tiles_h, step_n = p.ksplit / (p.tiles_w * p.tiles_h);
        const uint32_t stage_tx = (uint32_t)(b_bytes + nlive * box_bytes);
        uint32_t s = 0, ph = 0;
        for (int i = 0; i < npat; ++i) {
            if (lane == 0) {
                mbar_wait(&empty[s], ph ^ 1);
                mbar_expect_tx(&full[s], stage_tx);
            }
            __syncwarp();
            if (act) tma_load_4d(smem + s * p.stage_bytes + my_off, my_map, &full[s], my_c, tw * p.Wb * my_sw + my_dw,
                                 th * p.Hb * my_sh + my_dh, tn * p.Nb);
            tw += step_w;
            int carry = 0;
            if (tw >= p.tiles_w) { tw -= p.tiles_w; carry = 1; }
            th += step_h + carry;
            carry = 0;
            if (th >= p.tiles_h) { th -= p.tiles_h; carry = 1; }
            tn += step_n + carry;
            if (++s == (uint32_t)p.stages) { s = 0; ph ^= 1; }
        }
    } else if (warp == 1 && lane == 0) {
        // ---------------- MMA issuer ----------------
        constexpr uint32_t idesc = make_idesc(BN, true, true);
        const int ksteps = p.PK / 16;
        uint32_t s = 0, ph = 0;
        for (int i = 0; i < npat; ++i) {
            mbar_wait(&full[s], ph);
            fence_after();
            const uint32_t b = smem_u32(smem + s * p.stage_bytes);
            const uint32_t a = b + b_bytes;
            // MN-major operands: 64-wide MN atoms box_bytes apart (LBO), 8-pixel groups 1024 bytes apart (SBO),
            // one K step = 16 pixels = 2048 bytes
            const uint64_t db = make_desc(b, (uint32_t)box_bytes, 1024);
            for (int j = 0; j < nslices; ++j) {
                const uint64_t da = make_desc(a + j * 2 * box_bytes, (uint32_t)box_bytes, 1024);
                for (int ks = 0; ks < ksteps; ++ks)
                    umma(tmem_base + j * BN, da + (uint64_t)(ks * 128), db + (uint64_t)(ks * 128), idesc, (i > 0 || ks > 0) ? 1u : 0u);
            }
            umma_commit(&empty[s]);
            if (++s == (uint32_t)p.stages) { s = 0; ph ^= 1; }
        }
        umma_commit(tmem_full);
    } else if (warp >= 2) {
        // ---------------- epilogue: accumulators -> fp32 partials ----------------
        const int q = warp & 3;
        const int r = q * 32 + lane;                       // accumulator row = (box parity, channel)
        mbar_wait(tmem_full, 0);
        fence_after();
        // Each warp transposes its 32 accumulator rows through shared memory (the operand ring is idle now) so that a
        // store instruction writes 128 contiguous bytes of one partial row instead of 32 scattered 16-byte pieces.
        float* tb = reinterpret_cast<float*>(smem) + (warp - 2) * (32 * 33);
        for (int j = 0; j < nslices; ++j) {
#pragma unroll 1
            for (int c = 0; c < BN; c += 32) {
                uint32_t v[32];
                tmem_ld32_issue(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(j * BN + c), v);
                tmem_ld_wait();
                use_after_wait(v);
                if (npat == 0) {
#pragma unroll
                    for (int e = 0; e < 32; ++e) v[e] = 0u;
                }
                __syncwarp();
#pragma unroll
                for (int e = 0; e < 32; ++e) tb[lane * 33 + e] = __uint_as_float(v[e]);
                __syncwarp();
                if (co0 + c < p.Cout) {
#pragma unroll 4
                    for (int rr = 0; rr < 32; ++rr) {
                        const int row = q * 32 + rr;
                        const int bx = box0 + 2 * j + (row >> 6);
                        if (bx < p.nboxes)
                            p.ws[(((size_t)z * p.nboxes + bx) * 64 + (row & 63)) * p.Cout + co0 + c + lane] = tb[rr * 33 + lane];
                    }
                }
            }
        }
    }
    fence_before();
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        tmem_dealloc<TCOLS>(tmem_base);
    }
}

// Sum the split partials in a fixed order and transpose [box = (tap, chunk)][64 ci][Cout] -> [Cout][tap][Cin].
// grid = (nboxes * 8, Cout / 64), block = 256 = 16 column quads (float4 = 4 output channels) x 16 split lanes: a CTA
// owns 8 input channels x 64 output channels of one box; lane z sums splits z, z + 16, ... (8 independent 16-byte
// loads in flight per step, at most ceil(ksplit / 16) dependent steps), the lane sums are combined in a fixed order.
template <typename OutT>
__global__ void __launch_bounds__(256) k_wx_reduce(const float* __restrict__ ws, OutT* __restrict__ out, int ksplit, int nboxes,
                                                   int cchunks, int Cout, int Cin, int ntaps, float scale) {
    __shared__ float part[16][8][68];
    const int bx = blockIdx.x >> 3, i0 = (blockIdx.x & 7) * 8, co0 = blockIdx.y * 64;
    const int tap = bx / cchunks, cc = bx - tap * cchunks;
    const int tq = threadIdx.x & 15, zl = threadIdx.x >> 4;
    float4 s[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    const size_t zstride = (size_t)nboxes * 64 * Cout;
    const float* src = ws + ((size_t)bx * 64 + i0) * Cout + co0 + tq * 4;
    for (int zz = zl; zz < ksplit; zz += 16) {
        const float* q = src + (size_t)zz * zstride;
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const float4 v = __ldcg(reinterpret_cast<const float4*>(q + (size_t)i * Cout));
            s[i].x += v.x; s[i].y += v.y; s[i].z += v.z; s[i].w += v.w;
        }
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) *reinterpret_cast<float4*>(&part[zl][i][tq * 4]) = s[i];
    __syncthreads();
    for (int idx = threadIdx.x; idx < 512; idx += 256) {
        const int i = idx & 7, o = idx >> 3;
        const int ci = cc * 64 + i0 + i;
        if (ci < Cin) {
            float v = 0.0f;
#pragma unroll
            for (int z = 0; z < 16; ++z) v += part[z][i][o];
            v *= scale;
            OutT* dst = out + ((size_t)(co0 + o) * ntaps + tap) * Cin + ci;
            if constexpr (sizeof(OutT) == 4) *dst = v; else *dst = __float2bfloat16_rn(v);
        }
    }
}

// Combine the per-CTA statistics of k_fx into batch-norm mean / inverse std (+ running statistics), one warp per
// channel, fixed order, double accumulation (<= 148 partials per channel).
__global__ void __launch_bounds__(256) k_stats_finalize(const float* __restrict__ part, int nparts, int C, double count, float eps,
                                                        float momentum, float* __restrict__ mean, float* __restrict__ invstd,
                                                        float* __restrict__ running_mean, float* __restrict__ running_var,
                                                        long long* num_batches) {
    const int c = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
    if (num_batches && blockIdx.x == 0 && threadIdx.x == 0) *num_batches += 1;
    if (c >= C) return;
    double s = 0.0, q = 0.0;
    for (int b = lane; b < nparts; b += 32) {
        s += (double)part[(size_t)b * 2 * C + c];
        q += (double)part[(size_t)b * 2 * C + C + c];
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        s += __shfl_xor_sync(0xffffffffu, s, o);
        q += __shfl_xor_sync(0xffffffffu, q, o);
    }
    if (lane == 0) {
        const double mu = s / count;
        double var = q / count - mu * mu;
        if (var < 0.0) var = 0.0;
        mean[c] = (float)mu;
        invstd[c] = (float)(1.0 / sqrt(var + (double)eps));
        if (running_mean) {
            const double unb = count > 1.0 ? var * count / (count - 1.0) : var;
            running_mean[c] = (1.0f - momentum) * running_mean[c] + momentum * (float)mu;
            running_var[c] = (1.0f - momentum) * running_var[c] + momentum * (float)unb;
        }
    }
}

}  // namespace convx
}  // namespace lwc

using namespace lwc::convx;

// ---------------------------------------------------------------------------------------------- host side
namespace {

int g_sm_count = 0;
int sm_count() {
    if (!g_sm_count) {
        int dev = 0;
        cudaGetDevice(&dev);
        cudaDeviceGetAttribute(&g_sm_count, cudaDevAttrMultiProcessorCount, dev);
        if (g_sm_count <= 0) g_sm_count = 148;
    }
    return g_sm_count;
}

void ensure_context() {
    CUcontext ctx = nullptr;
    if (cuCtxGetCurrent(&ctx) != CUDA_SUCCESS || !ctx) {
        int dev = 0;
        cudaGetDevice(&dev);
        CUcontext prim = nullptr;
        if (cuDevicePrimaryCtxRetain(&prim, dev) == CUDA_SUCCESS) cuCtxSetCurrent(prim);
    }
}

// bf16 tensor map, 128-byte swizzle, zero fill out of bounds; estr = traversal strides (NULL = all 1).
int encode(CUtensorMap* map, const void* ptr, int rank, const cuuint64_t* gdim, const cuuint64_t* gstride_bytes,
           const cuuint32_t* box, const cuuint32_t* estr_in) {
    cuuint32_t estr[5] = {1, 1, 1, 1, 1};
    if (estr_in) for (int i = 0; i < rank; ++i) estr[i] = estr_in[i];
    CUresult r = cuTensorMapEncodeTiled(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, rank, const_cast<void*>(ptr), gdim, gstride_bytes,
                                        box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                                        CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS && getenv("LWC_DEBUG"))
        fprintf(stderr, "lwc_convx: cuTensorMapEncodeTiled rank %d -> %d (dims %llu %llu %llu %llu box %u %u %u %u)\n", rank, (int)r,
                (unsigned long long)gdim[0], (unsigned long long)gdim[1], rank > 2 ? (unsigned long long)gdim[2] : 0ULL,
                rank > 3 ? (unsigned long long)gdim[3] : 0ULL, box[0], box[1], rank > 2 ? box[2] : 0, rank > 3 ? box[3] : 0);
    return r == CUDA_SUCCESS ? 0 : LWC_EINVAL;
}

// Channels-last activation [N][H][W][C] (pixel pitch `cpitch` elements) as a 4-D map (C, W, H, N) whose boxes
// are 64 channels x (bw, bh, bn) traversed pixels with traversal stride (s, s) over W and H.
int encode_act(CUtensorMap* map, const void* ptr, int C, int cpitch, int W, int H, int N, int bw, int bh, int bn, int s) {
    cuuint64_t gd[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
    cuuint64_t gs[3] = {(cuuint64_t)cpitch * 2, (cuuint64_t)W * cpitch * 2, (cuuint64_t)H * W * cpitch * 2};
    cuuint32_t bx[4] = {64, (cuuint32_t)(bw * s), (cuuint32_t)(bh * s), (cuuint32_t)bn};
    cuuint32_t es[4] = {1, (cuuint32_t)s, (cuuint32_t)s, 1};
    if (bx[1] > 256 || bx[2] > 256 || bx[3] > 256) return LWC_EINVAL;
    return encode(map, ptr, 4, gd, gs, bx, es);
}

// Pixel patch (Wb, Hb, Nb) for an output extent (OW, OH, ON): rows = Wb * Hb * Nb <= max_rows, and, when `mult` > 1,
// a multiple of `mult` (the weight gradient's K steps).  Picks the best coverage, then the most rows, then the widest.
struct Patch { int Wb, Hb, Nb, rows, tiles_w, tiles_h, tiles_n; };
Patch pick_patch(int OW, int OH, int ON, int max_rows, int mult, int max_w) {
    Patch best = {1, 1, 1, 1, OW, OH, ON};
    double best_score = -1.0;
    for (int wb = 1; wb <= OW && wb <= max_rows && wb <= max_w; ++wb) {
        for (int hb = 1; hb <= OH && wb * hb <= max_rows && hb <= max_w; ++hb) {
            int nmax = max_rows / (wb * hb);
            if (nmax > 256) nmax = 256;
            for (int nb = 1; nb <= nmax && nb <= ON; ++nb) {
                const int rows = wb * hb * nb;
                if (mult > 1 && rows % mult) continue;
                const int tw = (OW + wb - 1) / wb, th = (OH + hb - 1) / hb, tn = (ON + nb - 1) / nb;
                const double eff = (double)OW * OH * ON / ((double)tw * wb * th * hb * tn * nb);
                const double score = eff * 1e6 + rows * 10.0 + wb * 0.01 + hb * 0.0001;
                if (score > best_score) { best_score = score; best = {wb, hb, nb, rows, tw, th, tn}; }
            }
        }
    }
    return best;
}

struct FxGeom {
    // input activation
    const void* x; int C, cpitch, W, H, N;       // channels visible to the map, pixel pitch, extent
    int window;                                  // 1: overlapping-window map (stem): W axis pitch = `wpitch` elements,
    int wpitch;                                  //    `cpitch` is then the ROW pitch in elements
    // weights: K-major [rows = Cout][Ktot] (b_mn = 0) or MN-major [Ktot][Cout] (b_mn = 1)
    const void* w; int Ktot; int b_mn; int b_ncols;   // b_ncols: columns of the MN-major matrix (0 = Cout)
    // output [ON][OH][OW][Cout] seen through (base, pitches): element strides in elements
    void* y; int OW, OH, ON, Cout; long long y_wpitch, y_hpitch, y_npitch;
    int stride;                                  // input traversal stride
    int ntaps, kchunks;
    short dw[MAXT], dh[MAXT]; int a_c0[MAXT], b_k0[MAXT], b_n0[MAXT];
    const float* bias; int relu; float* stats; int bn_hint;
    // batch-norm backward sums (see FxParams::bnb): x = the batch norm's input, same shape and layout as the output
    const void* bnx; const float *bn_gamma, *bn_beta, *bn_mean, *bn_invstd; int bnb_relu;
    // residual fork (see FxParams::add): addend has the output's shape and layout, mask is [pixels][Cout / 8] bytes or NULL
    const void* addend; const void* mask;
};

template <int BN, bool B_MN, bool FORK, bool REGSUM, bool TWO = false>
int launch_fx(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& tc_, const CUtensorMap& tx, const CUtensorMap& td,
              const FxParams& p, int smem, int grid, cudaStream_t st) {
    static bool attr = false;
    if (!attr) {
        LWC_CUDA_CHECK(cudaFuncSetAttribute(k_fx<BN, B_MN, FORK, REGSUM, TWO>, cudaFuncAttributeMaxDynamicSharedMemorySize, TWO ? SMEM_BUDGET / 2 - 1024 : SMEM_BUDGET));
        attr = true;
    }
    k_fx<BN, B_MN, FORK, REGSUM, TWO><<<grid, THREADS, smem, st>>>(ta, tb, tc_, tx, td, p);
    return 0;
}

int pick_bn(int Cout, int hint) {
    if ((hint == 64 || hint == 128 || hint == 256) && Cout % hint == 0) return hint;
    if (Cout % 256 == 0) return 256;
    if (Cout % 128 == 0) return 128;
    return 64;
}

// Halo mode is asked for with bn_hint = 1000 + BN (ops/conv.py times it against the per-tap variants).
bool halo_ok(const FxGeom& g, int* span_w, int* span_h, int* dw0, int* dh0) {
    if (g.kchunks != 1 || g.stride != 1 || g.ntaps < 2 || g.addend) return false;
    int lo_w = g.dw[0], hi_w = g.dw[0], lo_h = g.dh[0], hi_h = g.dh[0];
    for (int t = 1; t < g.ntaps; ++t) {
        if (g.a_c0[t] != g.a_c0[0]) return false;
        lo_w = g.dw[t] < lo_w ? g.dw[t] : lo_w; hi_w = g.dw[t] > hi_w ? g.dw[t] : hi_w;
        lo_h = g.dh[t] < lo_h ? g.dh[t] : lo_h; hi_h = g.dh[t] > hi_h ? g.dh[t] : hi_h;
    }
    *span_w = hi_w - lo_w + 1; *span_h = hi_h - lo_h + 1; *dw0 = lo_w; *dh0 = lo_h;
    return *span_w <= 5 && *span_h <= 5;
}

int run_fx(const FxGeom& g, cudaStream_t st, int* grid_out) {
    if (g.Cout % 64 || g.ntaps < 1 || g.ntaps > MAXT || g.kchunks < 1) return LWC_EINVAL;
    ensure_context();
    // bn_hint = BN + 1000 (halo variant) + 2000 (two CTAs per SM)
    const bool want_halo = (g.bn_hint / 1000) & 1;
    const bool two = (g.bn_hint / 1000) & 2;
    int BN = pick_bn(g.Cout, g.bn_hint % 1000);
    if ((g.bnx || g.addend) && BN == 256) BN = 128;   // the auxiliary rings take the shared memory of a 256-wide operand stage
    if (g.bnx && !g.stats) return LWC_EINVAL;
    if (g.mask && !g.addend) return LWC_EINVAL;
    FxParams p;
    memset(&p, 0, sizeof(p));
    int span_w = 1, span_h = 1, dw0 = 0, dh0 = 0;
    const bool halo = want_halo && halo_ok(g, &span_w, &span_h, &dw0, &dh0);
    if (want_halo && !halo) return LWC_EINVAL;
    Patch pt;
    if (halo) {
        // accumulator rows = 8 box rows x 16 box columns; the last span_w - 1 columns of every row are overhang
        pt.Wb = HALO_PW - (span_w - 1); pt.Hb = BM / HALO_PW; pt.Nb = 1; pt.rows = pt.Wb * pt.Hb;
        pt.tiles_w = (g.OW + pt.Wb - 1) / pt.Wb; pt.tiles_h = (g.OH + pt.Hb - 1) / pt.Hb; pt.tiles_n = g.ON;
        if (BN == 256) BN = 128;
    } else {
        pt = pick_patch(g.OW, g.OH, g.ON, BM, 1, 256 / g.stride);
    }
    p.N = g.Cout; p.ntaps = g.ntaps; p.kchunks = g.kchunks;
    p.Wb = pt.Wb; p.Hb = pt.Hb; p.Nb = pt.Nb; p.tiles_w = pt.tiles_w; p.tiles_h = pt.tiles_h;
    p.sw = g.stride; p.sh = g.stride; p.OW = g.OW; p.OH = g.OH; p.ON = g.ON;
    p.tiles_m = pt.tiles_w * pt.tiles_h * pt.tiles_n; p.tiles_n = g.Cout / BN; p.rows_per_tile = pt.rows;
    p.relu = g.relu; p.bias = g.bias; p.stats = g.stats;
    for (int t = 0; t < g.ntaps; ++t) { p.dw[t] = g.dw[t]; p.dh[t] = g.dh[t]; p.a_c0[t] = g.a_c0[t]; p.b_k0[t] = g.b_k0[t]; p.b_n0[t] = g.b_n0[t]; }
    // shared-memory plan: [operand ring | resident weights + halo ring][staging ring][x ring][addend ring][barriers]
    const int rings = 1 + (g.bnx ? 1 : 0) + (g.addend ? 1 : 0);
    const bool aux = rings > 1;
    const bool regsum = g.stats != nullptr;  // the variant with per-channel sums in registers
    const int fixed = 256;
    if (two && (BN == 256 || (g.addend && g.bnx) || (BN == 128 && (g.bnx || g.addend)))) return LWC_EINVAL;
    const int budget = two ? SMEM_BUDGET / 2 - 1024 : SMEM_BUDGET;
    int ncs = two ? 2 : 4, stages = 0, operand_bytes = 0;
    const int halo_rows = BM / HALO_PW + span_h - 1;
    if (halo) {
        const int w_bytes = g.ntaps * BN * BK * 2;
        // the shifted reads of the last tap run (span_h - 1) * 16 + span_w - 1 rows past the 128th row of the box
        const int a_stage = ((BM + (span_h - 1) * HALO_PW + span_w) * 128 + 1023) / 1024 * 1024;
        if (aux) ncs = 3;
        stages = (budget - fixed - w_bytes - rings * ncs * SUB_BYTES) / a_stage;
        if (stages < 2 && !aux) { ncs = 2; stages = (budget - fixed - w_bytes - rings * ncs * SUB_BYTES) / a_stage; }
        if (stages > 8) stages = 8;
        if (stages < 2) return LWC_ETOOBIG;
        p.halo = 1; p.halo_dw0 = dw0; p.halo_dh0 = dh0; p.a_off = w_bytes; p.a_stage = a_stage;
        p.halo_bytes = HALO_PW * halo_rows * 128;
        const char* bo = getenv("LWC_HALO_BO");
        p.halo_bo = bo ? atoi(bo) : 0;
        operand_bytes = w_bytes + stages * a_stage;
    } else {
        const int stage_bytes = BM * BK * 2 + BN * BK * 2;
        stages = (budget - fixed - rings * ncs * SUB_BYTES) / stage_bytes;
        if (stages < 3) {
            ncs = (aux && !two) ? 3 : 2;         // an auxiliary prefetch two chunks ahead needs >= 3 slots (one ahead: 2)
            stages = (budget - fixed - rings * ncs * SUB_BYTES) / stage_bytes;
        }
        if (stages > 8) stages = 8;
        if (stages < 2) return LWC_ETOOBIG;
        operand_bytes = stages * stage_bytes;
    }
    p.stages = stages; p.ncs = ncs;
    p.c_off = operand_bytes;
    p.x_off = p.c_off + ncs * SUB_BYTES;
    p.add_off = p.x_off + (g.bnx ? ncs * SUB_BYTES : 0);
    p.stat_off = p.add_off + (g.addend ? ncs * SUB_BYTES : 0);
    p.bar_off = p.stat_off;
    p.bnb = g.bnx ? 1 : 0; p.bnb_relu = g.bnb_relu;
    p.bn_gamma = g.bn_gamma; p.bn_beta = g.bn_beta; p.bn_mean = g.bn_mean; p.bn_invstd = g.bn_invstd;
    p.add = g.addend ? 1 : 0; p.mask = static_cast<const uint8_t*>(g.mask);
    const int smem = p.bar_off + 256;

    CUtensorMap ta, tb, tc_;
    int rc;
    if (g.window) {
        // overlapping windows: "pixel" w of the map starts wpitch elements after pixel w - 1 and spans 64 elements
        cuuint64_t gd[4] = {64, (cuuint64_t)g.W, (cuuint64_t)g.H, (cuuint64_t)g.N};
        cuuint64_t gs[3] = {(cuuint64_t)g.wpitch * 2, (cuuint64_t)g.cpitch * 2, (cuuint64_t)g.H * g.cpitch * 2};
        cuuint32_t bx[4] = {64, (cuuint32_t)pt.Wb, (cuuint32_t)pt.Hb, (cuuint32_t)pt.Nb};
        if (halo) { bx[1] = HALO_PW; bx[2] = (cuuint32_t)halo_rows; bx[3] = 1; }   // the stem's row taps: 16 x (8 + 3) windows
        rc = encode(&ta, g.x, 4, gd, gs, bx, nullptr);
    } else if (halo) {
        rc = encode_act(&ta, g.x, g.C, g.cpitch, g.W, g.H, g.N, HALO_PW, halo_rows, 1, 1);
    } else {
        rc = encode_act(&ta, g.x, g.C, g.cpitch, g.W, g.H, g.N, pt.Wb, pt.Hb, pt.Nb, g.stride);
    }
    if (rc) return rc;
    if (g.b_mn) {
        const cuuint64_t ncols = g.b_ncols ? g.b_ncols : g.Cout;
        cuuint64_t bd[2] = {ncols, (cuuint64_t)g.Ktot}, bs[1] = {ncols * 2};
        cuuint32_t bb[2] = {64, BK};
        rc = encode(&tb, g.w, 2, bd, bs, bb, nullptr);
    } else {
        cuuint64_t bd[2] = {(cuuint64_t)g.Ktot, (cuuint64_t)g.Cout}, bs[1] = {(cuuint64_t)g.Ktot * 2};
        cuuint32_t bb[2] = {BK, (cuuint32_t)BN};
        rc = encode(&tb, g.w, 2, bd, bs, bb, nullptr);
    }
    if (rc) return rc;
    // output-shaped maps: the output itself, the batch norm's input, the addend
    auto encode_out = [&](CUtensorMap* m, const void* ptr) {
        cuuint64_t cd[4] = {(cuuint64_t)g.Cout, (cuuint64_t)g.OW, (cuuint64_t)g.OH, (cuuint64_t)g.ON};
        cuuint64_t cs[3] = {(cuuint64_t)g.y_wpitch * 2, (cuuint64_t)g.y_hpitch * 2, (cuuint64_t)g.y_npitch * 2};
        cuuint32_t cb[4] = {64, (cuuint32_t)pt.Wb, (cuuint32_t)pt.Hb, (cuuint32_t)pt.Nb};
        return encode(m, ptr, 4, cd, cs, cb, nullptr);
    };
    rc = encode_out(&tc_, g.y);
    if (rc) return rc;
    CUtensorMap tx = tc_, td = tc_;
    if (g.bnx) {
        if (!g.bn_mean || !g.bn_invstd) return LWC_EINVAL;
        rc = encode_out(&tx, g.bnx);
        if (rc) return rc;
    }
    if (g.addend) {
        rc = encode_out(&td, g.addend);
        if (rc) return rc;
    }
    int grid = p.tiles_m * p.tiles_n;
    const int max_grid = sm_count() * (two ? 2 : 1);
    if (grid > max_grid) grid = max_grid;
    // with statistics (and with resident weights) a CTA must stay on one column of output-channel tiles
    // (tile = mt * tiles_n + nt, stride = grid)
    const bool pinned = g.stats || halo;
    if (pinned && grid % p.tiles_n) grid = grid > p.tiles_n ? grid / p.tiles_n * p.tiles_n : p.tiles_n;
    if (pinned && p.tiles_n > sm_count()) return LWC_ETOOBIG;
    if (grid_out) *grid_out = grid;
#define LWC_FX(BN_, MN_, FK_, RS_) launch_fx<BN_, MN_, FK_, RS_>(ta, tb, tc_, tx, td, p, smem, grid, st)
#define LWC_FX2(BN_, MN_, RS_) launch_fx<BN_, MN_, false, RS_, true>(ta, tb, tc_, tx, td, p, smem, grid, st)
    if (two && g.addend) {
        if (!g.b_mn) return LWC_EINVAL;
        rc = launch_fx<64, true, true, true, true>(ta, tb, tc_, tx, td, p, smem, grid, st);
    } else if (two) {
        if (BN == 128 && regsum) rc = g.b_mn ? LWC_EINVAL : LWC_FX2(128, false, true);
        else if (BN == 128) rc = g.b_mn ? LWC_FX2(128, true, false) : LWC_FX2(128, false, false);
        else if (regsum) rc = g.b_mn ? LWC_FX2(64, true, true) : LWC_FX2(64, false, true);
        else rc = g.b_mn ? LWC_FX2(64, true, false) : LWC_FX2(64, false, false);
    } else if (g.addend) {
        // the fork variant exists for the MN-major weight (data gradients) and 64- / 128-wide tiles only
        if (!g.b_mn) return LWC_EINVAL;
        rc = BN == 128 ? LWC_FX(128, true, true, true) : LWC_FX(64, true, true, true);
    } else if (regsum) {
        if (BN == 256) rc = g.b_mn ? LWC_FX(256, true, false, true) : LWC_FX(256, false, false, true);
        else if (BN == 128) rc = g.b_mn ? LWC_FX(128, true, false, true) : LWC_FX(128, false, false, true);
        else rc = g.b_mn ? LWC_FX(64, true, false, true) : LWC_FX(64, false, false, true);
    } else {
        if (BN == 256) rc = g.b_mn ? LWC_FX(256, true, false, false) : LWC_FX(256, false, false, false);
        else if (BN == 128) rc = g.b_mn ? LWC_FX(128, true, false, false) : LWC_FX(128, false, false, false);
        else rc = g.b_mn ? LWC_FX(64, true, false, false) : LWC_FX(64, false, false, false);
    }
#undef LWC_FX
#undef LWC_FX2
    if (rc) return rc;
    return (int)cudaGetLastError();
}

}  // namespace

// Number of per-CTA statistics partials a forward call may write (callers size the buffer [this][2][Cout]).
extern "C" int lwc_convx_max_partials(void) { return 2 * sm_count(); }   // two CTAs per SM in the densest variant

// Forward convolution, square kernel k (k * k <= 16), stride 1 or 2, symmetric padding `pad`.
//   x    : [N][H][W][Cx] bf16 channels-last; Cx % 8 == 0 (3-channel inputs are stored with 8 channels)
//   wt   : [Cout][k*k][Kc] bf16 with Kc = round_up(Cx, 64) (channels past Cx zero); Cout % 64 == 0
//   y    : [N][OH][OW][Cout] bf16, OH = (H + 2 pad - k) / stride + 1
//   bias : fp32 [Cout] or NULL; relu in {0, 1}
//   stats: [lwc_convx_max_partials()][2][Cout] fp32 or NULL; *nparts_out receives the number of partials written
//   bn_hint: 0 = automatic output-tile width, else 64 / 128 / 256
extern "C" int lwc_convx_forward(const void* x, const void* wt, void* y, int N, int H, int W, int Cx, int Cout, int k, int stride,
                                 int pad, const float* bias, int relu, float* stats, int* nparts_out, int bn_hint, void* stream) {
    if (!x || !wt || !y || N <= 0 || H <= 0 || W <= 0 || (Cx % 8) || (Cout % 64) || k < 1 || k * k > MAXT) return LWC_EINVAL;
    if (stride != 1 && stride != 2) return LWC_EINVAL;
    FxGeom g;
    memset(&g, 0, sizeof(g));
    const int OH = (H + 2 * pad - k) / stride + 1, OW = (W + 2 * pad - k) / stride + 1;
    if (OH <= 0 || OW <= 0) return LWC_EINVAL;
    const int Kc = (Cx + 63) / 64 * 64;
    g.x = x; g.C = Cx; g.cpitch = Cx; g.W = W; g.H = H; g.N = N;
    g.w = wt; g.Ktot = k * k * Kc; g.b_mn = 0;
    g.y = y; g.OW = OW; g.OH = OH; g.ON = N; g.Cout = Cout;
    g.y_wpitch = Cout; g.y_hpitch = (long long)OW * Cout; g.y_npitch = (long long)OH * OW * Cout;
    g.stride = stride; g.ntaps = k * k; g.kchunks = Kc / 64;
    g.bias = bias; g.relu = relu; g.stats = stats; g.bn_hint = bn_hint;
    if (k == 1 && stride == 1 && pad == 0) {
        // flatten the pixels: any 128 consecutive rows form a tile
        const long long M = (long long)N * H * W;
        if (M > 0x7fffffffLL) return LWC_ETOOBIG;
        g.W = (int)M; g.H = 1; g.N = 1; g.OW = (int)M; g.OH = 1; g.ON = 1;
        g.y_hpitch = M * Cout; g.y_npitch = M * Cout;
    }
    for (int a = 0; a < k; ++a)
        for (int b = 0; b < k; ++b) {
            const int t = a * k + b;
            g.dh[t] = (short)(a - pad); g.dw[t] = (short)(b - pad); g.a_c0[t] = 0; g.b_k0[t] = t * Kc;
        }
    return run_fx(g, (cudaStream_t)stream, nparts_out);
}

// 1x1 stride-1 data gradient with the forward weight as stored: dX[p][ci] = sum_co dY[p][co] * W[co][ci]
// (W is the MN-major B operand: no transposed copy).  Cin, Cout multiples of 64.
extern "C" int lwc_convx_dgrad1x1(const void* dy, const void* w, void* dx, long long M, int Cout, int Cin, int bn_hint, void* stream) {
    if (!dy || !w || !dx || M <= 0 || M > 0x7fffffffLL || (Cout % 64) || (Cin % 64)) return LWC_EINVAL;
    FxGeom g;
    memset(&g, 0, sizeof(g));
    g.x = dy; g.C = Cout; g.cpitch = Cout; g.W = (int)M; g.H = 1; g.N = 1;
    g.w = w; g.Ktot = Cout; g.b_mn = 1;
    g.y = dx; g.OW = (int)M; g.OH = 1; g.ON = 1; g.Cout = Cin;
    g.y_wpitch = Cin; g.y_hpitch = M * Cin; g.y_npitch = M * Cin;
    g.stride = 1; g.ntaps = 1; g.kchunks = Cout / 64; g.bn_hint = bn_hint;
    return run_fx(g, (cudaStream_t)stream, nullptr);
}

// Data gradient of a k x k / stride s / pad convolution:  dX[h][w] = sum_{(a,b)} W[a][b]^T dY[(h + pad - a) / s][(w + pad - b) / s]
// over the taps whose divisions are exact.  One launch per output parity class (s * s of them); classes without a tap
// are zero (dX is cleared up front when there is one, e.g. 1x1 / stride 2).  The forward weight is read as stored: for a
// tap it is a [Cout][Cin] matrix with Cin contiguous, i.e. the MN-major B operand of this GEMM -- no transposed copy.
//   dy : [N][OH][OW][Cout];  wt : [Cout][k*k][Cin] bf16 (the channels-last forward weight)
//   dx : [N][H][W][Cin];  Cin, Cout multiples of 64
extern "C" int lwc_convx_dgrad(const void* dy, const void* wt, void* dx, int N, int H, int W, int Cin, int Cout, int k, int stride,
                               int pad, int bn_hint, void* stream) {
    if (!dy || !wt || !dx || N <= 0 || (Cin % 64) || (Cout % 64) || k < 1 || k * k > MAXT) return LWC_EINVAL;
    if (stride != 1 && stride != 2) return LWC_EINVAL;
    const int OH = (H + 2 * pad - k) / stride + 1, OW = (W + 2 * pad - k) / stride + 1;
    cudaStream_t st = (cudaStream_t)stream;
    bool any_empty = false;
    for (int ph = 0; ph < stride; ++ph)
        for (int pw = 0; pw < stride; ++pw) {
            int n = 0;
            for (int a = 0; a < k; ++a)
                for (int b = 0; b < k; ++b)
                    if ((ph + pad - a) % stride == 0 && (pw + pad - b) % stride == 0) ++n;
            if (!n) any_empty = true;
        }
    if (any_empty) LWC_CUDA_CHECK(cudaMemsetAsync(dx, 0, (size_t)N * H * W * Cin * 2, st));
    for (int ph = 0; ph < stride; ++ph)
        for (int pw = 0; pw < stride; ++pw) {
            FxGeom g;
            memset(&g, 0, sizeof(g));
            int n = 0;
            for (int a = 0; a < k; ++a)
                for (int b = 0; b < k; ++b) {
                    const int nh = ph + pad - a, nw = pw + pad - b;
                    // C++ '%' and '/' truncate toward zero: exactness and the quotient are still right for negative nh
                    if (nh % stride || nw % stride) continue;
                    g.dh[n] = (short)(nh / stride); g.dw[n] = (short)(nw / stride); g.a_c0[n] = 0; g.b_k0[n] = 0;
                    g.b_n0[n] = (a * k + b) * Cin;
                    ++n;
                }
            if (!n) continue;
            const int CH = (H - ph + stride - 1) / stride, CW = (W - pw + stride - 1) / stride;   // pixels of this class
            if (CH <= 0 || CW <= 0) continue;
            g.x = dy; g.C = Cout; g.cpitch = Cout; g.W = OW; g.H = OH; g.N = N;
            g.w = wt; g.Ktot = Cout; g.b_mn = 1; g.b_ncols = k * k * Cin;
            g.y = (char*)dx + ((size_t)ph * W + pw) * Cin * 2; g.OW = CW; g.OH = CH; g.ON = N; g.Cout = Cin;
            g.y_wpitch = (long long)stride * Cin; g.y_hpitch = (long long)stride * W * Cin; g.y_npitch = (long long)H * W * Cin;
            g.stride = 1; g.ntaps = n; g.kchunks = Cout / 64; g.bn_hint = bn_hint;
            const int rc = run_fx(g, st, nullptr);
            if (rc) return rc;
        }
    return 0;
}

// Stride-1 data gradient whose output dz is the gradient of a batch-norm (+ ReLU) output: the same launch also sums
// S1 = sum d and S2 = sum d * x per channel (d = dz * [bn(x) > 0] with relu, else dz) into per-CTA partials
// [lwc_convx_max_partials()][2][Cin] -- the batch norm's backward then needs no reduction pass over dz and x
// (lwc_bn_backward_from_partials).  k = 1 / pad 0 uses the forward weight as stored, like lwc_convx_dgrad1x1.
//   x_bn : the batch norm's input [N][H][W][Cin] bf16; gamma / beta nullable; mean / invstd its batch statistics
extern "C" int lwc_convx_dgrad_bnstats(const void* dy, const void* wt, void* dx, int N, int H, int W, int Cin, int Cout, int k, int pad,
                                       const void* x_bn, const float* gamma, const float* beta, const float* mean, const float* invstd,
                                       int relu, float* partials, int* nparts_out, int bn_hint, void* stream) {
    if (!dy || !wt || !dx || !x_bn || !mean || !invstd || !partials || N <= 0 || (Cin % 64) || (Cout % 64) || k < 1 || k * k > MAXT)
        return LWC_EINVAL;
    FxGeom g;
    memset(&g, 0, sizeof(g));
    const int OH = H + 2 * pad - k + 1, OW = W + 2 * pad - k + 1;
    if (OH <= 0 || OW <= 0) return LWC_EINVAL;
    g.x = dy; g.C = Cout; g.cpitch = Cout; g.W = OW; g.H = OH; g.N = N;
    g.w = wt; g.Ktot = Cout; g.b_mn = 1; g.b_ncols = k * k * Cin;
    g.y = dx; g.OW = W; g.OH = H; g.ON = N; g.Cout = Cin;
    g.y_wpitch = Cin; g.y_hpitch = (long long)W * Cin; g.y_npitch = (long long)H * W * Cin;
    g.stride = 1; g.kchunks = Cout / 64; g.bn_hint = bn_hint;
    g.stats = partials; g.bnx = x_bn; g.bn_gamma = gamma; g.bn_beta = beta; g.bn_mean = mean; g.bn_invstd = invstd; g.bnb_relu = relu;
    if (k == 1 && pad == 0) {
        const long long M = (long long)N * H * W;
        if (M > 0x7fffffffLL) return LWC_ETOOBIG;
        g.W = (int)M; g.H = 1; g.N = 1; g.OW = (int)M; g.OH = 1; g.ON = 1;
        g.y_hpitch = M * Cin; g.y_npitch = M * Cin;
        g.ntaps = 1;
    } else {
        int n = 0;
        for (int a = 0; a < k; ++a)
            for (int b = 0; b < k; ++b) {
                g.dh[n] = (short)(pad - a); g.dw[n] = (short)(pad - b); g.b_n0[n] = (a * k + b) * Cin;
                ++n;
            }
        g.ntaps = n;
    }
    return run_fx(g, (cudaStream_t)stream, nparts_out);
}

// Residual fork: data gradient of the block's first 1x1 convolution with the skip connection's gradient added and the
// producing batch norm's ReLU mask applied in the epilogue,  dX = mask * (dY W + addend),  W the forward weight as stored
// ([Cout][Cin], MN-major B operand).  dX is then the gradient of that batch norm's (pre-masked) output AND of its
// residual input; with x_bn (that batch norm's input) the launch also sums S1 = sum dX, S2 = sum dX * x per channel
// into per-CTA partials for lwc_bn_backward_from_partials (relu = 0: the mask is already in dX).
//   dy [M][Cout], addend / dx / x_bn [M][Cin] bf16; mask [M][Cin / 8] bytes (bit j of byte b = channel 8 b + j) or NULL
extern "C" int lwc_convx_dgrad1x1_fork(const void* dy, const void* w, const void* addend, const void* mask, void* dx, long long M,
                                       int Cout, int Cin, const void* x_bn, const float* mean, const float* invstd, float* partials,
                                       int* nparts_out, int bn_hint, void* stream) {
    if (!dy || !w || !addend || !dx || M <= 0 || M > 0x7fffffffLL || (Cout % 64) || (Cin % 64)) return LWC_EINVAL;
    if (x_bn && (!mean || !invstd || !partials)) return LWC_EINVAL;
    FxGeom g;
    memset(&g, 0, sizeof(g));
    g.x = dy; g.C = Cout; g.cpitch = Cout; g.W = (int)M; g.H = 1; g.N = 1;
    g.w = w; g.Ktot = Cout; g.b_mn = 1;
    g.y = dx; g.OW = (int)M; g.OH = 1; g.ON = 1; g.Cout = Cin;
    g.y_wpitch = Cin; g.y_hpitch = M * Cin; g.y_npitch = M * Cin;
    g.stride = 1; g.ntaps = 1; g.kchunks = Cout / 64; g.bn_hint = bn_hint;
    g.addend = addend; g.mask = mask;
    if (x_bn) { g.stats = partials; g.bnx = x_bn; g.bn_mean = mean; g.bn_invstd = invstd; g.bnb_relu = 0; }
    return run_fx(g, (cudaStream_t)stream, nparts_out);
}

// The 7x7/2 ResNet stem on the 2x2 space-to-depth input (see lwc_normalize_u8 mode 1): a dense 4x4/1 convolution
// whose four kernel rows are the taps and whose 4 pixels x 16 channels kernel-row windows (128 bytes, overlapping,
// one pixel apart) are the K = 64 chunks.
//   xs : [N][OH+3][OW+3][16] bf16;  w4 : [Cout][4 rows][4 cols * 16 ch] bf16 (s2d_stem_weight, channels-last);  y : [N][OH][OW][Cout]
extern "C" int lwc_convx_stem_forward(const void* xs, const void* w4, void* y, int N, int OH, int OW, int Cout, float* stats,
                                      int* nparts_out, int bn_hint, void* stream) {
    if (!xs || !w4 || !y || N <= 0 || OH <= 0 || OW <= 0 || (Cout % 64)) return LWC_EINVAL;
    FxGeom g;
    memset(&g, 0, sizeof(g));
    g.x = xs; g.window = 1; g.wpitch = 16; g.cpitch = (OW + 3) * 16; g.W = OW; g.H = OH + 3; g.N = N; g.C = 64;
    g.w = w4; g.Ktot = 256; g.b_mn = 0;
    g.y = y; g.OW = OW; g.OH = OH; g.ON = N; g.Cout = Cout;
    g.y_wpitch = Cout; g.y_hpitch = (long long)OW * Cout; g.y_npitch = (long long)OH * OW * Cout;
    g.stride = 1; g.ntaps = 4; g.kchunks = 1; g.stats = stats; g.bn_hint = bn_hint;
    for (int a = 0; a < 4; ++a) { g.dh[a] = (short)a; g.dw[a] = 0; g.a_c0[a] = 0; g.b_k0[a] = a * 64; }
    return run_fx(g, (cudaStream_t)stream, nparts_out);
}

// Batch-norm statistics from the forward kernel's per-CTA partials.
extern "C" int lwc_convx_stats_finalize(const float* partials, int nparts, int C, long long count, float eps, float momentum,
                                        float* mean, float* invstd, float* running_mean, float* running_var,
                                        long long* num_batches_tracked, void* stream) {
    if (!partials || nparts <= 0 || C <= 0 || count <= 0 || !mean || !invstd) return LWC_EINVAL;
    k_stats_finalize<<<(C + 7) / 8, 256, 0, (cudaStream_t)stream>>>(partials, nparts, C, (double)count, eps, momentum, mean, invstd,
                                                                    running_mean, running_var, num_batches_tracked);
    return (int)cudaGetLastError();
}

namespace {

struct WxPlan { WxParams p; int BN; int smem; int ngroups; Patch pt; size_t ws_bytes; };

int plan_wx(int N, int H, int W, int Cx, int Cout, int k, int stride, int pad, int bn_hint, WxPlan* out) {
    if (N <= 0 || H <= 0 || W <= 0 || (Cx % 8) || (Cout % 64) || k < 1 || k * k > MAXT) return LWC_EINVAL;
    if (stride != 1 && stride != 2) return LWC_EINVAL;
    int OH = (H + 2 * pad - k) / stride + 1, OW = (W + 2 * pad - k) / stride + 1, ON = N;
    if (OH <= 0 || OW <= 0) return LWC_EINVAL;
    if (k == 1 && stride == 1 && pad == 0) {
        const long long M = (long long)N * H * W;
        if (M > 0x7fffffffLL) return LWC_ETOOBIG;
        OW = (int)M; OH = 1; ON = 1;
    }
    WxPlan& P = *out;
    memset(&P, 0, sizeof(P));
    WxParams& p = P.p;
    // hint = 1000 * (cap on accumulator slices per CTA, 0 = as many as tensor memory holds) + tile width
    const int nacc_cap = bn_hint / 1000;
    const int BN = pick_bn(Cout, bn_hint % 1000);
    p.Cout = Cout; p.cchunks = (Cx + 63) / 64; p.Cin_pad = p.cchunks * 64; p.ntaps = k * k; p.nboxes = p.ntaps * p.cchunks;
    const int nslices = (p.nboxes + 1) / 2;
    p.nacc = 512 / BN;
    if (nacc_cap > 0 && p.nacc > nacc_cap) p.nacc = nacc_cap;
    if (p.nacc > nslices) p.nacc = nslices;
    p.groups_per_co = (nslices + p.nacc - 1) / p.nacc;
    p.nacc = (nslices + p.groups_per_co - 1) / p.groups_per_co;          // balance the groups
    P.ngroups = p.groups_per_co * (Cout / BN);
    P.BN = BN;
    int stages = 0;
    Patch pt = {};
    for (int max_rows = 64; max_rows >= 16; max_rows /= 2) {
        pt = pick_patch(OW, OH, ON, max_rows, 16, 256 / stride);
        if (pt.rows % 16) continue;
        p.PK = pt.rows;
        p.stage_bytes = p.PK * 128 * (BN / 64 + 2 * p.nacc);
        stages = (SMEM_BUDGET - 2048) / p.stage_bytes;
        if (stages >= 3 || max_rows == 16) break;
    }
    if (pt.rows % 16 || stages < 2) return LWC_ETOOBIG;
    if (stages > 8) stages = 8;
    p.stages = stages;
    p.bar_off = stages * p.stage_bytes;
    P.smem = p.bar_off + 256 + 1024;
    p.Wb = pt.Wb; p.Hb = pt.Hb; p.Nb = pt.Nb; p.tiles_w = pt.tiles_w; p.tiles_h = pt.tiles_h;
    p.npatches = pt.tiles_w * pt.tiles_h * pt.tiles_n;
    p.sw = stride; p.sh = stride;
    int ks = sm_count() / P.ngroups;
    if (ks < 1) ks = 1;
    if (ks > p.npatches) ks = p.npatches;
    p.ksplit = ks;
    for (int a = 0; a < k; ++a)
        for (int b = 0; b < k; ++b) { p.dh[a * k + b] = (short)(a - pad); p.dw[a * k + b] = (short)(b - pad); }
    P.pt = pt;
    P.ws_bytes = (size_t)ks * p.nboxes * 64 * Cout * sizeof(float);
    return 0;
}

template <int BN>
int launch_wx(const CUtensorMap& tx, const CUtensorMap& td, const WxPlan& P, cudaStream_t st) {
    static bool attr = false;
    if (!attr) {
        LWC_CUDA_CHECK(cudaFuncSetAttribute(k_wx<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BUDGET));
        attr = true;
    }
    k_wx<BN><<<dim3(P.ngroups, P.p.ksplit), WX_THREADS, P.smem, st>>>(tx, td, P.p);
    return 0;
}

}  // namespace

extern "C" size_t lwc_convx_wgrad_workspace_bytes(int N, int H, int W, int Cx, int Cout, int k, int stride, int pad, int bn_hint) {
    WxPlan P;
    if (plan_wx(N, H, W, Cx, Cout, k, stride, pad, bn_hint, &P)) return 0;
    return P.ws_bytes;
}

// Weight gradient  dW[co][tap][ci] = scale * sum_p dY[p][co] * X[p * stride + tap - pad][ci].
//   x  : [N][H][W][Cx] bf16 (Cx % 8 == 0);  dy : [N][OH][OW][Cout] bf16
//   dw : [Cout][k*k][Cin_out] in out_dtype (LWC_BF16 / LWC_F32); Cin_out <= Cx (3 for an 8-channel-stored RGB input)
//   workspace : lwc_convx_wgrad_workspace_bytes(...) bytes of fp32 partials
extern "C" int lwc_convx_wgrad(const void* x, const void* dy, void* dw, float* workspace, int N, int H, int W, int Cx, int Cout,
                               int Cin_out, int k, int stride, int pad, int out_dtype, float scale, int bn_hint, void* stream) {
    if (!x || !dy || !dw || !workspace || Cin_out <= 0 || Cin_out > Cx) return LWC_EINVAL;
    if (out_dtype != LWC_F32 && out_dtype != LWC_BF16) return LWC_EINVAL;
    WxPlan P;
    int rc = plan_wx(N, H, W, Cx, Cout, k, stride, pad, bn_hint, &P);
    if (rc) return rc;
    ensure_context();
    cudaStream_t st = (cudaStream_t)stream;
    P.p.ws = workspace;
    int OH = (H + 2 * pad - k) / stride + 1, OW = (W + 2 * pad - k) / stride + 1, ON = N;
    int XW = W, XH = H, XN = N;
    if (k == 1 && stride == 1 && pad == 0) {
        const long long M = (long long)N * H * W;
        OW = (int)M; OH = 1; ON = 1; XW = (int)M; XH = 1; XN = 1;
    }
    CUtensorMap tx, td;
    rc = encode_act(&tx, x, Cx, Cx, XW, XH, XN, P.pt.Wb, P.pt.Hb, P.pt.Nb, stride);
    if (rc) return rc;
    rc = encode_act(&td, dy, Cout, Cout, OW, OH, ON, P.pt.Wb, P.pt.Hb, P.pt.Nb, 1);
    if (rc) return rc;
    if (P.BN == 256) rc = launch_wx<256>(tx, td, P, st);
    else if (P.BN == 128) rc = launch_wx<128>(tx, td, P, st);
    else rc = launch_wx<64>(tx, td, P, st);
    if (rc) return rc;
    const dim3 rg(P.p.nboxes * 8, Cout / 64), rb(256);
    if (out_dtype == LWC_F32)
        k_wx_reduce<float><<<rg, rb, 0, st>>>(workspace, (float*)dw, P.p.ksplit, P.p.nboxes, P.p.cchunks, Cout, Cin_out, P.p.ntaps, scale);
    else
        k_wx_reduce<__nv_bfloat16><<<rg, rb, 0, st>>>(workspace, (__nv_bfloat16*)dw, P.p.ksplit, P.p.nboxes, P.p.cchunks, Cout, Cin_out,
                                                      P.p.ntaps, scale);
    return (int)cudaGetLastError();
}
