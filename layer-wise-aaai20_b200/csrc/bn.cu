// Fused batch normalisation for channels-last activations ([M = N*H*W, C] row-major).
//
// Training forward : k_bn_stats (per-block partial sum / sum-of-squares, fp32)
//                    k_bn_finalize (deterministic Chan combination of the partials -> mean, invstd,
//                                   running-stat update)
//                    k_bn_apply   y = relu?( (x-mean)*invstd*gamma + beta (+ residual) )
// Training backward: k_bn_bwd_reduce (sum dy', sum dy'*xhat with dy' = dy * [y > 0] when ReLU is fused)
//                    k_bn_bwd_finalize (dgamma, dbeta)
//                    k_bn_bwd_apply  dx = gamma*invstd*(dy' - mean(dy') - xhat*mean(dy'*xhat)), dres = dy'
// Eval forward     : k_bn_apply with the running statistics.
//
// The stock path needs a statistics kernel, a transform kernel, a ReLU kernel and (at residual
// joins) an add kernel forward, and four more backward, each streaming the activation again
// (reference ops: nn.BatchNorm2d + nn.ReLU(inplace) + `out += residual`,
//  CIFAR10/dawn.py:23-28, IMAGENET/training/resnet.py:40-56,:75-95).  Here the activation is read twice and
// written once forward, and dy/x(/y) are read twice and dx written once backward.
//
// Thread mapping: a thread owns 8 consecutive channels (one 16-byte bf16 vector); TPR = min(C/8, 256)
// threads span a row, 256/TPR rows are processed per pass, so every warp access is contiguous.
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include "common.cuh"

namespace lwc {
namespace bn {

constexpr int THREADS = 256;
constexpr int VEC = 8;

struct Params {
    const void* x;
    void* y;              // forward output / backward: the saved forward output (ReLU mask)
    const void* res;      // forward: residual added before ReLU (nullable)
    const void* dy;       // backward
    void* dx;             // backward
    void* dres;           // backward: gradient for the residual branch (nullable)
    const float* gamma;
    const float* beta;
    float* mean;
    float* invstd;
    float* running_mean;
    float* running_var;
    float* dgamma;
    float* dbeta;
    float* partial;       // [nblk][2][C]
    long long* num_batches;   // nn.BatchNorm2d.num_batches_tracked, bumped by the finalize kernel (nullable)
    unsigned char* mask;      // ReLU mask, one bit per element ([M][C/8] bytes): written by the forward apply when
                              // non-NULL, read by the backward kernels instead of the saved output y (1/16 of its bytes)
    long long M;
    int C;
    int nblk;
    int rows_per_blk;
    int relu;
    float eps;
    float momentum;
};

template <typename T>
__device__ __forceinline__ void ld8(const T* p, float v[8]);
template <>
__device__ __forceinline__ void ld8<__nv_bfloat16>(const __nv_bfloat16* p, float v[8]) {
    const uint4 t = *reinterpret_cast<const uint4*>(p);
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(&w[i]);
        v[2 * i] = __low2float(h);
        v[2 * i + 1] = __high2float(h);
    }
}
template <>
__device__ __forceinline__ void ld8<__half>(const __half* p, float v[8]) {
    const uint4 t = *reinterpret_cast<const uint4*>(p);
    const uint32_t w[4] = {t.x, t.y, t.z, t.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const __half2 h = *reinterpret_cast<const __half2*>(&w[i]);
        v[2 * i] = __low2float(h);
        v[2 * i + 1] = __high2float(h);
    }
}
template <>
__device__ __forceinline__ void ld8<float>(const float* p, float v[8]) {
    const float4 a = *reinterpret_cast<const float4*>(p);
    const float4 b = *reinterpret_cast<const float4*>(p + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
template <typename T>
__device__ __forceinline__ void st8(T* p, const float v[8]);
template <>
__device__ __forceinline__ void st8<__nv_bfloat16>(__nv_bfloat16* p, const float v[8]) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
        w[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
}
template <>
__device__ __forceinline__ void st8<__half>(__half* p, const float v[8]) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        __half2 h = __floats2half2_rn(v[2 * i], v[2 * i + 1]);
        w[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
}
template <>
__device__ __forceinline__ void st8<float>(float* p, const float v[8]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}

// Packed 8-element vector kept in its storage format while loads are in flight (4 registers for
// 16-bit types), unpacked to fp32 only when consumed: lets a thread keep 4 rows x several tensors
// outstanding without blowing the register budget.
template <typename T> struct Raw { uint4 v; };
template <> struct Raw<float> { float4 a, b; };
template <typename T>
__device__ __forceinline__ void ldraw(const T* p, Raw<T>& r) { r.v = *reinterpret_cast<const uint4*>(p); }
template <>
__device__ __forceinline__ void ldraw<float>(const float* p, Raw<float>& r) {
    r.a = *reinterpret_cast<const float4*>(p);
    r.b = *reinterpret_cast<const float4*>(p + 4);
}
__device__ __forceinline__ void unpack(const Raw<__nv_bfloat16>& r, float v[8]) {
    const uint32_t w[4] = {r.v.x, r.v.y, r.v.z, r.v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {     // bf16 -> fp32 is a 16-bit shift
        v[2 * i] = __uint_as_float(w[i] << 16);
        v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
}
__device__ __forceinline__ void unpack(const Raw<__half>& r, float v[8]) {
    const uint32_t w[4] = {r.v.x, r.v.y, r.v.z, r.v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
        const __half2 h = *reinterpret_cast<const __half2*>(&w[i]);
        v[2 * i] = __low2float(h);
        v[2 * i + 1] = __high2float(h);
    }
}
__device__ __forceinline__ void unpack(const Raw<float>& r, float v[8]) {
    v[0] = r.a.x; v[1] = r.a.y; v[2] = r.a.z; v[3] = r.a.w; v[4] = r.b.x; v[5] = r.b.y; v[6] = r.b.z; v[7] = r.b.w;
}
constexpr int ROWS = 4;      // rows in flight per thread

struct Map {
    int tpr, rpp, tx, ry, cg;      // threads per row, rows per pass, my column / row lane, channel group
    bool active;
    long long row0, row1;
};
__device__ __forceinline__ Map make_map(const Params& p) {
    Map m;
    const int ncg = p.C / VEC;
    m.tpr = ncg < THREADS ? ncg : THREADS;
    m.rpp = THREADS / m.tpr;
    m.tx = threadIdx.x % m.tpr;
    m.ry = threadIdx.x / m.tpr;
    m.cg = blockIdx.y * m.tpr + m.tx;
    m.active = m.ry < m.rpp && m.cg < ncg;
    m.row0 = (long long)blockIdx.x * p.rows_per_blk;
    m.row1 = m.row0 + p.rows_per_blk;
    if (m.row1 > p.M) m.row1 = p.M;
    return m;
}

// Reduce 16 per-thread accumulators over the row lanes (ry) of a CTA; the ry == 0 lanes get the totals.
__device__ __forceinline__ void reduce_rows(const Map& m, float a[8], float b[8], float* sh) {
    // sh: [THREADS][16]
    float* mine = sh + threadIdx.x * 16;
#pragma unroll
    for (int k = 0; k < 8; ++k) { mine[k] = a[k]; mine[8 + k] = b[k]; }
    __syncthreads();
    if (m.active && m.ry == 0) {
        for (int r = 1; r < m.rpp; ++r) {
            const float* o = sh + (r * m.tpr + m.tx) * 16;
#pragma unroll
            for (int k = 0; k < 8; ++k) { a[k] += o[k]; b[k] += o[8 + k]; }
        }
    }
}

// ---- forward: partial statistics --------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(THREADS) k_bn_stats(Params p) {
    __shared__ float sh[THREADS * 16];
    const Map m = make_map(p);
    float s[8] = {0, 0, 0, 0, 0, 0, 0, 0}, q[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (m.active) {
        const T* x = reinterpret_cast<const T*>(p.x) + (size_t)m.cg * VEC;
        constexpr int U = 8;
        long long r = m.row0 + m.ry;
        for (; r + (long long)(U - 1) * m.rpp < m.row1; r += (long long)U * m.rpp) {
            Raw<T> raw[U];
#pragma unroll
            for (int u = 0; u < U; ++u) ldraw<T>(x + (size_t)(r + (long long)u * m.rpp) * p.C, raw[u]);
#pragma unroll
            for (int u = 0; u < U; ++u) {
                float v[8];
                unpack(raw[u], v);
#pragma unroll
                for (int k = 0; k < 8; ++k) { s[k] += v[k]; q[k] = fmaf(v[k], v[k], q[k]); }
            }
        }
        if (r < m.row1) {
            // fewer than U rows left: ONE batch of predicated loads (rows past the end re-read row r and are not
            // counted) instead of a serial one-load-per-round-trip tail of up to U - 1 iterations
            Raw<T> raw[U];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const long long ru = r + (long long)u * m.rpp;
                ldraw<T>(x + (size_t)(ru < m.row1 ? ru : r) * p.C, raw[u]);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (r + (long long)u * m.rpp >= m.row1) break;
                float v[8];
                unpack(raw[u], v);
#pragma unroll
                for (int k = 0; k < 8; ++k) { s[k] += v[k]; q[k] = fmaf(v[k], v[k], q[k]); }
            }
        }
    }
    reduce_rows(m, s, q, sh);
    if (m.active && m.ry == 0) {
        float* o = p.partial + (size_t)blockIdx.x * 2 * p.C + (size_t)m.cg * VEC;
#pragma unroll
        for (int k = 0; k < 8; ++k) { o[k] = s[k]; o[p.C + k] = q[k]; }
    }
}

// ---- forward: combine partials (fixed order -> deterministic) ------------------------------------------
constexpr int MAX_PART_PER_LANE = 20;     // plan_grid keeps nblk <= 32 * 20
constexpr int FIN_THREADS = 256;

// A CTA finalises 8 adjacent channels.  Lane = (partial row r = lane / 8, channel = lane % 8): a warp load reads
// four full 32-byte sectors of the [nblk][2][C] partial array (a warp-per-channel mapping touches a different
// sector per lane and fetches every sector eight times).  Thread (warp w, row r) owns partials b = 4 w + r (mod 32);
// its <= 40 loads are independent.  Everything is fp32 and index arithmetic 32-bit: with one CTA per 8 channels the
// kernel is a short dependent chain, and ncu showed the earlier fp64 / 64-bit version issue-bound (1.4 k
// instructions per warp, 13 us for C = 256).
__device__ __forceinline__ float fin_reduce(float (*sh)[8], int w, int lane, float v) {
    v += __shfl_xor_sync(0xffffffffu, v, 8);
    v += __shfl_xor_sync(0xffffffffu, v, 16);
    __syncthreads();
    if (lane < 8) sh[w][lane] = v;
    __syncthreads();
    float t = 0.0f;
#pragma unroll
    for (int k = 0; k < 8; ++k) t += sh[k][lane & 7];
    return t;
}

template <int N>
__device__ __forceinline__ float tree_sum(const float (&v)[N]) {
    float t[4] = {0.0f, 0.0f, 0.0f, 0.0f};
#pragma unroll
    for (int i = 0; i < N; ++i) t[i & 3] += v[i];
    return (t[0] + t[1]) + (t[2] + t[3]);
}

__device__ __forceinline__ void fin_load(const Params& p, int b0, int c, float (&sv)[MAX_PART_PER_LANE],
                                         float (&qv)[MAX_PART_PER_LANE]) {
    const int stride = 2 * p.C;
    const float* part = p.partial + c + b0 * stride;
#pragma unroll
    for (int i = 0; i < MAX_PART_PER_LANE; ++i) {
        const bool ok = b0 + 32 * i < p.nblk;
        sv[i] = ok ? part[i * 32 * stride] : 0.0f;
        qv[i] = ok ? part[i * 32 * stride + p.C] : 0.0f;
    }
}

__global__ void __launch_bounds__(FIN_THREADS) k_bn_finalize(Params p) {
    __shared__ float sh[8][8];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int c = blockIdx.x * 8 + (lane & 7);          // C % 8 == 0: always a live channel
    const int b0 = w * 4 + (lane >> 3);
    float sv[MAX_PART_PER_LANE], qv[MAX_PART_PER_LANE];
    fin_load(p, b0, c, sv, qv);
    const float s = fin_reduce(sh, w, lane, tree_sum(sv));
    const float n = (float)p.M;
    const float mean = s / n;
    // M2 = sum_b [ M2_b + n_b * (mean_b - mean)^2 ]   (Chan's combination against the global mean)
    const float n_full = (float)p.rows_per_blk, n_last = (float)(p.M - (long long)(p.nblk - 1) * p.rows_per_blk);
    const float i_full = 1.0f / n_full, i_last = 1.0f / n_last;
    float m2v[MAX_PART_PER_LANE];
#pragma unroll
    for (int i = 0; i < MAX_PART_PER_LANE; ++i) {
        const int b = b0 + 32 * i;
        const bool last = b == p.nblk - 1;
        const float nb = last ? n_last : n_full;
        const float mb = sv[i] * (last ? i_last : i_full);
        const float m2b = fmaxf(qv[i] - sv[i] * mb, 0.0f);
        const float d = mb - mean;
        m2v[i] = b < p.nblk ? m2b + nb * d * d : 0.0f;
    }
    const float m2 = fin_reduce(sh, w, lane, tree_sum(m2v));
    if (p.num_batches && blockIdx.x == 0 && threadIdx.x == 0) *p.num_batches += 1;
    if (threadIdx.x < 8) {
        const float var = m2 / n;
        p.mean[c] = mean;
        p.invstd[c] = 1.0f / sqrtf(var + p.eps);
        if (p.running_mean) {
            const float unb = p.M > 1 ? m2 / (n - 1.0f) : var;
            p.running_mean[c] = (1.0f - p.momentum) * p.running_mean[c] + p.momentum * mean;
            p.running_var[c] = (1.0f - p.momentum) * p.running_var[c] + p.momentum * unb;
        }
    }
}

// ---- forward: normalise (+ residual) (+ ReLU) ----------------------------------------------------------
// use_running: mean = running_mean, invstd = rsqrt(running_var + eps) (eval mode)
template <typename T>
__global__ void __launch_bounds__(THREADS) k_bn_apply(Params p, int use_running) {
    const Map m = make_map(p);
    if (!m.active) return;
    float a[8], b[8];
    const int c0 = m.cg * VEC;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const float mu = use_running ? p.running_mean[c0 + k] : p.mean[c0 + k];
        const float is = use_running ? rsqrtf(p.running_var[c0 + k] + p.eps) : p.invstd[c0 + k];
        const float g = p.gamma ? p.gamma[c0 + k] : 1.0f;
        a[k] = g * is;
        b[k] = (p.beta ? p.beta[c0 + k] : 0.0f) - mu * a[k];
    }
    const T* x = reinterpret_cast<const T*>(p.x) + c0;
    const T* rs = p.res ? reinterpret_cast<const T*>(p.res) + c0 : nullptr;
    T* y = reinterpret_cast<T*>(p.y) + c0;
    for (long long r = m.row0 + m.ry; r < m.row1; r += (long long)ROWS * m.rpp) {
        Raw<T> xr[ROWS], rr[ROWS];
#pragma unroll
        for (int u = 0; u < ROWS; ++u) {
            const long long ru = r + (long long)u * m.rpp;
            if (ru < m.row1) {
                ldraw<T>(x + (size_t)ru * p.C, xr[u]);
                if (rs) ldraw<T>(rs + (size_t)ru * p.C, rr[u]);
            }
        }
#pragma unroll
        for (int u = 0; u < ROWS; ++u) {
            const long long ru = r + (long long)u * m.rpp;
            if (ru >= m.row1) break;
            float v[8], w[8];
            unpack(xr[u], v);
            if (rs) unpack(rr[u], w);
#pragma unroll
            unsigned bits = 0;
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                float o = fmaf(v[k], a[k], b[k]);
                if (rs) o += w[k];
                bits |= (o > 0.0f ? 1u : 0u) << k;
                v[k] = p.relu ? fmaxf(o, 0.0f) : o;
            }
            st8<T>(y + (size_t)ru * p.C, v);
            if (p.mask) p.mask[(size_t)ru * (p.C >> 3) + m.cg] = (unsigned char)bits;
        }
    }
}

// ---- backward: partial reductions -----------------------------------------------------------------------
// Accumulates S1 = sum dy' and S2 = sum dy' * x per channel; the finalize kernel turns them into
// dbeta = S1 and dgamma = invstd * (S2 - mean * S1).  dy' = dy * [y > 0] when ReLU is fused; y is either the
// saved forward output (residual fused) or recomputed as fma(x, a, b) exactly like k_bn_apply.
template <typename T>
__global__ void __launch_bounds__(THREADS) k_bn_bwd_reduce(Params p) {
    __shared__ float sh[THREADS * 16];
    const Map m = make_map(p);
    float sd[8] = {0, 0, 0, 0, 0, 0, 0, 0}, sx[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (m.active) {
        const int c0 = m.cg * VEC;
        const bool have_mask = p.mask != nullptr;
        const bool have_y = !have_mask && p.y != nullptr;
        const bool recompute = p.relu && !have_y && !have_mask;
        const unsigned char* mk = have_mask ? p.mask + m.cg : nullptr;
        const size_t mstride = (size_t)(p.C >> 3);
        float fa[8], fb[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            fa[k] = (p.gamma ? p.gamma[c0 + k] : 1.0f) * p.invstd[c0 + k];
            fb[k] = (p.beta ? p.beta[c0 + k] : 0.0f) - p.mean[c0 + k] * fa[k];
        }
        const T* x = reinterpret_cast<const T*>(p.x) + c0;
        const T* dy = reinterpret_cast<const T*>(p.dy) + c0;
        const T* y = reinterpret_cast<const T*>(p.y) + c0;
        for (long long r = m.row0 + m.ry; r < m.row1; r += (long long)ROWS * m.rpp) {
            Raw<T> xr[ROWS], dr[ROWS], yr[ROWS];
            unsigned mb[ROWS];
#pragma unroll
            for (int u = 0; u < ROWS; ++u) {
                const long long ru = r + (long long)u * m.rpp;
                if (ru < m.row1) {
                    ldraw<T>(x + (size_t)ru * p.C, xr[u]);
                    ldraw<T>(dy + (size_t)ru * p.C, dr[u]);
                    if (p.relu && have_y) ldraw<T>(y + (size_t)ru * p.C, yr[u]);
                    if (p.relu && have_mask) mb[u] = mk[(size_t)ru * mstride];
                }
            }
#pragma unroll
            for (int u = 0; u < ROWS; ++u) {
                const long long ru = r + (long long)u * m.rpp;
                if (ru >= m.row1) break;
                float xv[8], dv[8], yv[8];
                unpack(xr[u], xv);
                unpack(dr[u], dv);
                if (p.relu && have_y) unpack(yr[u], yv);
#pragma unroll
                for (int k = 0; k < 8; ++k) {
                    float d = dv[k];
                    if (p.relu) {
                        if (have_mask) {
                            if (!((mb[u] >> k) & 1u)) d = 0.0f;
                        } else {
                            const float yy = recompute ? fmaf(xv[k], fa[k], fb[k]) : yv[k];
                            if (!(yy > 0.0f)) d = 0.0f;
                        }
                    }
                    sd[k] += d;
                    sx[k] = fmaf(d, xv[k], sx[k]);
                }
            }
        }
    }
    reduce_rows(m, sd, sx, sh);
    if (m.active && m.ry == 0) {
        float* o = p.partial + (size_t)blockIdx.x * 2 * p.C + (size_t)m.cg * VEC;
#pragma unroll
        for (int k = 0; k < 8; ++k) { o[k] = sd[k]; o[p.C + k] = sx[k]; }
    }
}

__global__ void __launch_bounds__(FIN_THREADS) k_bn_bwd_finalize(Params p) {
    __shared__ float sh[8][8];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int c = blockIdx.x * 8 + (lane & 7);
    float sv[MAX_PART_PER_LANE], qv[MAX_PART_PER_LANE];
    fin_load(p, w * 4 + (lane >> 3), c, sv, qv);
    const float sd = fin_reduce(sh, w, lane, tree_sum(sv));
    const float sx = fin_reduce(sh, w, lane, tree_sum(qv));
    if (threadIdx.x < 8) {
        p.dbeta[c] = sd;
        p.dgamma[c] = p.invstd[c] * (sx - p.mean[c] * sd);
    }
}

// dx = A*dy' + B*x + C  with  A = gamma*invstd,  B = -A*invstd*mean(dy'*xhat),  C = -B*mean - A*mean(dy')
template <typename T>
__global__ void __launch_bounds__(THREADS) k_bn_bwd_apply(Params p) {
    const Map m = make_map(p);
    if (!m.active) return;
    const int c0 = m.cg * VEC;
    const bool have_mask = p.mask != nullptr;
    const bool have_y = !have_mask && p.y != nullptr;
    const bool recompute = p.relu && !have_y && !have_mask;
    const unsigned char* mk = have_mask ? p.mask + m.cg : nullptr;
    const size_t mstride = (size_t)(p.C >> 3);
    float A[8], B[8], Cc[8], fb[8];
    const float invM = 1.0f / (float)p.M;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const float mu = p.mean[c0 + k], is = p.invstd[c0 + k];
        A[k] = (p.gamma ? p.gamma[c0 + k] : 1.0f) * is;
        fb[k] = (p.beta ? p.beta[c0 + k] : 0.0f) - mu * A[k];
        B[k] = -A[k] * is * (p.dgamma[c0 + k] * invM);
        Cc[k] = -B[k] * mu - A[k] * (p.dbeta[c0 + k] * invM);
    }
    const T* x = reinterpret_cast<const T*>(p.x) + c0;
    const T* dy = reinterpret_cast<const T*>(p.dy) + c0;
    const T* y = reinterpret_cast<const T*>(p.y) + c0;
    T* dx = reinterpret_cast<T*>(p.dx) + c0;
    T* dr = p.dres ? reinterpret_cast<T*>(p.dres) + c0 : nullptr;
    for (long long r = m.row0 + m.ry; r < m.row1; r += (long long)ROWS * m.rpp) {
        Raw<T> xr[ROWS], gr[ROWS], yr[ROWS];
        unsigned mb[ROWS];
#pragma unroll
        for (int u = 0; u < ROWS; ++u) {
            const long long ru = r + (long long)u * m.rpp;
            if (ru < m.row1) {
                ldraw<T>(x + (size_t)ru * p.C, xr[u]);
                ldraw<T>(dy + (size_t)ru * p.C, gr[u]);
                if (p.relu && have_y) ldraw<T>(y + (size_t)ru * p.C, yr[u]);
                if (p.relu && have_mask) mb[u] = mk[(size_t)ru * mstride];
            }
        }
#pragma unroll
        for (int u = 0; u < ROWS; ++u) {
            const long long ru = r + (long long)u * m.rpp;
            if (ru >= m.row1) break;
            float xv[8], dv[8], yv[8];
            unpack(xr[u], xv);
            unpack(gr[u], dv);
            if (p.relu && have_y) unpack(yr[u], yv);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                float d = dv[k];
                if (p.relu) {
                    if (have_mask) {
                        if (!((mb[u] >> k) & 1u)) d = 0.0f;
                    } else {
                        const float yy = recompute ? fmaf(xv[k], A[k], fb[k]) : yv[k];
                        if (!(yy > 0.0f)) d = 0.0f;
                    }
                }
                dv[k] = d;
                xv[k] = fmaf(A[k], d, fmaf(B[k], xv[k], Cc[k]));
            }
            st8<T>(dx + (size_t)ru * p.C, xv);
            if (dr) st8<T>(dr + (size_t)ru * p.C, dv);
        }
    }
}

}  // namespace bn
}  // namespace lwc

using namespace lwc::bn;

static void plan_grid(long long M, int C, int* nblk, int* rows_per_blk, dim3* grid, int ctas_per_sm = 2) {
    const int ncg = C / VEC;
    const int tpr = ncg < THREADS ? ncg : THREADS;
    const int rpp = THREADS / tpr;
    const int gy = (ncg + tpr - 1) / tpr;
    // 2 CTAs per SM (16 warps x 8-12 outstanding 16-byte loads each), at least 16 passes of rows per CTA
    long long target = (148LL * ctas_per_sm + gy - 1) / gy;
    long long rows = (M + target - 1) / target;
    const long long min_rows = 16LL * rpp;
    if (rows < min_rows) rows = min_rows;
    const long long max_blk = 32LL * 20;                 // k_bn_finalize holds <= 20 partials per lane
    if ((M + rows - 1) / rows > max_blk) rows = (M + max_blk - 1) / max_blk;
    rows = (rows + rpp - 1) / rpp * rpp;
    *rows_per_blk = (int)rows;
    *nblk = (int)((M + rows - 1) / rows);
    *grid = dim3(*nblk, gy, 1);
}

extern "C" size_t lwc_bn_workspace_bytes(long long M, int C) {
    int nblk, rows;
    dim3 g;
    plan_grid(M, C, &nblk, &rows, &g, 4);       // the largest partial count any pass uses
    return (size_t)nblk * 2 * C * sizeof(float);
}

#define LWC_BN_DISPATCH(dt, ...)                                               \
    switch (dt) {                                                              \
        case LWC_F32: { using T = float; __VA_ARGS__; } break;                 \
        case LWC_BF16: { using T = __nv_bfloat16; __VA_ARGS__; } break;        \
        case LWC_F16: { using T = __half; __VA_ARGS__; } break;                \
        default: return LWC_EINVAL;                                            \
    }

// Training (training=1): batch statistics are computed, saved in mean/invstd and folded into the running stats.
// Eval (training=0): running statistics are used.  res may be NULL; relu in {0,1}.
extern "C" int lwc_bn_forward_ex(const void* x, void* y, const void* res, const float* gamma, const float* beta,
                                 float* mean, float* invstd, float* running_mean, float* running_var, float* workspace,
                                 long long M, int C, int dtype, int training, int relu, float eps, float momentum,
                                 long long* num_batches_tracked, void* relu_mask, void* stream);

extern "C" int lwc_bn_forward(const void* x, void* y, const void* res, const float* gamma, const float* beta,
                              float* mean, float* invstd, float* running_mean, float* running_var, float* workspace,
                              long long M, int C, int dtype, int training, int relu, float eps, float momentum,
                              void* stream) {
    return lwc_bn_forward_ex(x, y, res, gamma, beta, mean, invstd, running_mean, running_var, workspace, M, C, dtype,
                             training, relu, eps, momentum, nullptr, nullptr, stream);
}

// num_batches_tracked (int64 device scalar, nullable): incremented by one in training mode inside the finalize
// kernel -- the module's counter costs no launch of its own.
// relu_mask (M * C / 8 bytes, nullable): one bit per output element, set where the pre-activation is positive; hand it
// to lwc_bn_backward_ex instead of y and the backward kernels read 1/16 of the bytes (matters when a residual was
// fused: without one the mask is recomputed from x and neither y nor a mask is needed).
extern "C" int lwc_bn_forward_ex(const void* x, void* y, const void* res, const float* gamma, const float* beta,
                                 float* mean, float* invstd, float* running_mean, float* running_var, float* workspace,
                                 long long M, int C, int dtype, int training, int relu, float eps, float momentum,
                                 long long* num_batches_tracked, void* relu_mask, void* stream) {
    if (!x || !y || M <= 0 || C <= 0 || (C % VEC)) return LWC_EINVAL;
    if (training && (!mean || !invstd || !workspace)) return LWC_EINVAL;
    if (!training && (!running_mean || !running_var)) return LWC_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    Params p;
    memset(&p, 0, sizeof(p));
    p.x = x; p.y = y; p.res = res; p.gamma = gamma; p.beta = beta; p.mean = mean; p.invstd = invstd;
    p.running_mean = running_mean; p.running_var = running_var; p.partial = workspace;
    p.M = M; p.C = C; p.relu = relu; p.eps = eps; p.momentum = momentum;
    p.num_batches = training ? num_batches_tracked : nullptr;
    p.mask = reinterpret_cast<unsigned char*>(relu_mask);
    dim3 grid;
    if (training) {
        // the pure-read statistics pass (50 registers) likes 4 CTAs per SM; the apply pass 2
        plan_grid(M, C, &p.nblk, &p.rows_per_blk, &grid, 4);
        LWC_BN_DISPATCH(dtype, k_bn_stats<T><<<grid, THREADS, 0, st>>>(p));
        k_bn_finalize<<<C / 8, FIN_THREADS, 0, st>>>(p);
    }
    plan_grid(M, C, &p.nblk, &p.rows_per_blk, &grid, 2);
    LWC_BN_DISPATCH(dtype, k_bn_apply<T><<<grid, THREADS, 0, st>>>(p, training ? 0 : 1));
    return (int)cudaGetLastError();
}

// y: the forward output (needed only when relu=1 and a residual was fused; NULL -> the ReLU mask is recomputed
// from x, gamma, beta); dres: gradient of the fused residual input (nullable).
extern "C" int lwc_bn_backward_ex(const void* dy, const void* x, const void* y, void* dx, void* dres, const float* gamma,
                                  const float* beta, const float* mean, const float* invstd, float* dgamma, float* dbeta,
                                  float* workspace, long long M, int C, int dtype, int relu, const void* relu_mask, void* stream);

extern "C" int lwc_bn_backward(const void* dy, const void* x, const void* y, void* dx, void* dres, const float* gamma,
                               const float* beta, const float* mean, const float* invstd, float* dgamma, float* dbeta, float* workspace,
                               long long M, int C, int dtype, int relu, void* stream) {
    return lwc_bn_backward_ex(dy, x, y, dx, dres, gamma, beta, mean, invstd, dgamma, dbeta, workspace, M, C, dtype, relu, nullptr,
                              stream);
}

// relu_mask: the bit mask written by lwc_bn_forward_ex (takes precedence over y).
extern "C" int lwc_bn_backward_ex(const void* dy, const void* x, const void* y, void* dx, void* dres, const float* gamma,
                                  const float* beta, const float* mean, const float* invstd, float* dgamma, float* dbeta,
                                  float* workspace, long long M, int C, int dtype, int relu, const void* relu_mask, void* stream) {
    if (!dy || !x || !dx || !mean || !invstd || !dgamma || !dbeta || !workspace) return LWC_EINVAL;
    if (relu && !y && !relu_mask && dres) return LWC_EINVAL;    // with a fused residual the mask needs y or the bit mask
    if (M <= 0 || C <= 0 || (C % VEC)) return LWC_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    Params p;
    memset(&p, 0, sizeof(p));
    p.x = x; p.y = const_cast<void*>(y); p.dy = dy; p.dx = dx; p.dres = dres; p.gamma = gamma; p.beta = beta;
    p.mean = const_cast<float*>(mean); p.invstd = const_cast<float*>(invstd);
    p.dgamma = dgamma; p.dbeta = dbeta; p.partial = workspace;
    p.M = M; p.C = C; p.relu = relu;
    p.mask = reinterpret_cast<unsigned char*>(const_cast<void*>(relu_mask));
    dim3 grid;
    plan_grid(M, C, &p.nblk, &p.rows_per_blk, &grid);
    LWC_BN_DISPATCH(dtype, k_bn_bwd_reduce<T><<<grid, THREADS, 0, st>>>(p));
    k_bn_bwd_finalize<<<C / 8, FIN_THREADS, 0, st>>>(p);
    LWC_BN_DISPATCH(dtype, k_bn_bwd_apply<T><<<grid, THREADS, 0, st>>>(p));
    return (int)cudaGetLastError();
}

// Training-mode normalisation with batch statistics computed elsewhere (the convolution epilogue,
// lwc_convx_stats_finalize): only the apply pass runs.  mean / invstd are the per-channel batch statistics.
extern "C" int lwc_bn_apply_stats(const void* x, void* y, const void* res, const float* gamma, const float* beta, const float* mean,
                                  const float* invstd, long long M, int C, int dtype, int relu, void* relu_mask, void* stream) {
    if (!x || !y || !mean || !invstd || M <= 0 || C <= 0 || (C % VEC)) return LWC_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    Params p;
    memset(&p, 0, sizeof(p));
    p.x = x; p.y = y; p.res = res; p.gamma = gamma; p.beta = beta;
    p.mean = const_cast<float*>(mean); p.invstd = const_cast<float*>(invstd);
    p.M = M; p.C = C; p.relu = relu;
    p.mask = reinterpret_cast<unsigned char*>(relu_mask);
    dim3 grid;
    plan_grid(M, C, &p.nblk, &p.rows_per_blk, &grid, 2);
    LWC_BN_DISPATCH(dtype, k_bn_apply<T><<<grid, THREADS, 0, st>>>(p, 0));
    return (int)cudaGetLastError();
}

// Batch-norm backward whose reduction sums were produced elsewhere (the data-gradient epilogue of the consuming
// convolution, lwc_convx_dgrad_bnstats): partials [nparts][2][C] = per-CTA (sum d, sum d * x).  Sums them in a fixed
// order (double), writes dbeta / dgamma and runs the apply pass.
namespace lwc { namespace bn {
__global__ void __launch_bounds__(256) k_bn_bwd_finalize_partials(const float* __restrict__ part, int nparts, int C,
                                                                  const float* __restrict__ mean, const float* __restrict__ invstd,
                                                                  float* __restrict__ dgamma, float* __restrict__ dbeta) {
    const int c = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
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
        dbeta[c] = (float)s;
        dgamma[c] = invstd[c] * (float)(q - (double)mean[c] * s);
    }
}
} }

extern "C" int lwc_bn_backward_from_partials(const void* dy, const void* x, void* dx, const float* gamma, const float* beta,
                                             const float* mean, const float* invstd, float* dgamma, float* dbeta,
                                             const float* partials, int nparts, long long M, int C, int dtype, int relu,
                                             void* stream) {
    if (!dy || !x || !dx || !mean || !invstd || !dgamma || !dbeta || !partials || nparts <= 0) return LWC_EINVAL;
    if (M <= 0 || C <= 0 || (C % VEC)) return LWC_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    k_bn_bwd_finalize_partials<<<(C + 7) / 8, 256, 0, st>>>(partials, nparts, C, mean, invstd, dgamma, dbeta);
    Params p;
    memset(&p, 0, sizeof(p));
    p.x = x; p.dy = dy; p.dx = dx; p.gamma = gamma; p.beta = beta;
    p.mean = const_cast<float*>(mean); p.invstd = const_cast<float*>(invstd);
    p.dgamma = dgamma; p.dbeta = dbeta;
    p.M = M; p.C = C; p.relu = relu;
    dim3 grid;
    plan_grid(M, C, &p.nblk, &p.rows_per_blk, &grid);
    LWC_BN_DISPATCH(dtype, k_bn_bwd_apply<T><<<grid, THREADS, 0, st>>>(p));
    return (int)cudaGetLastError();
}
