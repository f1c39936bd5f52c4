// Fused batch normalisation for channels-last activations ([M = N*H*W, C] row-major).
//
// Training forward : k_bn_stats (per-block partial sum / sum-of-squares, fp32)
//                    k_bn_finalize (deterministic Chan combination in double -> mean, invstd,
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
        long long r = m.row0 + m.ry;
        for (; r + 3LL * m.rpp < m.row1; r += 4LL * m.rpp) {
            float v0[8], v1[8], v2[8], v3[8];
            ld8<T>(x + (size_t)r * p.C, v0);
            ld8<T>(x + (size_t)(r + m.rpp) * p.C, v1);
            ld8<T>(x + (size_t)(r + 2 * m.rpp) * p.C, v2);
            ld8<T>(x + (size_t)(r + 3 * m.rpp) * p.C, v3);
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                s[k] += (v0[k] + v1[k]) + (v2[k] + v3[k]);
                q[k] = fmaf(v0[k], v0[k], q[k]);
                q[k] = fmaf(v1[k], v1[k], q[k]);
                q[k] = fmaf(v2[k], v2[k], q[k]);
                q[k] = fmaf(v3[k], v3[k], q[k]);
            }
        }
        for (; r < m.row1; r += m.rpp) {
            float v[8];
            ld8<T>(x + (size_t)r * p.C, v);
#pragma unroll
            for (int k = 0; k < 8; ++k) { s[k] += v[k]; q[k] = fmaf(v[k], v[k], q[k]); }
        }
    }
    reduce_rows(m, s, q, sh);
    if (m.active && m.ry == 0) {
        float* o = p.partial + (size_t)blockIdx.x * 2 * p.C + (size_t)m.cg * VEC;
#pragma unroll
        for (int k = 0; k < 8; ++k) { o[k] = s[k]; o[p.C + k] = q[k]; }
    }
}

// ---- forward: combine partials (one CTA = 32 channels x 8 lanes) --------------------------------------
__global__ void __launch_bounds__(256) k_bn_finalize(Params p) {
    const int c = blockIdx.x * 32 + (threadIdx.x >> 3);
    const int l = threadIdx.x & 7;
    double n = 0.0, mean = 0.0, m2 = 0.0;
    if (c < p.C) {
        for (int b = l; b < p.nblk; b += 8) {
            long long r0 = (long long)b * p.rows_per_blk, r1 = r0 + p.rows_per_blk;
            if (r1 > p.M) r1 = p.M;
            const double nb = (double)(r1 - r0);
            if (nb <= 0) continue;
            const float sb = p.partial[(size_t)b * 2 * p.C + c], qb = p.partial[(size_t)b * 2 * p.C + p.C + c];
            const double mb = (double)sb / nb;
            double m2b = (double)qb - (double)sb * mb;
            if (m2b < 0) m2b = 0;
            const double delta = mb - mean, tot = n + nb;
            mean += delta * nb / tot;
            m2 += m2b + delta * delta * n * nb / tot;
            n = tot;
        }
    }
    // fixed-order tree over the 8 lanes (deterministic)
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
        const double n2 = __shfl_xor_sync(0xffffffffu, n, o);
        const double mean2 = __shfl_xor_sync(0xffffffffu, mean, o);
        const double m22 = __shfl_xor_sync(0xffffffffu, m2, o);
        // combine in a symmetric way so both partners compute the same value
        const double tot = n + n2;
        if (tot > 0) {
            const double delta = mean2 - mean;
            const double nm = (n * mean + n2 * mean2) / tot;
            m2 = m2 + m22 + delta * delta * n * n2 / tot;
            mean = nm;
            n = tot;
        }
    }
    if (c < p.C && l == 0) {
        const double var = n > 0 ? m2 / n : 0.0;
        p.mean[c] = (float)mean;
        p.invstd[c] = (float)(1.0 / sqrt(var + (double)p.eps));
        if (p.running_mean) {
            const double unb = n > 1 ? m2 / (n - 1.0) : var;
            p.running_mean[c] = (float)((1.0 - p.momentum) * p.running_mean[c] + p.momentum * mean);
            p.running_var[c] = (float)((1.0 - p.momentum) * p.running_var[c] + p.momentum * unb);
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
    for (long long r = m.row0 + m.ry; r < m.row1; r += 2LL * m.rpp) {
        const long long r2 = r + m.rpp;
        const bool two = r2 < m.row1;
        float v0[8], v1[8], w0[8], w1[8];
        ld8<T>(x + (size_t)r * p.C, v0);
        if (two) ld8<T>(x + (size_t)r2 * p.C, v1);
        if (rs) {
            ld8<T>(rs + (size_t)r * p.C, w0);
            if (two) ld8<T>(rs + (size_t)r2 * p.C, w1);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            float o0 = fmaf(v0[k], a[k], b[k]), o1 = fmaf(v1[k], a[k], b[k]);
            if (rs) { o0 += w0[k]; o1 += w1[k]; }
            if (p.relu) { o0 = fmaxf(o0, 0.0f); o1 = fmaxf(o1, 0.0f); }
            v0[k] = o0; v1[k] = o1;
        }
        st8<T>(y + (size_t)r * p.C, v0);
        if (two) st8<T>(y + (size_t)r2 * p.C, v1);
    }
}

// ---- backward: partial reductions -----------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(THREADS) k_bn_bwd_reduce(Params p) {
    __shared__ float sh[THREADS * 16];
    const Map m = make_map(p);
    float sd[8] = {0, 0, 0, 0, 0, 0, 0, 0}, sx[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    if (m.active) {
        const int c0 = m.cg * VEC;
        float mu[8], is[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) { mu[k] = p.mean[c0 + k]; is[k] = p.invstd[c0 + k]; }
        const T* x = reinterpret_cast<const T*>(p.x) + c0;
        const T* dy = reinterpret_cast<const T*>(p.dy) + c0;
        const T* y = reinterpret_cast<const T*>(p.y) + c0;
        for (long long r = m.row0 + m.ry; r < m.row1; r += 2LL * m.rpp) {
            const long long r2 = r + m.rpp;
            const bool two = r2 < m.row1;
            float xv0[8], dv0[8], yv0[8], xv1[8], dv1[8], yv1[8];
            ld8<T>(x + (size_t)r * p.C, xv0);
            ld8<T>(dy + (size_t)r * p.C, dv0);
            if (p.relu) ld8<T>(y + (size_t)r * p.C, yv0);
            if (two) {
                ld8<T>(x + (size_t)r2 * p.C, xv1);
                ld8<T>(dy + (size_t)r2 * p.C, dv1);
                if (p.relu) ld8<T>(y + (size_t)r2 * p.C, yv1);
            }
#pragma unroll
            for (int k = 0; k < 8; ++k) {
                float d0 = dv0[k];
                if (p.relu && !(yv0[k] > 0.0f)) d0 = 0.0f;
                sd[k] += d0;
                sx[k] = fmaf(d0, (xv0[k] - mu[k]) * is[k], sx[k]);
                if (two) {
                    float d1 = dv1[k];
                    if (p.relu && !(yv1[k] > 0.0f)) d1 = 0.0f;
                    sd[k] += d1;
                    sx[k] = fmaf(d1, (xv1[k] - mu[k]) * is[k], sx[k]);
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

__global__ void __launch_bounds__(256) k_bn_bwd_finalize(Params p) {
    const int c = blockIdx.x * 32 + (threadIdx.x >> 3);
    const int l = threadIdx.x & 7;
    double sd = 0.0, sx = 0.0;
    if (c < p.C) {
        for (int b = l; b < p.nblk; b += 8) {
            sd += (double)p.partial[(size_t)b * 2 * p.C + c];
            sx += (double)p.partial[(size_t)b * 2 * p.C + p.C + c];
        }
    }
#pragma unroll
    for (int o = 1; o < 8; o <<= 1) {
        sd += __shfl_xor_sync(0xffffffffu, sd, o);
        sx += __shfl_xor_sync(0xffffffffu, sx, o);
    }
    if (c < p.C && l == 0) {
        p.dbeta[c] = (float)sd;
        p.dgamma[c] = (float)sx;
    }
}

template <typename T>
__global__ void __launch_bounds__(THREADS) k_bn_bwd_apply(Params p) {
    const Map m = make_map(p);
    if (!m.active) return;
    const int c0 = m.cg * VEC;
    float mu[8], is[8], a[8], md[8], mx[8];
    const float invM = 1.0f / (float)p.M;
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        mu[k] = p.mean[c0 + k];
        is[k] = p.invstd[c0 + k];
        a[k] = (p.gamma ? p.gamma[c0 + k] : 1.0f) * is[k];
        md[k] = p.dbeta[c0 + k] * invM;
        mx[k] = p.dgamma[c0 + k] * invM;
    }
    const T* x = reinterpret_cast<const T*>(p.x) + c0;
    const T* dy = reinterpret_cast<const T*>(p.dy) + c0;
    const T* y = reinterpret_cast<const T*>(p.y) + c0;
    T* dx = reinterpret_cast<T*>(p.dx) + c0;
    T* dr = p.dres ? reinterpret_cast<T*>(p.dres) + c0 : nullptr;
    for (long long r = m.row0 + m.ry; r < m.row1; r += 2LL * m.rpp) {
        const long long r2 = r + m.rpp;
        const bool two = r2 < m.row1;
        float xv0[8], dv0[8], yv0[8], xv1[8], dv1[8], yv1[8];
        ld8<T>(x + (size_t)r * p.C, xv0);
        ld8<T>(dy + (size_t)r * p.C, dv0);
        if (p.relu) ld8<T>(y + (size_t)r * p.C, yv0);
        if (two) {
            ld8<T>(x + (size_t)r2 * p.C, xv1);
            ld8<T>(dy + (size_t)r2 * p.C, dv1);
            if (p.relu) ld8<T>(y + (size_t)r2 * p.C, yv1);
        }
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            float d0 = dv0[k], d1 = dv1[k];
            if (p.relu) {
                if (!(yv0[k] > 0.0f)) d0 = 0.0f;
                if (!(yv1[k] > 0.0f)) d1 = 0.0f;
            }
            dv0[k] = d0; dv1[k] = d1;
            xv0[k] = a[k] * (d0 - md[k] - (xv0[k] - mu[k]) * is[k] * mx[k]);
            xv1[k] = a[k] * (d1 - md[k] - (xv1[k] - mu[k]) * is[k] * mx[k]);
        }
        st8<T>(dx + (size_t)r * p.C, xv0);
        if (dr) st8<T>(dr + (size_t)r * p.C, dv0);
        if (two) {
            st8<T>(dx + (size_t)r2 * p.C, xv1);
            if (dr) st8<T>(dr + (size_t)r2 * p.C, dv1);
        }
    }
}

}  // namespace bn
}  // namespace lwc

using namespace lwc::bn;

static void plan_grid(long long M, int C, int* nblk, int* rows_per_blk, dim3* grid) {
    const int ncg = C / VEC;
    const int tpr = ncg < THREADS ? ncg : THREADS;
    const int rpp = THREADS / tpr;
    const int gy = (ncg + tpr - 1) / tpr;
    // ~4 CTAs per SM, at least 4 passes of rows per CTA
    long long target = (148LL * 4 + gy - 1) / gy;
    long long rows = (M + target - 1) / target;
    const long long min_rows = 4LL * rpp;
    if (rows < min_rows) rows = min_rows;
    rows = (rows + rpp - 1) / rpp * rpp;
    *rows_per_blk = (int)rows;
    *nblk = (int)((M + rows - 1) / rows);
    *grid = dim3(*nblk, gy, 1);
}

extern "C" size_t lwc_bn_workspace_bytes(long long M, int C) {
    int nblk, rows;
    dim3 g;
    plan_grid(M, C, &nblk, &rows, &g);
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
extern "C" int lwc_bn_forward(const void* x, void* y, const void* res, const float* gamma, const float* beta,
                              float* mean, float* invstd, float* running_mean, float* running_var, float* workspace,
                              long long M, int C, int dtype, int training, int relu, float eps, float momentum,
                              void* stream) {
    if (!x || !y || M <= 0 || C <= 0 || (C % VEC)) return LWC_EINVAL;
    if (training && (!mean || !invstd || !workspace)) return LWC_EINVAL;
    if (!training && (!running_mean || !running_var)) return LWC_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    Params p;
    memset(&p, 0, sizeof(p));
    p.x = x; p.y = y; p.res = res; p.gamma = gamma; p.beta = beta; p.mean = mean; p.invstd = invstd;
    p.running_mean = running_mean; p.running_var = running_var; p.partial = workspace;
    p.M = M; p.C = C; p.relu = relu; p.eps = eps; p.momentum = momentum;
    dim3 grid;
    plan_grid(M, C, &p.nblk, &p.rows_per_blk, &grid);
    if (training) {
        LWC_BN_DISPATCH(dtype, k_bn_stats<T><<<grid, THREADS, 0, st>>>(p));
        k_bn_finalize<<<(C + 31) / 32, 256, 0, st>>>(p);
    }
    LWC_BN_DISPATCH(dtype, k_bn_apply<T><<<grid, THREADS, 0, st>>>(p, training ? 0 : 1));
    return (int)cudaGetLastError();
}

// y: the forward output (needed only when relu=1); dres: gradient of the fused residual input (nullable).
extern "C" int lwc_bn_backward(const void* dy, const void* x, const void* y, void* dx, void* dres, const float* gamma,
                               const float* mean, const float* invstd, float* dgamma, float* dbeta, float* workspace,
                               long long M, int C, int dtype, int relu, void* stream) {
    if (!dy || !x || !dx || !mean || !invstd || !dgamma || !dbeta || !workspace) return LWC_EINVAL;
    if (relu && !y) return LWC_EINVAL;
    if (M <= 0 || C <= 0 || (C % VEC)) return LWC_EINVAL;
    cudaStream_t st = (cudaStream_t)stream;
    Params p;
    memset(&p, 0, sizeof(p));
    p.x = x; p.y = const_cast<void*>(y); p.dy = dy; p.dx = dx; p.dres = dres; p.gamma = gamma;
    p.mean = const_cast<float*>(mean); p.invstd = const_cast<float*>(invstd);
    p.dgamma = dgamma; p.dbeta = dbeta; p.partial = workspace;
    p.M = M; p.C = C; p.relu = relu;
    dim3 grid;
    plan_grid(M, C, &p.nblk, &p.rows_per_blk, &grid);
    LWC_BN_DISPATCH(dtype, k_bn_bwd_reduce<T><<<grid, THREADS, 0, st>>>(p));
    k_bn_bwd_finalize<<<(C + 31) / 32, 256, 0, st>>>(p);
    LWC_BN_DISPATCH(dtype, k_bn_bwd_apply<T><<<grid, THREADS, 0, st>>>(p));
    return (int)cudaGetLastError();
}
