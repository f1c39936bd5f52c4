// Fused gradient compress -> exchange -> decompress for sm_100a (B200), NVLink peer stores.
//
// One "plan" = one bucket of gradient tensors.  Work is cut into chunks of LWC_CHUNK elements
// (one CTA each) listed in a device chunk table, so a bucket with hundreds of tiny tensors and
// a bucket with one 100M-element tensor both run as a handful of launches ("multi-tensor").
//
// Sparse methods (Top-k, Random-k, Threshold-v, adaptive threshold):
//   k_pre      : acc = grad + residual (error feedback), residual <- acc, first radix histogram
//                of the 31-bit selection key (|acc| bit pattern, or a counter-based hash for
//                Random-k) or a max-reduction (adaptive)
//   k_group    : per selection group: find the histogram bin holding the keep-th largest key
//   k_hist     : second / third radix level restricted to the current prefix
//   k_count    : per chunk number of keys > T and == T
//   k_group    : exclusive scan of the chunk counts (ordered, deterministic payload layout)
//   k_compact_send : ordered compaction of (index, value), residual <- 0 at sent entries,
//                coalesced stores of the packed run into the receive slot of EVERY rank
//                (local + peer-mapped pointers over NVLink), release flag by the last CTA
//   k_recv_sparse : acquire W-1 flags, each CTA owns one chunk of the dense output: binary-search
//                every source's sorted payload for its index range, accumulate in shared memory in
//                rank order (deterministic, no atomics), scale by 1/W, write the gradient
// Quantisers (TernGrad 2-bit, QSGD/dithering 8-bit + sign) and the dense path reuse k_pre for the
// scale (max / L2 norm) and then k_quant_send / k_recv_dense.
//
// Reference behaviour covered: CIFAR10/core.py:175-301 (compress + all_reduce + /W + copy back),
// IMAGENET/training/sparsified_ddp.py:403-414,:468-483 (error feedback, pack, expand).
#include <math.h>
#include <stdio.h>
#include <string.h>

#include <vector>

#include "common.cuh"

namespace lwc {

constexpr int kHist1 = 2048;     // key bits 30..20
constexpr int kHist2 = 1024;     // key bits 19..10
constexpr int kHist3 = 1024;     // key bits  9..0
constexpr int kStage = 2048;     // staged payload entries per CTA
constexpr uint32_t kNoKey = 0x80000000u;   // larger than every 31-bit key
constexpr unsigned long long kNoQuota = ~0ULL;

struct TensorRec {
    const void* grad;
    void* out;
    float* resid;
    const void* rnd;
    long long n;
    long long goff;     // offset of element 0 inside the group's index space
    int group;
    int uid;
};
struct ChunkRec {
    int tensor;
    int len;
    long long start;
};
struct GroupCfg {
    long long n;
    long long keep;
    long long cap;
    long long pay_off;
    int chunk_begin;
    int chunk_end;
    int uid;
    int pad;
};
struct GroupState {
    unsigned hist1[kHist1];
    unsigned hist2[kHist2];
    unsigned hist3[kHist3];
    unsigned prefix;
    unsigned thresh;
    unsigned maxbits;
    float scale;
    unsigned long long remaining;
    unsigned long long count_gt;
    unsigned long long eq_quota;
    unsigned long long gt_total;
    unsigned long long eq_total;
    unsigned long long kept;
    unsigned ticket[4];     // per phase: chunk CTAs of this group that finished (the last one runs the group step)
};
struct Counters {
    unsigned long long sent;
    unsigned long long dropped;
    unsigned long long timeouts;
    unsigned long long bytes;
    unsigned int done;      // CTA completion ticket of the send kernel
    unsigned int done_recv; // CTA completion ticket of the receive kernel (device-side step counter mode)
    unsigned long long dev_step;   // step counter kept on the device (CUDA-graph replays cannot change kernel arguments)
};

struct DevPlan {     // passed by value to kernels
    const TensorRec* tensors;
    const ChunkRec* chunks;
    const GroupCfg* gcfg;
    GroupState* gstate;
    unsigned* chunk_cnt;       // packed gt | eq<<16 per chunk
    unsigned long long* chunk_off;   // packed exclusive offsets gt | eq<<32
    float* chunk_part;         // per chunk partial (sum of squares)
    Counters* counters;
    int num_chunks;
    int num_groups;
    int method;
    int world;
    int rank;
    int qstates;
    float threshold;
    unsigned long long seed;
    unsigned long long step;
    unsigned long long* step_ptr;    // non-NULL: the step lives on the device (read at kernel start, bumped by the receiver)
    unsigned long long timeout_ns;   // bounded wait on peer flags
    // slot geometry
    unsigned long long slot_bytes;
    unsigned long long hdr_bytes;
    unsigned long long idx_off;     // byte offset of idx array inside a slot
    unsigned long long val_off;     // byte offset of val array
    unsigned long long code_off;    // byte offset of code area (dense / quantised)
    unsigned long long code_stride; // bytes per chunk in the code area
    char* region[LWC_MAX_WORLD];
    uint64_t* flags[LWC_MAX_WORLD];
    // NVLS variant: multicast address of this plan's region (one store lands in every rank's slot through the
    // NVSwitch); NULL -> one unicast store per destination
    char* mc_region;
    // shared-seed Random-k: the mask is a function of (seed, step, tensor, index) alone, so every rank selects the same
    // positions; only values travel and the receiver recomputes the positions (reference: sparsified_ddp.py:410-412)
    int values_only;
    // stand-alone selection: write (idx, val, count) here instead of into exchange slots
    uint32_t* direct_idx;
    float* direct_val;
    long long* direct_count;
};

// Per-rank pointer tables live in the kernel parameter block.  Indexing them with a run-time rank makes the compiler copy
// the whole block to local memory (k_compact_send had 1047 local loads); a compare chain over constant indices keeps the
// reads in the constant bank.
template <typename T>
__device__ __forceinline__ T pick_rank(T const (&tab)[LWC_MAX_WORLD], int r) {
    T x = tab[0];
#pragma unroll
    for (int i = 1; i < LWC_MAX_WORLD; ++i)
        if (r == i) x = tab[i];
    return x;
}

__device__ __forceinline__ char* slot_ptr(const DevPlan& p, int dst, int src) {
    const unsigned long long parity = p.step & 1ULL;
    return pick_rank(p.region, dst) + (parity * p.world + src) * p.slot_bytes;
}

__device__ __forceinline__ char* mc_slot_ptr(const DevPlan& p, int src) {
    const unsigned long long parity = p.step & 1ULL;
    return p.mc_region + (parity * p.world + src) * p.slot_bytes;
}
// multimem.st: a store to a multicast address is replicated by the switch to every GPU bound to it
__device__ __forceinline__ void mm_st_u32(void* a, uint32_t v) {
    asm volatile("multimem.st.relaxed.sys.global.u32 [%0], %1;" ::"l"(a), "r"(v) : "memory");
}
__device__ __forceinline__ void mm_st_f32(void* a, float v) {
    asm volatile("multimem.st.relaxed.sys.global.f32 [%0], %1;" ::"l"(a), "f"(v) : "memory");
}
__device__ __forceinline__ void mm_st_u64(void* a, unsigned long long v) {
    asm volatile("multimem.st.relaxed.sys.global.u64 [%0], %1;" ::"l"(a), "l"(v) : "memory");
}
__device__ __forceinline__ void mm_st_v4(void* a, float x, float y, float z, float w) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(a), "f"(x), "f"(y), "f"(z), "f"(w) : "memory");
}

// ---- selection key -----------------------------------------------------------------------------
__device__ __forceinline__ uint64_t group_salt(const DevPlan& p, int uid) {
    return mix64(p.seed + p.step * 0xD6E8FEB86659FD93ULL + (unsigned long long)uid * 0xA24BAED4963EE407ULL);
}
__device__ __forceinline__ uint32_t make_key(int method, float acc, const void* rnd, long long i,
                                             long long gidx, uint64_t salt) {
    if (method == LWC_RANDOMK) {
        if (rnd) return reinterpret_cast<const uint32_t*>(rnd)[i] & 0x7fffffffu;
        return (uint32_t)(mix64((uint64_t)gidx * 0x9E3779B97F4A7C15ULL + salt) >> 33);
    }
    if (method == LWC_ADAPTIVE) return __float_as_uint(fabsf(acc * 2.0f));
    return __float_as_uint(fabsf(acc));
}

// Source of the accumulated gradient for the passes after k_pre.
template <typename GradT>
__device__ __forceinline__ void load_acc(const TensorRec& t, long long start, int j, int len, float v[4]) {
    if (t.resid) {
        const float* b = t.resid + start;
        load_quad<float>(b, j, len, vec_ok(b), v);
    } else {
        const GradT* b = reinterpret_cast<const GradT*>(t.grad) + start;
        load_quad<GradT>(b, j, len, vec_ok(b), v);
    }
}

enum { STAT_NONE = 0, STAT_HIST = 1, STAT_MAX = 2, STAT_SUMSQ = 3 };
enum { G_SCAN1 = 1, G_SCAN2, G_SCAN3, G_SET_THRESH, G_SET_ADAPT, G_FIN_MAX, G_FIN_NORM, G_SCAN_COUNTS };
__device__ void group_tail(const DevPlan& p, int group, int phase, int mode);   // defined after the group step

// Selection threshold / tie quota of a group.  The radix-select methods leave them in the group state (G_SCAN3); the
// two threshold methods need no group step at all: V is a plan constant and max|g| is complete once k_pre has ended.
__device__ __forceinline__ unsigned group_threshold(const DevPlan& p, const GroupState& gs) {
    if (p.method == LWC_THRESHOLDV) return __float_as_uint(fabsf(p.threshold));
    if (p.method == LWC_ADAPTIVE) return gs.maxbits;
    return gs.thresh;
}
__device__ __forceinline__ unsigned long long group_quota(const DevPlan& p, const GroupState& gs) {
    return (p.method == LWC_THRESHOLDV || p.method == LWC_ADAPTIVE) ? kNoQuota : gs.eq_quota;
}


// ---- pass 1: error-feedback accumulate + first statistic -----------------------------------------------
template <typename GradT, int STAT>
__global__ void __launch_bounds__(kThreads) k_pre(DevPlan p) {
    if (p.step_ptr) p.step = *p.step_ptr;
    __shared__ unsigned sh_hist[STAT == STAT_HIST ? kHist1 : 1];
    __shared__ float sh_red[kThreads / 32];
    const ChunkRec c = p.chunks[blockIdx.x];
    const TensorRec t = p.tensors[c.tensor];
    const int tid = threadIdx.x;
    if (STAT == STAT_HIST) {
        for (int b = tid; b < kHist1; b += kThreads) sh_hist[b] = 0;
        __syncthreads();
    }
    const GradT* g = reinterpret_cast<const GradT*>(t.grad) + c.start;
    float* r = t.resid ? t.resid + c.start : nullptr;
    const bool gv = vec_ok(g), rv = r && vec_ok(r);
    const uint64_t salt = group_salt(p, p.gcfg[t.group].uid);
    float red = 0.0f;
#pragma unroll 2
    for (int it = 0; it < kQuadsPerThread; ++it) {
        const int j = (it * kThreads + tid) * 4;
        if (j >= c.len) break;
        float v[4];
        load_quad<GradT>(g, j, c.len, gv, v);
        if (r) {
            float e[4];
            load_quad<float>(r, j, c.len, rv, e);
#pragma unroll
            for (int k = 0; k < 4; ++k) v[k] += e[k];
            store_quad<float>(r, j, c.len, rv, v);
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (j + k >= c.len) break;
            if (STAT == STAT_HIST) {
                const long long i = c.start + j + k;
                const uint32_t key = make_key(p.method, v[k], t.rnd, i, t.goff + i, salt);
                atomicAdd(&sh_hist[key >> 20], 1u);
            } else if (STAT == STAT_MAX) {
                red = fmaxf(red, fabsf(v[k]));
            } else if (STAT == STAT_SUMSQ) {
                red = fmaf(v[k], v[k], red);
            }
        }
    }
    if (STAT == STAT_HIST) {
        __syncthreads();
        GroupState& gs = p.gstate[t.group];
        for (int b = tid; b < kHist1; b += kThreads) {
            const unsigned h = sh_hist[b];
            if (h) atomicAdd(&gs.hist1[b], h);
        }
    } else if (STAT == STAT_MAX || STAT == STAT_SUMSQ) {
        red = (STAT == STAT_MAX) ? warp_max(red) : warp_sum(red);
        if ((tid & 31) == 0) sh_red[tid >> 5] = red;
        __syncthreads();
        if (tid < 32) {
            float x = tid < kThreads / 32 ? sh_red[tid] : 0.0f;
            x = (STAT == STAT_MAX) ? warp_max(x) : warp_sum(x);
            if (tid == 0) {
                if (STAT == STAT_MAX) atomicMax(&p.gstate[t.group].maxbits, __float_as_uint(x));
                else p.chunk_part[blockIdx.x] = x;
            }
        }
    }
    if (STAT == STAT_HIST) group_tail(p, t.group, 0, G_SCAN1);
    else if (STAT == STAT_SUMSQ) group_tail(p, t.group, 0, G_FIN_NORM);
}

// ---- radix levels 2 and 3 -----------------------------------------------------------------------------
template <typename GradT, int LEVEL>
__global__ void __launch_bounds__(kThreads) k_hist(DevPlan p) {
    if (p.step_ptr) p.step = *p.step_ptr;
    __shared__ unsigned sh_hist[kHist2];
    const ChunkRec c = p.chunks[blockIdx.x];
    const TensorRec t = p.tensors[c.tensor];
    GroupState& gs = p.gstate[t.group];
    const unsigned prefix = gs.prefix;
    const int tid = threadIdx.x;
    for (int b = tid; b < kHist2; b += kThreads) sh_hist[b] = 0;
    __syncthreads();
    const uint64_t salt = group_salt(p, p.gcfg[t.group].uid);
    bool any = false;
#pragma unroll 2
    for (int it = 0; it < kQuadsPerThread; ++it) {
        const int j = (it * kThreads + tid) * 4;
        if (j >= c.len) break;
        float v[4];
        load_acc<GradT>(t, c.start, j, c.len, v);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (j + k >= c.len) break;
            const long long i = c.start + j + k;
            const uint32_t key = make_key(p.method, v[k], t.rnd, i, t.goff + i, salt);
            if (LEVEL == 2) {
                if ((key >> 20) == prefix) { atomicAdd(&sh_hist[(key >> 10) & 1023u], 1u); any = true; }
            } else {
                if ((key >> 10) == prefix) { atomicAdd(&sh_hist[key & 1023u], 1u); any = true; }
            }
        }
    }
    if (__syncthreads_or(any)) {
        unsigned* dst = (LEVEL == 2) ? gs.hist2 : gs.hist3;
        for (int b = tid; b < kHist2; b += kThreads) {
            const unsigned h = sh_hist[b];
            if (h) atomicAdd(&dst[b], h);
        }
    }
    group_tail(p, t.group, LEVEL == 2 ? 1 : 2, LEVEL == 2 ? G_SCAN2 : G_SCAN3);
}

// ---- per-group bookkeeping (one CTA per group) -----------------------------------------------------

__device__ __forceinline__ unsigned long long block_excl_scan_u64(unsigned long long v, unsigned long long* sh,
                                                                  unsigned long long* total) {
    // sh: kThreads/32 + 1 entries
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    unsigned long long x = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        unsigned long long y = __shfl_up_sync(0xffffffffu, x, o);
        if (lane >= o) x += y;
    }
    if (lane == 31) sh[w] = x;
    __syncthreads();
    if (w == 0) {
        unsigned long long s = lane < kThreads / 32 ? sh[lane] : 0ULL;
        unsigned long long xs = s;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            unsigned long long y = __shfl_up_sync(0xffffffffu, xs, o);
            if (lane >= o) xs += y;
        }
        if (lane < kThreads / 32) sh[lane] = xs - s;           // exclusive warp offsets
        if (lane == kThreads / 32 - 1) sh[kThreads / 32] = xs;  // grand total
    }
    __syncthreads();
    const unsigned long long res = x - v + sh[w];
    if (total) *total = sh[kThreads / 32];
    __syncthreads();
    return res;
}

// Find the bin b (scanning from the top) with S(b+1) < target <= S(b); returns via shared result.
template <int BINS>
__device__ void select_bin(const unsigned* hist, unsigned long long target, unsigned long long* sh_scan,
                           unsigned* out_bin, unsigned long long* out_above, unsigned* out_cnt) {
    constexpr int PER = BINS / kThreads;
    const int tid = threadIdx.x;
    // thread tid owns bins [BINS-(tid+1)*PER, BINS-tid*PER), visited from the top down
    unsigned local[PER];
    unsigned long long sum = 0;
#pragma unroll
    for (int k = 0; k < PER; ++k) {
        local[k] = __ldcg(hist + (BINS - 1 - (tid * PER + k)));
        sum += local[k];
    }
    unsigned long long total;
    unsigned long long above = block_excl_scan_u64(sum, sh_scan, &total);
    __shared__ unsigned s_bin;
    __shared__ unsigned long long s_above;
    __shared__ unsigned s_cnt;
    if (tid == 0) { const unsigned h0 = __ldcg(hist); s_bin = 0; s_above = total - h0; s_cnt = h0; }   // default: lowest bin
    __syncthreads();
    if (target > above && target <= above + sum) {
        unsigned long long a = above;
#pragma unroll
        for (int k = 0; k < PER; ++k) {
            if (target <= a + local[k]) {
                s_bin = BINS - 1 - (tid * PER + k);
                s_above = a;
                s_cnt = local[k];
                break;
            }
            a += local[k];
        }
    }
    __syncthreads();
    *out_bin = s_bin;
    *out_above = s_above;
    *out_cnt = s_cnt;
}

// Per-group bookkeeping, executed by all threads of one CTA.  It used to be a kernel of its own between the chunk
// passes (k_group, one CTA per group: four extra launches per bucket); now the chunk CTA that finishes a group's
// pass last runs it (group_tail), so a pass and its group step are one launch.
__device__ void group_step(const DevPlan& p, int g, int mode, unsigned long long* sh_scan) {
    const GroupCfg cfg = p.gcfg[g];
    GroupState& gs = p.gstate[g];
    const int tid = threadIdx.x;
    if (mode == G_SCAN1 || mode == G_SCAN2 || mode == G_SCAN3) {
        if (cfg.keep <= 0) {
            if (tid == 0 && mode == G_SCAN3) { gs.thresh = kNoKey; gs.count_gt = 0; gs.eq_quota = 0; }
            return;
        }
        unsigned bin, cnt;
        unsigned long long above;
        const unsigned long long target = (mode == G_SCAN1) ? (unsigned long long)cfg.keep : gs.remaining;
        if (mode == G_SCAN1) select_bin<kHist1>(gs.hist1, target, sh_scan, &bin, &above, &cnt);
        else if (mode == G_SCAN2) select_bin<kHist2>(gs.hist2, target, sh_scan, &bin, &above, &cnt);
        else select_bin<kHist3>(gs.hist3, target, sh_scan, &bin, &above, &cnt);
        if (tid == 0) {
            const unsigned long long cg = (mode == G_SCAN1 ? 0ULL : gs.count_gt) + above;
            gs.count_gt = cg;
            gs.remaining = target > above ? target - above : 1ULL;
            if (mode == G_SCAN1) gs.prefix = bin;
            else if (mode == G_SCAN2) gs.prefix = (gs.prefix << 10) | bin;
            else {
                gs.thresh = (gs.prefix << 10) | bin;
                // Top-k keeps every tie (reference: |g| >= thres); Random-k keeps exactly `keep`
                const unsigned long long k = (unsigned long long)cfg.keep;
                gs.eq_quota = (p.method == LWC_RANDOMK) ? (k > cg ? k - cg : 0ULL) : kNoQuota;
            }
        }
    } else if (mode == G_SET_THRESH) {
        if (tid == 0) { gs.thresh = __float_as_uint(fabsf(p.threshold)); gs.eq_quota = kNoQuota; }
    } else if (mode == G_SET_ADAPT) {
        if (tid == 0) { gs.thresh = gs.maxbits; gs.eq_quota = kNoQuota; }
    } else if (mode == G_FIN_MAX) {
        if (tid == 0) gs.scale = __uint_as_float(gs.maxbits);
    } else if (mode == G_FIN_NORM) {
        // fixed-order double accumulation of the chunk partials: run-to-run deterministic
        double acc = 0.0;
        for (int c = cfg.chunk_begin + tid; c < cfg.chunk_end; c += kThreads) acc += (double)__ldcg(p.chunk_part + c);
        __shared__ double sh_d[kThreads];
        sh_d[tid] = acc;
        __syncthreads();
        for (int o = kThreads / 2; o > 0; o >>= 1) {
            if (tid < o) sh_d[tid] += sh_d[tid + o];
            __syncthreads();
        }
        if (tid == 0) gs.scale = (float)sqrt(sh_d[0]);
    } else if (mode == G_SCAN_COUNTS) {
        unsigned long long run_gt = 0, run_eq = 0;
        for (int base = cfg.chunk_begin; base < cfg.chunk_end; base += kThreads) {
            const int c = base + tid;
            const unsigned packed = c < cfg.chunk_end ? __ldcg(p.chunk_cnt + c) : 0u;
            const unsigned long long v = (unsigned long long)(packed & 0xffffu) | ((unsigned long long)(packed >> 16) << 32);
            unsigned long long tot;
            const unsigned long long ex = block_excl_scan_u64(v, sh_scan, &tot);
            if (c < cfg.chunk_end) {
                const unsigned long long gt = run_gt + (ex & 0xffffffffULL);
                const unsigned long long eq = run_eq + (ex >> 32);
                p.chunk_off[c] = gt | (eq << 32);
            }
            run_gt += tot & 0xffffffffULL;
            run_eq += tot >> 32;
        }
        if (tid == 0) {
            const unsigned long long quota = group_quota(p, gs);
            unsigned long long want = run_gt + (run_eq < quota ? run_eq : quota);
            unsigned long long kept = want < (unsigned long long)cfg.cap ? want : (unsigned long long)cfg.cap;
            gs.gt_total = run_gt;
            gs.eq_total = run_eq;
            gs.kept = kept;
            atomicAdd(&p.counters->sent, kept);
            atomicAdd(&p.counters->dropped, want - kept);
        }
    }
}

// Ticket per (group, phase): the CTA that completes the group's pass runs the group step on the finished statistics.
__device__ void group_tail(const DevPlan& p, int group, int phase, int mode) {
    __shared__ unsigned long long sh_tail_scan[kThreads / 32 + 1];
    __shared__ int s_last;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();                        // one cumulative fence per CTA (the barrier ordered the CTA's writes before it)
        const GroupCfg& cfg = p.gcfg[group];
        const unsigned nchunks = (unsigned)(cfg.chunk_end - cfg.chunk_begin);
        s_last = atomicAdd(&p.gstate[group].ticket[phase], 1u) == nchunks - 1 ? 1 : 0;
    }
    __syncthreads();
    if (s_last) {
        __threadfence();
        group_step(p, group, mode, sh_tail_scan);
    }
}

__global__ void __launch_bounds__(kThreads) k_group(DevPlan p, int mode) {
    __shared__ unsigned long long sh_scan[kThreads / 32 + 1];
    group_step(p, blockIdx.x, mode, sh_scan);
}

// ---- per chunk counts of keys > T and == T -----------------------------------------------------------
template <typename GradT>
__global__ void __launch_bounds__(kThreads) k_count(DevPlan p) {
    if (p.step_ptr) p.step = *p.step_ptr;
    __shared__ unsigned sh[kThreads / 32];
    const ChunkRec c = p.chunks[blockIdx.x];
    const TensorRec t = p.tensors[c.tensor];
    const unsigned T = group_threshold(p, p.gstate[t.group]);
    const uint64_t salt = group_salt(p, p.gcfg[t.group].uid);
    const int tid = threadIdx.x;
    unsigned cnt = 0;   // gt in low 16 bits, eq in high 16 bits (a thread sees at most 32 elements)
#pragma unroll 2
    for (int it = 0; it < kQuadsPerThread; ++it) {
        const int j = (it * kThreads + tid) * 4;
        if (j >= c.len) break;
        float v[4];
        load_acc<GradT>(t, c.start, j, c.len, v);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (j + k >= c.len) break;
            if (v[k] == 0.0f && !p.values_only) continue;
            const long long i = c.start + j + k;
            const uint32_t key = make_key(p.method, v[k], t.rnd, i, t.goff + i, salt);
            cnt += (key > T) ? 1u : (key == T ? 0x10000u : 0u);
        }
    }
    cnt = warp_sum_u(cnt);
    if ((tid & 31) == 0) sh[tid >> 5] = cnt;
    __syncthreads();
    if (tid == 0) {
        unsigned s = 0;
#pragma unroll
        for (int w = 0; w < kThreads / 32; ++w) s += sh[w];
        p.chunk_cnt[blockIdx.x] = s;
    }
    group_tail(p, t.group, 3, G_SCAN_COUNTS);
}

// Publish "this rank's payload for `step` is complete" to every destination.  Called by all
// threads of every CTA of the sending kernel after their payload stores.
__device__ __forceinline__ void publish_when_last(const DevPlan& p) {
    if (p.direct_idx) return;
    // The CTA barrier orders every thread's payload stores before thread 0's fence; the fence is cumulative, so ONE
    // system-scope fence per CTA publishes them all (a fence in every thread cost a third of k_compact_send's time:
    // 841 k membar.sys for 3288 CTAs, profiles/ncu_exchange_r2.csv).  The multicast (multimem.st) variant keeps the
    // per-thread fence it was validated with on 8 GPUs; the unicast path was re-validated on real GPUs after the change.
    if (p.mc_region) __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        const unsigned ticket = atomicAdd(&p.counters->done, 1u);
        if (ticket == gridDim.x - 1) {
            __threadfence_system();
            p.counters->done = 0;
            for (int d = 0; d < p.world; ++d) st_release_sys(pick_rank(p.flags, d) + p.rank, p.step + 1);
        }
    }
}

// ---- ordered compaction + residual update + stores to every rank ---------------------------------------
template <typename GradT>
__global__ void __launch_bounds__(kThreads, 3) k_compact_send(DevPlan p) {
    if (p.step_ptr) p.step = *p.step_ptr;
    __shared__ uint32_t st_idx[kStage];
    __shared__ float st_val[kStage];
    __shared__ unsigned sh_warp[kThreads / 32];
    const ChunkRec c = p.chunks[blockIdx.x];
    const TensorRec t = p.tensors[c.tensor];
    const GroupCfg cfg = p.gcfg[t.group];
    const GroupState& gs = p.gstate[t.group];
    const unsigned T = group_threshold(p, gs);
    const unsigned long long quota = group_quota(p, gs);
    const unsigned long long cap = (unsigned long long)cfg.cap;
    const uint64_t salt = group_salt(p, cfg.uid);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    const unsigned long long off = p.chunk_off[blockIdx.x];
    const unsigned long long gt_off = off & 0xffffffffULL, eq_off = off >> 32;
    const unsigned packed = p.chunk_cnt[blockIdx.x];
    const unsigned long long cgt = packed & 0xffffu, ceq = packed >> 16;
    auto pos_of = [&](unsigned long long g, unsigned long long e) {
        return g + (e < quota ? e : quota);
    };
    unsigned long long pos_begin = pos_of(gt_off, eq_off);
    unsigned long long pos_end = pos_of(gt_off + cgt, eq_off + ceq);
    if (pos_begin > cap) pos_begin = cap;
    if (pos_end > cap) pos_end = cap;
    const unsigned n_out = (unsigned)(pos_end - pos_begin);
    const bool staged = n_out <= (unsigned)kStage;

    // header: the CTA owning the group's first chunk publishes the entry count to every rank
    if (p.direct_idx) {
        if ((int)blockIdx.x == cfg.chunk_begin && tid == 0) p.direct_count[t.group] = (long long)gs.kept;
    } else if ((int)blockIdx.x == cfg.chunk_begin && p.mc_region) {
        if (tid == 0) mm_st_u64(reinterpret_cast<long long*>(mc_slot_ptr(p, p.rank)) + t.group, (unsigned long long)gs.kept);
    } else if ((int)blockIdx.x == cfg.chunk_begin && tid < p.world) {
        long long* hdr = reinterpret_cast<long long*>(slot_ptr(p, tid, p.rank));
        hdr[t.group] = (long long)gs.kept;
    }
    // Each warp owns a contiguous 1024-element range (8 coalesced iterations of 128 elements); the values and
    // predicate bits stay in registers, so the gradient is read once and the block synchronises twice.
    constexpr int kWarps = kThreads / 32;
    constexpr int kPerWarp = kChunk / kWarps;
    constexpr int kIters = kPerWarp / 128;
    float v[kIters][4];
    unsigned fl[kIters], excl[kIters];
    if (n_out > 0) {
        unsigned run_base = 0;        // packed (gt | eq << 16) count of this warp before iteration `it`
        // all eight loads of the warp's range go out before the first key is computed (the kernel is latency-bound)
#pragma unroll
        for (int it = 0; it < kIters; ++it) {
            const int j = warp * kPerWarp + it * 128 + lane * 4;
            v[it][0] = v[it][1] = v[it][2] = v[it][3] = 0.0f;
            if (j < c.len) load_acc<GradT>(t, c.start, j, c.len, v[it]);
        }
#pragma unroll
        for (int it = 0; it < kIters; ++it) {
            const int j = warp * kPerWarp + it * 128 + lane * 4;
            unsigned f = 0;           // bit k: key > T, bit 4+k: key == T
            if (j < c.len) {
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (j + k >= c.len || (v[it][k] == 0.0f && !p.values_only)) continue;
                    const long long i = c.start + j + k;
                    const uint32_t key = make_key(p.method, v[it][k], t.rnd, i, t.goff + i, salt);
                    f |= (key > T) ? (1u << k) : (key == T ? (16u << k) : 0u);
                }
            }
            fl[it] = f;
            const unsigned mine = __popc(f & 0xfu) | (__popc(f >> 4) << 16);
            const unsigned incl = warp_incl_scan(mine);
            excl[it] = run_base + incl - mine;
            run_base += __shfl_sync(0xffffffffu, incl, 31);
        }
        if (lane == 0) sh_warp[warp] = run_base;
    }
    __syncthreads();
    if (n_out > 0) {
        unsigned wbase = 0;
#pragma unroll
        for (int w = 0; w < kWarps; ++w)
            if (w < warp) wbase += sh_warp[w];
#pragma unroll
        for (int it = 0; it < kIters; ++it) {
            if (!fl[it]) continue;
            const int j = warp * kPerWarp + it * 128 + lane * 4;
            unsigned run = wbase + excl[it];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const bool is_gt = (fl[it] >> k) & 1u, is_eq = (fl[it] >> (4 + k)) & 1u;
                if (!(is_gt || is_eq)) continue;
                const unsigned long long g_before = gt_off + (run & 0xffffu);
                const unsigned long long e_before = eq_off + (run >> 16);
                const unsigned long long pos = pos_of(g_before, e_before);
                const bool keep = pos < cap && (is_gt || e_before < quota);
                run += is_gt ? 1u : 0x10000u;
                if (!keep) continue;
                const long long i = c.start + j + k;
                const uint32_t gidx = (uint32_t)(t.goff + i);
                if (staged) {
                    st_idx[pos - pos_begin] = gidx;           // (unused in values-only mode)
                    st_val[pos - pos_begin] = v[it][k];
                } else if (p.direct_idx) {
                    p.direct_idx[cfg.pay_off + pos] = gidx;
                    p.direct_val[cfg.pay_off + pos] = v[it][k];
                } else if (p.mc_region) {
                    char* s = mc_slot_ptr(p, p.rank);
                    if (!p.values_only) mm_st_u32(reinterpret_cast<uint32_t*>(s + p.idx_off) + cfg.pay_off + pos, gidx);
                    mm_st_f32(reinterpret_cast<float*>(s + p.val_off) + cfg.pay_off + pos, v[it][k]);
                } else {
                    for (int d = 0; d < p.world; ++d) {
                        char* s = slot_ptr(p, d, p.rank);
                        if (!p.values_only) reinterpret_cast<uint32_t*>(s + p.idx_off)[cfg.pay_off + pos] = gidx;
                        reinterpret_cast<float*>(s + p.val_off)[cfg.pay_off + pos] = v[it][k];
                    }
                }
                if (t.resid) t.resid[i] = 0.0f;     // sent exactly: nothing left behind
            }
        }
    }
    __syncthreads();
    if (n_out > 0 && staged && p.mc_region && !p.direct_idx) {
        char* s = mc_slot_ptr(p, p.rank);
        uint32_t* di = reinterpret_cast<uint32_t*>(s + p.idx_off) + cfg.pay_off + pos_begin;
        float* dv = reinterpret_cast<float*>(s + p.val_off) + cfg.pay_off + pos_begin;
        for (unsigned e = tid; e < n_out; e += kThreads) {
            if (!p.values_only) mm_st_u32(di + e, st_idx[e]);
            mm_st_f32(dv + e, st_val[e]);
        }
    } else if (n_out > 0 && staged) {
        const int ndst = p.direct_idx ? 1 : p.world;
        for (int d = 0; d < ndst; ++d) {
            char* s = p.direct_idx ? nullptr : slot_ptr(p, d, p.rank);
            uint32_t* di = (p.direct_idx ? p.direct_idx : reinterpret_cast<uint32_t*>(s + p.idx_off)) + cfg.pay_off + pos_begin;
            float* dv = (p.direct_val ? p.direct_val : reinterpret_cast<float*>(s + p.val_off)) + cfg.pay_off + pos_begin;
            for (unsigned e = tid; e < n_out; e += kThreads) {
                if (!p.values_only) di[e] = st_idx[e];
                dv[e] = st_val[e];
            }
        }
    }
    publish_when_last(p);
}

// ---- fail closed: a flag wait that ran out poisons the step --------------------------------------------------
// One word per device.  A receiver (or the dense all-reduce) whose peer did not show up within the time-out sets it;
// the fused SGD kernels read it (LwcSgdDesc.skip_flag) and leave weights and momentum untouched, so a stalled peer
// costs skipped updates instead of updates computed from stale slots.  The host reads / clears it at its next
// natural synchronisation point (GradientExchanger.check_health).
__device__ unsigned int g_lwc_poison = 0;

// ---- wait for every source's flag (bounded spin) -------------------------------------------------------
__device__ __forceinline__ void wait_sources(const DevPlan& p) {
    if (threadIdx.x < p.world && (int)threadIdx.x != p.rank) {
        const uint64_t* f = pick_rank(p.flags, p.rank) + threadIdx.x;
        const uint64_t want = p.step + 1;
        const uint64_t t0 = global_timer_ns();
        unsigned spins = 0;
        while (ld_acquire_sys(f) < want) {
            if ((++spins & 63u) == 0 && global_timer_ns() - t0 > p.timeout_ns) {
                atomicAdd(&p.counters->timeouts, 1ULL);
                atomicExch(&g_lwc_poison, 1u);
                break;
            }
            __nanosleep(64);
        }
    }
    __syncthreads();
}

// ---- receiver, shared-seed Random-k (values only): every CTA owns one chunk of the dense output.  It recomputes which
// of its elements were selected and at which payload position from the hash keys, the threshold and the chunk offsets of
// this rank's OWN send of the same step (all ranks selected the same positions), then sums that position of every
// source's value array.  No index ever crosses the link: 4 bytes per kept element instead of 8.
template <typename OutT>
__global__ void __launch_bounds__(kThreads) k_recv_values(DevPlan p) {
    if (p.step_ptr) p.step = *p.step_ptr;
    __shared__ unsigned sh_warp[kThreads / 32];
    const ChunkRec c = p.chunks[blockIdx.x];
    const TensorRec t = p.tensors[c.tensor];
    const GroupCfg cfg = p.gcfg[t.group];
    const GroupState& gs = p.gstate[t.group];
    const unsigned T = group_threshold(p, gs);
    const unsigned long long quota = group_quota(p, gs);
    const unsigned long long cap = (unsigned long long)cfg.cap;
    const uint64_t salt = group_salt(p, cfg.uid);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const unsigned long long off = p.chunk_off[blockIdx.x];
    const unsigned long long gt_off = off & 0xffffffffULL, eq_off = off >> 32;
    constexpr int kWarps = kThreads / 32;
    constexpr int kPerWarp = kChunk / kWarps;
    constexpr int kIters = kPerWarp / 128;
    unsigned fl[kIters], excl[kIters];
    unsigned run_base = 0;
#pragma unroll
    for (int it = 0; it < kIters; ++it) {
        const int j = warp * kPerWarp + it * 128 + lane * 4;
        unsigned f = 0;
        if (j < c.len) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (j + k >= c.len) continue;
                const long long i = c.start + j + k;
                const uint32_t key = make_key(p.method, 0.0f, t.rnd, i, t.goff + i, salt);
                f |= (key > T) ? (1u << k) : (key == T ? (16u << k) : 0u);
            }
        }
        fl[it] = f;
        const unsigned mine = __popc(f & 0xfu) | (__popc(f >> 4) << 16);
        const unsigned incl = warp_incl_scan(mine);
        excl[it] = run_base + incl - mine;
        run_base += __shfl_sync(0xffffffffu, incl, 31);
    }
    if (lane == 0) sh_warp[warp] = run_base;
    wait_sources(p);                      // (contains the block barrier that publishes sh_warp)
    unsigned wbase = 0;
#pragma unroll
    for (int w = 0; w < kWarps; ++w)
        if (w < warp) wbase += sh_warp[w];
    const float inv = 1.0f / (float)p.world;
    OutT* o = reinterpret_cast<OutT*>(t.out) + c.start;
    const bool ov = vec_ok(o);
#pragma unroll
    for (int it = 0; it < kIters; ++it) {
        const int j = warp * kPerWarp + it * 128 + lane * 4;
        if (j >= c.len) continue;
        float v[4] = {0.0f, 0.0f, 0.0f, 0.0f};
        unsigned run = wbase + excl[it];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const bool is_gt = (fl[it] >> k) & 1u, is_eq = (fl[it] >> (4 + k)) & 1u;
            if (!(is_gt || is_eq)) continue;
            const unsigned long long g_before = gt_off + (run & 0xffffu);
            const unsigned long long e_before = eq_off + (run >> 16);
            const unsigned long long pos = g_before + (e_before < quota ? e_before : quota);
            const bool keep = pos < cap && (is_gt || e_before < quota);
            run += is_gt ? 1u : 0x10000u;
            if (!keep) continue;
            float a = 0.0f;
            for (int src = 0; src < p.world; ++src)          // rank order: every replica sums in the same order
                a += reinterpret_cast<const float*>(slot_ptr(p, p.rank, src) + p.val_off)[cfg.pay_off + pos];
            v[k] = a * inv;
        }
        store_quad<OutT>(o, j, c.len, ov, v);
    }
    if (p.step_ptr) {      // the last CTA to finish advances the device-side step counter
        __syncthreads();
        if (threadIdx.x == 0) {
            const unsigned ticket = atomicAdd(&p.counters->done_recv, 1u);
            if (ticket == gridDim.x - 1) {
                p.counters->done_recv = 0;
                __threadfence();
                *p.step_ptr = p.step + 1;
            }
        }
    }
}

// ---- receiver, sparse: every CTA owns one chunk of the dense output --------------------------------------
template <typename OutT>
__global__ void __launch_bounds__(kThreads) k_recv_sparse(DevPlan p) {
    if (p.step_ptr) p.step = *p.step_ptr;
    __shared__ float acc[kChunk];
    __shared__ int sh_rng[2];
    const ChunkRec c = p.chunks[blockIdx.x];
    const TensorRec t = p.tensors[c.tensor];
    const GroupCfg cfg = p.gcfg[t.group];
    const int tid = threadIdx.x;
    for (int j = tid; j < kChunk; j += kThreads) acc[j] = 0.0f;
    wait_sources(p);
    const unsigned long long lo = (unsigned long long)(t.goff + c.start), hi = lo + c.len;
    for (int src = 0; src < p.world; ++src) {
        const char* s = slot_ptr(p, p.rank, src);
        long long kept = reinterpret_cast<const long long*>(s)[t.group];
        if (kept > cfg.cap) kept = cfg.cap;
        const uint32_t* idx = reinterpret_cast<const uint32_t*>(s + p.idx_off) + cfg.pay_off;
        const float* val = reinterpret_cast<const float*>(s + p.val_off) + cfg.pay_off;
        if (tid < 2) {     // lower_bound(lo) and lower_bound(hi)
            const unsigned long long key = tid == 0 ? lo : hi;
            long long a = 0, b = kept;
            while (a < b) {
                const long long m = (a + b) >> 1;
                if ((unsigned long long)idx[m] < key) a = m + 1; else b = m;
            }
            sh_rng[tid] = (int)a;
        }
        __syncthreads();
        const int e0 = sh_rng[0], e1 = sh_rng[1];
        for (int e = e0 + tid; e < e1; e += kThreads) acc[idx[e] - (uint32_t)lo] += val[e];
        __syncthreads();
    }
    const float inv = 1.0f / (float)p.world;
    OutT* o = reinterpret_cast<OutT*>(t.out) + c.start;
    const bool ov = vec_ok(o);
    for (int it = 0; it < kQuadsPerThread; ++it) {
        const int j = (it * kThreads + tid) * 4;
        if (j >= c.len) break;
        float v[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = (j + k < c.len) ? acc[j + k] * inv : 0.0f;
        store_quad<OutT>(o, j, c.len, ov, v);
    }
    if (p.step_ptr) {      // the last CTA to finish advances the device-side step counter
        __syncthreads();
        if (threadIdx.x == 0) {
            const unsigned ticket = atomicAdd(&p.counters->done_recv, 1u);
            if (ticket == gridDim.x - 1) {
                p.counters->done_recv = 0;
                __threadfence();
                *p.step_ptr = p.step + 1;
            }
        }
    }
}

// ---- quantisers and the dense path -------------------------------------------------------------------
// Code area of chunk ci inside a slot: code_off + ci * code_stride.
//   dense    : raw GradT values
//   TernGrad : 2 bits per element (0 = zero, 1 = +s, 2 = -s), 4 elements per byte
//   dithering: one level byte per element followed by one sign bit per element
__device__ __forceinline__ void quad_uniforms(const DevPlan& p, const TensorRec& t, long long i, long long gidx,
                                              uint64_t salt, float u[4]) {
    if (t.rnd) {
        const float* r = reinterpret_cast<const float*>(t.rnd);
#pragma unroll
        for (int k = 0; k < 4; ++k) u[k] = (i + k < t.n) ? r[i + k] : 0.0f;
        return;
    }
    const uint64_t q = (uint64_t)gidx >> 2;
    const uint4 w = philox4x32_10(make_uint4((uint32_t)q, (uint32_t)(q >> 32), (uint32_t)p.step, (uint32_t)(p.step >> 32)),
                                  make_uint2((uint32_t)salt, (uint32_t)(salt >> 32)));
    u[0] = u32_to_unit(w.x); u[1] = u32_to_unit(w.y); u[2] = u32_to_unit(w.z); u[3] = u32_to_unit(w.w);
}

template <typename GradT>
__global__ void __launch_bounds__(kThreads) k_quant_send(DevPlan p) {
    if (p.step_ptr) p.step = *p.step_ptr;
    const ChunkRec c = p.chunks[blockIdx.x];
    const TensorRec t = p.tensors[c.tensor];
    const GroupCfg cfg = p.gcfg[t.group];
    const GroupState& gs = p.gstate[t.group];
    const int tid = threadIdx.x, lane = tid & 31;
    const float scale = p.method == LWC_TERNGRAD ? __uint_as_float(gs.maxbits) : gs.scale;   // max|g| is final after k_pre
    const float qf = (float)p.qstates;
    const uint64_t salt = group_salt(p, cfg.uid);
    const unsigned long long coff = p.code_off + (unsigned long long)blockIdx.x * p.code_stride;

    if ((int)blockIdx.x == cfg.chunk_begin && p.mc_region) {
        if (tid == 0) mm_st_f32(reinterpret_cast<float*>(mc_slot_ptr(p, p.rank)) + t.group, scale);
    } else if ((int)blockIdx.x == cfg.chunk_begin && tid < p.world) {
        float* hdr = reinterpret_cast<float*>(slot_ptr(p, tid, p.rank));
        hdr[t.group] = scale;
    }
    for (int it = 0; it < kQuadsPerThread; ++it) {
        const int q = it * kThreads + tid;
        const int j = q * 4;
        if (it * kThreads * 4 >= c.len) break;      // whole warp-row beyond the chunk (uniform)
        float v[4] = {0.f, 0.f, 0.f, 0.f};
        if (j < c.len) load_acc<GradT>(t, c.start, j, c.len, v);
        if (p.method == LWC_NONE) {
            if (j < c.len) {
                if (p.mc_region && sizeof(GradT) == 4 && j + 4 <= c.len) {
                    mm_st_v4(reinterpret_cast<float*>(mc_slot_ptr(p, p.rank) + coff) + j, v[0], v[1], v[2], v[3]);
                } else {
                    for (int d = 0; d < p.world; ++d) {
                        GradT* dst = reinterpret_cast<GradT*>(slot_ptr(p, d, p.rank) + coff);
                        store_quad<GradT>(dst, j, c.len, true, v);
                    }
                }
            }
            continue;
        }
        float u[4] = {1.f, 1.f, 1.f, 1.f};
        if (j < c.len) quad_uniforms(p, t, c.start + j, t.goff + c.start + j, salt, u);
        float deq[4];
        unsigned code = 0, sign = 0;
        if (p.method == LWC_TERNGRAD) {
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float a = fabsf(v[k]);
                const bool b = (scale > 0.0f) && (u[k] < __fdiv_rn(a, scale)) && (j + k < c.len);
                const unsigned cde = b ? (v[k] < 0.0f ? 2u : 1u) : 0u;
                code |= cde << (2 * k);
                deq[k] = b ? (v[k] < 0.0f ? -scale : scale) : 0.0f;
            }
            // four lanes -> one 32-bit word (16 elements)
            unsigned word = code << (8 * (lane & 3));
            word |= __shfl_xor_sync(0xffffffffu, word, 1);
            word |= __shfl_xor_sync(0xffffffffu, word, 2);
            if ((lane & 3) == 0 && j < c.len) {
                if (p.mc_region) {
                    mm_st_u32(reinterpret_cast<uint32_t*>(mc_slot_ptr(p, p.rank) + coff) + (q >> 2), word);
                } else {
                    for (int d = 0; d < p.world; ++d)
                        reinterpret_cast<uint32_t*>(slot_ptr(p, d, p.rank) + coff)[q >> 2] = word;
                }
            }
        } else {   // LWC_DITHERING
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                float level = 0.0f;
                if (scale > 0.0f && j + k < c.len)
                    level = floorf(__fadd_rn(__fmul_rn(__fdiv_rn(fabsf(v[k]), scale), qf), u[k]));
                level = fminf(level, 255.0f);
                code |= ((unsigned)level) << (8 * k);
                sign |= (v[k] < 0.0f ? 1u : 0u) << k;
                const float mag = __fmul_rn(scale, __fdiv_rn(level, qf));
                deq[k] = v[k] < 0.0f ? -mag : (v[k] > 0.0f ? mag : 0.0f);
            }
            // eight lanes -> one 32-bit word of sign bits (32 elements)
            unsigned sw = sign << (4 * (lane & 7));
            sw |= __shfl_xor_sync(0xffffffffu, sw, 1);
            sw |= __shfl_xor_sync(0xffffffffu, sw, 2);
            sw |= __shfl_xor_sync(0xffffffffu, sw, 4);
            if (j < c.len) {
                if (p.mc_region) {
                    char* base = mc_slot_ptr(p, p.rank) + coff;
                    mm_st_u32(reinterpret_cast<uint32_t*>(base) + q, code);
                    if ((lane & 7) == 0) mm_st_u32(reinterpret_cast<uint32_t*>(base + kChunk) + (q >> 3), sw);
                } else {
                    for (int d = 0; d < p.world; ++d) {
                        char* base = slot_ptr(p, d, p.rank) + coff;
                        reinterpret_cast<uint32_t*>(base)[q] = code;
                        if ((lane & 7) == 0) reinterpret_cast<uint32_t*>(base + kChunk)[q >> 3] = sw;
                    }
                }
            }
        }
        if (t.resid && j < c.len) {
            float e[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) e[k] = v[k] - deq[k];
            float* r = t.resid + c.start;
            store_quad<float>(r, j, c.len, vec_ok(r), e);
        }
    }
    publish_when_last(p);
}

template <typename GradT, typename OutT>
__global__ void __launch_bounds__(kThreads) k_recv_dense(DevPlan p) {
    if (p.step_ptr) p.step = *p.step_ptr;
    const ChunkRec c = p.chunks[blockIdx.x];
    const TensorRec t = p.tensors[c.tensor];
    const int tid = threadIdx.x;
    const float qf = (float)p.qstates;
    const unsigned long long coff = p.code_off + (unsigned long long)blockIdx.x * p.code_stride;
    wait_sources(p);
    const float inv = 1.0f / (float)p.world;
    OutT* o = reinterpret_cast<OutT*>(t.out) + c.start;
    const bool ov = vec_ok(o);
    for (int it = 0; it < kQuadsPerThread; ++it) {
        const int q = it * kThreads + tid;
        const int j = q * 4;
        if (j >= c.len) break;
        float sum[4] = {0.f, 0.f, 0.f, 0.f};
        for (int src = 0; src < p.world; ++src) {
            const char* s = slot_ptr(p, p.rank, src);
            const char* base = s + coff;
            if (p.method == LWC_NONE) {
                float v[4];
                load_quad<GradT>(reinterpret_cast<const GradT*>(base), j, c.len, true, v);
#pragma unroll
                for (int k = 0; k < 4; ++k) sum[k] += v[k];
            } else {
                const float scale = reinterpret_cast<const float*>(s)[t.group];
                if (p.method == LWC_TERNGRAD) {
                    const unsigned code = reinterpret_cast<const uint8_t*>(base)[q];
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const unsigned cde = (code >> (2 * k)) & 3u;
                        sum[k] += cde == 1u ? scale : (cde == 2u ? -scale : 0.0f);
                    }
                } else {
                    const unsigned code = reinterpret_cast<const uint32_t*>(base)[q];
                    const unsigned sb = (reinterpret_cast<const uint8_t*>(base + kChunk)[q >> 1] >> (4 * (q & 1))) & 0xfu;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const float level = (float)((code >> (8 * k)) & 0xffu);
                        const float mag = __fmul_rn(scale, __fdiv_rn(level, qf));
                        sum[k] += ((sb >> k) & 1u) ? -mag : mag;
                    }
                }
            }
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) sum[k] *= inv;
        store_quad<OutT>(o, j, c.len, ov, sum);
    }
    if (p.step_ptr) {      // the last CTA to finish advances the device-side step counter
        __syncthreads();
        if (threadIdx.x == 0) {
            const unsigned ticket = atomicAdd(&p.counters->done_recv, 1u);
            if (ticket == gridDim.x - 1) {
                p.counters->done_recv = 0;
                __threadfence();
                *p.step_ptr = p.step + 1;
            }
        }
    }
}

}  // namespace lwc

// =====================================================================================================
// Host side
// =====================================================================================================
using namespace lwc;

// Upper bound of every wait on a peer's flag (default 10 s).  A rank whose peer died or hung leaves its kernels
// after this long, counts the event (stats()[2]) and lets the host decide -- it never spins forever.
static unsigned long long g_flag_timeout_ns = 10000000000ULL;
extern "C" int lwc_set_flag_timeout_ms(unsigned int ms) {
    g_flag_timeout_ns = (unsigned long long)(ms ? ms : 1) * 1000000ULL;
    return 0;
}

struct LwcPlan {
    DevPlan dev;
    bool device_step = false;
    void* h_pinned = nullptr;      // pinned staging of the tensor table (graph-capturable uploads)
    int num_tensors;
    int grad_dtype, out_dtype;
    bool any_resid;
    bool bound;
    size_t state_bytes;
    std::vector<TensorRec> h_tensors;
    void* d_tensors = nullptr;
    void* d_chunks = nullptr;
    void* d_gcfg = nullptr;
    void* d_gstate = nullptr;
    void* d_chunk_cnt = nullptr;
    void* d_chunk_off = nullptr;
    void* d_chunk_part = nullptr;
    void* d_counters = nullptr;
};

static size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }
static size_t dtype_size(int dt) { return dt == LWC_F32 ? 4 : 2; }

extern "C" int lwc_plan_create(LwcPlan** out, const LwcPlanDesc* d) {
    if (!out || !d || d->num_tensors <= 0 || d->num_groups <= 0) return LWC_EINVAL;
    if (d->world < 1 || d->world > LWC_MAX_WORLD || d->rank < 0 || d->rank >= d->world) return LWC_EINVAL;
    if (d->method < LWC_NONE || d->method > LWC_DITHERING) return LWC_EINVAL;
    if (d->method == LWC_DITHERING && (d->qstates < 1 || d->qstates > 255)) return LWC_EINVAL;
    LwcPlan* P = new LwcPlan();
    P->num_tensors = d->num_tensors;
    P->grad_dtype = d->grad_dtype;
    P->out_dtype = d->out_dtype;
    P->bound = false;
    P->any_resid = false;
    const int G = d->num_groups;
    std::vector<GroupCfg> gcfg(G);
    for (int g = 0; g < G; ++g) {
        gcfg[g] = GroupCfg{0, d->keep ? d->keep[g] : 0, d->cap ? d->cap[g] : 0, 0, 0, 0, 0, 0};
    }
    std::vector<ChunkRec> chunks;
    P->h_tensors.resize(d->num_tensors);
    int prev_group = -1;
    for (int t = 0; t < d->num_tensors; ++t) {
        const int g = d->group_of[t];
        if (g < prev_group || g >= G || g > prev_group + 1) { delete P; return LWC_EINVAL; }
        TensorRec& r = P->h_tensors[t];
        r.grad = d->grad_dev[t];
        r.out = d->out_dev[t];
        r.resid = (d->resid_dev && d->method != LWC_NONE) ? d->resid_dev[t] : nullptr;
        r.rnd = nullptr;
        r.n = d->numel[t];
        r.goff = gcfg[g].n;
        r.group = g;
        r.uid = d->uid ? d->uid[t] : t;
        if (r.resid) P->any_resid = true;
        if (g != prev_group) {
            gcfg[g].chunk_begin = (int)chunks.size();
            gcfg[g].uid = r.uid;
            prev_group = g;
        }
        for (long long s = 0; s < r.n; s += kChunk) {
            ChunkRec c;
            c.tensor = t;
            c.start = s;
            c.len = (int)((r.n - s) < kChunk ? (r.n - s) : kChunk);
            chunks.push_back(c);
        }
        gcfg[g].n += r.n;
        gcfg[g].chunk_end = (int)chunks.size();
        if (gcfg[g].n > 0xffffffffLL) { delete P; return LWC_ETOOBIG; }
    }
    if (prev_group != G - 1 || chunks.empty()) { delete P; return LWC_EINVAL; }
    long long pay = 0;
    for (int g = 0; g < G; ++g) {
        gcfg[g].pay_off = pay;
        if (gcfg[g].cap > gcfg[g].n) gcfg[g].cap = gcfg[g].n;
        pay += (gcfg[g].cap + 3) / 4 * 4;
    }
    DevPlan& D = P->dev;
    memset(&D, 0, sizeof(D));
    D.num_chunks = (int)chunks.size();
    D.num_groups = G;
    D.method = d->method;
    D.world = d->world;
    D.rank = d->rank;
    D.qstates = d->qstates;
    D.threshold = d->threshold;
    D.seed = d->seed;
    const bool sparse = d->method >= LWC_TOPK && d->method <= LWC_ADAPTIVE;
    D.hdr_bytes = align_up((size_t)G * 8, 256);
    if (sparse) {
        D.idx_off = D.hdr_bytes;
        D.val_off = D.idx_off + align_up((size_t)pay * 4, 256);
        D.slot_bytes = D.val_off + align_up((size_t)pay * 4, 256);
    } else {
        D.code_off = D.hdr_bytes;
        D.code_stride = d->method == LWC_NONE ? (size_t)kChunk * dtype_size(d->grad_dtype)
                        : d->method == LWC_TERNGRAD ? kChunk / 4 : kChunk + kChunk / 8;
        D.slot_bytes = D.code_off + (size_t)D.num_chunks * D.code_stride;
    }
    P->state_bytes = sizeof(GroupState) * (size_t)G;
    cudaError_t e = cudaSuccess;
    auto alloc = [&](void** p, size_t bytes) { if (e == cudaSuccess) e = cudaMalloc(p, bytes ? bytes : 16); };
    alloc(&P->d_tensors, sizeof(TensorRec) * P->h_tensors.size());
    alloc(&P->d_chunks, sizeof(ChunkRec) * chunks.size());
    alloc(&P->d_gcfg, sizeof(GroupCfg) * G);
    alloc(&P->d_gstate, P->state_bytes);
    alloc(&P->d_chunk_cnt, sizeof(unsigned) * chunks.size());
    alloc(&P->d_chunk_off, sizeof(unsigned long long) * chunks.size());
    alloc(&P->d_chunk_part, sizeof(float) * chunks.size());
    alloc(&P->d_counters, sizeof(Counters));
    if (e == cudaSuccess) e = cudaMemcpy(P->d_tensors, P->h_tensors.data(), sizeof(TensorRec) * P->h_tensors.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(P->d_chunks, chunks.data(), sizeof(ChunkRec) * chunks.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemcpy(P->d_gcfg, gcfg.data(), sizeof(GroupCfg) * G, cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMemset(P->d_gstate, 0, P->state_bytes);
    if (e == cudaSuccess) e = cudaMemset(P->d_counters, 0, sizeof(Counters));
    if (e != cudaSuccess) { lwc_plan_destroy(P); return (int)e; }
    D.tensors = (const TensorRec*)P->d_tensors;
    D.chunks = (const ChunkRec*)P->d_chunks;
    D.gcfg = (const GroupCfg*)P->d_gcfg;
    D.gstate = (GroupState*)P->d_gstate;
    D.chunk_cnt = (unsigned*)P->d_chunk_cnt;
    D.chunk_off = (unsigned long long*)P->d_chunk_off;
    D.chunk_part = (float*)P->d_chunk_part;
    D.counters = (Counters*)P->d_counters;
    *out = P;
    return 0;
}

extern "C" int lwc_plan_destroy(LwcPlan* P) {
    if (!P) return 0;
    cudaFree(P->d_tensors); cudaFree(P->d_chunks); cudaFree(P->d_gcfg); cudaFree(P->d_gstate);
    cudaFree(P->d_chunk_cnt); cudaFree(P->d_chunk_off); cudaFree(P->d_chunk_part); cudaFree(P->d_counters);
    if (P->h_pinned) cudaFreeHost(P->h_pinned);
    delete P;
    return 0;
}

extern "C" size_t lwc_plan_slot_bytes(const LwcPlan* P) { return P ? (size_t)P->dev.slot_bytes : 0; }
extern "C" size_t lwc_plan_region_bytes(const LwcPlan* P) {
    return P ? (size_t)P->dev.slot_bytes * 2 * P->dev.world : 0;
}

extern "C" int lwc_plan_bind(LwcPlan* P, void* const* region_dev, void* const* flags_dev) {
    if (!P || !region_dev || !flags_dev) return LWC_EINVAL;
    for (int d = 0; d < P->dev.world; ++d) {
        if (!region_dev[d] || !flags_dev[d]) return LWC_EINVAL;
        P->dev.region[d] = (char*)region_dev[d];
        P->dev.flags[d] = (uint64_t*)flags_dev[d];
    }
    P->bound = true;
    return 0;
}

// NVLS variant: mc_region_dev is the multicast mapping of the SAME symmetric region that lwc_plan_bind received
// as unicast pointers (all ranks bound to one multicast object at identical offsets).  NULL switches back.
extern "C" int lwc_plan_bind_multicast(LwcPlan* P, void* mc_region_dev) {
    if (!P) return LWC_EINVAL;
    P->dev.mc_region = (char*)mc_region_dev;
    return 0;
}

// Shared-seed Random-k ships values only (4 bytes per kept element; indices are implied by the hash).  Call on every rank
// alike, before the first send; injected random keys (lwc_plan_set_random) switch it off again.
extern "C" int lwc_plan_set_values_only(LwcPlan* P, int enable) {
    if (!P) return LWC_EINVAL;
    if (enable && (P->dev.method != LWC_RANDOMK || P->dev.direct_idx)) return LWC_EINVAL;
    P->dev.values_only = enable ? 1 : 0;
    return 0;
}

extern "C" int lwc_plan_set_random(LwcPlan* P, const void* const* rand_dev) {
    if (!P) return LWC_EINVAL;
    if (rand_dev) P->dev.values_only = 0;
    for (int t = 0; t < P->num_tensors; ++t) P->h_tensors[t].rnd = rand_dev ? rand_dev[t] : nullptr;
    LWC_CUDA_CHECK(cudaMemcpy(P->d_tensors, P->h_tensors.data(), sizeof(TensorRec) * P->h_tensors.size(),
                              cudaMemcpyHostToDevice));
    return 0;
}

#define LWC_DISPATCH_GRAD(dt, ...)                                            \
    switch (dt) {                                                             \
        case LWC_F32: { using GradT = float; __VA_ARGS__; } break;            \
        case LWC_BF16: { using GradT = __nv_bfloat16; __VA_ARGS__; } break;   \
        case LWC_F16: { using GradT = __half; __VA_ARGS__; } break;           \
        default: return LWC_EINVAL;                                           \
    }

extern "C" int lwc_plan_send(LwcPlan* P, uint64_t step, void* stream_) {
    if (!P || !P->bound) return LWC_ESTATE;
    cudaStream_t st = (cudaStream_t)stream_;
    DevPlan D = P->dev;
    D.step = step;
    D.step_ptr = P->device_step ? &reinterpret_cast<Counters*>(P->d_counters)->dev_step : nullptr;
    D.timeout_ns = g_flag_timeout_ns;
    const int NC = D.num_chunks, G = D.num_groups, m = D.method;
    LWC_CUDA_CHECK(cudaMemsetAsync(P->d_gstate, 0, P->state_bytes, st));
    // Every pass is one launch: the group bookkeeping that used to sit between the passes as a kernel of its own runs in
    // the chunk CTA that finishes the group last (group_tail).  Top-k / Random-k: 5 kernels, was 10.
    switch (m) {
        case LWC_TOPK:
        case LWC_RANDOMK:
            LWC_DISPATCH_GRAD(P->grad_dtype, k_pre<GradT, STAT_HIST><<<NC, kThreads, 0, st>>>(D));
            LWC_DISPATCH_GRAD(P->grad_dtype, k_hist<GradT, 2><<<NC, kThreads, 0, st>>>(D));
            LWC_DISPATCH_GRAD(P->grad_dtype, k_hist<GradT, 3><<<NC, kThreads, 0, st>>>(D));
            break;
        case LWC_THRESHOLDV:
            if (P->any_resid) { LWC_DISPATCH_GRAD(P->grad_dtype, k_pre<GradT, STAT_NONE><<<NC, kThreads, 0, st>>>(D)); }
            break;
        case LWC_ADAPTIVE:
        case LWC_TERNGRAD:
            LWC_DISPATCH_GRAD(P->grad_dtype, k_pre<GradT, STAT_MAX><<<NC, kThreads, 0, st>>>(D));
            break;
        case LWC_DITHERING:
            LWC_DISPATCH_GRAD(P->grad_dtype, k_pre<GradT, STAT_SUMSQ><<<NC, kThreads, 0, st>>>(D));
            break;
        default: break;
    }
    if (m >= LWC_TOPK && m <= LWC_ADAPTIVE) {
        LWC_DISPATCH_GRAD(P->grad_dtype, k_count<GradT><<<NC, kThreads, 0, st>>>(D));
        LWC_DISPATCH_GRAD(P->grad_dtype, k_compact_send<GradT><<<NC, kThreads, 0, st>>>(D));
    } else {
        LWC_DISPATCH_GRAD(P->grad_dtype, k_quant_send<GradT><<<NC, kThreads, 0, st>>>(D));
    }
    (void)G;
    return (int)cudaGetLastError();
}

template <typename GradT>
static int launch_recv_dense(const DevPlan& D, int out_dtype, cudaStream_t st) {
    switch (out_dtype) {
        case LWC_F32: k_recv_dense<GradT, float><<<D.num_chunks, kThreads, 0, st>>>(D); break;
        case LWC_BF16: k_recv_dense<GradT, __nv_bfloat16><<<D.num_chunks, kThreads, 0, st>>>(D); break;
        case LWC_F16: k_recv_dense<GradT, __half><<<D.num_chunks, kThreads, 0, st>>>(D); break;
        default: return LWC_EINVAL;
    }
    return 0;
}

extern "C" int lwc_plan_recv(LwcPlan* P, uint64_t step, void* stream_) {
    if (!P || !P->bound) return LWC_ESTATE;
    cudaStream_t st = (cudaStream_t)stream_;
    DevPlan D = P->dev;
    D.step = step;
    D.step_ptr = P->device_step ? &reinterpret_cast<Counters*>(P->d_counters)->dev_step : nullptr;
    D.timeout_ns = g_flag_timeout_ns;
    const int m = D.method;
    if (m == LWC_RANDOMK && D.values_only) {
        switch (P->out_dtype) {
            case LWC_F32: k_recv_values<float><<<D.num_chunks, kThreads, 0, st>>>(D); break;
            case LWC_BF16: k_recv_values<__nv_bfloat16><<<D.num_chunks, kThreads, 0, st>>>(D); break;
            case LWC_F16: k_recv_values<__half><<<D.num_chunks, kThreads, 0, st>>>(D); break;
            default: return LWC_EINVAL;
        }
    } else if (m >= LWC_TOPK && m <= LWC_ADAPTIVE) {
        switch (P->out_dtype) {
            case LWC_F32: k_recv_sparse<float><<<D.num_chunks, kThreads, 0, st>>>(D); break;
            case LWC_BF16: k_recv_sparse<__nv_bfloat16><<<D.num_chunks, kThreads, 0, st>>>(D); break;
            case LWC_F16: k_recv_sparse<__half><<<D.num_chunks, kThreads, 0, st>>>(D); break;
            default: return LWC_EINVAL;
        }
    } else {
        int rc = 0;
        LWC_DISPATCH_GRAD(P->grad_dtype, rc = launch_recv_dense<GradT>(D, P->out_dtype, st));
        if (rc) return rc;
    }
    return (int)cudaGetLastError();
}

extern "C" int lwc_plan_stats(LwcPlan* P, uint64_t out[4], void* stream_) {
    if (!P || !out) return LWC_EINVAL;
    Counters c;
    cudaStream_t st = (cudaStream_t)stream_;
    LWC_CUDA_CHECK(cudaMemcpyAsync(&c, P->d_counters, sizeof(c), cudaMemcpyDeviceToHost, st));
    LWC_CUDA_CHECK(cudaStreamSynchronize(st));
    out[0] = c.sent; out[1] = c.dropped; out[2] = c.timeouts;
    const bool sparse = P->dev.method >= LWC_TOPK && P->dev.method <= LWC_ADAPTIVE;
    out[3] = sparse ? c.sent * (P->dev.values_only ? 4 : 8) : (uint64_t)(P->dev.slot_bytes - P->dev.hdr_bytes);
    return 0;
}


// =====================================================================================================
// Stand-alone exact Top-k selection of one fp32 vector, scratch supplied by the caller.
// =====================================================================================================
namespace lwc {
struct TopkHeader {
    TensorRec t;
    GroupCfg g;
    Counters c;
    GroupState gs;
};
__host__ __device__ inline size_t topk_align(size_t x) { return (x + 255) / 256 * 256; }
struct TopkLayout {
    size_t chunks, cnt, off, total;
    __host__ __device__ explicit TopkLayout(long long n) {
        const size_t nc = (size_t)((n + kChunk - 1) / kChunk);
        chunks = topk_align(sizeof(TopkHeader));
        cnt = chunks + topk_align(nc * sizeof(ChunkRec));
        off = cnt + topk_align(nc * sizeof(unsigned));
        total = off + topk_align(nc * sizeof(unsigned long long));
    }
};
__global__ void k_topk_init(char* ws, const float* x, long long n, long long keep, long long cap) {
    const TopkLayout L(n);
    const int nc = (int)((n + kChunk - 1) / kChunk);
    TopkHeader* H = reinterpret_cast<TopkHeader*>(ws);
    ChunkRec* ch = reinterpret_cast<ChunkRec*>(ws + L.chunks);
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < nc) {
        const long long s = (long long)i * kChunk;
        ch[i].tensor = 0;
        ch[i].start = s;
        ch[i].len = (int)((n - s) < kChunk ? (n - s) : kChunk);
    }
    if (i == 0) {
        H->t = TensorRec{x, nullptr, nullptr, nullptr, n, 0, 0, 0};
        H->g = GroupCfg{n, keep, cap < n ? cap : n, 0, 0, nc, 0, 0};
        H->c = Counters{0, 0, 0, 0, 0, 0};
    }
}
}  // namespace lwc

extern "C" size_t lwc_topk_workspace_bytes(int64_t n) {
    if (n <= 0) return 0;
    return TopkLayout(n).total;
}

// Shared by the public entry point (caller-provided scratch) and any variant that brings its own.
int lwc_topk_select_impl(const float* x, int64_t n, int64_t keep, int64_t cap, uint32_t* idx_out, float* val_out,
                         int64_t* count_out, void* ws_, cudaStream_t st) {
    if (!x || !idx_out || !val_out || !count_out || !ws_ || n <= 0 || keep < 0 || cap <= 0) return LWC_EINVAL;
    if (n > 0xffffffffLL) return LWC_ETOOBIG;
    char* ws = (char*)ws_;
    const TopkLayout L(n);
    const int nc = (int)((n + kChunk - 1) / kChunk);
    TopkHeader* H = reinterpret_cast<TopkHeader*>(ws);
    LWC_CUDA_CHECK(cudaMemsetAsync(&H->gs, 0, sizeof(GroupState), st));
    k_topk_init<<<(nc + 255) / 256, 256, 0, st>>>(ws, x, n, keep, cap);
    DevPlan D;
    memset(&D, 0, sizeof(D));
    D.tensors = &H->t;
    D.chunks = reinterpret_cast<const ChunkRec*>(ws + L.chunks);
    D.gcfg = &H->g;
    D.gstate = &H->gs;
    D.chunk_cnt = reinterpret_cast<unsigned*>(ws + L.cnt);
    D.chunk_off = reinterpret_cast<unsigned long long*>(ws + L.off);
    D.chunk_part = nullptr;
    D.counters = &H->c;
    D.num_chunks = nc;
    D.num_groups = 1;
    D.method = LWC_TOPK;
    D.world = 1;
    D.rank = 0;
    D.direct_idx = idx_out;
    D.direct_val = val_out;
    D.direct_count = reinterpret_cast<long long*>(count_out);
    k_pre<float, STAT_HIST><<<nc, kThreads, 0, st>>>(D);
    k_hist<float, 2><<<nc, kThreads, 0, st>>>(D);
    k_hist<float, 3><<<nc, kThreads, 0, st>>>(D);
    k_count<float><<<nc, kThreads, 0, st>>>(D);
    k_compact_send<float><<<nc, kThreads, 0, st>>>(D);
    return (int)cudaGetLastError();
}

extern "C" int lwc_topk_select_f32(const float* x_dev, int64_t n, int64_t keep, int64_t cap, uint32_t* idx_out_dev,
                                   float* val_out_dev, int64_t* count_out_dev, void* workspace_dev, void* stream) {
    return lwc_topk_select_impl(x_dev, n, keep, cap, idx_out_dev, val_out_dev, count_out_dev, workspace_dev,
                                (cudaStream_t)stream);
}

// =====================================================================================================
// Top-k selection for callers that must not allocate device memory.
//
// Who needs this: code that runs where a device allocation is off the table -- a host thread inside a
// serving or training loop whose memory pool already owns all of HBM (so cudaMalloc would fail or steal
// from the pool), a plug-in that is handed streams and pointers but is forbidden from touching the
// allocator, or a latency-critical section that cannot afford cudaMalloc's implicit device-wide
// synchronisation.  Such a caller has nowhere to get the selection scratch (histograms, per-chunk
// counters, chunk table) from, so this variant keeps it in a fixed array inside the library image.
//
// The scratch lives in device memory of every GPU that loads the library, for the life of the process
// (LWC_TOPK_NOALLOC_SCRATCH_BYTES each).  It bounds the problem size: n <= LWC_TOPK_NOALLOC_MAX_N.
// =====================================================================================================
static __device__ __align__(256) unsigned char g_topk_noalloc_scratch[LWC_TOPK_NOALLOC_SCRATCH_BYTES];

namespace {
constexpr int kMaxDevices = 64;
void* g_noalloc_scratch_ptr[kMaxDevices] = {nullptr};   // device address of the array, per device
}  // namespace

// In host code `g_topk_noalloc_scratch` names a host-side shadow of the symbol, not device memory: the
// real address has to be asked for, once per device, while that device is current.
static int noalloc_scratch_for_current_device(void** out) {
    int dev = -1;
    LWC_CUDA_CHECK(cudaGetDevice(&dev));
    if (dev < 0 || dev >= kMaxDevices) return LWC_EINVAL;
    if (!g_noalloc_scratch_ptr[dev]) {
        void* ptr = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&ptr, g_topk_noalloc_scratch);
        if (e != cudaSuccess) return (int)e;
        if (!ptr) return LWC_ESTATE;
        g_noalloc_scratch_ptr[dev] = ptr;
    }
    *out = g_noalloc_scratch_ptr[dev];
    return 0;
}

extern "C" int lwc_topk_select_f32_noalloc(const float* x_dev, int64_t n, int64_t keep, int64_t cap,
                                           uint32_t* idx_out_dev, float* val_out_dev, int64_t* count_out_dev,
                                           void* stream) {
    if (n <= 0) return LWC_EINVAL;
    if (n > (int64_t)LWC_TOPK_NOALLOC_MAX_N || TopkLayout(n).total > (size_t)LWC_TOPK_NOALLOC_SCRATCH_BYTES)
        return LWC_ETOOBIG;      // "size beyond the supported maximum" (see lwc_error_string)
    void* scratch = nullptr;
    const int rc = noalloc_scratch_for_current_device(&scratch);
    if (rc) return rc;
    // lwc_topk_select_impl re-initialises everything it reads before use (selection state is memset,
    // header / chunk table / counters are rewritten by k_topk_init, per-chunk counts are rewritten for
    // every chunk of this call), so whatever the previous call left in the array is harmless.
    return lwc_topk_select_impl(x_dev, n, keep, cap, idx_out_dev, val_out_dev, count_out_dev, scratch,
                                (cudaStream_t)stream);
}

// =====================================================================================================
// Dense two-shot all-reduce over peer memory (the `--compress none` / DDP path for large dense models).
//
// Every rank's gradient arena lives in the peer-mapped heap.  Rank r owns slice r: it loads that slice
// from all W arenas (P2P loads over NVLink), sums in rank order, scales by 1/W and stores the result
// into slice r of all W arenas (P2P stores) -- reduce-scatter and all-gather in one kernel, 2(W-1)/W of
// the data per direction instead of the (W-1)x of the one-shot slots.  Two flag rounds: READY ("my
// gradients are final", written at kernel start) and DONE ("my slice has landed everywhere").
// Reference op: per-tensor dist.all_reduce + /W (CIFAR10/core.py:316-319; train_imagenet_nv.py:382-386).
// =====================================================================================================
namespace lwc {
struct DevAR {
    char* arena[LWC_MAX_WORLD];
    uint64_t* flags[LWC_MAX_WORLD];     // per rank: [2][LWC_MAX_WORLD] (ready row, done row)
    unsigned long long* counters;       // [0] ticket, [1] timeouts   (local)
    long long off, n;                   // element range [off, off + n) of the arena
    int world, rank;
    unsigned long long step;
    unsigned long long timeout_ns;
};

__device__ __forceinline__ void ar_wait(const DevAR& p, int row) {
    if (threadIdx.x < p.world && (int)threadIdx.x != p.rank) {
        const uint64_t* f = pick_rank(p.flags, p.rank) + row * LWC_MAX_WORLD + threadIdx.x;
        const uint64_t want = p.step + 1;
        const uint64_t t0 = global_timer_ns();
        unsigned spins = 0;
        while (ld_acquire_sys(f) < want) {
            if ((++spins & 63u) == 0 && global_timer_ns() - t0 > p.timeout_ns) {
                atomicAdd(p.counters + 1, 1ULL);
                atomicExch(&g_lwc_poison, 1u);
                break;
            }
            __nanosleep(64);
        }
    }
    __syncthreads();
}

template <typename T>
__global__ void __launch_bounds__(512) k_allreduce_two_shot(DevAR p) {
    constexpr int V = 16 / sizeof(T);         // elements per 16-byte vector
    if (blockIdx.x == 0 && threadIdx.x < p.world)
        st_release_sys(pick_rank(p.flags, (int)threadIdx.x) + 0 * LWC_MAX_WORLD + p.rank, p.step + 1);     // READY
    ar_wait(p, 0);
    // slice of this rank, in vectors
    const long long nvec = (p.n + V - 1) / V;                 // arena padding makes the tail vector addressable
    const long long per = (nvec + p.world - 1) / p.world;
    const long long v0 = per * p.rank, v1 = min(nvec, v0 + per);
    const float inv = 1.0f / (float)p.world;
    for (long long v = v0 + (long long)blockIdx.x * blockDim.x + threadIdx.x; v < v1; v += (long long)gridDim.x * blockDim.x) {
        const size_t byte = ((size_t)p.off * sizeof(T)) + (size_t)v * 16;
        float acc[V];
#pragma unroll
        for (int k = 0; k < V; ++k) acc[k] = 0.0f;
        uint4 raw[LWC_MAX_WORLD];
#pragma unroll
        for (int s = 0; s < LWC_MAX_WORLD; ++s)
            if (s < p.world) raw[s] = *reinterpret_cast<const uint4*>(p.arena[s] + byte);
#pragma unroll
        for (int s = 0; s < LWC_MAX_WORLD; ++s) {
            if (s >= p.world) break;
            const T* e = reinterpret_cast<const T*>(&raw[s]);
#pragma unroll
            for (int k = 0; k < V; ++k) acc[k] += DT<T>::ld(e + k);
        }
        uint4 outv;
        T* o = reinterpret_cast<T*>(&outv);
#pragma unroll
        for (int k = 0; k < V; ++k) DT<T>::st(o + k, acc[k] * inv);
#pragma unroll
        for (int d = 0; d < LWC_MAX_WORLD; ++d)
            if (d < p.world) *reinterpret_cast<uint4*>(p.arena[d] + byte) = outv;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();                 // cumulative: covers the whole CTA's stores (ordered by the barrier)
        const unsigned long long ticket = atomicAdd(p.counters, 1ULL);
        if (ticket == gridDim.x - 1) {
            __threadfence_system();
            p.counters[0] = 0;
            for (int d = 0; d < p.world; ++d) st_release_sys(pick_rank(p.flags, d) + 1 * LWC_MAX_WORLD + p.rank, p.step + 1);   // DONE
        }
    }
}

__global__ void k_allreduce_wait_done(DevAR p) { ar_wait(p, 1); }
}  // namespace lwc

// arena_dev[r] / flags_dev[r]: rank r's arena base and flag block (2 * LWC_MAX_WORLD uint64, zeroed) as seen from
// this rank; counters_dev: 2 local uint64 (zeroed).  The arena allocation must be padded to a multiple of 16 bytes
// past off + n.  In-place: on return (stream order) every arena holds the average.
extern "C" int lwc_dense_allreduce(void* const* arena_dev, void* const* flags_dev, void* counters_dev, int64_t off,
                                   int64_t n, int dtype, int world, int rank, uint64_t step, void* stream) {
    if (!arena_dev || !flags_dev || !counters_dev || n <= 0 || world < 1 || world > LWC_MAX_WORLD || rank < 0 || rank >= world)
        return LWC_EINVAL;
    const size_t esz = dtype == LWC_F32 ? 4 : 2;
    if (((size_t)off * esz) & 15) return LWC_EINVAL;
    DevAR p;
    memset(&p, 0, sizeof(p));
    for (int r = 0; r < world; ++r) {
        if (!arena_dev[r] || !flags_dev[r]) return LWC_EINVAL;
        p.arena[r] = (char*)arena_dev[r];
        p.flags[r] = (uint64_t*)flags_dev[r];
    }
    p.counters = (unsigned long long*)counters_dev;
    p.off = off; p.n = n; p.world = world; p.rank = rank; p.step = step; p.timeout_ns = g_flag_timeout_ns;
    cudaStream_t st = (cudaStream_t)stream;
    const long long nvec = ((long long)n * (long long)esz + 15) / 16;
    const long long per = (nvec + world - 1) / world;
    long long blocks = (per + 511) / 512;
    if (blocks > 148 * 2) blocks = 148 * 2;
    if (blocks < 1) blocks = 1;
    switch (dtype) {
        case LWC_F32: k_allreduce_two_shot<float><<<(int)blocks, 512, 0, st>>>(p); break;
        case LWC_BF16: k_allreduce_two_shot<__nv_bfloat16><<<(int)blocks, 512, 0, st>>>(p); break;
        case LWC_F16: k_allreduce_two_shot<__half><<<(int)blocks, 512, 0, st>>>(p); break;
        default: return LWC_EINVAL;
    }
    k_allreduce_wait_done<<<1, 32, 0, st>>>(p);
    return (int)cudaGetLastError();
}

// Device address of this device's poison word (hand it to LwcSgdDesc.skip_flag), its value, and a reset.
extern "C" int lwc_poison_ptr(void** out_dev) {
    if (!out_dev) return LWC_EINVAL;
    return (int)cudaGetSymbolAddress(out_dev, g_lwc_poison);
}
extern "C" int lwc_poison_read(unsigned int* out_host) {
    if (!out_host) return LWC_EINVAL;
    return (int)cudaMemcpyFromSymbol(out_host, g_lwc_poison, sizeof(unsigned int));
}
extern "C" int lwc_poison_clear(void) {
    const unsigned int z = 0;
    return (int)cudaMemcpyToSymbol(g_lwc_poison, &z, sizeof(z));
}

// Re-point the *input* gradient of every tensor of a plan (outputs, residuals and geometry unchanged).
// Lets autograd hand over freshly allocated gradient tensors without an accumulate/copy pass: the
// plan reads them where they are and writes the averaged gradient into its stable output arena.
// The table upload is ordered on `stream` before the kernels that follow.
extern "C" int lwc_plan_update_grads(LwcPlan* P, const void* const* grad_dev, void* stream) {
    if (!P || !grad_dev) return LWC_EINVAL;
    for (int t = 0; t < P->num_tensors; ++t) {
        if (!grad_dev[t]) return LWC_EINVAL;
        P->h_tensors[t].grad = grad_dev[t];
    }
    const size_t bytes = sizeof(TensorRec) * P->h_tensors.size();
    cudaStream_t st = (cudaStream_t)stream;
    cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
    cudaStreamIsCapturing(st, &cap);
    if (cap != cudaStreamCaptureStatusNone) {
        // inside a CUDA-graph capture the source must outlive the graph: stage in pinned memory owned by the plan
        // (every replay re-uploads the same table, which is what the captured kernels expect)
        if (!P->h_pinned) return LWC_ESTATE;       // lwc_plan_use_device_step() provides it before any capture
        memcpy(P->h_pinned, P->h_tensors.data(), bytes);
        LWC_CUDA_CHECK(cudaMemcpyAsync(P->d_tensors, P->h_pinned, bytes, cudaMemcpyHostToDevice, st));
    } else {
        LWC_CUDA_CHECK(cudaMemcpyAsync(P->d_tensors, P->h_tensors.data(), bytes, cudaMemcpyHostToDevice, st));
    }
    return 0;
}

// Graph-safe mode: the step counter lives on the device (kernels read it at start, the receiver bumps it), so a
// captured send/recv sequence can be replayed; `step` arguments are then ignored.  Call outside any capture.
extern "C" int lwc_plan_use_device_step(LwcPlan* P, int enable, uint64_t start_step) {
    if (!P) return LWC_EINVAL;
    if (enable && !P->h_pinned) LWC_CUDA_CHECK(cudaMallocHost(&P->h_pinned, sizeof(TensorRec) * P->h_tensors.size()));
    Counters* c = reinterpret_cast<Counters*>(P->d_counters);
    const unsigned long long v = start_step;
    LWC_CUDA_CHECK(cudaMemcpy(&c->dev_step, &v, sizeof(v), cudaMemcpyHostToDevice));
    P->device_step = enable != 0;
    return 0;
}

// =====================================================================================================
// Peer-memory broadcast: rank `root` stores a staging buffer into every rank's copy over NVLink (plain peer stores,
// 16 bytes per thread), the last CTA publishes a step flag; the other ranks run a one-CTA kernel that waits for it
// (bounded spin -> poison word, never a hang).  Staging is double-buffered by step parity, so the root may be one
// broadcast ahead of a receiver.  Replaces the coalesced NCCL broadcasts of the DDP wrappers: parameters + buffers at
// wrap time and the batch-norm buffers before every training forward
// (IMAGENET/training/sparsified_ddp.py:199-203,:366-375; ddp.py:190-194,:373-377).
// =====================================================================================================
namespace lwc {
struct DevBcast {
    char* buf[LWC_MAX_WORLD];           // per rank: staging [2][nbytes_cap]
    uint64_t* flags[LWC_MAX_WORLD];     // per rank: LWC_MAX_WORLD words, index = root
    unsigned int* ticket;               // local
    long long nbytes, cap;
    int world, rank, root;
    unsigned long long step, timeout_ns;
};

__global__ void __launch_bounds__(512) k_bcast_push(DevBcast p) {
    const size_t par = (size_t)(p.step & 1ULL) * (size_t)p.cap;
    const uint4* src = reinterpret_cast<const uint4*>(pick_rank(p.buf, p.rank) + par);
    const long long nvec = (p.nbytes + 15) / 16;
    for (long long v = (long long)blockIdx.x * blockDim.x + threadIdx.x; v < nvec; v += (long long)gridDim.x * blockDim.x) {
        const uint4 x = src[v];
#pragma unroll
        for (int d = 0; d < LWC_MAX_WORLD; ++d)
            if (d < p.world && d != p.rank) reinterpret_cast<uint4*>(p.buf[d] + par)[v] = x;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();                 // cumulative: covers the whole CTA's stores (ordered by the barrier)
        const unsigned t = atomicAdd(p.ticket, 1u);
        if (t == gridDim.x - 1) {
            __threadfence_system();
            *p.ticket = 0;
            for (int d = 0; d < p.world; ++d)
                if (d != p.rank) st_release_sys(pick_rank(p.flags, d) + p.root, p.step + 1);
        }
    }
}

__global__ void k_bcast_wait(DevBcast p) {
    if (threadIdx.x == 0) {
        const uint64_t* f = pick_rank(p.flags, p.rank) + p.root;
        const uint64_t want = p.step + 1;
        const uint64_t t0 = global_timer_ns();
        unsigned spins = 0;
        while (ld_acquire_sys(f) < want) {
            if ((++spins & 63u) == 0 && global_timer_ns() - t0 > p.timeout_ns) {
                atomicExch(&g_lwc_poison, 1u);
                break;
            }
            __nanosleep(64);
        }
    }
}
}  // namespace lwc

// buf_dev[r] / flags_dev[r]: rank r's staging base (2 * cap bytes, cap a multiple of 16) and flag block
// (LWC_MAX_WORLD uint64, zeroed) as seen from this rank; ticket_dev: one local uint32 (zeroed).  The root has filled
// its own staging half (step & 1) before the call; on return (stream order) every other rank's half holds the same
// nbytes.  `step` must advance by one per broadcast of this (buffers, root) pair.
extern "C" int lwc_peer_broadcast(void* const* buf_dev, void* const* flags_dev, void* ticket_dev, long long nbytes, long long cap,
                                  int world, int rank, int root, uint64_t step, void* stream) {
    if (!buf_dev || !flags_dev || !ticket_dev || nbytes <= 0 || nbytes > cap || (cap & 15) || world < 1 || world > LWC_MAX_WORLD ||
        rank < 0 || rank >= world || root < 0 || root >= world)
        return LWC_EINVAL;
    DevBcast p;
    memset(&p, 0, sizeof(p));
    for (int r = 0; r < world; ++r) {
        if (!buf_dev[r] || !flags_dev[r]) return LWC_EINVAL;
        p.buf[r] = (char*)buf_dev[r];
        p.flags[r] = (uint64_t*)flags_dev[r];
    }
    p.ticket = (unsigned int*)ticket_dev;
    p.nbytes = nbytes; p.cap = cap; p.world = world; p.rank = rank; p.root = root; p.step = step; p.timeout_ns = g_flag_timeout_ns;
    cudaStream_t st = (cudaStream_t)stream;
    if (world == 1) return 0;
    if (rank == root) {
        long long blocks = ((nbytes + 15) / 16 + 511) / 512;
        if (blocks > 148) blocks = 148;
        if (blocks < 1) blocks = 1;
        k_bcast_push<<<(int)blocks, 512, 0, st>>>(p);
    } else {
        k_bcast_wait<<<1, 32, 0, st>>>(p);
    }
    return (int)cudaGetLastError();
}
