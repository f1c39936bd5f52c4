/*
 * lwc_b200.h -- C ABI of the layerwise_b200 native library (sm_100a).
 *
 * All entry points return 0 on success, a positive cudaError_t on a CUDA failure and a
 * negative LWC_E* code on an argument error.  Pointers named *_dev are device pointers.
 * Streams are passed as void* (cudaStream_t).  Nothing here synchronises the device
 * unless stated.
 *
 * Reference citations name the file:line of aritra-dutta/layer-wise-aaai20 whose
 * behaviour the entry point covers.
 */
#ifndef LWC_B200_H_
#define LWC_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LWC_MAX_WORLD 8
#define LWC_CHUNK 8192            /* elements handled by one CTA */

/* error codes */
#define LWC_EINVAL   (-1)
#define LWC_ENOMEM   (-2)
#define LWC_ETOOBIG  (-3)
#define LWC_ESTATE   (-4)

/* dtypes */
#define LWC_F32  0
#define LWC_BF16 1
#define LWC_F16  2

/* methods (CIFAR10/core.py:178-215) */
#define LWC_NONE       0
#define LWC_TOPK       1   /* core.py:178-183 */
#define LWC_RANDOMK    2   /* core.py:184-188, sparsified_ddp.py:409-413 */
#define LWC_THRESHOLDV 3   /* core.py:189-193 */
#define LWC_ADAPTIVE   4   /* core.py:194-199 */
#define LWC_TERNGRAD   5   /* core.py:200-206 */
#define LWC_DITHERING  6   /* core.py:207-213 (QSGD) */

/* ------------------------------------------------------------------------------------------
 * Device memory that can be shared between processes (peer receive slots, gradient arenas).
 * cudaMalloc-backed so that cudaIpc handles can be taken (torch's caching allocator cannot
 * promise that).  Stand-in for the c10d-internal buffers behind dist._queue_reduction
 * (IMAGENET/training/sparsified_ddp.py:460).
 * ---------------------------------------------------------------------------------------- */
int lwc_malloc(void** out_dev, size_t bytes);
int lwc_free(void* dev);
int lwc_memset_async(void* dev, int value, size_t bytes, void* stream);
int lwc_ipc_get_handle(void* dev, void* handle_out_64B);
/* A handle names the driver's whole underlying block; the importer adds the exporter's offset inside that block
 * (lwc_ipc_base_offset) to the base that lwc_ipc_open_handle returns.  Opens are reference-counted per handle. */
int lwc_ipc_base_offset(void* dev, uint64_t* offset_out);
int lwc_ipc_open_handle(const void* handle_64B, void** out_dev);
int lwc_ipc_close_handle(void* dev);
int lwc_enable_peer_access(int peer_device);           /* idempotent */
const char* lwc_error_string(int code);

/* ------------------------------------------------------------------------------------------
 * Exchange plan: one bucket of gradient tensors that is compressed, exchanged and
 * decompressed together.  "Layer-wise" = one selection group per tensor; "entire-model" =
 * one group spanning all tensors (CIFAR10/core.py:175-225 vs :227-301).
 * ---------------------------------------------------------------------------------------- */
typedef struct LwcPlan LwcPlan;

typedef struct {
    int32_t  num_tensors;
    int32_t  num_groups;
    int32_t  method;          /* LWC_* */
    int32_t  grad_dtype;      /* dtype of grad_dev[] */
    int32_t  out_dtype;       /* dtype of out_dev[]  */
    int32_t  world;           /* ranks in the exchange (1..LWC_MAX_WORLD) */
    int32_t  rank;
    int32_t  qstates;         /* LWC_DITHERING levels (1..255) */
    float    threshold;       /* LWC_THRESHOLDV */
    uint64_t seed;            /* RNG key; same on all ranks for shared masks */
    /* per tensor */
    const void* const* grad_dev;
    void* const*       out_dev;
    float* const*      resid_dev;   /* error-feedback residuals, or NULL entries / NULL array */
    const int64_t*     numel;
    const int32_t*     group_of;    /* group id of each tensor, non-decreasing */
    const int32_t*     uid;         /* stable id per tensor for RNG keying */
    /* per group */
    const int64_t*     keep;        /* Top-k / Random-k: number of elements to keep */
    const int64_t*     cap;         /* payload capacity (entries) */
} LwcPlanDesc;

int lwc_plan_create(LwcPlan** out, const LwcPlanDesc* desc);
int lwc_plan_destroy(LwcPlan* plan);
/* bytes of one receive slot (one source rank, one parity) and of the whole region
 * (2 parities x world slots) a rank must provide for this plan */
size_t lwc_plan_slot_bytes(const LwcPlan* plan);
size_t lwc_plan_region_bytes(const LwcPlan* plan);
/* region_dev[d] = base of rank d's receive region for this plan, flags_dev[d] = rank d's
 * flag array (LWC_MAX_WORLD uint64 per plan, zero-initialised).  For rank==d these are
 * local allocations; for the others peer-mapped pointers. */
int lwc_plan_bind(LwcPlan* plan, void* const* region_dev, void* const* flags_dev);
/* optional: uniforms in [0,1) (quantisers) or integer keys (Random-k) injected per tensor
 * for bit-exact tests; NULL restores the counter-based generator */
/* Shared-seed Random-k: ship values only (4 bytes per kept element instead of (index, value) pairs); every rank selects
 * the same positions, the receiver recomputes them from the hash keys (IMAGENET/training/sparsified_ddp.py:410-412
 * packs k values the same way, relying on identical randperm streams). */
int lwc_plan_set_values_only(LwcPlan* plan, int enable);
int lwc_plan_set_random(LwcPlan* plan, const void* const* rand_dev);
/* Sender side: error-feedback accumulate, select / quantise, residual update, pack,
 * store into every rank's slot, release flags.  step must increase by 1 per call. */
int lwc_plan_send(LwcPlan* plan, uint64_t step, void* stream);
/* Receiver side: acquire the flags of step, scatter-add / dequantise all sources in rank
 * order, scale by 1/world, write out_dev[].  The optimiser step is a separate launch (lwc_sgd_step). */
int lwc_plan_recv(LwcPlan* plan, uint64_t step, void* stream);
/* counters: [0]=entries sent last step, [1]=entries dropped by capacity clamp,
 * [2]=flag wait timeouts, [3]=payload bytes stored per destination */
int lwc_plan_stats(LwcPlan* plan, uint64_t out[4], void* stream);

/* Dense two-shot all-reduce over peer memory, in place on gradient arenas that live in the
 * peer-mapped heap: rank r sums slice r of all arenas in rank order, scales by 1/world and stores it
 * into every arena.  arena_dev[r] / flags_dev[r] = rank r's arena base and flag block
 * (2 * LWC_MAX_WORLD uint64, zero-initialised); counters_dev = 2 local uint64 (zeroed).
 * (CIFAR10/core.py:316-319; IMAGENET/training/train_imagenet_nv.py:382-386; ddp.py:434-475) */
int lwc_dense_allreduce(void* const* arena_dev, void* const* flags_dev, void* counters_dev, int64_t off,
                        int64_t n, int dtype, int world, int rank, uint64_t step, void* stream);

/* ------------------------------------------------------------------------------------------
 * Stand-alone exact Top-k selection of one fp32 vector (the K1 kernel family without the
 * exchange): threshold = keep-th largest |x|, every |x| >= threshold and != 0 is emitted in
 * index order, at most cap entries.  workspace_dev must hold lwc_topk_workspace_bytes(n)
 * bytes of scratch.  (CIFAR10/core.py:178-183)
 * ---------------------------------------------------------------------------------------- */
size_t lwc_topk_workspace_bytes(int64_t n);
int lwc_topk_select_f32(const float* x_dev, int64_t n, int64_t keep, int64_t cap,
                        uint32_t* idx_out_dev, float* val_out_dev, int64_t* count_out_dev,
                        void* workspace_dev, void* stream);

/* No counterpart in the reference (it never manages device memory itself): the same exact Top-k
 * selection as lwc_topk_select_f32 for callers that cannot allocate device memory -- code running
 * under a memory pool that already owns all of HBM, plug-ins that are only handed pointers and
 * streams, or sections where cudaMalloc's device-wide synchronisation is unacceptable.  The scratch
 * is a fixed array inside the library (LWC_TOPK_NOALLOC_SCRATCH_BYTES of device memory on every GPU
 * that loads the library, held until the process exits), hence the size limit:
 * n > LWC_TOPK_NOALLOC_MAX_N returns LWC_ETOOBIG.  Inputs and outputs are still the caller's.
 * The call performs no cudaMalloc / cudaFree and uses no allocating library routine.
 *  - The first call on each device looks up the array's device address (that device must be
 *    current); make that first call before any stream capture begins.
 *  - All calls on a device share the one array: two calls whose kernels may overlap on the GPU
 *    (different streams, no ordering between them) race.  Serialise them.
 *  - Nothing has to be cleared by the caller; the entry point resets what must start at zero. */
#define LWC_TOPK_NOALLOC_MAX_N          (1 << 24)          /* 16 Mi elements */
#define LWC_TOPK_NOALLOC_SCRATCH_BYTES  (96 * 1024)
int lwc_topk_select_f32_noalloc(const float* x_dev, int64_t n, int64_t keep, int64_t cap,
                                uint32_t* idx_out_dev, float* val_out_dev, int64_t* count_out_dev,
                                void* stream);

/* ------------------------------------------------------------------------------------------
 * Fused multi-tensor SGD (torch.optim.SGD semantics: CIFAR10/torch_backend.py:132-143,
 * IMAGENET/training/train_imagenet_nv.py:188-191,:441; fp16 master copies
 * IMAGENET/training/fp16util.py:103-139 folded in).
 * ---------------------------------------------------------------------------------------- */
typedef struct {
    int32_t num_tensors;
    int32_t grad_dtype;      /* dtype of grads */
    int32_t model_dtype;     /* dtype of model copies (ignored when model_dev == NULL) */
    float lr, momentum, dampening, weight_decay, grad_scale;
    int32_t nesterov, first_step;
    const float* lr_dev;     /* optional: learning rate read from device memory (CUDA-graph replay) */
    int32_t param_dtype;     /* LWC_F32 (default): fp32 weights + momentum.  LWC_BF16 / LWC_F16: weights, momentum and
                              * gradients all in that type (training without fp32 masters); model copies unused */
    const unsigned int* skip_flag;   /* optional device word: non-zero -> the step is skipped (lwc_poison_ptr: set by an
                                      * exchange whose peer timed out, so stale gradients are never applied) */
} LwcSgdDesc;
/* table_dev: device array of num_tensors LwcSgdTensor; chunk table built by lwc_sgd_plan */
typedef struct { void* param; const void* grad; void* mom; void* model; int64_t n; } LwcSgdTensor;
typedef struct LwcSgdPlan LwcSgdPlan;
int lwc_sgd_plan_create(LwcSgdPlan** out, const LwcSgdTensor* tensors_host, int32_t num_tensors);
int lwc_sgd_plan_destroy(LwcSgdPlan* plan);
int lwc_sgd_step(LwcSgdPlan* plan, const LwcSgdDesc* desc, void* stream);
/* Diagnostics (no reference counterpart): cycles per tcgen05.mma (M = 128, N = n, K = 16, bf16, both operands in shared
 * memory) issued back to back by one thread per SM, for K-major / MN-major operand layouts.  tools/mma_rate_probe.py. */
int lwc_probe_mma_rate(int a_mn, int b_mn, int n, int iters, double* cycles_per_mma);

/* Peer-memory broadcast of a staging buffer from `root` into every rank's copy (NVLink peer stores + step flag;
 * receivers wait with a bound).  Stands in for the coalesced NCCL broadcasts of the reference's DDP wrappers
 * (IMAGENET/training/sparsified_ddp.py:199-203,:366-375; ddp.py:190-194,:373-377). */
int lwc_peer_broadcast(void* const* buf_dev, void* const* flags_dev, void* ticket_dev, long long nbytes, long long cap,
                       int world, int rank, int root, uint64_t step, void* stream);
/* Fail-closed word of this device: set by receivers / the dense all-reduce when a peer's flag wait ran out (no
 * reference counterpart beyond the try/except of IMAGENET/training/train_imagenet_nv.py:704-716). */
int lwc_poison_ptr(void** out_dev);
int lwc_poison_read(unsigned int* out_host);
int lwc_poison_clear(void);

/* Re-point the input gradients of a plan (autograd's freshly written tensors are consumed in place; outputs
 * stay in the arena).  Stream-ordered table upload.  (no accumulate pass: cf. core.py:177 `layer.grad.data`) */
int lwc_plan_update_grads(LwcPlan* plan, const void* const* grad_dev, void* stream);
/* Upper bound (ms, default 10000) of every in-kernel wait on a peer flag; expiry is counted in
 * lwc_plan_stats()[2] and the kernel proceeds -- a dead peer cannot hang a GPU. */
int lwc_set_flag_timeout_ms(unsigned int ms);
/* Graph-safe mode: keep the step counter on the device (read by every kernel, advanced by the receiver) so a
 * captured send/recv sequence can be replayed; gradient-table uploads made during capture use pinned staging. */
int lwc_plan_use_device_step(LwcPlan* plan, int enable, uint64_t start_step);
/* NVLS variant: multicast mapping of the region given to lwc_plan_bind (NULL switches back to unicast). */
int lwc_plan_bind_multicast(LwcPlan* plan, void* mc_region_dev);

/* ------------------------------------------------------------------------------------------
 * bf16 GEMM on tcgen05 / TMEM / TMA: C[M,N] (+)= A * B^T (+bias)(ReLU).  a_mn / b_mn select MN-major
 * operands (A stored [K][M], B stored [K][N]) so forward, dgrad and wgrad of a linear / 1x1-conv layer
 * need no transposes.  M,N >= 256 run on CTA pairs (cta_group::2, 256x256 tiles).
 * (nn.Linear / 1x1 nn.Conv2d: CIFAR10/alexnet.py:31-36, CIFAR10/vgg16.py:24-30, IMAGENET/training/resnet.py:110)
 * ---------------------------------------------------------------------------------------- */
size_t lwc_gemm_workspace_bytes(int M, int N, int splits);
int lwc_gemm_pick_splits(int M, int N, int K);
int lwc_gemm_bf16(const void* A, const void* B, void* C, int M, int N, int K, int lda, int ldb, int ldc,
                  int a_mn, int b_mn, int out_dtype, const float* bias, int relu, int accumulate,
                  int splits, float* ws, void* stream);

/* ------------------------------------------------------------------------------------------
 * Fused batch-norm (+ residual) (+ ReLU) on channels-last activations [M = N*H*W, C], C % 8 == 0.
 * (nn.BatchNorm2d + nn.ReLU + `out += residual`: CIFAR10/dawn.py:23-28, IMAGENET/training/resnet.py:40-95)
 * ---------------------------------------------------------------------------------------- */
size_t lwc_bn_workspace_bytes(long long M, int C);
int lwc_bn_forward(const void* x, void* y, const void* res, const float* gamma, const float* beta,
                   float* mean, float* invstd, float* running_mean, float* running_var, float* workspace,
                   long long M, int C, int dtype, int training, int relu, float eps, float momentum, void* stream);
/* same, and in training mode also increments *num_batches_tracked (int64 device scalar, nullable) inside the
 * statistics-finalize kernel, so nn.BatchNorm2d's counter costs no launch of its own.  relu_mask (M*C/8 bytes,
 * nullable): one bit per output element, set where the pre-activation is positive -- lwc_bn_backward_ex reads it
 * instead of the saved output y (1/16 of the bytes; needed only when a residual was fused). */
int lwc_bn_forward_ex(const void* x, void* y, const void* res, const float* gamma, const float* beta,
                      float* mean, float* invstd, float* running_mean, float* running_var, float* workspace,
                      long long M, int C, int dtype, int training, int relu, float eps, float momentum,
                      long long* num_batches_tracked, void* relu_mask, void* stream);
int lwc_bn_backward_ex(const void* dy, const void* x, const void* y, void* dx, void* dres, const float* gamma,
                       const float* beta, const float* mean, const float* invstd, float* dgamma, float* dbeta,
                       float* workspace, long long M, int C, int dtype, int relu, const void* relu_mask, void* stream);
int lwc_bn_backward(const void* dy, const void* x, const void* y, void* dx, void* dres, const float* gamma,
                    const float* beta, const float* mean, const float* invstd, float* dgamma, float* dbeta,
                    float* workspace, long long M, int C, int dtype, int relu, void* stream);

/* Persistent tcgen05 implicit-GEMM convolution, bf16 NHWC, stride 1, kh in {1, 3} ("same" padding), Cin % 64 == 0,
 * Cout % 64 == 0 (stands in for the cuDNN convolutions under IMAGENET/training/resnet.py:28-31 and
 * CIFAR10/dawn.py:34-41).  wt is [Cout][kh*kh][Cin] (b_mn = 0) or, for 1x1 data gradients, the forward weight
 * [Cin_fwd][Cout_fwd] read MN-major (b_mn = 1).  bias (fp32 per Cout) and relu are optional epilogue stages;
 * stats, when non-NULL, receives per-M-tile partial batch-norm sums: [tiles_m][2][Cout] floats with tiles_m from
 * lwc_conv_tiles_m (which also reports the rows covered by one tile). */
int lwc_conv_tiles_m(int Nimg, int H, int W, int kh, int* rows_per_tile_out);
int lwc_conv_forward(const void* x, const void* wt, void* y, int Nimg, int H, int W, int Cin, int Cout, int kh,
                     int b_mn, const float* bias, int relu, float* stats, void* stream);
/* 1x1 only: Y = conv(X) + addend (bf16, shape of Y), the addend tile prefetched by TMA: the data gradient at a
 * residual fork, dX = dY * W + d_identity, in one kernel. */
int lwc_conv1x1_add(const void* x, const void* wt, const void* addend, void* y, int Nimg, int H, int W, int Cin,
                    int Cout, int b_mn, void* stream);
/* Same, and outputs whose bit in relu_mask ([M][Cout/8] bytes, the layout lwc_bn_forward_ex writes) is 0 are zeroed:
 * Y = (conv(X) + addend) * mask.  The gradient then reaches the producing block already multiplied with that block's
 * ReLU mask, so its batch-norm backward neither reads a mask nor writes a residual gradient of its own.
 * addend may be NULL; Cout % 128 == 0. */
int lwc_conv1x1_add_mask(const void* x, const void* wt, const void* addend, const void* relu_mask, void* y, int Nimg,
                         int H, int W, int Cin, int Cout, int b_mn, void* stream);

/* Weight gradient of the ResNet stem (7x7/2 pad 3, 3 -> 64 channels; IMAGENET/training/resnet.py:100) on tcgen05:
 * dy [N][OH][OW][64] bf16, xs = the 2x2 space-to-depth input [N][OH+3][OW+3][16] bf16 written by lwc_normalize_u8
 * (mode 1); out [64][7][7][3] (channels-last [64,3,7,7]) in out_dtype (LWC_F32 / LWC_BF16) = scale * dW.
 * OH even, OW % 16 == 0, OW <= 144.  workspace: lwc_stem_wgrad_workspace_bytes(N, OH) bytes of device memory. */
size_t lwc_stem_wgrad_workspace_bytes(int N, int OH);
int lwc_stem_wgrad(const void* dy, const void* xs, void* out, float* workspace, int N, int OH, int OW, int out_dtype,
                   float scale, void* stream);

/* Tap-table tcgen05 convolution family (csrc/convx.cu): every convolution of the model zoo on one persistent
 * implicit-GEMM kernel (forward, data gradient) plus a split-pixel weight-gradient kernel.  bf16 channels-last.
 * (nn.Conv2d forward/backward through cuDNN: IMAGENET/training/resnet.py:18-21,59-66,100-101,122-126;
 * CIFAR10/dawn.py:23-35; CIFAR10/alexnet.py:15-27; CIFAR10/vgg16.py:69-82.)
 *   forward : x [N][H][W][Cx] (Cx % 8 == 0), wt [Cout][k*k][round_up(Cx,64)], y [N][OH][OW][Cout]; k*k <= 16,
 *             stride 1 or 2; bias fp32 / relu optional; stats (nullable) receives per-CTA batch-norm partials
 *             [lwc_convx_max_partials()][2][Cout] and *nparts_out how many of them are valid; bn_hint = 0 (auto),
 *             64, 128 or 256 output channels per tile.
 *   dgrad1x1: dX = dY * W with the forward weight as stored (no transposed copy).
 *   dgrad   : general k / stride / pad; wt = the forward weight [Cout][k*k][Cin] as stored (read MN-major per tap).
 *   stem    : the 7x7/2 stem as a 4-row-tap convolution over overlapping 128-byte windows of the space-to-depth input.
 *   wgrad   : dW [Cout][k*k][Cin_out] = scale * dY^T X, fp32 partials in `workspace` reduced in a fixed order. */
int lwc_convx_max_partials(void);
int lwc_convx_forward(const void* x, const void* wt, void* y, int N, int H, int W, int Cx, int Cout, int k, int stride,
                      int pad, const float* bias, int relu, float* stats, int* nparts_out, int bn_hint, void* stream);
int lwc_convx_dgrad1x1(const void* dy, const void* w, void* dx, long long M, int Cout, int Cin, int bn_hint, void* stream);
int lwc_convx_dgrad(const void* dy, const void* wt, void* dx, int N, int H, int W, int Cin, int Cout, int k, int stride,
                    int pad, int bn_hint, void* stream);
int lwc_convx_stem_forward(const void* xs, const void* w4, void* y, int N, int OH, int OW, int Cout, float* stats,
                           int* nparts_out, int bn_hint, void* stream);
size_t lwc_convx_wgrad_workspace_bytes(int N, int H, int W, int Cx, int Cout, int k, int stride, int pad, int bn_hint);
int lwc_convx_wgrad(const void* x, const void* dy, void* dw, float* workspace, int N, int H, int W, int Cx, int Cout,
                    int Cin_out, int k, int stride, int pad, int out_dtype, float scale, int bn_hint, void* stream);
/* Stride-1 data gradient that also produces the batch-norm BACKWARD reduction sums of the layer whose output it
 * differentiates (per-CTA partials of sum d and sum d * x), and the batch-norm backward that consumes them
 * (IMAGENET/training/resnet.py:72-92: conv -> bn -> relu -> conv chains). */
int lwc_convx_dgrad_bnstats(const void* dy, const void* wt, void* dx, int N, int H, int W, int Cin, int Cout, int k, int pad,
                            const void* x_bn, const float* gamma, const float* beta, const float* mean, const float* invstd,
                            int relu, float* partials, int* nparts_out, int bn_hint, void* stream);
/* Residual fork (IMAGENET/training/resnet.py:72-94, `out += identity; relu`): dX = mask * (dY W + addend) for the block's
 * first 1x1 convolution -- addend = the skip connection's gradient, mask = the producing batch norm's one-bit ReLU mask
 * ([M][Cin / 8] bytes, nullable) -- and, with x_bn, that batch norm's backward sums of dX into per-CTA partials.
 * bn_hint 1000 + BN on lwc_convx_forward / lwc_convx_dgrad selects the halo variant (64 input channels, stride 1):
 * resident weights, one TMA box per tile, taps read from shifted start addresses. */
int lwc_convx_dgrad1x1_fork(const void* dy, const void* w, const void* addend, const void* mask, void* dx, long long M,
                            int Cout, int Cin, const void* x_bn, const float* mean, const float* invstd, float* partials,
                            int* nparts_out, int bn_hint, void* stream);
int lwc_bn_backward_from_partials(const void* dy, const void* x, void* dx, const float* gamma, const float* beta,
                                  const float* mean, const float* invstd, float* dgamma, float* dbeta,
                                  const float* partials, int nparts, long long M, int C, int dtype, int relu, void* stream);
/* Batch-norm mean / inverse std (+ running statistics, + num_batches_tracked) from those partials, and the
 * apply-only batch-norm pass that consumes them (IMAGENET/training/resnet.py:72-92 conv -> bn -> relu chains). */
int lwc_convx_stats_finalize(const float* partials, int nparts, int C, long long count, float eps, float momentum,
                             float* mean, float* invstd, float* running_mean, float* running_var,
                             long long* num_batches_tracked, void* stream);
int lwc_bn_apply_stats(const void* x, void* y, const void* res, const float* gamma, const float* beta, const float* mean,
                       const float* invstd, long long M, int C, int dtype, int relu, void* relu_mask, void* stream);

/* Small model ops (csrc/misc.cu).  eltwise op: 0 = relu(x), 1 = alpha * x (Mul(0.125), CIFAR10/torch_backend.py:72-77),
 * 2 = ReLU backward dy * (aux > 0) (nn.ReLU after biased layers, CIFAR10/alexnet.py:31-36, vgg16.py:24-30).
 * Dropout (alexnet.py:30,33; vgg16.py:26,29): Philox4x32-10 keyed by (seed, offset + group), one mask bit per element.
 * Adaptive average pooling on channels-last tensors (IMAGENET/training/resnet.py:109, CIFAR10/vgg16.py:22).
 * Accuracy (torch_backend.py:88-90; train_imagenet_nv.py:679-689): per-sample argmax-correct bytes and/or
 * accumulated top-1 / top-5 hit counters (float[2]) -- no host synchronisation. */
int lwc_eltwise(const void* x, const void* aux, void* y, long long n, int op, float alpha, int dtype, void* stream);
int lwc_dropout_forward(const void* x, void* y, void* mask, long long n, float p, unsigned long long seed,
                        unsigned long long offset, int dtype, void* stream);
int lwc_dropout_backward(const void* dy, const void* mask, void* dx, long long n, float p, int dtype, void* stream);
int lwc_adaptive_avgpool(const void* in, void* out, int N, int IH, int IW, int OH, int OW, int C, int backward, int dtype,
                         void* stream);
int lwc_accuracy(const void* logits, const void* target_i64, void* correct, float* counters, int N, int C, int dtype,
                 void* stream);

/* Max pooling, NHWC, byte argmax forward + gather backward (CIFAR10/dawn.py:51,:70; resnet.py:104). */
int lwc_maxpool_forward(const void* x, void* y, void* idx, int N, int H, int W, int C, int k, int stride,
                        int pad, int dtype, void* stream);
int lwc_maxpool_backward(const void* dy, const void* idx, void* dx, int N, int H, int W, int C, int k,
                         int stride, int pad, int dtype, void* stream);

/* uint8 NCHW -> normalised channels-last (mode 0) or 2x2 space-to-depth with zero border (mode 1)
 * (IMAGENET/training/dataloader.py:88-93). */
int lwc_normalize_u8(const void* x_u8_nchw, void* y, int N, int H, int W, int c_out, int mode,
                     const float* mean3, const float* std3, int dtype, void* stream);

/* Cross entropy: per-sample loss + log-sum-exp; backward = (softmax - onehot) * gscale * g[n]
 * (CIFAR10/dawn.py:85; IMAGENET/training/train_imagenet_nv.py:187). */
int lwc_cross_entropy_forward(const void* logits, const void* target_i64, float* loss, float* lse, int N, int C,
                              int dtype, void* stream);
int lwc_cross_entropy_backward(const void* logits, const void* target_i64, const float* lse,
                               const float* grad_per_sample, void* dlogits, int N, int C, int dtype,
                               float gscale, void* stream);

#ifdef __cplusplus
}
#endif
#endif  /* LWC_B200_H_ */
