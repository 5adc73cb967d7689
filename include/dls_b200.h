/* dls_b200.h -- C ABI of libdlsb200.so, the B200 (sm_100a) device layer that replaces
 * Compute.scala's OpenCL trait + Memory + OpenCLKernelBuilder underneath the `Tensors` trait.
 *
 * Citations are file:line into the reference tree:
 *   OpenCL  = Compute.scala/OpenCL/src/main/scala/com/thoughtworks/compute/OpenCL.scala
 *   Tensors = Compute.scala/Tensors/src/main/scala/com/thoughtworks/compute/Tensors.scala
 *   KB      = Compute.scala/OpenCLKernelBuilder/src/main/scala/com/thoughtworks/compute/OpenCLKernelBuilder.scala
 *   Trees   = Compute.scala/Trees/src/main/scala/com/thoughtworks/compute/Trees.scala
 *   TW      = plugins-TensorWeights/src/main/scala/com/thoughtworks/deeplearning/plugins/TensorWeights.scala
 *
 * Conventions
 *   - every entry returns an int status (DLS_OK == 0); no exceptions cross the ABI; the message of the last
 *     failure on the calling thread is dls_last_error() (replaces OpenCL.Exceptions.fromErrorCode, OpenCL:143-312,
 *     and the program build log, OpenCL:885-915).
 *   - handles are opaque and reference counted; retain/release are thread safe (replaces clRetain/Release*).
 *   - all device work is asynchronous on a stream borrowed from the context's pool (replaces `dispatch`,
 *     OpenCL:1298-1329); ordering between launches is expressed with dls_event dependencies exactly like the
 *     reference's `waitingEvents` (Tensors:1235, :1246).  Buffers may be released while kernels that use them
 *     are still in flight (stream-ordered free).
 *   - plain pointers and sizes only: no torch / C++ types in any signature.
 *   - there is NO CPU fallback: every compute entry fails with DLS_ERR_NO_DEVICE when no sm_100 device is usable.
 */
#ifndef DLS_B200_H
#define DLS_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DLS_ABI_VERSION 1

typedef struct dls_ctx dls_ctx;     /* one per device: stream pool, caching allocator, plan cache     */
typedef struct dls_buf dls_buf;     /* refcounted device buffer of fp32 (OpenCL.DeviceBuffer, OpenCL:633-717) */
typedef struct dls_event dls_event; /* refcounted completion event (OpenCL.Event, OpenCL:565-612)      */

enum dls_status {
  DLS_OK = 0,
  DLS_ERR_INVALID = 1,      /* bad argument / malformed tree (IllegalArgumentException in the reference) */
  DLS_ERR_NO_DEVICE = 2,    /* no CUDA device / not sm_100 (OpenCL.Exceptions.DeviceNotFound)           */
  DLS_ERR_OOM = 3,          /* MemObjectAllocationFailure / OutOfResources                               */
  DLS_ERR_CUDA = 4,         /* any other CUDA runtime error; text in dls_last_error()                    */
  DLS_ERR_UNSUPPORTED = 5,  /* tree exceeds an interpreter limit (rank > 8, > 65535 leaves, ...)         */
  DLS_ERR_COMM = 6          /* NCCL failure                                                              */
};

enum dls_rng_hash {         /* Tensors:98-117 */
  DLS_HASH_WANG = 0,        /* WangHashingRandomNumberGenerator (the cpu/gpu singletons' default)        */
  DLS_HASH_LCG = 1          /* LcgRandomNumberGenerator                                                  */
};

enum dls_gemm_mode {        /* contraction precision for the tensor-core path                            */
  DLS_GEMM_FP32 = 0,        /* CUDA-core fp32 (bit-comparable with the unrolled elementwise chain)       */
  DLS_GEMM_TF32 = 1,
  DLS_GEMM_BF16 = 2
};

typedef struct dls_ctx_opts {
  int32_t device;               /* CUDA ordinal (replaces deviceIds / UseAll*Devices, OpenCL:314-374)      */
  int32_t streams_per_device;   /* replaces numberOfCommandQueuesPerDevice (OpenCL:425; cpu.scala uses 5) */
  int32_t rng_hash;             /* enum dls_rng_hash                                                       */
  int32_t gemm_mode;            /* enum dls_gemm_mode                                                      */
  uint64_t allocator_cap_bytes; /* 0 = unlimited: free blocks cached above this are returned to the driver */
  int32_t plan_cache_capacity;  /* 0 = default 4096 (replaces kernelCacheBuilder, Tensors:1139-1150)       */
  int32_t reserved;
} dls_ctx_opts;

typedef struct dls_stats {
  uint64_t kernel_launches;     /* kernels of this library launched so far                                 */
  uint64_t plan_cache_hits;     /* Tensors:1165 kernelCache.getIfPresent hits                              */
  uint64_t plan_cache_misses;
  uint64_t bytes_in_use;        /* device bytes held by live dls_buf                                       */
  uint64_t bytes_cached;        /* device bytes parked in the allocator's free lists                       */
  uint64_t driver_allocs;       /* cudaMalloc calls (the reference does one clCreateBuffer per launch)     */
  uint64_t h2d_bytes;
  uint64_t d2h_bytes;
} dls_stats;

/* ---- library ------------------------------------------------------------------------------------------ */
int dls_abi_version(void);
const char *dls_last_error(void);                 /* thread local; never NULL                              */

/* ---- context: replaces the lazy platform/device/context/queue-pool chain OpenCL:340-374, :1376-1393, :427-438 */
int dls_ctx_create(const dls_ctx_opts *opts, dls_ctx **out);
void dls_ctx_destroy(dls_ctx *ctx);               /* replaces the monadicClose chain OpenCL:1335, :443-446, Tensors:1159-1161 */
int dls_ctx_sync(dls_ctx *ctx);                   /* wait for every stream of the pool                     */
int dls_ctx_stats(dls_ctx *ctx, dls_stats *out);
int dls_ctx_trim(dls_ctx *ctx);                   /* return cached free blocks to the driver               */
void *dls_ctx_stream(dls_ctx *ctx, int index);    /* cudaStream_t of pool slot `index` (for event timing / NCCL plumbing) */

/* ---- buffers: replaces allocateBuffer OpenCL:1399, allocateBufferFrom OpenCL:1415, DeviceBuffer OpenCL:633-717 */
int dls_buf_alloc(dls_ctx *ctx, size_t n_floats, dls_buf **out);
int dls_buf_from_host(dls_ctx *ctx, const float *src, size_t n_floats, dls_buf **out, dls_event **done);
void dls_buf_retain(dls_buf *buf);                /* clRetainMemObject  OpenCL:637 */
void dls_buf_release(dls_buf *buf);               /* clReleaseMemObject OpenCL:641 */
size_t dls_buf_length(const dls_buf *buf);        /* floats (DeviceBuffer.length OpenCL:683-686)           */
void *dls_buf_device_ptr(const dls_buf *buf);     /* raw device pointer (interop with NCCL / torch only)   */
/* non-blocking read-back: replaces DeviceBuffer.toHostBuffer OpenCL:698-715 + nclEnqueueReadBuffer OpenCL:1224.
 * `dst` must stay valid until `*done` completes (or pass done == NULL for a blocking read). */
int dls_buf_read(dls_ctx *ctx, dls_buf *buf, size_t offset, float *dst, size_t n_floats,
                 dls_event *const *deps, int n_deps, dls_event **done);
int dls_buf_write(dls_ctx *ctx, dls_buf *buf, size_t offset, const float *src, size_t n_floats, dls_event **done);

/* Upload on the context's COPY stream (new functionality: a stream of its own, so the next batch travels over PCIe while
 * the current step computes; the reference uploads on the compute queue, Tensors:445-463).  The copy stream first waits
 * for `deps` (e.g. the event after the last consumer of `buf`); `*done` completes when the data is in `buf`.  Nothing
 * else is ordered: consumers must wait for `*done` (dls_stream_wait_event, or the deps of a launch).  `src` should be
 * page-locked (dls_host_alloc) and must stay valid until `*done`. */
int dls_buf_upload_async(dls_ctx *ctx, dls_buf *buf, size_t offset, const float *src, size_t n_floats,
                         dls_event *const *deps, int n_deps, dls_event **done);
/* make pool stream `stream_index` wait for `ev` before running anything queued after this call (clEnqueueBarrierWithWaitList) */
int dls_stream_wait_event(dls_ctx *ctx, int stream_index, dls_event *ev);

/* page-locked host staging memory: replaces Memory.allocate / MemoryUtil.memAllocFloat (Memory.scala:184-208) for
 * buffers that feed dls_buf_write / dls_buf_read at full PCIe rate without a bounce copy. */
int dls_host_alloc(size_t n_floats, float **out);
void dls_host_free(float *ptr);

/* ---- events: replaces Event.retain/release OpenCL:568-574, waitForComplete OpenCL:603-605, clSetEventCallback OpenCL:1251 */
void dls_event_retain(dls_event *ev);
void dls_event_release(dls_event *ev);
int dls_event_wait(dls_event *ev);
int dls_event_query(dls_event *ev);               /* DLS_OK when complete, 1 when still running            */
typedef void (*dls_event_cb)(void *user, int status);
int dls_event_on_complete(dls_event *ev, dls_event_cb cb, void *user);  /* cb must not call launch APIs (cf. OpenCL:378-392) */
int dls_event_elapsed_ms(dls_event *start, dls_event *end, float *ms);
int dls_event_record(dls_ctx *ctx, int stream_index, dls_event **out); /* marker on a pool stream (timing)  */

/* ---- expression trees: replaces enqueueClosure Tensors:1163-1264 (alpha-conversion Trees:266-268, kernelCache lookup
 * Tensors:1165/:1258, code generation KB:135-221, clBuildProgram OpenCL:917-932, Kernel.enqueue OpenCL:788-844).
 * `tree` is the serialised closure in wire format v1 (see DESIGN.md "Tree wire format"); `params[i]` is the buffer of
 * the i-th ArrayParameter in first-visit order (the alpha-normalised numbering of Trees:807-822).
 * The output buffer is freshly allocated from the caching allocator (`allocateBuffer(shape.product)`, Tensors:1217). */
enum dls_launch_flags {
  DLS_LAUNCH_DEFAULT = 0,
  DLS_LAUNCH_REDUCE_SUM = 1     /* do not materialise: *out is a 1-float buffer holding the sum of all elements */
};
int dls_launch_tree(dls_ctx *ctx, const uint8_t *tree, size_t tree_len, dls_buf *const *params, int n_params,
                    dls_event *const *deps, int n_deps, uint32_t flags, dls_buf **out, dls_event **done);
/* test back door, replaces `kernelCache.getIfPresent(closure)` (TensorsSpec:48-52) */
int dls_plan_cache_contains(dls_ctx *ctx, const uint8_t *tree, size_t tree_len, int *found);
/* describe how a tree lowers (kernel family, program length, leaf modes) into `text` -- diagnostics only */
int dls_plan_describe(dls_ctx *ctx, const uint8_t *tree, size_t tree_len, char *text, size_t cap);

/* ---- reductions: replaces Tensor.sum = reduce(PlusPrograms) Tensors:673-766 (kernels Tensors:313-392) */
int dls_sum(dls_ctx *ctx, dls_buf *in, size_t n_floats, dls_event *const *deps, int n_deps, dls_buf **out,
            dls_event **done);
/* out[o, i] = sum_k in[o, k, i] over a row-major [outer, k, inner] view: the batch-axis gradient sums that
 * TensorLayers.sumAs unrolls on the host (TL:124-162) */
int dls_sum_axis(dls_ctx *ctx, dls_buf *in, size_t outer, size_t k, size_t inner, dls_event *const *deps, int n_deps,
                 dls_buf **out, dls_event **done);
/* The full sum above for callers that may not allocate (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  No counterpart in the reference.
 * out[0] = sum of in[0 .. n_floats) into a buffer the caller already owns; the grid scratch lives in a file-scope
 * __device__ array (8448 bytes per GPU that loads the library, held until exit), so a call makes no cudaMalloc,
 * cudaFree, event or allocator call.  n_floats above DLS_SUM_STATIC_MAX_FLOATS is rejected with DLS_ERR_INVALID.
 * Every call on a device shares that array: calls ordered on one stream are safe, calls that overlap on the GPU
 * race.  The first call on a device resolves the array's address; make it before opening a capture.  Same lane
 * arithmetic as the allocating entry, so on devices with at most 256 SMs the two results are bit-identical. */
#define DLS_SUM_STATIC_MAX_FLOATS 268435456u      /* 2^28 floats = 1 GiB of input                            */
int dls_sum_static(dls_ctx *ctx, dls_buf *in, size_t n_floats, dls_buf *out);

/* ---- RNG: replaces the `random` / `random_normal` programs Tensors:398-443 and their launches Tensors:479-524 */
int dls_random(dls_ctx *ctx, size_t n_floats, uint32_t seed, dls_buf **out, dls_event **done);
int dls_random_normal(dls_ctx *ctx, size_t n_floats, uint32_t seed, dls_buf **out, dls_event **done);

/* ---- weights: fused `data - lr * delta` in place over a weight buffer; replaces (data - delta).doCache + buffer swap
 * TW:63-72 when the delta is already materialised (gradient arena after the all-reduce). */
int dls_sgd_update(dls_ctx *ctx, dls_buf *weight, dls_buf *delta, size_t n_floats, float learning_rate,
                   dls_event *const *deps, int n_deps, dls_event **done);

/* ---- stateful optimizers behind TensorOptimizerApi.delta (TW:155-161, Weights:46-51; the reference has learning rate /
 * L1 / L2 for scalars only, CumulativeFloatLayersSpec:18-47): weight, state and materialised gradient updated IN PLACE by
 * one kernel.  `hyper` is a device buffer of six floats (device memory, so that a captured step replays with the values
 * the host wrote before THIS replay -- Adam's bias corrections change every step):
 *     momentum = {lr, mu, weight_decay, nesterov(0|1), 0, 0}
 *     adam     = {lr / (1 - beta1^t), beta1, beta2, eps, weight_decay, 1 / (1 - beta2^t)}
 * The rounding order is fixed (csrc/aux_kernels.cu) and restated by oracle/optim_oracle.py.  In-place, so only valid once
 * every launch that reads the old weight has been issued (end of the step; harness `defer_updates`). */
enum dls_optimizer_kind { DLS_OPT_MOMENTUM = 0, DLS_OPT_ADAM = 1 };
int dls_optimizer_step(dls_ctx *ctx, int kind, dls_buf *weight, dls_buf *state1, dls_buf *state2, dls_buf *grad,
                       size_t n_floats, dls_buf *hyper);

/* ---- dense contraction C[M,N] = A[M,K] * B[K,N] (row-major, fp32 in/out); the lowering target for the unrolled
 * matmul trees of TensorsSpec:337-344 / :371-383 and the im2col convolution of TensorsSpec:144-210. */
int dls_matmul(dls_ctx *ctx, dls_buf *a, dls_buf *b, size_t m, size_t n, size_t k, int transpose_a, int transpose_b,
               dls_event *const *deps, int n_deps, dls_buf **out, dls_event **done);

/* Same contraction with an explicit precision and a fused epilogue, out = act(op(A) op(B) + bias): the body of a custom
 * TensorLayer forward (`TensorLayer(forward)`, TL:98-108) such as tanh(W x + b) in one kernel.  mode = -1 takes the
 * context's gemm_mode.  BF16 / TF32 run on the tcgen05 tensor cores with fp32 accumulation; stated tolerance against the
 * fp32 chain: 1e-3 relative (TF32), 4e-3 relative (BF16). */
enum dls_bias_mode { DLS_BIAS_NONE = 0, DLS_BIAS_ROW = 1 /* bias[m]: left-aligned broadcast, Tensors:208-222 */, DLS_BIAS_COL = 2 /* bias[n] */ };
enum dls_activation { DLS_ACT_NONE = 0, DLS_ACT_TANH = 1, DLS_ACT_RELU = 2 /* Tensor.max(x, fill 0) */ };
typedef struct dls_gemm_opts {
  int32_t mode;        /* enum dls_gemm_mode, or -1 = context default */
  int32_t bias_mode;   /* enum dls_bias_mode   */
  int32_t activation;  /* enum dls_activation  */
  int32_t flags;       /* DLS_GEMM_EMIT_OPERAND_COPY | (enum dls_gate << 8) */
} dls_gemm_opts;
/* flags bit 0: the kernel also writes the result once more in the mode's operand type (bf16), kept as the output buffer's
 * operand copy, so a contraction that consumes the result next starts without a staging pass (BF16 mode; ignored elsewhere). */
#define DLS_GEMM_EMIT_OPERAND_COPY 1
/* flags bits 8-9, dls_matmul_gated: out = (op(A) op(B) [+ bias, activation]) * act'(gate), gate an [M, N] fp32 buffer -- the
 * delta a custom layer's backward sends down, `delta * (1 - y * y)` (tanh; FloatLayers.scala derivative rules restated for
 * tensors) or `delta * min(y * FLT_MAX, 1)` (= 1[y > 0], relu), evaluated in the epilogue of the contraction that produces
 * `delta` instead of by a second elementwise launch over the materialised product.  Same nodes, one rounding each. */
enum dls_gate { DLS_GATE_NONE = 0, DLS_GATE_TANH = 1, DLS_GATE_RELU = 2 };
int dls_matmul_ex(dls_ctx *ctx, dls_buf *a, dls_buf *b, dls_buf *bias, size_t m, size_t n, size_t k, int transpose_a,
                  int transpose_b, const dls_gemm_opts *opts, dls_event *const *deps, int n_deps, dls_buf **out,
                  dls_event **done);
int dls_matmul_gated(dls_ctx *ctx, dls_buf *a, dls_buf *b, dls_buf *bias, dls_buf *gate, size_t m, size_t n, size_t k,
                     int transpose_a, int transpose_b, const dls_gemm_opts *opts, dls_event *const *deps, int n_deps,
                     dls_buf **out, dls_event **done);
/* weight[M,N] -= learning_rate * (op(A) op(B)), in place, in ONE kernel: the contraction is a weight gradient (TL:124-162 sums
 * over the batch) and its epilogue is the update `data - delta * learningRate` of TW:63-72 (same two roundings), so the
 * gradient never exists in HBM -- north_star's "gradient accumulation and the optimizer step fused into one pass over each
 * weight buffer".  BF16 mode also leaves the new weight's bf16 operand copy.  In place, so only valid once every launch that
 * reads the old weight has been issued (the caller orders a layer's backward-data pass first).  Tensor-core modes only
 * (mode = -1: the context's gemm_mode); DLS_ERR_UNSUPPORTED otherwise: materialise the gradient and use dls_sgd_update. */
int dls_matmul_sgd(dls_ctx *ctx, dls_buf *a, dls_buf *b, dls_buf *weight, size_t m, size_t n, size_t k, int transpose_a,
                   int transpose_b, float learning_rate, int mode);

/* ---- 'same' 3x3 convolution of NHWC tensors with [3,3,Cin,Cout] weights, without an im2col matrix.  The reference spells
 * a convolution as nine `translate`d views, `join` and an unrolled matmul (TensorsSpec:144-210 `convolute`,
 * benchmarks.scala:412-632); these three entries are the lowering target of that composite and of its two gradients:
 *   forward           out[b,y,x,f] = act(bias[f] + sum_{ky,kx,c} xs[b, y-(ky-1), x-(kx-1), c] * w[ky,kx,c,f])
 *   backward_data     dx[b,y,x,c]  = sum_{ky,kx,f} dz[b, y+(ky-1), x+(kx-1), f] * w[ky,kx,c,f]
 *   backward_weights  dw[ky,kx,c,f] = sum_{b,y,x} xs[b, y-(ky-1), x-(kx-1), c] * dz[b,y,x,f]
 * with zero padding outside the image.  xs is the [batch,h,w,cin] input; with in_stride = s > 1 it is read in place from a
 * [batch, h*s, w*s, cin] buffer as xs[b,y,x,c] = x[b, s*y, s*x, c] -- the nearest-neighbour `scale` view (Tensors:950-965)
 * that precedes the convolution, never materialised; dz_stride is the mirror image for the delta of a down-sampling that
 * FOLLOWS the layer (its zeros are neither stored nor, in the first-layer kernels, visited).  Tensor-core modes only (mode = -1 takes the context's gemm_mode):
 * operands rounded to bf16 / tf32 like every contraction of the mode, fp32 accumulation; same stated tolerances as
 * dls_matmul_ex.  Cin <= 4 (first layer) runs on direct kernels (warp-level mma over a shared-memory patch), everything else as an implicit GEMM on the
 * tcgen05 tensor cores.  DLS_ERR_UNSUPPORTED (ask dls_conv3x3_supported first) means: compose it from views + dls_matmul. */
enum dls_conv_pass { DLS_CONV_FORWARD = 0, DLS_CONV_BACKWARD_DATA = 1, DLS_CONV_BACKWARD_WEIGHTS = 2 };
typedef struct dls_conv_opts {
  int32_t mode;        /* enum dls_gemm_mode, or -1 = context default */
  int32_t activation;  /* enum dls_activation (forward only)          */
  int32_t in_stride;   /* 0 / 1 = dense input; s > 1 = input sampled at stride s (forward, backward_weights) */
  int32_t dz_stride;   /* backward passes: 0 / 1 = dense delta; s > 1 = dz is [batch, h/s, w/s, cout] and stands for a full-size
                          delta that is zero except at (s*y, s*x): what a `scale` down-sampling after the layer hands back.
                          forward: 0 = nothing; s >= 1 = a hint that the value is consumed next by a contraction (s = 1) or by
                          a convolution behind a `scale` down-sampling by s: the kernel also leaves that consumer's bf16
                          operand copy (BF16 mode, cout % 64 == 0), so it starts without a staging pass */
} dls_conv_opts;
int dls_conv3x3_supported(dls_ctx *ctx, int pass, size_t h, size_t w, size_t cin, size_t cout, int mode);
int dls_conv3x3_forward(dls_ctx *ctx, dls_buf *x, dls_buf *w, dls_buf *bias, size_t batch, size_t h, size_t wd, size_t cin,
                        size_t cout, const dls_conv_opts *opts, dls_event *const *deps, int n_deps, dls_buf **out,
                        dls_event **done);
int dls_conv3x3_backward_data(dls_ctx *ctx, dls_buf *dz, dls_buf *w, size_t batch, size_t h, size_t wd, size_t cin, size_t cout,
                              const dls_conv_opts *opts, dls_event *const *deps, int n_deps, dls_buf **out, dls_event **done);
/* backward_data with the activation derivative of the layer BELOW in its epilogue (enum dls_gate, as dls_matmul_gated):
 * out = dx * act'(gate[b, gs*y, gs*x, c]), gate = that layer's pinned output [batch, h*gs, w*gs, cin], gs = the factor of the
 * `scale` down-sampling between the two layers (1: none) -- i.e. the (sampled) delta of the layer below in one kernel. */
int dls_conv3x3_backward_data_gated(dls_ctx *ctx, dls_buf *dz, dls_buf *w, dls_buf *gate, int gate_mode, int gate_stride,
                                    size_t batch, size_t h, size_t wd, size_t cin, size_t cout, const dls_conv_opts *opts,
                                    dls_event *const *deps, int n_deps, dls_buf **out, dls_event **done);
int dls_conv3x3_backward_weights(dls_ctx *ctx, dls_buf *x, dls_buf *dz, size_t batch, size_t h, size_t wd, size_t cin,
                                 size_t cout, const dls_conv_opts *opts, dls_event *const *deps, int n_deps, dls_buf **out,
                                 dls_event **done);

/* backward_weights that also returns the bias gradient sum_{b,y,x} dz[b,y,x,f] (fp32 sums of the unrounded delta) from the same
 * pass over dz.  First-layer kernels (Cin <= 4) only: DLS_ERR_UNSUPPORTED elsewhere (use dls_sum_axis). */
int dls_conv3x3_backward_weights_bias(dls_ctx *ctx, dls_buf *x, dls_buf *dz, size_t batch, size_t h, size_t wd, size_t cin,
                                      size_t cout, const dls_conv_opts *opts, dls_event *const *deps, int n_deps,
                                      dls_buf **out, dls_buf **out_bias, dls_event **done);

/* ---- data-parallel exchange (new functionality: the reference has no collective) */
int dls_comm_unique_id(uint8_t *id128);                              /* 128-byte NCCL unique id (rank 0)   */
int dls_comm_init(dls_ctx *ctx, const uint8_t *id128, int rank, int world_size);
int dls_allreduce_sum(dls_ctx *ctx, dls_buf *buf, size_t n_floats, dls_event *const *deps, int n_deps,
                      dls_event **done);
/* one bucket: every buffer SUM-all-reduced in place by one grouped NCCL launch on the exchange stream, overlapping the
 * compute stream until a later launch touches one of the buffers (single-stream contexts; recordable) */
int dls_allreduce_sum_many(dls_ctx *ctx, dls_buf *const *bufs, const size_t *n_floats, int n_bufs);
/* The end of a data-parallel step as ONE kernel over NVLink peer memory (csrc/peer.cu).  dls_peer_init (collective, after
 * dls_comm_init, ranks on one node) allocates this rank's peer-visible staging area of `staging_floats` x 2 and maps
 * every peer's (cudaIpc handles all-gathered over the communicator).  dls_peer_sgd_step then does, in one launch:
 *   for every i with exchange[i] != 0: grads[i] = sum over the ranks, added in rank order (bit-identical on every rank);
 *   weights[i] = weights[i] - learning_rate * grads[i], in place (TW:64's `data - delta` with the optimizer hook of
 *   TW:155-161 folded in); segments with exchange[i] == 0 are updated from their local (already reduced) gradient.
 * weights == NULL: the sums are written back into grads[i] instead (all-reduce only, for stateful optimizers).
 * Without any exchanged segment it is a multi-tensor SGD and needs neither communicator nor dls_peer_init.
 * In place, so only valid at the end of the step; recordable.  Every rank must pass the same sizes in the same order.
 * A peer that does not arrive within 20 s makes the next dls_ctx_sync / dls_buf_read return DLS_ERR_COMM. */
int dls_peer_init(dls_ctx *ctx, size_t staging_floats);
int dls_peer_ready(dls_ctx *ctx);                                    /* 1 after a successful dls_peer_init */
int dls_peer_sgd_step(dls_ctx *ctx, dls_buf *const *weights, dls_buf *const *grads, const size_t *n_floats,
                      const int *exchange, int n, float learning_rate);
int dls_comm_destroy(dls_ctx *ctx);

/* ---- captured steps (new functionality).  The reference dispatches every tensor materialisation separately:
 * clSetKernelArg + clEnqueueNDRangeKernel per launch (Tensors:1203-1251, OpenCL:826) and one clCreateBuffer per output
 * (Tensors:1217).  Between dls_graph_begin and dls_graph_end every launch of this library is RECORDED instead of
 * executed (CUDA stream capture); dls_graph_launch replays the whole step with one driver call.  Buffers allocated
 * while recording come from an arena owned by the graph; host transfers, completion events and dls_ctx_sync are
 * rejected while a capture is open; single-stream contexts only.  See csrc/graph.cu for the full contract. */
typedef struct dls_graph dls_graph;
int dls_graph_begin(dls_ctx *ctx);
int dls_graph_end(dls_ctx *ctx, dls_graph **out);
int dls_graph_abort(dls_ctx *ctx);                /* drop an open capture (error paths); no-op when none is open */
int dls_graph_launch(dls_graph *graph);           /* asynchronous, on the context's stream */
int dls_graph_info(const dls_graph *graph, uint64_t *kernel_nodes, uint64_t *nodes, uint64_t *arena_bytes);
void dls_graph_retain(dls_graph *graph);
void dls_graph_release(dls_graph *graph);
/* device-to-device copy of the first n_floats (stream ordered, recordable): a captured step uses it to write a swapped
 * weight (TW:63-72) back into the buffer the next replay reads */
int dls_buf_copy(dls_ctx *ctx, dls_buf *dst, dls_buf *src, size_t n_floats);

#ifdef __cplusplus
}
#endif
#endif /* DLS_B200_H */
