// static_scratch.cu -- dls_sum_static: the full sum of dls_sum (N3, aux_kernels.cu) for callers that may not allocate.
//
// No counterpart in the reference.  Code under CUDA graph capture, tight
// serving loops and harnesses that forbid cudaMalloc cannot use dls_sum, which takes its one-float result from the
// context's allocator.  This entry point writes into a buffer the caller already owns and keeps its grid scratch (one
// partial per CTA + the arrival counter) in a file-scope __device__ array, so a call issues stream waits, one 4-byte
// memset node and one kernel -- no cudaMalloc, cudaFree, event creation or allocator traffic.
//
// Limits and sharing:
//   * the array is kStaticScratchBytes (8448 B) of device memory per GPU that loads the library, held until exit;
//   * at most DLS_SUM_STATIC_MAX_FLOATS (2^28) input floats per call: kStaticPartials CTAs x kStaticFoldPerBlock
//     floats each, which keeps every lane's serial fold no longer than the one dls_sum is tested with;
//   * every call on a device shares the array.  Calls ordered on one stream (a single-stream context, the default)
//     are safe; calls that overlap on the GPU (two contexts on one device, two streams of a multi-stream context,
//     two captured steps replayed on different streams) race on the partials and the counter and give wrong sums;
//   * the array keeps the previous call's contents: the arrival counter is cleared on the stream before each launch
//     (an aborted launch cannot poison the next one); the partials are fully overwritten before they are read.
#include <algorithm>
#include <mutex>

#include "common.h"
#include "reduce.cuh"

namespace dls {

constexpr int kStaticPartials = 2048;                 // CTAs one launch may use (148 x 8 = 1184 on B200)
constexpr size_t kStaticFoldPerBlock = 131072;        // floats one CTA folds at most at the size limit
constexpr int kStaticCounterSlot = kStaticPartials;   // the counter sits on its own 256-byte line behind the partials
constexpr int kStaticScratchFloats = kStaticPartials + 64;
constexpr size_t kStaticScratchBytes = kStaticScratchFloats * sizeof(float);
static_assert((size_t)kStaticPartials * kStaticFoldPerBlock == DLS_SUM_STATIC_MAX_FLOATS, "header and scratch disagree");
static_assert(kStaticScratchBytes == 8448, "the size stated in include/dls_b200.h, DESIGN.md and INTEGRATION.md");

// In HOST code this name is only a stand-in for the symbol (never a device address): the address the kernel gets
// comes from cudaGetSymbolAddress below.
static __device__ __align__(256) float g_sum_static_scratch[kStaticScratchFloats];

// Same lane / CTA arithmetic as sum_kernel (aux_kernels.cu:12), so equal grids give bit-identical sums.
__global__ void __launch_bounds__(kBlock) sum_static_kernel(const float *__restrict__ in, size_t n, ReduceArgs R) {
  const size_t n4 = n / 4;
  const float4 *in4 = reinterpret_cast<const float4 *>(in);
  const size_t stride = (size_t)gridDim.x * kBlock;
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
  size_t i = (size_t)blockIdx.x * kBlock + threadIdx.x;
  for (; i + 3 * stride < n4; i += 4 * stride) {
    float4 v0 = __ldcs(in4 + i), v1 = __ldcs(in4 + i + stride), v2 = __ldcs(in4 + i + 2 * stride),
           v3 = __ldcs(in4 + i + 3 * stride);
    a0 += (v0.x + v0.y) + (v0.z + v0.w);
    a1 += (v1.x + v1.y) + (v1.z + v1.w);
    a2 += (v2.x + v2.y) + (v2.z + v2.w);
    a3 += (v3.x + v3.y) + (v3.z + v3.w);
  }
  for (; i < n4; i += stride) {
    float4 v = __ldcs(in4 + i);
    a0 += (v.x + v.y) + (v.z + v.w);
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) a1 += in[n4 * 4 + threadIdx.x];
  grid_sum_finish((a0 + a1) + (a2 + a3), R);
}

// Set-up, once per device, with that device current: resolve the symbol to its address on THIS device.
constexpr int kStaticMaxDevices = 64;
static std::mutex g_static_mu;
static float *g_static_ptr[kStaticMaxDevices] = {};

static int static_scratch_for(int device, float **out) {
  DLS_CHECK(device >= 0 && device < kStaticMaxDevices, DLS_ERR_UNSUPPORTED, "dls_sum_static: device ordinal out of range");
  std::lock_guard<std::mutex> lk(g_static_mu);
  if (!g_static_ptr[device]) {
    float *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress((void **)&p, g_sum_static_scratch);
    if (e != cudaSuccess) return cuda_fail(e, "cudaGetSymbolAddress(g_sum_static_scratch)", __FILE__, __LINE__);
    DLS_CHECK(p != nullptr, DLS_ERR_CUDA, "dls_sum_static: the scratch symbol resolved to a NULL device address");
    g_static_ptr[device] = p;
  }
  *out = g_static_ptr[device];
  return DLS_OK;
}

}  // namespace dls

using namespace dls;

extern "C" int dls_sum_static(dls_ctx *ctx, dls_buf *in, size_t n_floats, dls_buf *out) {
  DLS_CHECK(ctx && in && out, DLS_ERR_INVALID, "dls_sum_static: NULL argument");
  DLS_CHECK(n_floats <= (size_t)DLS_SUM_STATIC_MAX_FLOATS, DLS_ERR_INVALID,
            "dls_sum_static: n_floats exceeds DLS_SUM_STATIC_MAX_FLOATS (268435456), the most the static scratch "
            "array is sized for; use dls_sum");
  DLS_CHECK(n_floats <= in->n, DLS_ERR_INVALID, "dls_sum_static: length exceeds the input buffer");
  DLS_CHECK(out->n >= 1, DLS_ERR_INVALID, "dls_sum_static: the output buffer must hold at least one float");
  DLS_CHECK(out != in && out->ptr != in->ptr, DLS_ERR_INVALID, "dls_sum_static: the output must not alias the input");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  for (int k = 0; k < kShadows; ++k)   // dropping stale operand copies would go through the allocator
    DLS_CHECK(out->shadow[k].ptr == nullptr, DLS_ERR_INVALID,
              "dls_sum_static: the output buffer carries staged operand copies (it was a contraction operand); "
              "pass a buffer that has not been one");
  DLS_CHECK(ctx->n_streams == 1 || out->ready != nullptr, DLS_ERR_UNSUPPORTED,
            "dls_sum_static: in a multi-stream context the output buffer must have been written once before "
            "(its ready event would have to be created here)");
  DLS_CUDA(cudaSetDevice(ctx->device));
  float *scratch = nullptr;
  int rc = static_scratch_for(ctx->device, &scratch);
  if (rc) return rc;
  const int stream = ctx->n_streams == 1 ? 0 : out->last_stream;
  cudaStream_t st = ctx->streams[stream];
  dls_buf *both[2] = {in, out};
  rc = wait_inputs(ctx, stream, both, 2, nullptr, 0);
  if (!rc) rc = wait_readers(ctx, stream, out);
  if (rc) return rc;
  if (n_floats == 0) {
    DLS_CUDA(cudaMemsetAsync(out->ptr, 0, sizeof(float), st));
  } else {
    unsigned int *counter = reinterpret_cast<unsigned int *>(scratch + kStaticCounterSlot);
    DLS_CUDA(cudaMemsetAsync(counter, 0, sizeof(unsigned int), st));
    ReduceArgs R{scratch, counter, out->ptr};
    const size_t want = (n_floats / 4 + (size_t)kBlock * 4 - 1) / ((size_t)kBlock * 4);
    const size_t cap = std::min<size_t>((size_t)ctx->sm_count * 8, (size_t)kStaticPartials);
    const int grid = (int)std::max<size_t>(1, std::min(want, cap));
    sum_static_kernel<<<grid, kBlock, 0, st>>>(in->ptr, n_floats, R);
    DLS_CUDA(cudaGetLastError());
    ctx->stats.kernel_launches++;
  }
  // the bookkeeping of finish_output without its allocator / event-creation branches
  in->block.stream_mask |= 1u << stream;
  out->block.stream_mask |= 1u << stream;
  out->last_stream = stream;
  if (ctx->n_streams > 1) DLS_CUDA(cudaEventRecord(out->ready, st));
  return DLS_OK;
}
