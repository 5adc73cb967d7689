// matmul.cu -- dense contraction C[M,N] = op(A) * op(B), row-major fp32 in / fp32 out.
// DLS_GEMM_FP32: CUDA-core tiled kernel (this file).  The tcgen05 / TMEM tensor-core path (TF32 / BF16 inputs,
// fp32 accumulate) lives in gemm_tc.cu and is selected by dls_ctx_opts.gemm_mode.
// Lowering target for the unrolled matmul trees (TensorsSpec:337-344, :371-383) and im2col convolution (:144-210).
#include "common.h"

namespace dls {

constexpr int TM = 64, TN = 64, TK = 16;

// 64x64 output tile per 256-thread block, 4x4 register tile per thread, k ascending per output element so the
// result equals the reference's left fold ((a0*b0 + a1*b1) + a2*b2) + ... bit for bit when started from the first term.
template <bool TA, bool TB>
__global__ void __launch_bounds__(256) sgemm_kernel(const float *__restrict__ A, const float *__restrict__ B,
                                                    float *__restrict__ C, int M, int N, int K) {
  __shared__ float As[TK][TM + 4];
  __shared__ float Bs[TK][TN + 4];
  const int tx = threadIdx.x % 16, ty = threadIdx.x / 16;
  const int m0 = blockIdx.y * TM, n0 = blockIdx.x * TN;
  float acc[4][4];
  bool first = true;
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int k0 = 0; k0 < K; k0 += TK) {
    for (int idx = threadIdx.x; idx < TM * TK; idx += 256) {
      int kk, mm;
      if (TA) { mm = idx % TM; kk = idx / TM; } else { kk = idx % TK; mm = idx / TK; }
      int gm = m0 + mm, gk = k0 + kk;
      float v = 0.f;
      if (gm < M && gk < K) v = TA ? A[(size_t)gk * M + gm] : A[(size_t)gm * K + gk];
      As[kk][mm] = v;
    }
    for (int idx = threadIdx.x; idx < TN * TK; idx += 256) {
      int kk, nn;
      if (TB) { kk = idx % TK; nn = idx / TK; } else { nn = idx % TN; kk = idx / TN; }
      int gn = n0 + nn, gk = k0 + kk;
      float v = 0.f;
      if (gn < N && gk < K) v = TB ? B[(size_t)gn * K + gk] : B[(size_t)gk * N + gn];
      Bs[kk][nn] = v;
    }
    __syncthreads();
    const int kmax = min(TK, K - k0);
    for (int kk = 0; kk < kmax; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          float p = __fmul_rn(a[i], b[j]);
          acc[i][j] = first ? p : __fadd_rn(acc[i][j], p);
        }
      first = false;
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    int gm = m0 + ty * 4 + i;
    if (gm >= M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int gn = n0 + tx * 4 + j;
      if (gn < N) C[(size_t)gm * N + gn] = acc[i][j];
    }
  }
}

int launch_sgemm(dls_ctx *ctx, int stream, const float *A, const float *B, float *C, int M, int N, int K, bool ta, bool tb) {
  dim3 grid((N + TN - 1) / TN, (M + TM - 1) / TM);
  cudaStream_t st = ctx->streams[stream];
  if (!ta && !tb) sgemm_kernel<false, false><<<grid, 256, 0, st>>>(A, B, C, M, N, K);
  else if (ta && !tb) sgemm_kernel<true, false><<<grid, 256, 0, st>>>(A, B, C, M, N, K);
  else if (!ta && tb) sgemm_kernel<false, true><<<grid, 256, 0, st>>>(A, B, C, M, N, K);
  else sgemm_kernel<true, true><<<grid, 256, 0, st>>>(A, B, C, M, N, K);
  DLS_CUDA(cudaGetLastError());
  ctx->stats.kernel_launches++;
  return DLS_OK;
}

}  // namespace dls

using namespace dls;

extern "C" int dls_matmul(dls_ctx *ctx, dls_buf *a, dls_buf *b, size_t m, size_t n, size_t k, int transpose_a, int transpose_b,
                          dls_event *const *deps, int n_deps, dls_buf **out, dls_event **done) {
  DLS_CHECK(ctx && a && b && out, DLS_ERR_INVALID, "dls_matmul: NULL argument");
  DLS_CHECK(m * k <= a->n && k * n <= b->n, DLS_ERR_INVALID, "dls_matmul: operand shorter than its shape");
  DLS_CHECK(m < (1u << 31) && n < (1u << 31) && k < (1u << 31), DLS_ERR_UNSUPPORTED, "dls_matmul: dimension too large");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CUDA(cudaSetDevice(ctx->device));
  const int stream = pick_stream(ctx);
  dls_buf *ins[2] = {a, b};
  int rc = wait_inputs(ctx, stream, ins, 2, deps, n_deps);
  if (rc) return rc;
  dls_buf *o = nullptr;
  rc = new_buf(ctx, m * n, stream, &o);
  if (rc) return rc;
  if (m * n != 0) {
    if (k == 0) {
      cudaMemsetAsync(o->ptr, 0, m * n * sizeof(float), ctx->streams[stream]);
    } else {
      rc = launch_sgemm(ctx, stream, a->ptr, b->ptr, o->ptr, (int)m, (int)n, (int)k, transpose_a != 0, transpose_b != 0);
    }
  }
  if (!rc) rc = finish_output(ctx, stream, &o, 1, ins, 2, done);
  if (rc) {
    dls_buf_release(o);
    return rc;
  }
  *out = o;
  return DLS_OK;
}
