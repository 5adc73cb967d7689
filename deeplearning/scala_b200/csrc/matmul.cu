// matmul.cu -- dense contraction C[M,N] = op(A) * op(B), row-major fp32 in / fp32 out.
// DLS_GEMM_FP32: CUDA-core tiled kernel (this file).  The tcgen05 / TMEM tensor-core path (TF32 / BF16 inputs,
// fp32 accumulate) lives in gemm_tc.cu and is selected by dls_ctx_opts.gemm_mode.
// Lowering target for the unrolled matmul trees (TensorsSpec:337-344, :371-383) and im2col convolution (:144-210).
#include "common.h"
#include "gemm_tc.h"

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

// out = act(out + bias) in place: the fused epilogue of the tensor-core kernel for the CUDA-core fp32 mode.
__global__ void __launch_bounds__(256) bias_act_kernel(float *__restrict__ c, size_t m, size_t n, const float *__restrict__ bias,
                                                       int bias_mode, int act) {
  const size_t total = m * n;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    float x = c[i];
    if (bias_mode == 1) x = __fadd_rn(x, __ldg(bias + i / n));
    else if (bias_mode == 2) x = __fadd_rn(x, __ldg(bias + i % n));
    if (act == 1) x = tanhf(x);
    else if (act == 2) x = fmaxf(x, 0.f);
    c[i] = x;
  }
}

int launch_bias_act(dls_ctx *ctx, int stream, float *c, size_t m, size_t n, const float *bias, int bias_mode, int act) {
  if (m * n == 0 || (bias_mode == 0 && act == 0)) return DLS_OK;
  const int grid = (int)std::min<size_t>((m * n + 255) / 256, (size_t)ctx->sm_count * 8);
  bias_act_kernel<<<grid, 256, 0, ctx->streams[stream]>>>(c, m, n, bias, bias_mode, act);
  DLS_CUDA(cudaGetLastError());
  ctx->stats.kernel_launches++;
  return DLS_OK;
}

// c *= act'(gate) in place (fp32 parity mode; the tensor-core kernel does it in its epilogue): same nodes, same roundings.
__global__ void __launch_bounds__(256) gate_kernel(float *__restrict__ c, size_t total, const float *__restrict__ gate, int mode) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const float h = __ldg(gate + i);
    c[i] = mode == 1 ? __fmul_rn(c[i], __fsub_rn(1.0f, __fmul_rn(h, h))) : __fmul_rn(c[i], fminf(__fmul_rn(h, 3.4028234663852886e38f), 1.0f));
  }
}

int launch_gate(dls_ctx *ctx, int stream, float *c, size_t total, const float *gate, int mode, float) {
  DLS_CHECK(mode != 3, DLS_ERR_UNSUPPORTED, "the in-place update epilogue exists in the tensor-core modes only");
  if (total == 0 || mode == 0) return DLS_OK;
  const int grid = (int)std::min<size_t>((total + 255) / 256, (size_t)ctx->sm_count * 8);
  gate_kernel<<<grid, 256, 0, ctx->streams[stream]>>>(c, total, gate, mode);
  DLS_CUDA(cudaGetLastError());
  ctx->stats.kernel_launches++;
  return DLS_OK;
}

}  // namespace dls

using namespace dls;

namespace {

// fp32 operand -> what the tensor-core kernel wants.  BF16: a bf16 copy with the same orientation (both major-nesses
// are native to the kernel).  TF32: an fp32 copy rounded to tf32 (round to nearest; the tensor core would truncate), K-major
// (transposed when the operand is stored MN-major), rows padded to 16 bytes.  Copies are cached on the buffer (dls_buf::shadow) until it is written again.
//   rows x cols       : the stored row-major shape
//   contraction_is_row: the contraction index is the ROW index, i.e. the operand is MN-major as stored
int stage_operand(dls_ctx *ctx, int stream, int mode, dls_buf *buf, long long rows, long long cols, bool contraction_is_row,
                  const void **ptr, long long *pitch, bool *mn_major, Block *private_blk = nullptr) {
  const float *src = buf->ptr;
  const int slot = mode == DLS_GEMM_BF16 ? 0 : 1;
  long long no_rows = 0, no_cols = 0;
  Block &sh = private_blk ? *private_blk : buf->shadow[slot];
  long long &tag_r = private_blk ? no_rows : buf->shadow_rows[slot];
  long long &tag_c = private_blk ? no_cols : buf->shadow_cols[slot];
  // TF32 operands stored MN-major are read in place through the 32-byte-atom swizzle (gemm_tc.cu); DLS_TF32_MN=0 goes back to
  // the transposing staging pass that made them K-major
  static const bool tf32_mn = [] {
    const char *e = getenv("DLS_TF32_MN");
    return !(e && atoi(e) == 0);
  }();
  const bool transposed = mode == DLS_GEMM_TF32 && contraction_is_row && !tf32_mn;
  const long long p = mode == DLS_GEMM_BF16 ? ((cols + 7) & ~7LL) : transposed ? ((rows + 3) & ~3LL) : ((cols + 3) & ~3LL);
  // slot 1 holds either the transposed or the re-pitched fp32 copy; the sign of shadow_rows tells which
  const long long tag_rows = transposed ? -rows : rows;
  // a staged copy made outside the capture that is open now (or inside one that has ended) holds data the graph being
  // recorded will not reproduce on replay: only copies staged in the current arena count as fresh
  // a dense bf16 copy (pitch == cols, cols % 8 == 0) is the same memory under every [rows, cols] reading of the buffer
  const bool same_flat = mode == DLS_GEMM_BF16 && sh.ptr && tag_c > 0 && tag_c % 8 == 0 && cols % 8 == 0 && tag_r > 0 &&
                         tag_r * tag_c == rows * cols;
  if (!(sh.ptr && ((tag_r == tag_rows && tag_c == cols) || same_flat) && sh.pool == ctx->capture_pool)) {
    if (sh.ptr) free_block(ctx, sh);
    int rc;
    if (mode == DLS_GEMM_BF16) {
      rc = alloc_block(ctx, (size_t)(rows * p) * 2, stream, &sh);
      if (!rc) rc = launch_to_bf16(ctx, stream, src, sh.ptr, rows, cols, p);
    } else if (transposed) {
      rc = alloc_block(ctx, (size_t)(cols * p) * 4, stream, &sh);
      if (!rc) rc = launch_transpose_f32(ctx, stream, src, (float *)sh.ptr, (int)rows, (int)cols, p);
    } else {
      rc = alloc_block(ctx, (size_t)(rows * p) * 4, stream, &sh);
      if (!rc) rc = launch_to_tf32(ctx, stream, src, (float *)sh.ptr, rows, cols, p);
    }
    if (rc) {
      if (sh.ptr) free_block(ctx, sh);
      return rc;
    }
    tag_r = tag_rows;
    tag_c = cols;
  } else {
    // reused from another stream of the pool: wait for the staging kernel that produced the copy (and, being
    // conservative, for everything else those streams have queued)
    int rc = wait_streams(ctx, stream, sh.stream_mask);
    if (rc) return rc;
    sh.stream_mask |= 1u << stream;
  }
  *ptr = sh.ptr;
  *pitch = p;
  *mn_major = (mode == DLS_GEMM_BF16 || tf32_mn) ? contraction_is_row : false;
  return DLS_OK;
}

}  // namespace

namespace dls {

// c[M, N] = act(op(a) op(b) + bias) on `stream` (inputs already waited for): the body of dls_matmul_ex, also the
// target of trees that the lowering recognised as one dense contraction (Plan::contraction).
int run_matmul(dls_ctx *ctx, int stream, int mode, dls_buf *a, dls_buf *b, dls_buf *bias, size_t m, size_t n, size_t k,
               bool transpose_a, bool transpose_b, int bias_mode, int activation, dls_buf *c, dls_buf *gate, int gate_mode,
               Block *operand_copy, float gate_scale) {
  if (m * n == 0) return DLS_OK;
  const float *bias_ptr = bias_mode ? bias->ptr : nullptr;
  const float *gate_ptr = gate_mode ? gate->ptr : nullptr;
  if (k == 0) {
    DLS_CUDA(cudaMemsetAsync(c->ptr, 0, m * n * sizeof(float), ctx->streams[stream]));
    int rc = launch_bias_act(ctx, stream, c->ptr, m, n, bias_ptr, bias_mode, activation);
    if (!rc) rc = launch_gate(ctx, stream, c->ptr, m * n, gate_ptr, gate_mode, gate_scale);
    return rc;
  }
  if (mode == DLS_GEMM_FP32) {
    int rc = launch_sgemm(ctx, stream, a->ptr, b->ptr, c->ptr, (int)m, (int)n, (int)k, transpose_a, transpose_b);
    if (!rc) rc = launch_bias_act(ctx, stream, c->ptr, m, n, bias_ptr, bias_mode, activation);
    if (!rc) rc = launch_gate(ctx, stream, c->ptr, m * n, gate_ptr, gate_mode, gate_scale);
    return rc;
  }
  const void *pa, *pb;
  long long lda, ldb;
  bool a_mn, b_mn;
  // a is stored [K][M] when transposed (contraction index = row index), else [M][K]
  int rc = stage_operand(ctx, stream, mode, a, transpose_a ? k : m, transpose_a ? m : k, transpose_a, &pa, &lda, &a_mn);
  // b is stored [N][K] when transposed, else [K][N] (contraction index = row index); the same buffer on both sides with
  // two different interpretations must not evict the first staged copy
  Block own;
  if (!rc) rc = stage_operand(ctx, stream, mode, b, transpose_b ? n : k, transpose_b ? k : n, !transpose_b, &pb, &ldb, &b_mn,
                              a == b ? &own : nullptr);
  GemmEpilogue epi;
  epi.bias = bias_ptr;
  epi.bias_mode = bias_mode;
  epi.act = activation;
  epi.gate = gate_ptr;
  epi.gate_mode = gate_mode;
  epi.gate_scale = gate_scale;
  if (!rc && operand_copy && mode == DLS_GEMM_BF16) {
    // the result's own bf16 operand copy, written by the epilogue: [m][n rounded up to 8] like stage_operand's
    epi.copy_ld = ((long long)n + 7) & ~7LL;
    rc = alloc_block(ctx, (size_t)m * epi.copy_ld * 2, stream, operand_copy);
    epi.copy = operand_copy->ptr;
  }
  if (!rc) rc = launch_gemm_tc(ctx, stream, mode, pa, a_mn, lda, pb, b_mn, ldb, c->ptr, (long long)n, (int)m, (int)n, (int)k, epi);
  if (own.ptr) free_block(ctx, own);
  return rc;
}

}  // namespace dls

extern "C" int dls_matmul_gated(dls_ctx *ctx, dls_buf *a, dls_buf *b, dls_buf *bias, dls_buf *gate, size_t m, size_t n, size_t k,
                                int transpose_a, int transpose_b, const dls_gemm_opts *opts, dls_event *const *deps, int n_deps,
                                dls_buf **out, dls_event **done) {
  DLS_CHECK(ctx && a && b && out, DLS_ERR_INVALID, "dls_matmul: NULL argument");
  DLS_CHECK(m * k <= a->n && k * n <= b->n, DLS_ERR_INVALID, "dls_matmul: operand shorter than its shape");
  DLS_CHECK(m < (1u << 31) && n < (1u << 31) && k < (1u << 31), DLS_ERR_UNSUPPORTED, "dls_matmul: dimension too large");
  dls_gemm_opts o{};
  o.mode = -1;
  if (opts) o = *opts;
  const int mode = o.mode < 0 ? ctx->opts.gemm_mode : o.mode;
  DLS_CHECK(mode == DLS_GEMM_FP32 || mode == DLS_GEMM_TF32 || mode == DLS_GEMM_BF16, DLS_ERR_INVALID, "dls_matmul: bad gemm mode");
  DLS_CHECK(o.bias_mode >= 0 && o.bias_mode <= 2 && o.activation >= 0 && o.activation <= 2, DLS_ERR_INVALID,
            "dls_matmul: bad epilogue options");
  DLS_CHECK(o.bias_mode == 0 || (bias && bias->n >= (o.bias_mode == 1 ? m : n)), DLS_ERR_INVALID,
            "dls_matmul: bias buffer missing or too short");
  const int gate_mode = (o.flags >> 8) & 3;
  DLS_CHECK(gate_mode <= 2 && (gate_mode == 0 || (gate && gate->n >= m * n)), DLS_ERR_INVALID,
            "dls_matmul_gated: gate buffer missing or shorter than the product");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CUDA(cudaSetDevice(ctx->device));
  const int stream = pick_stream(ctx);
  dls_buf *ins[4] = {a, b, o.bias_mode ? bias : nullptr, gate_mode ? gate : nullptr};
  int rc = wait_inputs(ctx, stream, ins, 4, deps, n_deps);
  if (rc) return rc;
  dls_buf *c = nullptr;
  rc = new_buf(ctx, m * n, stream, &c);
  if (rc) return rc;
  Block copy;
  rc = run_matmul(ctx, stream, mode, a, b, bias, m, n, k, transpose_a != 0, transpose_b != 0, o.bias_mode, o.activation, c, gate,
                  gate_mode, (o.flags & DLS_GEMM_EMIT_OPERAND_COPY) ? &copy : nullptr, 0.f);
  if (!rc) rc = finish_output(ctx, stream, &c, 1, ins, 4, done);
  if (rc) {
    if (copy.ptr) free_block(ctx, copy);
    dls_buf_release(c);
    return rc;
  }
  if (copy.ptr) {               // after finish_output (which drops the stale shadows of a written buffer)
    copy.stream_mask |= 1u << stream;
    c->shadow[0] = copy;
    c->shadow_rows[0] = (long long)m;
    c->shadow_cols[0] = (long long)n;
  }
  *out = c;
  return DLS_OK;
}

extern "C" int dls_matmul_sgd(dls_ctx *ctx, dls_buf *a, dls_buf *b, dls_buf *weight, size_t m, size_t n, size_t k, int transpose_a,
                              int transpose_b, float learning_rate, int mode) {
  DLS_CHECK(ctx && a && b && weight, DLS_ERR_INVALID, "dls_matmul_sgd: NULL argument");
  DLS_CHECK(weight != a && weight != b, DLS_ERR_INVALID, "dls_matmul_sgd: the weight is one of the operands");
  DLS_CHECK(m * k <= a->n && k * n <= b->n && m * n <= weight->n, DLS_ERR_INVALID, "dls_matmul_sgd: a buffer is shorter than its shape");
  DLS_CHECK(m < (1u << 31) && n < (1u << 31) && k < (1u << 31) && k > 0 && m * n > 0, DLS_ERR_UNSUPPORTED, "dls_matmul_sgd: bad dimension");
  if (mode < 0) mode = ctx->opts.gemm_mode;
  DLS_CHECK(mode == DLS_GEMM_TF32 || mode == DLS_GEMM_BF16, DLS_ERR_UNSUPPORTED,
            "dls_matmul_sgd: tensor-core modes only (materialise the gradient and use dls_sgd_update otherwise)");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CUDA(cudaSetDevice(ctx->device));
  const int stream = ctx->n_streams == 1 ? 0 : weight->last_stream;
  dls_buf *ins[3] = {a, b, weight};
  int rc = wait_inputs(ctx, stream, ins, 3, nullptr, 0);
  if (!rc) rc = wait_readers(ctx, stream, weight);   // in place: every launch that still reads the old weight comes first
  if (rc) return rc;
  Block copy;
  rc = run_matmul(ctx, stream, mode, a, b, nullptr, m, n, k, transpose_a != 0, transpose_b != 0, 0, 0, weight, weight, 3, &copy,
                  learning_rate);
  if (!rc) rc = finish_output(ctx, stream, &weight, 1, ins, 2, nullptr);
  if (rc) {
    if (copy.ptr) free_block(ctx, copy);
    return rc;
  }
  if (copy.ptr) {               // the new weight's operand copy: forward and backward-data of the next step need no staging pass
    copy.stream_mask |= 1u << stream;
    weight->shadow[0] = copy;
    weight->shadow_rows[0] = (long long)m;
    weight->shadow_cols[0] = (long long)n;
  }
  return DLS_OK;
}

extern "C" int dls_matmul_ex(dls_ctx *ctx, dls_buf *a, dls_buf *b, dls_buf *bias, size_t m, size_t n, size_t k, int transpose_a,
                             int transpose_b, const dls_gemm_opts *opts, dls_event *const *deps, int n_deps, dls_buf **out,
                             dls_event **done) {
  return dls_matmul_gated(ctx, a, b, bias, nullptr, m, n, k, transpose_a, transpose_b, opts, deps, n_deps, out, done);
}

extern "C" int dls_matmul(dls_ctx *ctx, dls_buf *a, dls_buf *b, size_t m, size_t n, size_t k, int transpose_a, int transpose_b,
                          dls_event *const *deps, int n_deps, dls_buf **out, dls_event **done) {
  return dls_matmul_ex(ctx, a, b, nullptr, m, n, k, transpose_a, transpose_b, nullptr, deps, n_deps, out, done);
}
