// gemm_tc.h -- internal interface of the tcgen05 / TMEM contraction kernel (gemm_tc.cu).
#pragma once
#include "common.h"

namespace dls {

struct GemmEpilogue {
  const float *bias = nullptr;
  int bias_mode = 0;   // 0 none, 1 bias[row] (feature-major layers: left-aligned broadcast, Tensors:208-222), 2 bias[col]
  int act = 0;         // 0 none, 1 tanh, 2 relu (= Tensor.max(x, fill 0))
};

// D[M,N] fp32 row-major (pitch ldc) = A * B (+ epilogue).
//   kind  : DLS_GEMM_BF16 (operands are bf16 in HBM) or DLS_GEMM_TF32 (operands are fp32 in HBM, read as tf32)
//   a_mn  : false -> A stored [M][K] (K contiguous, pitch lda elements); true -> stored [K][M] (M contiguous)
//   b_mn  : false -> B stored [N][K] (K contiguous, pitch ldb);          true -> stored [K][N] (N contiguous)
// Operand rows must start 16-byte aligned (pitch * element size % 16 == 0); TF32 operands must be K-major.
int launch_gemm_tc(dls_ctx *ctx, int stream, int kind, const void *A, bool a_mn, long long lda, const void *B, bool b_mn,
                   long long ldb, float *C, long long ldc, int M, int N, int K, const GemmEpilogue &epi);

// fp32 [rows, cols] dense -> bf16 [rows, pitch] (pitch % 8 == 0, pad columns zero), round to nearest even.
int launch_to_bf16(dls_ctx *ctx, int stream, const float *in, void *out, long long rows, long long cols, long long pitch);

// fp32 [rows, cols] dense -> tf32-rounded (RNA) fp32 [rows, pitch] (pitch % 4 == 0, pad columns zero)
int launch_to_tf32(dls_ctx *ctx, int stream, const float *in, float *out, long long rows, long long cols, long long pitch);

// fp32 [rows, cols] dense -> tf32-rounded fp32 [cols, pitch] (transpose; pitch % 4 == 0)
int launch_transpose_f32(dls_ctx *ctx, int stream, const float *in, float *out, int rows, int cols, long long pitch);

// matmul.cu
int run_matmul(dls_ctx *ctx, int stream, int mode, dls_buf *a, dls_buf *b, dls_buf *bias, size_t m, size_t n, size_t k,
               bool transpose_a, bool transpose_b, int bias_mode, int activation, dls_buf *c);
int launch_sgemm(dls_ctx *ctx, int stream, const float *A, const float *B, float *C, int M, int N, int K, bool ta, bool tb);

}  // namespace dls
