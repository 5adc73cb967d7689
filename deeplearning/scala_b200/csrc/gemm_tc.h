// gemm_tc.h -- internal interface of the tcgen05 / TMEM contraction kernel (gemm_tc.cu).
#pragma once
#include "common.h"

namespace dls {

struct GemmEpilogue {
  const float *bias = nullptr;
  int bias_mode = 0;   // 0 none, 1 bias[row] (feature-major layers: left-aligned broadcast, Tensors:208-222), 2 bias[col]
  int act = 0;         // 0 none, 1 tanh, 2 relu (= Tensor.max(x, fill 0))
  // gate: the result is multiplied by the activation derivative evaluated on `gate` ([M, N] fp32, same layout as the output):
  //   1: x * (1 - g * g)  (`delta * (1 - y * y)`, tanh)      2: x * min(g * FLT_MAX, 1)  (`delta * 1[y > 0]`, relu)
  // i.e. the elementwise tree a layer's backward would otherwise launch on the materialised product, node for node.
  //   3: g - x * gate_scale  (`data - delta * learningRate`, TW:63-72: the contraction IS a weight gradient and the epilogue
  //      applies it; gate = the weight, which is also the output buffer -- every element is read and written by one thread)
  const float *gate = nullptr;
  int gate_mode = 0;
  float gate_scale = 0.f;
  // copy: a second copy of the result rounded to the contraction mode's operand type ([M][copy_ld] bf16), so the contraction
  // that consumes the result next does not need a staging pass (it becomes the output buffer's operand shadow)
  void *copy = nullptr;
  long long copy_ld = 0;
};

// Geometry of the implicit-GEMM convolution modes of gemm_tc_kernel (zero for a plain GEMM).
struct ConvGeom {
  int H = 0, W = 0;               // spatial size of the NHWC tensor whose pixels are tiled
  int bw = 0, bh = 0, bb = 0;     // TMA box of one pixel tile: bb images x bh rows x bw (= W) columns
  int tiles_per_img = 0;          // H / bh when bb == 1
  int kc_per_tap = 0;             // CONV 1: k blocks per tap (padded channels / BLOCK_K)
  int tap_sign = 1;               // CONV 1: +1 forward (reads x[y - oy, x - ox]), -1 backward-data (reads dz[y + oy, x + ox])
  int n_mma = 0;                  // UMMA N (multiple of 16, <= 256) and rows of the weight box
  int b_bytes = 0;                // bytes of one B stage (n_mma x 128)
  int cblks = 0;                  // CONV 2: 64-channel blocks per tap
  int gate_stride = 1;            // CONV 1: the epilogue gate is [images, H*gs, W*gs, N] read at (gs*y, gs*x) (a layer's pinned
                                  // full-size output under the `scale` view that feeds this convolution)
  // CONV 1 epilogue: bf16 operand copies of the result for the convolution that consumes it next (no staging pass there).
  //   copy_plain: [images, H, W, ld] as is;  copy_down (s = copy_s): only pixels (s*y', s*x') -> [images, H/s, W/s, ld] (the
  //   input of a layer behind a `scale` down-sampling);  copy_up (s = copy_s): [images, H*s, W*s, ld] with the result at
  //   (s*y, s*x) and zeros elsewhere (a sampled delta as backward-data reads it).  ld = N (no channel padding).
  void *copy_plain = nullptr, *copy_down = nullptr, *copy_up = nullptr;
  int copy_s = 1;
  int px_stride = 1;              // CONV 2: the k pixels are every px_stride-th pixel of the activation (a sampled delta)
};

// Implicit-GEMM 'same' 3x3 convolution on the tensor cores, out[P, N] fp32 (P = batch*H*W pixels, row-major) =
//   sum over taps t = (ky, kx) and channels k of act[b, y - s*(ky-1), x - s*(kx-1), k] * wts[t][n][k]   (s = tap_sign)
// act: staged NHWC operand [batch][H][W][kp] (bf16, or tf32-rounded fp32), kp % BLOCK_K == 0 (64 bf16 / 32 tf32);
// wts: staged [9][N][kp] in the same element type.  Forward: act = x, wts[t][f][c] = w[t][c][f], s = +1.
// Backward-data: act = dz, wts[t][c][f] = w[t][c][f], s = -1.
bool conv_tc_supported(int kind, int H, int W, int k_channels, int n);
bool conv_tiling(int H, int W, int pixels, ConvGeom *g);
int launch_conv_tc(dls_ctx *ctx, int stream, int kind, const void *act, int batch, int H, int W, int kp, const void *wts, int N,
                   float *C, int tap_sign, const GemmEpilogue &epi, int gate_stride = 1, void *copy_plain = nullptr,
                   void *copy_down = nullptr, void *copy_up = nullptr, int copy_s = 1);

// Weight gradient dw[9*cin, cout] fp32 (= [3,3,cin,cout]) = sum over pixels of act[p - tap][c] * dz[p][f]: contraction over the
// pixels, both operands read MN-major from their staged NHWC copies (act [batch][H][W][cxp], dz [batch*H*W][cdp]; bf16: 64-element atoms, tf32: 32-element atoms through the 32-byte-atom swizzle;
// channels padded to 64), A through tap-shifted 4-D boxes; split over the SMs along the pixels + one reduce launch.
bool conv_dw_tc_supported(int kind, int H, int W, int cin, int cout);
// dzs > 1: dz is the COMPACT sampled delta [batch * H/dzs * W/dzs][cdp]; only those pixels are contracted, and the activation
// boxes are read with TMA element strides (dzs along x and y).
int launch_conv_dw_tc(dls_ctx *ctx, int stream, int kind, const void *act, int cxp, const void *dz, int cdp, int batch, int H, int W,
                      int cin, int cout, float *dw, int dzs = 1);

// conv.cu
bool conv_direct_supported(int cin, int cout);
int launch_conv_direct_fwd(dls_ctx *ctx, int stream, int mode, const float *x, const float *w, const float *bias, float *out, int B,
                           int H, int W, int cin, int cout, int act, int stride);
int launch_conv_direct_dw(dls_ctx *ctx, int stream, int mode, const float *x, const float *dz, float *dw, int B, int H, int W, int cin,
                          int cout, int stride, int dzs, float *dbias = nullptr);
int launch_conv_stage_act(dls_ctx *ctx, int stream, int mode, const float *in, void *out, long long pixels, int H, int W, int C, int Cp,
                          int stride, int up);
int launch_conv_pack_weights(dls_ctx *ctx, int stream, int mode, const float *w, void *wf, void *wb, int cin, int cout, int cinp,
                             int coutp);

// D[M,N] fp32 row-major (pitch ldc) = A * B (+ epilogue).
//   kind  : DLS_GEMM_BF16 (operands are bf16 in HBM) or DLS_GEMM_TF32 (operands are fp32 in HBM, read as tf32)
//   a_mn  : false -> A stored [M][K] (K contiguous, pitch lda elements); true -> stored [K][M] (M contiguous)
//   b_mn  : false -> B stored [N][K] (K contiguous, pitch ldb);          true -> stored [K][N] (N contiguous)
// Operand rows must start 16-byte aligned (pitch * element size % 16 == 0); MN-major TF32 operands are read through the 32-byte-atom swizzle.
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
               bool transpose_a, bool transpose_b, int bias_mode, int activation, dls_buf *c, dls_buf *gate = nullptr,
               int gate_mode = 0, Block *operand_copy = nullptr, float gate_scale = 0.f);
int launch_gate(dls_ctx *ctx, int stream, float *c, size_t total, const float *gate, int mode, float scale = 0.f);
int launch_sgemm(dls_ctx *ctx, int stream, const float *A, const float *B, float *C, int M, int N, int K, bool ta, bool tb);

}  // namespace dls
