// conv.cu -- 'same' 3x3 convolution of NHWC tensors with [3,3,Cin,Cout] weights WITHOUT an im2col matrix.
//
// The reference can only spell a convolution as a composition of views and elementwise trees (TensorsSpec:144-210
// `convolute`: nine `translate`d copies, `join`, then the unrolled matmul; benchmarks.scala:412-632).  Round 1 lowered that
// composition literally: one gather launch wrote the [B*H*W, 9*Cin] fp32 im2col matrix, a staging pass re-wrote it as bf16,
// the tensor-core GEMM read it, and backward scattered a [B*H*W, 9*Cin] delta back with a nine-leaf gather.  Here the
// three passes of a convolution layer are kernels of their own:
//   forward / backward-data : implicit GEMM on the tcgen05 tensor cores (gemm_tc.cu, CONV = 1): the A tile of a tap is one
//                             4-D TMA box of the staged NHWC activation, shifted by the tap, zero-filled outside the image;
//   backward-weights        : implicit GEMM contracting over pixels (gemm_tc.cu, CONV = 2), split over the SMs;
//   first layer (Cin <= 4)  : direct CUDA-core kernels -- 6..16-byte pixel rows are below TMA's 16-byte granule, and the
//                             layer is HBM bound by two orders of magnitude (27 FMA per 128 output bytes).
// Index convention (TensorsSpec:144-210): out[b,y,x,f] = bias[f] + sum_{ky,kx,c} x[b, y-(ky-1), x-(kx-1), c] * w[ky,kx,c,f].
// Arithmetic: operands rounded once to the context's contraction precision (bf16 RNE / tf32 RNA, like every other
// contraction of the mode), fp32 accumulation.
#include <cuda_bf16.h>

#include "common.h"
#include "gemm_tc.h"

namespace dls {
namespace {

__device__ __forceinline__ float round_tf32_rna(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}
// MODE: DLS_GEMM_TF32 (1) / DLS_GEMM_BF16 (2): what the tensor cores would see of the operand
template <int MODE>
__device__ __forceinline__ float round_operand(float x) {
  if (MODE == DLS_GEMM_TF32) return round_tf32_rna(x);
  if (MODE == DLS_GEMM_BF16) return __bfloat162float(__float2bfloat16_rn(x));
  return x;
}

// ---- staging: fp32 NHWC [B, H*s, W*s, C] sampled at stride s (the `scale` down-sampling view, Tensors:950-965, read in
// place) -> operand copy [B, H, W, Cp] (bf16, or tf32-rounded fp32), pad channels zero.  One thread per 4 channels.
template <bool TF32>
__global__ void __launch_bounds__(256) conv_stage_act_kernel(const float *__restrict__ in, void *__restrict__ out, long long pixels,
                                                             int H, int W, int C, int Cp, int stride) {
  const int groups = Cp >> 2;
  const long long total = pixels * groups;
  const bool vec = (C & 3) == 0;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long p = i / groups;
    const int c = (int)(i - p * groups) << 2;
    const int x = (int)(p % W);
    const long long q = p / W;
    const int y = (int)(q % H);
    const long long b = q / H;
    const float *src = in + (((b * H + y) * stride) * (long long)(W * stride) + (long long)x * stride) * C + c;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (vec) {
      if (c < C) v = __ldg(reinterpret_cast<const float4 *>(src));
    } else {
      if (c < C) v.x = __ldg(src);
      if (c + 1 < C) v.y = __ldg(src + 1);
      if (c + 2 < C) v.z = __ldg(src + 2);
      if (c + 3 < C) v.w = __ldg(src + 3);
    }
    if (TF32) {
      v.x = round_tf32_rna(v.x);
      v.y = round_tf32_rna(v.y);
      v.z = round_tf32_rna(v.z);
      v.w = round_tf32_rna(v.w);
      reinterpret_cast<float4 *>(out)[i] = v;
    } else {
      __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
      uint2 o;
      o.x = *reinterpret_cast<uint32_t *>(&lo);
      o.y = *reinterpret_cast<uint32_t *>(&hi);
      reinterpret_cast<uint2 *>(out)[i] = o;
    }
  }
}

// ---- weights [9][Cin][Cout] fp32 -> both per-tap operand layouts in one launch:
//   wf[t][f][c], c < Cinp  (forward:        K = input channels contiguous)
//   wb[t][c][f], f < Coutp (backward-data:  K = output channels contiguous)
template <bool TF32>
__global__ void __launch_bounds__(256) conv_pack_weights_kernel(const float *__restrict__ w, void *__restrict__ wf, void *__restrict__ wb,
                                                                int Cin, int Cout, int Cinp, int Coutp) {
  const int nf = 9 * Cout * Cinp, nb = 9 * Cin * Coutp;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < nf + nb; i += gridDim.x * blockDim.x) {
    float v = 0.f;
    if (i < nf) {
      const int c = i % Cinp, f = (i / Cinp) % Cout, t = i / (Cinp * Cout);
      if (c < Cin) v = __ldg(w + ((long long)t * Cin + c) * Cout + f);
      if (TF32) reinterpret_cast<float *>(wf)[i] = round_tf32_rna(v);
      else reinterpret_cast<__nv_bfloat16 *>(wf)[i] = __float2bfloat16_rn(v);
    } else {
      const int j = i - nf;
      const int f = j % Coutp, c = (j / Coutp) % Cin, t = j / (Coutp * Cin);
      if (f < Cout) v = __ldg(w + ((long long)t * Cin + c) * Cout + f);
      if (TF32) reinterpret_cast<float *>(wb)[j] = round_tf32_rna(v);
      else reinterpret_cast<__nv_bfloat16 *>(wb)[j] = __float2bfloat16_rn(v);
    }
  }
}

// ---- first layer, direct ---------------------------------------------------------------------------------------------------
// A CTA owns a band of R output rows of one image: the (R + 2) x (W + 2) x CIN input patch (zero ring = the 'same' padding,
// operands rounded on the way in) sits in shared memory; lane = output channel, so every weight lives in a register, every
// input value is a shared-memory broadcast and every output pixel leaves as one coalesced 128-byte store.
constexpr int kDirectWarps = 8;

template <int CIN, int MODE>
__device__ __forceinline__ void direct_fill_patch(float *sx, const float *__restrict__ x, long long b, int y0, int R, int H, int W,
                                                  int stride) {
  const int row_len = (W + 2) * CIN;
  const int total = (R + 2) * row_len;
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int r = i / row_len, rem = i - r * row_len;
    const int col = rem / CIN, c = rem - col * CIN;
    const int gy = y0 + r - 1, gx = col - 1;
    float v = 0.f;
    // the source is [B, H*stride, W*stride, CIN] sampled at `stride` (the `scale` down-sampling view read in place)
    if (gy >= 0 && gy < H && gx >= 0 && gx < W)
      v = round_operand<MODE>(__ldg(x + (((b * H + gy) * stride) * (long long)(W * stride) + (long long)gx * stride) * CIN + c));
    sx[i] = v;
  }
}

template <int CIN, int MODE>
__global__ void __launch_bounds__(kDirectWarps * 32) conv_direct_fwd_kernel(const float *__restrict__ x, const float *__restrict__ w,
                                                                            const float *__restrict__ bias, float *__restrict__ out,
                                                                            int B, int H, int W, int Cout, int act, int R, int stride) {
  extern __shared__ float sx[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float wr[9 * CIN];
#pragma unroll
  for (int k = 0; k < 9 * CIN; ++k) wr[k] = lane < Cout ? round_operand<MODE>(__ldg(w + k * Cout + lane)) : 0.f;
  const float bv = (bias && lane < Cout) ? __ldg(bias + lane) : 0.f;
  const int bands = (H + R - 1) / R;
  const int row_len = (W + 2) * CIN;
  for (long long t = blockIdx.x; t < (long long)B * bands; t += gridDim.x) {
    const long long b = t / bands;
    const int y0 = (int)(t - b * bands) * R;
    const int rows = min(R, H - y0);
    __syncthreads();                          // the previous band has been consumed
    direct_fill_patch<CIN, MODE>(sx, x, b, y0, R, H, W, stride);
    __syncthreads();
    for (int p = warp; p < rows * W; p += kDirectWarps) {
      const int yy = p / W, xx = p - yy * W;
      float a0 = 0.f, a1 = 0.f;
#pragma unroll
      for (int ky = 0; ky < 3; ++ky) {
        // tap (ky, kx) reads input (y - ky + 1, x - kx + 1): patch row yy - ky + 2, patch column xx - kx + 2
        const float *prow = sx + (yy - ky + 2) * row_len;
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          const float *pp = prow + (xx - kx + 2) * CIN;
#pragma unroll
          for (int c = 0; c < CIN; ++c) {
            const int k = (ky * 3 + kx) * CIN + c;
            if (k & 1) a1 = fmaf(pp[c], wr[k], a1);
            else a0 = fmaf(pp[c], wr[k], a0);
          }
        }
      }
      float v = __fadd_rn(__fadd_rn(a0, a1), bv);
      if (act == 1) v = tanhf(v);
      else if (act == 2) v = fmaxf(v, 0.f);
      if (lane < Cout) out[((b * H + y0 + yy) * (long long)W + xx) * Cout + lane] = v;
    }
  }
}

// Weight gradient of the first layer: dW[ky,kx,c,f] = sum over pixels of x[b, y-(ky-1), x-(kx-1), c] * dz[b,y,x,f].
// Same patch; lane f keeps the 9*CIN partial sums of its output channel in registers over every pixel its warp visits, the
// warps of a CTA are added through shared memory, and each CTA writes one [9*CIN][32] slab that conv_direct_dw_reduce sums
// in a fixed order (deterministic).
template <int CIN, int MODE>
__global__ void __launch_bounds__(kDirectWarps * 32) conv_direct_dw_kernel(const float *__restrict__ x, const float *__restrict__ dz,
                                                                           float *__restrict__ partial, int B, int H, int W, int Cout,
                                                                           int R, int stride) {
  extern __shared__ float sx[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float acc[9 * CIN];
#pragma unroll
  for (int k = 0; k < 9 * CIN; ++k) acc[k] = 0.f;
  const int bands = (H + R - 1) / R;
  const int row_len = (W + 2) * CIN;
  for (long long t = blockIdx.x; t < (long long)B * bands; t += gridDim.x) {
    const long long b = t / bands;
    const int y0 = (int)(t - b * bands) * R;
    const int rows = min(R, H - y0);
    __syncthreads();
    direct_fill_patch<CIN, MODE>(sx, x, b, y0, R, H, W, stride);
    __syncthreads();
    for (int p = warp; p < rows * W; p += kDirectWarps) {
      const int yy = p / W, xx = p - yy * W;
      const float d = lane < Cout ? round_operand<MODE>(__ldg(dz + ((b * H + y0 + yy) * (long long)W + xx) * Cout + lane)) : 0.f;
#pragma unroll
      for (int ky = 0; ky < 3; ++ky) {
        const float *prow = sx + (yy - ky + 2) * row_len;
#pragma unroll
        for (int kx = 0; kx < 3; ++kx) {
          const float *pp = prow + (xx - kx + 2) * CIN;
#pragma unroll
          for (int c = 0; c < CIN; ++c) acc[(ky * 3 + kx) * CIN + c] = fmaf(pp[c], d, acc[(ky * 3 + kx) * CIN + c]);
        }
      }
    }
  }
  __syncthreads();
  float *red = sx;                            // [warps][9*CIN][32]; the launcher sizes the buffer for both uses
#pragma unroll
  for (int k = 0; k < 9 * CIN; ++k) red[(warp * 9 * CIN + k) * 32 + lane] = acc[k];
  __syncthreads();
  for (int i = threadIdx.x; i < 9 * CIN * 32; i += blockDim.x) {
    float s = red[i];
    for (int wv = 1; wv < kDirectWarps; ++wv) s = __fadd_rn(s, red[wv * 9 * CIN * 32 + i]);
    partial[(long long)blockIdx.x * 9 * CIN * 32 + i] = s;
  }
}

__global__ void __launch_bounds__(256) conv_direct_dw_reduce_kernel(const float *__restrict__ partial, float *__restrict__ dw, int slabs,
                                                                    int K, int Cout) {
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < K * Cout; i += gridDim.x * blockDim.x) {
    const int k = i / Cout, f = i - k * Cout;
    float s = 0.f;
    for (int g = 0; g < slabs; ++g) s = __fadd_rn(s, __ldg(partial + ((long long)g * K + k) * 32 + f));
    dw[i] = s;
  }
}

int direct_rows(int H, int W, int cin) {
  // band height: a patch of (R + 2) x (W + 2) x cin floats, a few KB, and enough bands to fill the machine
  int r = 4;
  while (r > 1 && (long long)(r + 2) * (W + 2) * cin * 4 > 24 * 1024) r >>= 1;
  return r < H ? r : H;
}

}  // namespace

bool conv_direct_supported(int cin, int cout) { return cin >= 1 && cin <= 4 && cout >= 1 && cout <= 32; }

template <int MODE>
static int launch_direct_fwd_mode(dls_ctx *ctx, cudaStream_t st, const float *x, const float *w, const float *bias, float *out, int B,
                                  int H, int W, int cin, int cout, int act, int stride) {
  const int R = direct_rows(H, W, cin);
  const size_t smem = (size_t)(R + 2) * (W + 2) * cin * sizeof(float);
  const long long bands = (long long)B * ((H + R - 1) / R);
  const int grid = (int)std::min<long long>(bands, (long long)ctx->sm_count * 8);
  switch (cin) {
    case 1: conv_direct_fwd_kernel<1, MODE><<<grid, kDirectWarps * 32, smem, st>>>(x, w, bias, out, B, H, W, cout, act, R, stride); break;
    case 2: conv_direct_fwd_kernel<2, MODE><<<grid, kDirectWarps * 32, smem, st>>>(x, w, bias, out, B, H, W, cout, act, R, stride); break;
    case 3: conv_direct_fwd_kernel<3, MODE><<<grid, kDirectWarps * 32, smem, st>>>(x, w, bias, out, B, H, W, cout, act, R, stride); break;
    default: conv_direct_fwd_kernel<4, MODE><<<grid, kDirectWarps * 32, smem, st>>>(x, w, bias, out, B, H, W, cout, act, R, stride); break;
  }
  DLS_CUDA(cudaGetLastError());
  ctx->stats.kernel_launches++;
  return DLS_OK;
}

int launch_conv_direct_fwd(dls_ctx *ctx, int stream, int mode, const float *x, const float *w, const float *bias, float *out, int B,
                           int H, int W, int cin, int cout, int act, int stride) {
  cudaStream_t st = ctx->streams[stream];
  if (mode == DLS_GEMM_TF32) return launch_direct_fwd_mode<DLS_GEMM_TF32>(ctx, st, x, w, bias, out, B, H, W, cin, cout, act, stride);
  if (mode == DLS_GEMM_BF16) return launch_direct_fwd_mode<DLS_GEMM_BF16>(ctx, st, x, w, bias, out, B, H, W, cin, cout, act, stride);
  return launch_direct_fwd_mode<DLS_GEMM_FP32>(ctx, st, x, w, bias, out, B, H, W, cin, cout, act, stride);
}

template <int MODE>
static int launch_direct_dw_mode(dls_ctx *ctx, int stream, const float *x, const float *dz, float *dw, int B, int H, int W, int cin,
                                 int cout, int stride) {
  cudaStream_t st = ctx->streams[stream];
  const int R = direct_rows(H, W, cin);
  const int K = 9 * cin;
  const size_t patch = (size_t)(R + 2) * (W + 2) * cin, red = (size_t)kDirectWarps * K * 32;
  const size_t smem = std::max(patch, red) * sizeof(float);
  const long long bands = (long long)B * ((H + R - 1) / R);
  const int grid = (int)std::min<long long>(bands, (long long)ctx->sm_count * 4);
  Block scratch;
  int rc = alloc_block(ctx, (size_t)grid * K * 32 * sizeof(float), stream, &scratch);
  if (rc) return rc;
  float *partial = (float *)scratch.ptr;
  switch (cin) {
    case 1: conv_direct_dw_kernel<1, MODE><<<grid, kDirectWarps * 32, smem, st>>>(x, dz, partial, B, H, W, cout, R, stride); break;
    case 2: conv_direct_dw_kernel<2, MODE><<<grid, kDirectWarps * 32, smem, st>>>(x, dz, partial, B, H, W, cout, R, stride); break;
    case 3: conv_direct_dw_kernel<3, MODE><<<grid, kDirectWarps * 32, smem, st>>>(x, dz, partial, B, H, W, cout, R, stride); break;
    default: conv_direct_dw_kernel<4, MODE><<<grid, kDirectWarps * 32, smem, st>>>(x, dz, partial, B, H, W, cout, R, stride); break;
  }
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) {
    conv_direct_dw_reduce_kernel<<<(K * cout + 255) / 256, 256, 0, st>>>(partial, dw, grid, K, cout);
    e = cudaGetLastError();
  }
  ctx->stats.kernel_launches += 2;
  free_block(ctx, scratch);
  if (e != cudaSuccess) return cuda_fail(e, "conv_direct_dw_kernel", __FILE__, __LINE__);
  return DLS_OK;
}

int launch_conv_direct_dw(dls_ctx *ctx, int stream, int mode, const float *x, const float *dz, float *dw, int B, int H, int W, int cin,
                          int cout, int stride) {
  if (mode == DLS_GEMM_TF32) return launch_direct_dw_mode<DLS_GEMM_TF32>(ctx, stream, x, dz, dw, B, H, W, cin, cout, stride);
  if (mode == DLS_GEMM_BF16) return launch_direct_dw_mode<DLS_GEMM_BF16>(ctx, stream, x, dz, dw, B, H, W, cin, cout, stride);
  return launch_direct_dw_mode<DLS_GEMM_FP32>(ctx, stream, x, dz, dw, B, H, W, cin, cout, stride);
}

int launch_conv_stage_act(dls_ctx *ctx, int stream, int mode, const float *in, void *out, long long pixels, int H, int W, int C, int Cp,
                          int stride) {
  const long long total = pixels * (Cp / 4);
  if (total == 0) return DLS_OK;
  const int grid = (int)std::min<long long>((total + 255) / 256, (long long)ctx->sm_count * 16);
  if (mode == DLS_GEMM_TF32) conv_stage_act_kernel<true><<<grid, 256, 0, ctx->streams[stream]>>>(in, out, pixels, H, W, C, Cp, stride);
  else conv_stage_act_kernel<false><<<grid, 256, 0, ctx->streams[stream]>>>(in, out, pixels, H, W, C, Cp, stride);
  DLS_CUDA(cudaGetLastError());
  ctx->stats.kernel_launches++;
  return DLS_OK;
}

int launch_conv_pack_weights(dls_ctx *ctx, int stream, int mode, const float *w, void *wf, void *wb, int cin, int cout, int cinp, int coutp) {
  const int total = 9 * cout * cinp + 9 * cin * coutp;
  const int grid = std::min((total + 255) / 256, ctx->sm_count * 8);
  if (mode == DLS_GEMM_TF32) conv_pack_weights_kernel<true><<<grid, 256, 0, ctx->streams[stream]>>>(w, wf, wb, cin, cout, cinp, coutp);
  else conv_pack_weights_kernel<false><<<grid, 256, 0, ctx->streams[stream]>>>(w, wf, wb, cin, cout, cinp, coutp);
  DLS_CUDA(cudaGetLastError());
  ctx->stats.kernel_launches++;
  return DLS_OK;
}

}  // namespace dls

using namespace dls;

namespace {

inline int pad_channels(int mode, int c) {
  const int q = mode == DLS_GEMM_TF32 ? 32 : 64;   // one 128-byte swizzle row of K
  return (c + q - 1) / q * q;
}
inline size_t elem_bytes(int mode) { return mode == DLS_GEMM_TF32 ? 4 : 2; }

// Staged NHWC operand copy of `buf` read as [batch, h*stride, w*stride, c] and sampled at `stride` (shadow slot 2, kept
// until the buffer is written again: the forward pass and the weight gradient share the copy of an activation, the
// backward-data pass and the weight gradient share the copy of a delta).
int staged_act(dls_ctx *ctx, int stream, int mode, dls_buf *buf, int batch, int h, int w, int c, int stride, const void **ptr, int *cp) {
  Block &sh = buf->shadow[2];
  const long long tag_r = ((long long)mode << 56) | ((long long)stride << 48) | ((long long)h << 24) | (long long)w;
  const long long tag_c = ((long long)batch << 24) | (long long)c;
  const int p = pad_channels(mode, c);
  if (!(sh.ptr && buf->shadow_rows[2] == tag_r && buf->shadow_cols[2] == tag_c && sh.pool == ctx->capture_pool)) {
    if (sh.ptr) free_block(ctx, sh);
    const long long pixels = (long long)batch * h * w;
    int rc = alloc_block(ctx, (size_t)pixels * p * elem_bytes(mode), stream, &sh);
    if (!rc) rc = launch_conv_stage_act(ctx, stream, mode, buf->ptr, sh.ptr, pixels, h, w, c, p, stride);
    if (rc) {
      if (sh.ptr) free_block(ctx, sh);
      return rc;
    }
    buf->shadow_rows[2] = tag_r;
    buf->shadow_cols[2] = tag_c;
  } else {
    int rc = wait_streams(ctx, stream, sh.stream_mask);
    if (rc) return rc;
    sh.stream_mask |= 1u << stream;
  }
  *ptr = sh.ptr;
  *cp = p;
  return DLS_OK;
}

// Both per-tap operand layouts of a [3,3,cin,cout] weight in one block (shadow slot 3): wf first, then wb.
int staged_weights(dls_ctx *ctx, int stream, int mode, dls_buf *w, int cin, int cout, const void **wf, const void **wb) {
  Block &sh = w->shadow[3];
  const long long tag_r = ((long long)mode << 56) | (long long)cin, tag_c = cout;
  const int cinp = pad_channels(mode, cin), coutp = pad_channels(mode, cout);
  const size_t nf = ((size_t)9 * cout * cinp * elem_bytes(mode) + 1023) & ~(size_t)1023, nb = (size_t)9 * cin * coutp * elem_bytes(mode);
  if (!(sh.ptr && w->shadow_rows[3] == tag_r && w->shadow_cols[3] == tag_c && sh.pool == ctx->capture_pool)) {
    if (sh.ptr) free_block(ctx, sh);
    int rc = alloc_block(ctx, nf + nb, stream, &sh);
    if (!rc) rc = launch_conv_pack_weights(ctx, stream, mode, w->ptr, sh.ptr, (char *)sh.ptr + nf, cin, cout, cinp, coutp);
    if (rc) {
      if (sh.ptr) free_block(ctx, sh);
      return rc;
    }
    w->shadow_rows[3] = tag_r;
    w->shadow_cols[3] = tag_c;
  } else {
    int rc = wait_streams(ctx, stream, sh.stream_mask);
    if (rc) return rc;
    sh.stream_mask |= 1u << stream;
  }
  *wf = sh.ptr;
  *wb = (char *)sh.ptr + nf;
  return DLS_OK;
}

int conv_mode(dls_ctx *ctx, const dls_conv_opts *opts) {
  const int m = (opts && opts->mode >= 0) ? opts->mode : ctx->opts.gemm_mode;
  return m;
}

bool fits_int(size_t v) { return v > 0 && v < (1u << 24); }

}  // namespace

extern "C" {

int dls_conv3x3_supported(dls_ctx *ctx, int pass, size_t h, size_t w, size_t cin, size_t cout, int mode) {
  if (!ctx) return 0;
  if (mode < 0) mode = ctx->opts.gemm_mode;
  if (mode != DLS_GEMM_TF32 && mode != DLS_GEMM_BF16) return 0;
  if (!fits_int(h) || !fits_int(w) || !fits_int(cin) || !fits_int(cout)) return 0;
  if (conv_direct_supported((int)cin, (int)cout)) return pass == DLS_CONV_FORWARD || pass == DLS_CONV_BACKWARD_WEIGHTS;
  if (pass == DLS_CONV_FORWARD) return conv_tc_supported(mode, (int)h, (int)w, pad_channels(mode, (int)cin), (int)cout) ? 1 : 0;
  if (pass == DLS_CONV_BACKWARD_DATA) return conv_tc_supported(mode, (int)h, (int)w, pad_channels(mode, (int)cout), (int)cin) ? 1 : 0;
  if (pass == DLS_CONV_BACKWARD_WEIGHTS) return conv_dw_tc_supported(mode, (int)h, (int)w, (int)cin, (int)cout) ? 1 : 0;
  return 0;
}

int dls_conv3x3_forward(dls_ctx *ctx, dls_buf *x, dls_buf *w, dls_buf *bias, size_t batch, size_t h, size_t wd, size_t cin, size_t cout,
                        const dls_conv_opts *opts, dls_event *const *deps, int n_deps, dls_buf **out, dls_event **done) {
  DLS_CHECK(ctx && x && w && out, DLS_ERR_INVALID, "dls_conv3x3_forward: NULL argument");
  const int stride = (opts && opts->in_stride > 1) ? opts->in_stride : 1;
  const int act = opts ? opts->activation : 0;
  DLS_CHECK(act >= 0 && act <= 2, DLS_ERR_INVALID, "dls_conv3x3_forward: bad activation");
  DLS_CHECK(fits_int(batch) && fits_int(h) && fits_int(wd) && fits_int(cin) && fits_int(cout), DLS_ERR_INVALID,
            "dls_conv3x3_forward: bad shape");
  DLS_CHECK(batch * h * stride * wd * stride * cin <= x->n && 9 * cin * cout <= w->n && (!bias || bias->n >= cout), DLS_ERR_INVALID,
            "dls_conv3x3_forward: a buffer is shorter than its shape");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  const int mode = conv_mode(ctx, opts);
  DLS_CHECK(dls_conv3x3_supported(ctx, DLS_CONV_FORWARD, h, wd, cin, cout, mode), DLS_ERR_UNSUPPORTED,
            "dls_conv3x3_forward: shape / mode not covered (compose the convolution from translate, join and matmul instead)");
  DLS_CUDA(cudaSetDevice(ctx->device));
  const int stream = pick_stream(ctx);
  dls_buf *ins[3] = {x, w, bias};
  int rc = wait_inputs(ctx, stream, ins, 3, deps, n_deps);
  if (rc) return rc;
  dls_buf *o = nullptr;
  rc = new_buf(ctx, batch * h * wd * cout, stream, &o);
  if (rc) return rc;
  if (conv_direct_supported((int)cin, (int)cout)) {
    rc = launch_conv_direct_fwd(ctx, stream, mode, x->ptr, w->ptr, bias ? bias->ptr : nullptr, o->ptr, (int)batch, (int)h, (int)wd,
                                (int)cin, (int)cout, act, stride);
  } else {
    const void *xa, *wf, *wb;
    int cp;
    rc = staged_act(ctx, stream, mode, x, (int)batch, (int)h, (int)wd, (int)cin, stride, &xa, &cp);
    if (!rc) rc = staged_weights(ctx, stream, mode, w, (int)cin, (int)cout, &wf, &wb);
    GemmEpilogue epi;
    epi.bias = bias ? bias->ptr : nullptr;
    epi.bias_mode = bias ? 2 : 0;
    epi.act = act;
    if (!rc) rc = launch_conv_tc(ctx, stream, mode, xa, (int)batch, (int)h, (int)wd, cp, wf, (int)cout, o->ptr, +1, epi);
  }
  if (!rc) rc = finish_output(ctx, stream, &o, 1, ins, 3, done);
  if (rc) {
    dls_buf_release(o);
    return rc;
  }
  *out = o;
  return DLS_OK;
}

int dls_conv3x3_backward_data(dls_ctx *ctx, dls_buf *dz, dls_buf *w, size_t batch, size_t h, size_t wd, size_t cin, size_t cout,
                              const dls_conv_opts *opts, dls_event *const *deps, int n_deps, dls_buf **out, dls_event **done) {
  DLS_CHECK(ctx && dz && w && out, DLS_ERR_INVALID, "dls_conv3x3_backward_data: NULL argument");
  DLS_CHECK(fits_int(batch) && fits_int(h) && fits_int(wd) && fits_int(cin) && fits_int(cout), DLS_ERR_INVALID,
            "dls_conv3x3_backward_data: bad shape");
  DLS_CHECK(batch * h * wd * cout <= dz->n && 9 * cin * cout <= w->n, DLS_ERR_INVALID,
            "dls_conv3x3_backward_data: a buffer is shorter than its shape");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  const int mode = conv_mode(ctx, opts);
  DLS_CHECK(dls_conv3x3_supported(ctx, DLS_CONV_BACKWARD_DATA, h, wd, cin, cout, mode), DLS_ERR_UNSUPPORTED,
            "dls_conv3x3_backward_data: shape / mode not covered");
  DLS_CUDA(cudaSetDevice(ctx->device));
  const int stream = pick_stream(ctx);
  dls_buf *ins[2] = {dz, w};
  int rc = wait_inputs(ctx, stream, ins, 2, deps, n_deps);
  if (rc) return rc;
  dls_buf *o = nullptr;
  rc = new_buf(ctx, batch * h * wd * cin, stream, &o);
  if (rc) return rc;
  const void *da, *wf, *wb;
  int cp;
  rc = staged_act(ctx, stream, mode, dz, (int)batch, (int)h, (int)wd, (int)cout, 1, &da, &cp);
  if (!rc) rc = staged_weights(ctx, stream, mode, w, (int)cin, (int)cout, &wf, &wb);
  // dx[b,y,x,c] = sum_{ky,kx,f} dz[b, y+(ky-1), x+(kx-1), f] * w[ky,kx,c,f]: the same kernel with the taps mirrored
  if (!rc) rc = launch_conv_tc(ctx, stream, mode, da, (int)batch, (int)h, (int)wd, cp, wb, (int)cin, o->ptr, -1, GemmEpilogue());
  if (!rc) rc = finish_output(ctx, stream, &o, 1, ins, 2, done);
  if (rc) {
    dls_buf_release(o);
    return rc;
  }
  *out = o;
  return DLS_OK;
}

int dls_conv3x3_backward_weights(dls_ctx *ctx, dls_buf *x, dls_buf *dz, size_t batch, size_t h, size_t wd, size_t cin, size_t cout,
                                 const dls_conv_opts *opts, dls_event *const *deps, int n_deps, dls_buf **out, dls_event **done) {
  DLS_CHECK(ctx && x && dz && out, DLS_ERR_INVALID, "dls_conv3x3_backward_weights: NULL argument");
  const int stride = (opts && opts->in_stride > 1) ? opts->in_stride : 1;
  DLS_CHECK(fits_int(batch) && fits_int(h) && fits_int(wd) && fits_int(cin) && fits_int(cout), DLS_ERR_INVALID,
            "dls_conv3x3_backward_weights: bad shape");
  DLS_CHECK(batch * h * stride * wd * stride * cin <= x->n && batch * h * wd * cout <= dz->n, DLS_ERR_INVALID,
            "dls_conv3x3_backward_weights: a buffer is shorter than its shape");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  const int mode = conv_mode(ctx, opts);
  DLS_CHECK(dls_conv3x3_supported(ctx, DLS_CONV_BACKWARD_WEIGHTS, h, wd, cin, cout, mode), DLS_ERR_UNSUPPORTED,
            "dls_conv3x3_backward_weights: shape / mode not covered");
  DLS_CUDA(cudaSetDevice(ctx->device));
  const int stream = pick_stream(ctx);
  dls_buf *ins[2] = {x, dz};
  int rc = wait_inputs(ctx, stream, ins, 2, deps, n_deps);
  if (rc) return rc;
  dls_buf *o = nullptr;
  rc = new_buf(ctx, 9 * cin * cout, stream, &o);
  if (rc) return rc;
  if (conv_direct_supported((int)cin, (int)cout)) {
    rc = launch_conv_direct_dw(ctx, stream, mode, x->ptr, dz->ptr, o->ptr, (int)batch, (int)h, (int)wd, (int)cin, (int)cout, stride);
  } else {
    const void *xa, *da;
    int cxp, cdp;
    rc = staged_act(ctx, stream, mode, x, (int)batch, (int)h, (int)wd, (int)cin, stride, &xa, &cxp);
    if (!rc) rc = staged_act(ctx, stream, mode, dz, (int)batch, (int)h, (int)wd, (int)cout, 1, &da, &cdp);
    if (!rc) rc = launch_conv_dw_tc(ctx, stream, mode, xa, cxp, da, cdp, (int)batch, (int)h, (int)wd, (int)cin, (int)cout, o->ptr);
  }
  if (!rc) rc = finish_output(ctx, stream, &o, 1, ins, 2, done);
  if (rc) {
    dls_buf_release(o);
    return rc;
  }
  *out = o;
  return DLS_OK;
}

}  // extern "C"
