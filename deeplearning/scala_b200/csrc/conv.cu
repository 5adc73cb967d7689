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

#include <climits>

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

// ---- staging: fp32 NHWC -> operand copy [B, H, W, Cp] (bf16, or tf32-rounded fp32), pad channels zero.  One thread per 4
// channels.  Two read patterns besides the dense one, both views the reference would evaluate lazily (Tensors:950-965):
//   stride > 1: the source is [B, H*stride, W*stride, C] sampled at `stride` (a `scale` down-sampling read in place);
//   up > 1:     the source is [B, H/up, W/up, C] and stands for a tensor that is zero except at (up*y, up*x) -- the delta
//               that a down-sampling layer hands back (its `scale` up, masked to the sampled positions).
template <bool TF32>
__global__ void __launch_bounds__(256) conv_stage_act_kernel(const float *__restrict__ in, void *__restrict__ out, long long pixels,
                                                             int H, int W, int C, int Cp, int stride, int up) {
  // one thread per 8 channels (bf16: one 16-byte store; tf32: two), two 128-bit loads when the source channels allow
  const int groups = Cp >> 3;
  const long long total = pixels * groups;
  const bool vec = (C & 3) == 0;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long p = i / groups;
    const int c = (int)(i - p * groups) << 3;
    const int x = (int)(p % W);
    const long long q = p / W;
    const int y = (int)(q % H);
    const long long b = q / H;
    const float *src;
    bool live = true;
    if (up > 1) {
      live = (y % up) == 0 && (x % up) == 0;
      src = in + ((b * (H / up) + y / up) * (long long)(W / up) + x / up) * C + c;
    } else {
      src = in + (((b * H + y) * stride) * (long long)(W * stride) + (long long)x * stride) * C + c;
    }
    float v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) v[k] = 0.f;
    if (live) {
      if (vec) {
        if (c < C) {
          const float4 q0 = __ldg(reinterpret_cast<const float4 *>(src));
          v[0] = q0.x; v[1] = q0.y; v[2] = q0.z; v[3] = q0.w;
        }
        if (c + 4 < C) {
          const float4 q1 = __ldg(reinterpret_cast<const float4 *>(src) + 1);
          v[4] = q1.x; v[5] = q1.y; v[6] = q1.z; v[7] = q1.w;
        }
      } else {
#pragma unroll
        for (int k = 0; k < 8; ++k)
          if (c + k < C) v[k] = __ldg(src + k);
      }
    }
    if (TF32) {
      float4 o0 = make_float4(round_tf32_rna(v[0]), round_tf32_rna(v[1]), round_tf32_rna(v[2]), round_tf32_rna(v[3]));
      float4 o1 = make_float4(round_tf32_rna(v[4]), round_tf32_rna(v[5]), round_tf32_rna(v[6]), round_tf32_rna(v[7]));
      reinterpret_cast<float4 *>(out)[2 * i] = o0;
      reinterpret_cast<float4 *>(out)[2 * i + 1] = o1;
    } else {
      __nv_bfloat162 p0 = __floats2bfloat162_rn(v[0], v[1]), p1 = __floats2bfloat162_rn(v[2], v[3]);
      __nv_bfloat162 p2 = __floats2bfloat162_rn(v[4], v[5]), p3 = __floats2bfloat162_rn(v[6], v[7]);
      uint4 o;
      o.x = *reinterpret_cast<uint32_t *>(&p0);
      o.y = *reinterpret_cast<uint32_t *>(&p1);
      o.z = *reinterpret_cast<uint32_t *>(&p2);
      o.w = *reinterpret_cast<uint32_t *>(&p3);
      reinterpret_cast<uint4 *>(out)[i] = o;
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
// Cin <= 4: 6..16-byte pixel rows are below TMA's 16-byte granule and the layer is HBM bound by two orders of magnitude, so
// it gets kernels of its own.  A CTA owns a band of R output rows of one image: the (R + 2) x (W + 2) input patch (zero ring
// = the 'same' padding, channels padded to 4, operands rounded to the contraction mode on the way in) sits in shared memory.
// The arithmetic is warp-level mma.sync.m16n8k8 (tf32 inputs, fp32 accumulate; a bf16-rounded value is a valid tf32, so one
// instruction serves both modes): with the CUDA cores the layer needs 27 FFMA + 27 shared-memory loads per pixel and warp and
// is issue bound at 8x its HBM time; as 16-pixel x 32-channel x 40-tap MMA tiles it needs ~6 instructions per pixel.
// k index of the implicit im2col row: k = tap * 4 + c (c = 3 and taps 9.. are zero).
constexpr int kDirectWarps = 8;
constexpr int kDirectK = 40;
constexpr int kSlabRows = 37;                // 36 im2col rows of the weight gradient + one row for the bias gradient                  // 9 taps x 4 padded channels, rounded up to 5 k-steps of 8

__device__ __forceinline__ void mma_tf32_16x8x8(float (&d)[4], const uint32_t (&a)[4], const uint32_t (&b)[2]) {
  asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b[0]), "r"(b[1]));
}

// patch[r][col][4]: r = 0 .. R+1 <-> input row y0 + r - 1, col = 0 .. W+1 <-> input column col - 1.  The source is
// [B, H*stride, W*stride, CIN] sampled at `stride` (the `scale` down-sampling view read in place).
template <int CIN, int MODE>
__device__ __forceinline__ void direct_fill_patch(float *sx, const float *__restrict__ x, long long b, int y0, int R, int H, int W,
                                                  int stride) {
  const int row_len = (W + 2) * 4;
  const int total = (R + 2) * row_len;
  if (stride == 1 && ((W * CIN) & 3) == 0 && ((uintptr_t)x & 15) == 0) {
    // dense source whose image rows are whole float4s: every thread's 128-bit loads are independent (one round trip to
    // memory for the whole patch instead of one per element); the ring, the pad channel and rows outside the image are
    // zeroed first
    for (int i = threadIdx.x; i < total; i += blockDim.x) sx[i] = 0.f;
    __syncthreads();
    const int n4 = (W * CIN) >> 2;
    for (int i = threadIdx.x; i < (R + 2) * n4; i += blockDim.x) {
      const int r = i / n4, q = i - r * n4;
      const int gy = y0 + r - 1;
      if (gy < 0 || gy >= H) continue;
      const float4 v = __ldg(reinterpret_cast<const float4 *>(x + ((b * H + gy) * (long long)W) * CIN) + q);
      const float e[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int flat = 4 * q + k, px = flat / CIN, c = flat - px * CIN;
        sx[(r * (W + 2) + px + 1) * 4 + c] = round_operand<MODE>(e[k]);
      }
    }
    return;
  }
  for (int i = threadIdx.x; i < total; i += blockDim.x) {
    const int r = i / row_len, rem = i - r * row_len;
    const int col = rem >> 2, c = rem & 3;
    const int gy = y0 + r - 1, gx = col - 1;
    float v = 0.f;
    if (c < CIN && gy >= 0 && gy < H && gx >= 0 && gx < W)
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
  const int g = lane >> 2, t = lane & 3;
  // B fragments of the whole [40 x 32] weight matrix stay in registers: b0 = (k = 8 ks + t, n = 8 nt + g), b1 = (k + 4, n)
  uint32_t bw[kDirectK / 8][4][2];
#pragma unroll
  for (int ks = 0; ks < kDirectK / 8; ++ks)
#pragma unroll
    for (int nt = 0; nt < 4; ++nt)
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int tap = 2 * ks + j, n = nt * 8 + g;
        float v = 0.f;
        if (tap < 9 && t < CIN && n < Cout) v = round_operand<MODE>(__ldg(w + (tap * CIN + t) * Cout + n));
        bw[ks][nt][j] = __float_as_uint(v);
      }
  float bv[4][2];
#pragma unroll
  for (int nt = 0; nt < 4; ++nt)
#pragma unroll
    for (int j = 0; j < 2; ++j) bv[nt][j] = (bias && nt * 8 + 2 * t + j < Cout) ? __ldg(bias + nt * 8 + 2 * t + j) : 0.f;
  const int bands = (H + R - 1) / R;
  const int prow = (W + 2) * 4;
  for (long long band = blockIdx.x; band < (long long)B * bands; band += gridDim.x) {
    const long long b = band / bands;
    const int y0 = (int)(band - b * bands) * R;
    const int rows = min(R, H - y0);
    __syncthreads();                          // the previous band has been consumed
    direct_fill_patch<CIN, MODE>(sx, x, b, y0, R, H, W, stride);
    __syncthreads();
    const int npx = rows * W;
    for (int tile = warp; tile * 16 < npx; tile += kDirectWarps) {
      // A fragment rows: pixels p0 = 16 tile + g and p0 + 8 of the band
      const int p0 = tile * 16 + g, p1 = p0 + 8;
      const int q0 = min(p0, npx - 1), q1 = min(p1, npx - 1);
      const int yy0 = q0 / W, xx0 = q0 - yy0 * W, yy1 = q1 / W, xx1 = q1 - yy1 * W;
      const float *a0p = sx + ((yy0 + 2) * (W + 2) + xx0 + 2) * 4 + t;
      const float *a1p = sx + ((yy1 + 2) * (W + 2) + xx1 + 2) * 4 + t;
      float acc[4][4];
#pragma unroll
      for (int nt = 0; nt < 4; ++nt)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[nt][j] = 0.f;
#pragma unroll
      for (int ks = 0; ks < kDirectK / 8; ++ks) {
        uint32_t a[4];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int tap = 2 * ks + j;             // tap (ky, kx) reads input (y - ky + 1, x - kx + 1)
          if (tap < 9) {
            const int off = -((tap / 3) * prow + (tap % 3) * 4);
            a[2 * j] = __float_as_uint(a0p[off]);
            a[2 * j + 1] = __float_as_uint(a1p[off]);
          } else {
            a[2 * j] = a[2 * j + 1] = 0u;
          }
        }
#pragma unroll
        for (int nt = 0; nt < 4; ++nt) mma_tf32_16x8x8(acc[nt], a, bw[ks][nt]);
      }
      // C fragment: (pixel p0, channels 8 nt + 2 t, + 1), (pixel p1, same channels)
      float *o0 = out + ((b * H + y0) * (long long)W + p0) * Cout;
      float *o1 = out + ((b * H + y0) * (long long)W + p1) * Cout;
#pragma unroll
      for (int nt = 0; nt < 4; ++nt) {
        const int n = nt * 8 + 2 * t;
        float v[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          v[j] = __fadd_rn(acc[nt][j], bv[nt][j & 1]);
          if (act == 1) v[j] = tanhf(v[j]);
          else if (act == 2) v[j] = fmaxf(v[j], 0.f);
        }
        if ((Cout & 1) == 0) {
          if (p0 < npx && n < Cout) *reinterpret_cast<float2 *>(o0 + n) = make_float2(v[0], v[1]);
          if (p1 < npx && n < Cout) *reinterpret_cast<float2 *>(o1 + n) = make_float2(v[2], v[3]);
        } else {
          if (p0 < npx && n < Cout) o0[n] = v[0];
          if (p0 < npx && n + 1 < Cout) o0[n + 1] = v[1];
          if (p1 < npx && n < Cout) o1[n] = v[2];
          if (p1 < npx && n + 1 < Cout) o1[n + 1] = v[3];
        }
      }
    }
  }
}

// Weight gradient of the first layer: dW[ky,kx,c,f] = sum over pixels of x[b, y-(ky-1), x-(kx-1), c] * dz[b,y,x,f], as
// D[k, f] += A[k, pixel] * B[pixel, f] with 8 pixels per mma.  `dzs` > 1: the delta arrives SAMPLED -- dz is
// [B, H/dzs, W/dzs, Cout] and stands for a full-size delta that is zero except at (dzs*y, dzs*x), the gradient of a `scale`
// down-sampling that follows the layer (Tensors:950-965) -- so only those pixels are visited.  Every warp keeps its
// [48 x 32] partial in registers over all pixels it visits; the warps of a CTA are added through shared memory and each
// CTA writes one [36][32] slab that conv_direct_dw_reduce sums in a fixed order (deterministic).
template <int CIN, int MODE>
__global__ void __launch_bounds__(kDirectWarps * 32) conv_direct_dw_kernel(const float *__restrict__ x, const float *__restrict__ dz,
                                                                           float *__restrict__ partial, int B, int H, int W, int Cout,
                                                                           int R, int stride, int dzs) {
  extern __shared__ float sx[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int g = lane >> 2, t = lane & 3;
  const int prow = (W + 2) * 4;
  float acc[3][4][4];
  float db[4] = {0.f, 0.f, 0.f, 0.f};         // sum of the (unrounded) delta over this lane's pixels, channel 8 nt + g
#pragma unroll
  for (int mt = 0; mt < 3; ++mt)
#pragma unroll
    for (int nt = 0; nt < 4; ++nt)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[mt][nt][j] = 0.f;
  // A fragment rows m = 16 mt + g (+ 8): tap = m / 4, channel = m % 4; offset of that tap relative to the pixel's patch cell
  int aoff[3][2];
#pragma unroll
  for (int mt = 0; mt < 3; ++mt)
#pragma unroll
    for (int h = 0; h < 2; ++h) {
      const int m = mt * 16 + g + 8 * h, tap = m >> 2;
      aoff[mt][h] = tap < 9 ? -((tap / 3) * prow + (tap % 3) * 4) + (m & 3) : INT_MIN;
    }
  const int bands = (H + R - 1) / R;
  const int Hs = H / dzs, Ws = W / dzs;
  for (long long band = blockIdx.x; band < (long long)B * bands; band += gridDim.x) {
    const long long b = band / bands;
    const int y0 = (int)(band - b * bands) * R;   // R % dzs == 0, so y0 % dzs == 0
    const int rows = min(R, H - y0);
    __syncthreads();
    direct_fill_patch<CIN, MODE>(sx, x, b, y0, R, H, W, stride);
    __syncthreads();
    const int nps = (rows / dzs) * Ws;            // delta-carrying pixels of the band
    const float *dzb = dz + ((b * Hs + y0 / dzs) * (long long)Ws) * Cout;
    // two groups of 8 pixels per trip: all 16 delta loads of a trip are issued before the first mma needs one
    for (int grp0 = warp; grp0 * 8 < nps; grp0 += 2 * kDirectWarps) {
      int pq[2][2];
      bool ok[2][2];
      uint32_t bf[2][4][2];
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int grp = grp0 + u * kDirectWarps;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int q = grp * 8 + t + 4 * h;          // k columns of A / rows of B: pixels grp * 8 + t and + 4
          ok[u][h] = q < nps;
          pq[u][h] = ok[u][h] ? q : 0;
        }
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            const int n = nt * 8 + g;
            float v = 0.f;
            if (ok[u][h] && n < Cout) v = __ldg(dzb + (long long)pq[u][h] * Cout + n);
            db[nt] += v;
            bf[u][nt][h] = __float_as_uint(v);
          }
      }
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        if ((grp0 + u * kDirectWarps) * 8 >= nps) break;
        const float *ap[2];
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          const int ys = pq[u][h] / Ws, xs = pq[u][h] - ys * Ws;
          ap[h] = sx + ((ys * dzs + 2) * (W + 2) + xs * dzs + 2) * 4;
        }
#pragma unroll
        for (int nt = 0; nt < 4; ++nt)
#pragma unroll
          for (int h = 0; h < 2; ++h) bf[u][nt][h] = __float_as_uint(round_operand<MODE>(__uint_as_float(bf[u][nt][h])));
#pragma unroll
        for (int mt = 0; mt < 3; ++mt) {
          uint32_t a[4];
#pragma unroll
          for (int h = 0; h < 2; ++h)    // a0 = (m, pixel t), a1 = (m + 8, pixel t), a2 = (m, pixel t + 4), a3 = (m + 8, pixel t + 4)
#pragma unroll
            for (int r = 0; r < 2; ++r) a[2 * h + r] = aoff[mt][r] != INT_MIN ? __float_as_uint(ap[h][aoff[mt][r]]) : 0u;
#pragma unroll
          for (int nt = 0; nt < 4; ++nt) mma_tf32_16x8x8(acc[mt][nt], a, bf[u][nt]);
        }
      }
    }
  }
  __syncthreads();
  // C fragment: c0 = (m = 16 mt + g, f = 8 nt + 2 t), c1 = (m, f + 1), c2 = (m + 8, f), c3 = (m + 8, f + 1); rows m < 36 only
  float *red = sx;                            // [warps][36][32]; the launcher sizes the buffer for both uses
#pragma unroll
  for (int mt = 0; mt < 3; ++mt)
#pragma unroll
    for (int nt = 0; nt < 4; ++nt)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int m = mt * 16 + g + 8 * (j >> 1), f = nt * 8 + 2 * t + (j & 1);
        if (m < 36) red[(warp * kSlabRows + m) * 32 + f] = acc[mt][nt][j];
      }
  // bias gradient: fold the four lanes (t) that share a channel, lane t == 0 holds the warp's sum for channel 8 nt + g
#pragma unroll
  for (int nt = 0; nt < 4; ++nt) {
    float v = db[nt];
    v += __shfl_xor_sync(0xffffffffu, v, 1);
    v += __shfl_xor_sync(0xffffffffu, v, 2);
    if (t == 0) red[(warp * kSlabRows + 36) * 32 + nt * 8 + g] = v;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < kSlabRows * 32; i += blockDim.x) {
    float s = red[i];
    for (int wv = 1; wv < kDirectWarps; ++wv) s = __fadd_rn(s, red[wv * kSlabRows * 32 + i]);
    partial[(long long)blockIdx.x * kSlabRows * 32 + i] = s;
  }
}

// dw[(tap * cin + c) * cout + f] = sum over slabs of partial[slab][tap * 4 + c][f]: one warp per output, the lanes take
// every 32nd slab and are folded by a fixed shuffle tree (deterministic for a given grid).
__global__ void __launch_bounds__(256) conv_direct_dw_reduce_kernel(const float *__restrict__ partial, float *__restrict__ dw,
                                                                    float *__restrict__ dbias, int slabs, int cin, int cout) {
  const int lane = threadIdx.x & 31;
  const int o = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const int n_dw = 9 * cin * cout;
  if (o >= n_dw + (dbias ? cout : 0)) return;
  const float *src;
  if (o < n_dw) {
    const int f = o % cout, c = (o / cout) % cin, tap = o / (cout * cin);
    src = partial + (tap * 4 + c) * 32 + f;
  } else {
    src = partial + 36 * 32 + (o - n_dw);            // row 36 of every slab: the bias gradient
  }
  float s = 0.f;
  for (int sl = lane; sl < slabs; sl += 32) s = __fadd_rn(s, __ldg(src + (long long)sl * kSlabRows * 32));
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) s = __fadd_rn(s, __shfl_down_sync(0xffffffffu, s, d));
  if (lane == 0) {
    if (o < n_dw) dw[o] = s;
    else dbias[o - n_dw] = s;
  }
}

int direct_rows(int H, int W, int dzs, int B, int sm_count) {
  // band height.  The kernels are latency bound (one patch fill, one barrier, then a few mma tiles per warp), so a band is
  // as tall as fits -- the whole image when its (H + 2) x (W + 2) x 4 float patch stays below 40 KB (18.5 KB at 32 x 32:
  // 64 forward tiles / 32 sampled-delta groups per band instead of 16 / 8, and no halo rows read twice) -- as long as
  // there is still a band for every SM; a multiple of the delta's sampling factor
  auto patch = [&](int r) { return (long long)(r + 2) * (W + 2) * 16; };
  int r = H;
  while (r > 1 && (patch(r) > 40 * 1024 || ((long long)B * ((H + r - 1) / r) < sm_count && r > 8))) r = (r + 1) / 2;
  if (dzs > 1) r = std::max(dzs, r / dzs * dzs);
  return r;
}

}  // namespace

bool conv_direct_supported(int cin, int cout) { return cin >= 1 && cin <= 4 && cout >= 1 && cout <= 32; }

template <int MODE>
static int launch_direct_fwd_mode(dls_ctx *ctx, cudaStream_t st, const float *x, const float *w, const float *bias, float *out, int B,
                                  int H, int W, int cin, int cout, int act, int stride) {
  const int R = direct_rows(H, W, 1, B, ctx->sm_count);
  const size_t smem = (size_t)(R + 2) * (W + 2) * 4 * sizeof(float);
  const long long bands = (long long)B * ((H + R - 1) / R);
  const int grid = (int)std::min<long long>(bands, (long long)ctx->sm_count * 4);
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
static int launch_direct_dw_mode(dls_ctx *ctx, int stream, const float *x, const float *dz, float *dw, float *dbias, int B, int H, int W,
                                 int cin, int cout, int stride, int dzs) {
  cudaStream_t st = ctx->streams[stream];
  const int R = direct_rows(H, W, dzs, B, ctx->sm_count);
  const size_t patch = (size_t)(R + 2) * (W + 2) * 4, red = (size_t)kDirectWarps * kSlabRows * 32;
  const size_t smem = std::max(patch, red) * sizeof(float);
  const long long bands = (long long)B * ((H + R - 1) / R);
  const int grid = (int)std::min<long long>(bands, (long long)ctx->sm_count * 2);
  Block scratch;
  int rc = alloc_block(ctx, (size_t)grid * kSlabRows * 32 * sizeof(float), stream, &scratch);
  if (rc) return rc;
  float *partial = (float *)scratch.ptr;
  switch (cin) {
    case 1: conv_direct_dw_kernel<1, MODE><<<grid, kDirectWarps * 32, smem, st>>>(x, dz, partial, B, H, W, cout, R, stride, dzs); break;
    case 2: conv_direct_dw_kernel<2, MODE><<<grid, kDirectWarps * 32, smem, st>>>(x, dz, partial, B, H, W, cout, R, stride, dzs); break;
    case 3: conv_direct_dw_kernel<3, MODE><<<grid, kDirectWarps * 32, smem, st>>>(x, dz, partial, B, H, W, cout, R, stride, dzs); break;
    default: conv_direct_dw_kernel<4, MODE><<<grid, kDirectWarps * 32, smem, st>>>(x, dz, partial, B, H, W, cout, R, stride, dzs); break;
  }
  cudaError_t e = cudaGetLastError();
  if (e == cudaSuccess) {
    conv_direct_dw_reduce_kernel<<<((9 * cin + 1) * cout * 32 + 255) / 256, 256, 0, st>>>(partial, dw, dbias, grid, cin, cout);
    e = cudaGetLastError();
  }
  ctx->stats.kernel_launches += 2;
  free_block(ctx, scratch);
  if (e != cudaSuccess) return cuda_fail(e, "conv_direct_dw_kernel", __FILE__, __LINE__);
  return DLS_OK;
}

int launch_conv_direct_dw(dls_ctx *ctx, int stream, int mode, const float *x, const float *dz, float *dw, int B, int H, int W, int cin,
                          int cout, int stride, int dzs, float *dbias) {
  if (mode == DLS_GEMM_TF32) return launch_direct_dw_mode<DLS_GEMM_TF32>(ctx, stream, x, dz, dw, dbias, B, H, W, cin, cout, stride, dzs);
  if (mode == DLS_GEMM_BF16) return launch_direct_dw_mode<DLS_GEMM_BF16>(ctx, stream, x, dz, dw, dbias, B, H, W, cin, cout, stride, dzs);
  return launch_direct_dw_mode<DLS_GEMM_FP32>(ctx, stream, x, dz, dw, dbias, B, H, W, cin, cout, stride, dzs);
}

int launch_conv_stage_act(dls_ctx *ctx, int stream, int mode, const float *in, void *out, long long pixels, int H, int W, int C, int Cp,
                          int stride, int up) {
  const long long total = pixels * (Cp / 8);
  if (total == 0) return DLS_OK;
  const int grid = (int)std::min<long long>((total + 255) / 256, (long long)ctx->sm_count * 16);
  if (mode == DLS_GEMM_TF32) conv_stage_act_kernel<true><<<grid, 256, 0, ctx->streams[stream]>>>(in, out, pixels, H, W, C, Cp, stride, up);
  else conv_stage_act_kernel<false><<<grid, 256, 0, ctx->streams[stream]>>>(in, out, pixels, H, W, C, Cp, stride, up);
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

// The operand copy of `buf` read as [batch, h, w, c] (sampled at `stride` / zero-filled by `up`) written by the kernel that
// PRODUCED the buffer instead of by a staging pass: allocate it before the launch, attach it afterwards (finish_output drops
// the shadows of a written buffer) under the tag staged_act will look for.
long long act_tag_r(int mode, int up, int stride, int h, int w) {
  return ((long long)mode << 60) | ((long long)up << 56) | ((long long)stride << 48) | ((long long)h << 24) | (long long)w;
}
long long act_tag_c(int batch, int c) { return ((long long)batch << 24) | (long long)c; }

int alloc_emitted(dls_ctx *ctx, int stream, long long pixels, int c, Block *blk) {
  return alloc_block(ctx, (size_t)pixels * c * 2, stream, blk);
}
void attach_emitted(dls_ctx *ctx, int stream, dls_buf *buf, int slot, Block &blk, long long tag_r, long long tag_c) {
  if (!blk.ptr) return;
  if (buf->shadow[slot].ptr) free_block(ctx, buf->shadow[slot]);
  blk.stream_mask |= 1u << stream;
  buf->shadow[slot] = blk;
  buf->shadow_rows[slot] = tag_r;
  buf->shadow_cols[slot] = tag_c;
  blk = Block();
}

// Staged NHWC operand copy of `buf` read as [batch, h*stride, w*stride, c] and sampled at `stride` (shadow slot 2, kept
// until the buffer is written again: the forward pass and the weight gradient share the copy of an activation, the
// backward-data pass and the weight gradient share the copy of a delta).
int staged_act(dls_ctx *ctx, int stream, int mode, dls_buf *buf, int batch, int h, int w, int c, int stride, int up, const void **ptr,
               int *cp, int slot = 2) {
  Block &sh = buf->shadow[slot];
  const long long tag_r = act_tag_r(mode, up, stride, h, w);
  const long long tag_c = act_tag_c(batch, c);
  const int p = pad_channels(mode, c);
  {
    // a dense bf16 copy of the whole buffer left by the kernel that produced it (DLS_GEMM_EMIT_OPERAND_COPY or a convolution
    // epilogue, slot 0, [rows][cols] with cols % 8 == 0) is this operand when no channel padding and no resampling is needed
    Block &flat = buf->shadow[0];
    if (mode == DLS_GEMM_BF16 && stride == 1 && up == 1 && p == c && flat.ptr && flat.pool == ctx->capture_pool &&
        buf->shadow_cols[0] > 0 && buf->shadow_cols[0] % 8 == 0 &&
        buf->shadow_rows[0] * buf->shadow_cols[0] == (long long)batch * h * w * c) {
      int rc = wait_streams(ctx, stream, flat.stream_mask);
      if (rc) return rc;
      flat.stream_mask |= 1u << stream;
      *ptr = flat.ptr;
      *cp = p;
      return DLS_OK;
    }
  }
  if (!(sh.ptr && buf->shadow_rows[slot] == tag_r && buf->shadow_cols[slot] == tag_c && sh.pool == ctx->capture_pool)) {
    if (sh.ptr) free_block(ctx, sh);
    const long long pixels = (long long)batch * h * w;
    int rc = alloc_block(ctx, (size_t)pixels * p * elem_bytes(mode), stream, &sh);
    if (!rc) rc = launch_conv_stage_act(ctx, stream, mode, buf->ptr, sh.ptr, pixels, h, w, c, p, stride, up);
    if (rc) {
      if (sh.ptr) free_block(ctx, sh);
      return rc;
    }
    buf->shadow_rows[slot] = tag_r;
    buf->shadow_cols[slot] = tag_c;
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
// first-layer kernels: the (R + 2) x (W + 2) x 4 patch of even a one-row band must fit the default shared-memory window
bool use_direct(size_t cin, size_t cout, size_t w) { return conv_direct_supported((int)cin, (int)cout) && w <= 512; }

}  // namespace

extern "C" {

int dls_conv3x3_supported(dls_ctx *ctx, int pass, size_t h, size_t w, size_t cin, size_t cout, int mode) {
  if (!ctx) return 0;
  if (mode < 0) mode = ctx->opts.gemm_mode;
  if (mode != DLS_GEMM_TF32 && mode != DLS_GEMM_BF16) return 0;
  if (!fits_int(h) || !fits_int(w) || !fits_int(cin) || !fits_int(cout)) return 0;
  if (use_direct(cin, cout, w)) return pass == DLS_CONV_FORWARD || pass == DLS_CONV_BACKWARD_WEIGHTS;
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
  Block emitted;
  int emit_stride = 0;
  if (use_direct(cin, cout, wd)) {
    rc = launch_conv_direct_fwd(ctx, stream, mode, x->ptr, w->ptr, bias ? bias->ptr : nullptr, o->ptr, (int)batch, (int)h, (int)wd,
                                (int)cin, (int)cout, act, stride);
  } else {
    const void *xa, *wf, *wb;
    int cp;
    rc = staged_act(ctx, stream, mode, x, (int)batch, (int)h, (int)wd, (int)cin, stride, 1, &xa, &cp);
    if (!rc) rc = staged_weights(ctx, stream, mode, w, (int)cin, (int)cout, &wf, &wb);
    GemmEpilogue epi;
    epi.bias = bias ? bias->ptr : nullptr;
    epi.bias_mode = bias ? 2 : 0;
    epi.act = act;
    // opts->dz_stride (forward: "the value will be consumed sampled at this stride"): s = 1 leaves the dense bf16 copy of the
    // result (slot 0: the next contraction or convolution starts without a staging pass), s > 1 the copy of the pixels
    // (s*y, s*x) only -- the staged input of a convolution behind a `scale` down-sampling (slot 2)
    const int os = opts ? opts->dz_stride : 0;
    if (!rc && mode == DLS_GEMM_BF16 && os >= 1 && cout % 64 == 0 && h % os == 0 && wd % os == 0)
      rc = alloc_emitted(ctx, stream, (long long)batch * (h / os) * (wd / os), (int)cout, &emitted);
    if (!rc)
      rc = launch_conv_tc(ctx, stream, mode, xa, (int)batch, (int)h, (int)wd, cp, wf, (int)cout, o->ptr, +1, epi, 1,
                          os == 1 ? emitted.ptr : nullptr, os > 1 ? emitted.ptr : nullptr, nullptr, os > 1 ? os : 1);
    emit_stride = os;
  }
  if (!rc) rc = finish_output(ctx, stream, &o, 1, ins, 3, done);
  if (rc) {
    if (emitted.ptr) free_block(ctx, emitted);
    dls_buf_release(o);
    return rc;
  }
  if (emitted.ptr && emit_stride == 1) {
    emitted.stream_mask |= 1u << stream;
    o->shadow[0] = emitted;
    o->shadow_rows[0] = (long long)(batch * h * wd);
    o->shadow_cols[0] = (long long)cout;
  } else if (emitted.ptr) {
    attach_emitted(ctx, stream, o, 2, emitted, act_tag_r(mode, 1, emit_stride, (int)h / emit_stride, (int)wd / emit_stride),
                   act_tag_c((int)batch, (int)cout));
  }
  *out = o;
  return DLS_OK;
}

int dls_conv3x3_backward_data_gated(dls_ctx *ctx, dls_buf *dz, dls_buf *w, dls_buf *gate, int gate_mode, int gate_stride, size_t batch,
                                    size_t h, size_t wd, size_t cin, size_t cout, const dls_conv_opts *opts, dls_event *const *deps,
                                    int n_deps, dls_buf **out, dls_event **done) {
  DLS_CHECK(ctx && dz && w && out, DLS_ERR_INVALID, "dls_conv3x3_backward_data: NULL argument");
  if (gate_stride < 1) gate_stride = 1;
  DLS_CHECK(gate_mode >= 0 && gate_mode <= 2 && (gate_mode == 0 || (gate && gate->n >= batch * h * gate_stride * wd * gate_stride * cin)),
            DLS_ERR_INVALID, "dls_conv3x3_backward_data_gated: gate missing or shorter than [batch, h*gs, w*gs, cin]");
  DLS_CHECK(fits_int(batch) && fits_int(h) && fits_int(wd) && fits_int(cin) && fits_int(cout), DLS_ERR_INVALID,
            "dls_conv3x3_backward_data: bad shape");
  const int dzs = (opts && opts->dz_stride > 1) ? opts->dz_stride : 1;
  DLS_CHECK(h % dzs == 0 && wd % dzs == 0, DLS_ERR_INVALID, "dls_conv3x3_backward_data: dz_stride must divide the image size");
  DLS_CHECK(batch * (h / dzs) * (wd / dzs) * cout <= dz->n && 9 * cin * cout <= w->n, DLS_ERR_INVALID,
            "dls_conv3x3_backward_data: a buffer is shorter than its shape");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  const int mode = conv_mode(ctx, opts);
  DLS_CHECK(dls_conv3x3_supported(ctx, DLS_CONV_BACKWARD_DATA, h, wd, cin, cout, mode), DLS_ERR_UNSUPPORTED,
            "dls_conv3x3_backward_data: shape / mode not covered");
  DLS_CUDA(cudaSetDevice(ctx->device));
  const int stream = pick_stream(ctx);
  dls_buf *ins[3] = {dz, w, gate_mode ? gate : nullptr};
  int rc = wait_inputs(ctx, stream, ins, 3, deps, n_deps);
  if (rc) return rc;
  dls_buf *o = nullptr;
  rc = new_buf(ctx, batch * h * wd * cin, stream, &o);
  if (rc) return rc;
  const void *da, *wf, *wb;
  int cp;
  rc = staged_act(ctx, stream, mode, dz, (int)batch, (int)h, (int)wd, (int)cout, 1, dzs, &da, &cp);
  if (!rc) rc = staged_weights(ctx, stream, mode, w, (int)cin, (int)cout, &wf, &wb);
  // dx[b,y,x,c] = sum_{ky,kx,f} dz[b, y+(ky-1), x+(kx-1), f] * w[ky,kx,c,f]: the same kernel with the taps mirrored
  GemmEpilogue epi;
  epi.gate = gate_mode ? gate->ptr : nullptr;
  epi.gate_mode = gate_mode;
  // gated: the result IS the delta of the layer below, whose two backward passes consume it as a convolution operand --
  // dense (gate_stride 1: one copy, slot 0) or sampled at gate_stride (compact copy for its weight gradient, slot 4, and the
  // zero-filled full-size copy for its backward-data pass, slot 2), written by this epilogue instead of two staging passes
  Block plain, upcopy;
  const bool emit = mode == DLS_GEMM_BF16 && gate_mode && cin % 64 == 0;
  if (!rc && emit) rc = alloc_emitted(ctx, stream, (long long)batch * h * wd, (int)cin, &plain);
  if (!rc && emit && gate_stride > 1)
    rc = alloc_emitted(ctx, stream, (long long)batch * h * gate_stride * wd * gate_stride, (int)cin, &upcopy);
  if (!rc)
    rc = launch_conv_tc(ctx, stream, mode, da, (int)batch, (int)h, (int)wd, cp, wb, (int)cin, o->ptr, -1, epi, gate_stride, plain.ptr,
                        nullptr, upcopy.ptr, gate_stride);
  if (!rc) rc = finish_output(ctx, stream, &o, 1, ins, 3, done);
  if (rc) {
    if (plain.ptr) free_block(ctx, plain);
    if (upcopy.ptr) free_block(ctx, upcopy);
    dls_buf_release(o);
    return rc;
  }
  if (plain.ptr && gate_stride == 1) {
    plain.stream_mask |= 1u << stream;
    o->shadow[0] = plain;
    o->shadow_rows[0] = (long long)(batch * h * wd);
    o->shadow_cols[0] = (long long)cin;
  } else if (plain.ptr) {
    attach_emitted(ctx, stream, o, 4, plain, act_tag_r(mode, 1, 1, (int)h, (int)wd), act_tag_c((int)batch, (int)cin));
    attach_emitted(ctx, stream, o, 2, upcopy, act_tag_r(mode, gate_stride, 1, (int)h * gate_stride, (int)wd * gate_stride),
                   act_tag_c((int)batch, (int)cin));
  }
  *out = o;
  return DLS_OK;
}

int dls_conv3x3_backward_data(dls_ctx *ctx, dls_buf *dz, dls_buf *w, size_t batch, size_t h, size_t wd, size_t cin, size_t cout,
                              const dls_conv_opts *opts, dls_event *const *deps, int n_deps, dls_buf **out, dls_event **done) {
  return dls_conv3x3_backward_data_gated(ctx, dz, w, nullptr, 0, 1, batch, h, wd, cin, cout, opts, deps, n_deps, out, done);
}

int dls_conv3x3_backward_weights_bias(dls_ctx *ctx, dls_buf *x, dls_buf *dz, size_t batch, size_t h, size_t wd, size_t cin, size_t cout,
                                      const dls_conv_opts *opts, dls_event *const *deps, int n_deps, dls_buf **out,
                                      dls_buf **out_bias, dls_event **done) {
  DLS_CHECK(ctx && x && dz && out, DLS_ERR_INVALID, "dls_conv3x3_backward_weights: NULL argument");
  DLS_CHECK(!out_bias || use_direct(cin, cout, wd), DLS_ERR_UNSUPPORTED,
            "dls_conv3x3_backward_weights_bias: only the first-layer kernel (Cin <= 4) sums the bias gradient on the way");
  const int stride = (opts && opts->in_stride > 1) ? opts->in_stride : 1;
  DLS_CHECK(fits_int(batch) && fits_int(h) && fits_int(wd) && fits_int(cin) && fits_int(cout), DLS_ERR_INVALID,
            "dls_conv3x3_backward_weights: bad shape");
  const int dzs = (opts && opts->dz_stride > 1) ? opts->dz_stride : 1;
  DLS_CHECK(h % dzs == 0 && wd % dzs == 0, DLS_ERR_INVALID, "dls_conv3x3_backward_weights: dz_stride must divide the image size");
  DLS_CHECK(batch * h * stride * wd * stride * cin <= x->n && batch * (h / dzs) * (wd / dzs) * cout <= dz->n, DLS_ERR_INVALID,
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
  dls_buf *o = nullptr, *ob = nullptr;
  rc = new_buf(ctx, 9 * cin * cout, stream, &o);
  if (!rc && out_bias) rc = new_buf(ctx, cout, stream, &ob);
  if (rc) {
    if (o) dls_buf_release(o);
    return rc;
  }
  if (use_direct(cin, cout, wd)) {
    rc = launch_conv_direct_dw(ctx, stream, mode, x->ptr, dz->ptr, o->ptr, (int)batch, (int)h, (int)wd, (int)cin, (int)cout, stride, dzs,
                               ob ? ob->ptr : nullptr);
  } else {
    const void *xa, *da;
    int cxp, cdp;
    rc = staged_act(ctx, stream, mode, x, (int)batch, (int)h, (int)wd, (int)cin, stride, 1, &xa, &cxp);
    // a sampled delta is contracted over its own pixels only: staged compact (slot 4; backward-data keeps the zero-filled
    // full-size copy in slot 2), the activation read through TMA element strides
    ConvGeom probe;
    const bool compact = dzs > 1 && dzs <= 8 && conv_tiling((int)h / dzs, (int)wd / dzs, mode == DLS_GEMM_TF32 ? 32 : 64, &probe) &&
                         probe.bw * dzs <= 256 &&
                         probe.bh * dzs <= 256;
    if (!rc && compact) rc = staged_act(ctx, stream, mode, dz, (int)batch, (int)h / dzs, (int)wd / dzs, (int)cout, 1, 1, &da, &cdp, 4);
    else if (!rc) rc = staged_act(ctx, stream, mode, dz, (int)batch, (int)h, (int)wd, (int)cout, 1, dzs, &da, &cdp);
    if (!rc) rc = launch_conv_dw_tc(ctx, stream, mode, xa, cxp, da, cdp, (int)batch, (int)h, (int)wd, (int)cin, (int)cout, o->ptr,
                                    compact ? dzs : 1);
  }
  dls_buf *outs[2] = {o, ob};
  if (!rc) rc = finish_output(ctx, stream, outs, ob ? 2 : 1, ins, 2, done);
  if (rc) {
    dls_buf_release(o);
    if (ob) dls_buf_release(ob);
    return rc;
  }
  *out = o;
  if (out_bias) *out_bias = ob;
  return DLS_OK;
}

int dls_conv3x3_backward_weights(dls_ctx *ctx, dls_buf *x, dls_buf *dz, size_t batch, size_t h, size_t wd, size_t cin, size_t cout,
                                 const dls_conv_opts *opts, dls_event *const *deps, int n_deps, dls_buf **out, dls_event **done) {
  return dls_conv3x3_backward_weights_bias(ctx, x, dz, batch, h, wd, cin, cout, opts, deps, n_deps, out, nullptr, done);
}

}  // extern "C"
