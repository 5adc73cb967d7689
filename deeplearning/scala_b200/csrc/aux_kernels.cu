// aux_kernels.cu -- the remaining hand-written sm_100a kernels:
//   N3  sum_kernel / sum_axis_*   full and single-axis reductions (warp shuffles + block tree, one launch)
//   N5  random_kernel / random_normal_kernel   counter-based RNG, bit-compatible with Tensors:98-117, :398-443
//   N6  sgd_kernel                fused in-place `data - lr * delta` (TW:63-72 without the buffer swap)
//   N6b optimizer_kernel<kind>    momentum SGD / Adam: weight, state and gradient in one pass
#include "common.h"
#include "reduce.cuh"

namespace dls {

// ---- N3: full sum (replaces reduce(PlusPrograms), Tensors:673-766) ---------------------------------------------------
__global__ void __launch_bounds__(kBlock) sum_kernel(const float *__restrict__ in, size_t n, ReduceArgs R) {
  const size_t n4 = n / 4;
  const float4 *in4 = reinterpret_cast<const float4 *>(in);
  const size_t stride = (size_t)gridDim.x * kBlock;
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
  size_t i = (size_t)blockIdx.x * kBlock + threadIdx.x;
  // four independent 128-bit loads in flight per lane
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

int launch_sum(dls_ctx *ctx, int stream, const float *in, size_t n, float *out) {
  cudaStream_t st = ctx->streams[stream];
  if (n == 0) {
    DLS_CUDA(cudaMemsetAsync(out, 0, sizeof(float), st));
    return DLS_OK;
  }
  ReduceArgs R{ctx->red_partials[stream], ctx->red_counter[stream], out};
  size_t want = (n / 4 + (size_t)kBlock * 4 - 1) / ((size_t)kBlock * 4);
  size_t cap = (size_t)ctx->sm_count * 8;
  if (cap > (size_t)kReduceMaxBlocks) cap = kReduceMaxBlocks;
  int grid = (int)std::max<size_t>(1, std::min(want, cap));
  sum_kernel<<<grid, kBlock, 0, st>>>(in, n, R);
  DLS_CUDA(cudaGetLastError());
  ctx->stats.kernel_launches++;
  return DLS_OK;
}

// ---- N3: one-axis sum over a row-major [outer, k, inner] view ---------------------------------------------------------
// inner > 1: one lane per (outer, inner) column, lanes coalesced along `inner`, k summed in ascending order -- the same
// left fold ((t0 + t1) + t2) + ... the reference builds on the host (TL:124-162), so results are bit-identical to it.
__global__ void __launch_bounds__(kBlock) sum_axis_cols_kernel(const float *__restrict__ in, float *__restrict__ out,
                                                               size_t outer, size_t k, size_t inner) {
  const size_t total = outer * inner;
  for (size_t c = (size_t)blockIdx.x * kBlock + threadIdx.x; c < total; c += (size_t)gridDim.x * kBlock) {
    const size_t o = c / inner, i = c - o * inner;
    const float *p = in + o * k * inner + i;
    float acc = __ldg(p);
    size_t j = 1;
    for (; j + 4 <= k; j += 4) {
      float v0 = __ldg(p + j * inner), v1 = __ldg(p + (j + 1) * inner), v2 = __ldg(p + (j + 2) * inner),
            v3 = __ldg(p + (j + 3) * inner);
      acc = __fadd_rn(__fadd_rn(__fadd_rn(__fadd_rn(acc, v0), v1), v2), v3);
    }
    for (; j < k; ++j) acc = __fadd_rn(acc, __ldg(p + j * inner));
    out[c] = acc;
  }
}

// inner == 1: one warp per row, lanes strided along k, shuffle tree (order differs from the host fold: <= 1e-5 rel).
__global__ void __launch_bounds__(kBlock) sum_axis_rows_kernel(const float *__restrict__ in, float *__restrict__ out,
                                                               size_t rows, size_t k) {
  const size_t warps = (size_t)gridDim.x * (kBlock / 32);
  const int lane = threadIdx.x & 31;
  for (size_t r = (size_t)blockIdx.x * (kBlock / 32) + (threadIdx.x >> 5); r < rows; r += warps) {
    const float *p = in + r * k;
    float acc = 0.f;
    for (size_t j = lane; j < k; j += 32) acc += __ldg(p + j);
    acc = warp_sum(acc);
    if (lane == 0) out[r] = acc;
  }
}

// few columns, long k (bias gradients: [B*H*W, F] summed over rows): stage 1 sums `chunk` consecutive rows per lane into
// partial[outer, chunks, inner]; stage 2 is the column kernel over the chunks.  Lanes stay coalesced along `inner`.
__global__ void __launch_bounds__(kBlock) sum_axis_chunks_kernel(const float *__restrict__ in, float *__restrict__ partial,
                                                                 size_t outer, size_t k, size_t inner, size_t chunks, size_t chunk) {
  const size_t total = outer * chunks * inner;
  for (size_t t = (size_t)blockIdx.x * kBlock + threadIdx.x; t < total; t += (size_t)gridDim.x * kBlock) {
    const size_t i = t % inner, c = (t / inner) % chunks, o = t / (inner * chunks);
    const size_t j0 = c * chunk, j1 = min(k, j0 + chunk);
    const float *p = in + (o * k + j0) * inner + i;
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
    size_t j = j0;
    for (; j + 4 <= j1; j += 4, p += 4 * inner) {
      a0 += __ldg(p);
      a1 += __ldg(p + inner);
      a2 += __ldg(p + 2 * inner);
      a3 += __ldg(p + 3 * inner);
    }
    for (; j < j1; ++j, p += inner) a0 += __ldg(p);
    partial[t] = (a0 + a1) + (a2 + a3);
  }
}

// few, short columns and a long k (inner <= 256): a block owns a slab of consecutive rows and reads it as ONE contiguous
// run (thread t -> element t of the slab, so column = t % inner), keeps one partial per (row-lane, column), folds the
// row-lanes through shared memory and writes partial[outer][block][column].  Launched twice: slabs -> gridDim.x partials,
// then one block per outer index over those partials.
__global__ void __launch_bounds__(kBlock) sum_axis_slab_kernel(const float *__restrict__ in, float *__restrict__ out, size_t k,
                                                               size_t inner, size_t rows_per_block) {
  __shared__ float part[kBlock];
  const size_t o = blockIdx.y;
  const int lanes = kBlock / (int)inner;                 // row-lanes per block (>= 1)
  const int col = threadIdx.x % (int)inner, rl = threadIdx.x / (int)inner;
  const size_t r0 = (size_t)blockIdx.x * rows_per_block, r1 = min(k, r0 + rows_per_block);
  float a0 = 0.f, a1 = 0.f;
  if (rl < lanes) {
    const float *p = in + (o * k + r0) * inner + col;
    size_t r = r0 + rl;
    for (; r + lanes < r1; r += 2 * (size_t)lanes) {
      a0 += __ldg(p + (r - r0) * inner);
      a1 += __ldg(p + (r - r0 + lanes) * inner);
    }
    if (r < r1) a0 += __ldg(p + (r - r0) * inner);
  }
  part[threadIdx.x] = a0 + a1;
  __syncthreads();
  if (rl == 0) {
    float s = part[col];
    for (int l = 1; l < lanes; ++l) s += part[l * (int)inner + col];
    out[(o * gridDim.x + blockIdx.x) * inner + col] = s;
  }
}

// Tall narrow slabs with inner % 4 == 0 (bias gradients: [batch * H * W, channels]): a thread owns 4 consecutive columns
// and walks rows with four independent 128-bit loads in flight; the block's row-lanes are folded through shared memory.
__global__ void __launch_bounds__(kBlock) sum_axis_slab4_kernel(const float *__restrict__ in, float *__restrict__ out, size_t k,
                                                                size_t inner, size_t rows_per_block) {
  __shared__ float4 part[kBlock];
  const size_t o = blockIdx.y;
  const int inner4 = (int)(inner / 4);
  const int lanes = kBlock / inner4;                      // row-lanes per block (>= 1)
  const int c4 = threadIdx.x % inner4, rl = threadIdx.x / inner4;
  const size_t r0 = (size_t)blockIdx.x * rows_per_block, r1 = min(k, r0 + rows_per_block);
  float4 a[4];
#pragma unroll
  for (int u = 0; u < 4; ++u) a[u] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (rl < lanes) {
    const float4 *p = reinterpret_cast<const float4 *>(in + o * k * inner) + c4;
    size_t r = r0 + rl;
    for (; r + 3 * (size_t)lanes < r1; r += 4 * (size_t)lanes) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = __ldcs(p + (r + (size_t)u * lanes) * inner4);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        a[u].x += v[u].x;
        a[u].y += v[u].y;
        a[u].z += v[u].z;
        a[u].w += v[u].w;
      }
    }
    for (; r < r1; r += lanes) {
      const float4 v = __ldcs(p + r * inner4);
      a[0].x += v.x;
      a[0].y += v.y;
      a[0].z += v.z;
      a[0].w += v.w;
    }
  }
  part[threadIdx.x] = make_float4((a[0].x + a[1].x) + (a[2].x + a[3].x), (a[0].y + a[1].y) + (a[2].y + a[3].y),
                                  (a[0].z + a[1].z) + (a[2].z + a[3].z), (a[0].w + a[1].w) + (a[2].w + a[3].w));
  __syncthreads();
  if (rl == 0) {
    float4 s = part[c4];
    for (int l = 1; l < lanes; ++l) {
      const float4 v = part[l * inner4 + c4];
      s.x += v.x;
      s.y += v.y;
      s.z += v.z;
      s.w += v.w;
    }
    reinterpret_cast<float4 *>(out + (o * gridDim.x + blockIdx.x) * inner)[c4] = s;
  }
}

// Second stage of the slab reduction: `rows` partial rows of `inner` columns per outer index; one warp per column.
__global__ void __launch_bounds__(kBlock) sum_axis_finish_kernel(const float *__restrict__ in, float *__restrict__ out, size_t rows,
                                                                 size_t inner) {
  const size_t o = blockIdx.y;
  const size_t col = (size_t)blockIdx.x * (kBlock / 32) + (threadIdx.x >> 5);
  if (col >= inner) return;
  const float *p = in + o * rows * inner + col;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  size_t r = threadIdx.x & 31;
  for (; r + 96 < rows; r += 128) {
    s0 += __ldg(p + r * inner);
    s1 += __ldg(p + (r + 32) * inner);
    s2 += __ldg(p + (r + 64) * inner);
    s3 += __ldg(p + (r + 96) * inner);
  }
  for (; r < rows; r += 32) s0 += __ldg(p + r * inner);
  const float s = warp_sum((s0 + s1) + (s2 + s3));
  if ((threadIdx.x & 31) == 0) out[o * inner + col] = s;
}

// One-launch forms for the bias gradients of a training step (outer == 1, [k, inner] summed over its rows), where a second
// launch costs as much as the reduction itself.
// (a) small matrices (a few thousand elements per 32-column group): a block owns 32 columns, its 8 warps stride the rows,
//     one shared-memory fold.
__global__ void __launch_bounds__(kBlock) sum_axis_small_kernel(const float *__restrict__ in, float *__restrict__ out, size_t k,
                                                                size_t inner) {
  __shared__ float part[kBlock / 32][33];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const size_t col = (size_t)blockIdx.x * 32 + lane;
  float a0 = 0.f, a1 = 0.f;
  if (col < inner) {
    size_t r = w;
    for (; r + kBlock / 32 < k; r += 2 * (kBlock / 32)) {
      a0 += __ldg(in + r * inner + col);
      a1 += __ldg(in + (r + kBlock / 32) * inner + col);
    }
    if (r < k) a0 += __ldg(in + r * inner + col);
  }
  part[w][lane] = a0 + a1;
  __syncthreads();
  if (w == 0 && col < inner) {
    float s = part[0][lane];
#pragma unroll
    for (int i = 1; i < kBlock / 32; ++i) s += part[i][lane];
    out[col] = s;
  }
}

// (b) tall slabs with inner % 4 == 0 and inner <= 256: sum_axis_slab4's first stage, and the block that draws the last
//     ticket folds the per-block partial rows in a fixed order (deterministic), so there is no finish launch.
__global__ void __launch_bounds__(kBlock) sum_axis_slab4_ticket_kernel(const float *__restrict__ in, float *__restrict__ partial,
                                                                       float *__restrict__ out, unsigned int *counter, size_t k,
                                                                       size_t inner, size_t rows_per_block) {
  __shared__ float4 part[kBlock];
  __shared__ bool is_last;
  const int inner4 = (int)(inner / 4);
  const int lanes = kBlock / inner4;
  const int c4 = threadIdx.x % inner4, rl = threadIdx.x / inner4;
  const size_t r0 = (size_t)blockIdx.x * rows_per_block, r1 = min(k, r0 + rows_per_block);
  // eight independent 128-bit loads in flight per thread and trip; rows past the slab are predicated off, so a slab of up to
  // 8 x lanes rows is ONE round trip to memory whatever its exact height
  auto fold8 = [&](const float4 *base, size_t first, size_t limit, size_t step, bool cg) {
    float4 a[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) a[u] = make_float4(0.f, 0.f, 0.f, 0.f);
    for (size_t r = first; r < limit; r += 8 * step) {
      float4 v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const size_t row = r + (size_t)u * step;
        v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row < limit) v[u] = cg ? __ldcg(base + row * inner4) : __ldcs(base + row * inner4);
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        a[u & 3].x += v[u].x;
        a[u & 3].y += v[u].y;
        a[u & 3].z += v[u].z;
        a[u & 3].w += v[u].w;
      }
    }
    return make_float4((a[0].x + a[1].x) + (a[2].x + a[3].x), (a[0].y + a[1].y) + (a[2].y + a[3].y),
                       (a[0].z + a[1].z) + (a[2].z + a[3].z), (a[0].w + a[1].w) + (a[2].w + a[3].w));
  };
  auto fold_lanes = [&](float4 mine, float *dst) {   // row-lanes of the block -> one row, in lane order
    part[threadIdx.x] = mine;
    __syncthreads();
    if (rl == 0) {
      float4 s = part[c4];
      for (int l = 1; l < lanes; ++l) {
        const float4 v = part[l * inner4 + c4];
        s.x += v.x;
        s.y += v.y;
        s.z += v.z;
        s.w += v.w;
      }
      reinterpret_cast<float4 *>(dst)[c4] = s;
    }
  };
  float4 mine = make_float4(0.f, 0.f, 0.f, 0.f);
  if (rl < lanes) mine = fold8(reinterpret_cast<const float4 *>(in) + c4, r0 + rl, r1, (size_t)lanes, false);
  fold_lanes(mine, partial + (size_t)blockIdx.x * inner);
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) is_last = atomicAdd(counter, 1u) == gridDim.x - 1;
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  mine = make_float4(0.f, 0.f, 0.f, 0.f);
  if (rl < lanes) mine = fold8(reinterpret_cast<const float4 *>(partial) + c4, (size_t)rl, (size_t)gridDim.x, (size_t)lanes, true);
  __syncthreads();
  fold_lanes(mine, out);
  if (threadIdx.x == 0) *counter = 0u;   // ready for the next launch on this stream
}

int launch_sum_axis(dls_ctx *ctx, int stream, const float *in, size_t outer, size_t k, size_t inner, float *out) {
  cudaStream_t st = ctx->streams[stream];
  if (outer * inner == 0) return DLS_OK;
  if (k == 0) {
    DLS_CUDA(cudaMemsetAsync(out, 0, outer * inner * sizeof(float), st));
    return DLS_OK;
  }
  size_t cap = (size_t)ctx->sm_count * 8;
  const size_t lanes_wanted = (size_t)ctx->sm_count * kBlock * 4;
  if (outer == 1 && inner > 1 && k * std::min<size_t>(inner, 32) <= 16384 && inner <= 4096) {
    sum_axis_small_kernel<<<(unsigned)((inner + 31) / 32), kBlock, 0, st>>>(in, out, k, inner);
  } else if (outer == 1 && inner % 4 == 0 && inner <= kBlock && k >= 512 && ((uintptr_t)in & 15) == 0 && ((uintptr_t)out & 15) == 0) {
    // one block per SM: the last block's fold over the per-block partials stays a couple of trips
    size_t blocks = std::max<size_t>(1, std::min<size_t>((size_t)ctx->sm_count, (k + 63) / 64));
    const size_t rows_per_block = (k + blocks - 1) / blocks;
    blocks = (k + rows_per_block - 1) / rows_per_block;
    Block scratch;
    int rc = alloc_block(ctx, blocks * inner * sizeof(float), stream, &scratch);
    if (rc) return rc;
    sum_axis_slab4_ticket_kernel<<<(unsigned)blocks, kBlock, 0, st>>>(in, (float *)scratch.ptr, out, ctx->red_counter[stream], k, inner,
                                                                     rows_per_block);
    free_block(ctx, scratch);
  } else if (inner == 1 && k >= 64) {
    size_t want = (outer + kBlock / 32 - 1) / (kBlock / 32);
    sum_axis_rows_kernel<<<(int)std::max<size_t>(1, std::min(want, cap)), kBlock, 0, st>>>(in, out, outer, k);
  } else if (inner <= kBlock && k >= 512 && outer * inner * 8 <= lanes_wanted && outer < 65536) {
    size_t blocks = std::min<size_t>((size_t)ctx->sm_count * 4 / std::max<size_t>(1, std::min<size_t>(outer, 4)), (k + 63) / 64);
    blocks = std::max<size_t>(1, blocks);
    const size_t rows_per_block = (k + blocks - 1) / blocks;
    blocks = (k + rows_per_block - 1) / rows_per_block;
    Block scratch;
    int rc = alloc_block(ctx, outer * blocks * inner * sizeof(float), stream, &scratch);
    if (rc) return rc;
    if (inner % 4 == 0 && ((uintptr_t)in & 15) == 0)
      sum_axis_slab4_kernel<<<dim3((unsigned)blocks, (unsigned)outer), kBlock, 0, st>>>(in, (float *)scratch.ptr, k, inner, rows_per_block);
    else
      sum_axis_slab_kernel<<<dim3((unsigned)blocks, (unsigned)outer), kBlock, 0, st>>>(in, (float *)scratch.ptr, k, inner, rows_per_block);
    sum_axis_finish_kernel<<<dim3((unsigned)((inner + kBlock / 32 - 1) / (kBlock / 32)), (unsigned)outer), kBlock, 0, st>>>(
        (const float *)scratch.ptr, out, blocks, inner);
    free_block(ctx, scratch);
    ctx->stats.kernel_launches++;
  } else if (outer * inner * 8 <= lanes_wanted && k >= 256) {
    // two stages: enough chunks to fill the machine, at least 32 rows each
    size_t chunks = std::min(lanes_wanted / (outer * inner), (k + 31) / 32);
    const size_t chunk = (k + chunks - 1) / chunks;
    chunks = (k + chunk - 1) / chunk;
    Block scratch;
    int rc = alloc_block(ctx, outer * chunks * inner * sizeof(float), stream, &scratch);
    if (rc) return rc;
    size_t want = (outer * chunks * inner + kBlock - 1) / kBlock;
    sum_axis_chunks_kernel<<<(int)std::max<size_t>(1, std::min(want, cap)), kBlock, 0, st>>>(in, (float *)scratch.ptr, outer, k, inner,
                                                                                          chunks, chunk);
    want = (outer * inner + kBlock - 1) / kBlock;
    sum_axis_cols_kernel<<<(int)std::max<size_t>(1, std::min(want, cap)), kBlock, 0, st>>>((const float *)scratch.ptr, out, outer, chunks,
                                                                                        inner);
    free_block(ctx, scratch);   // stream ordered: the next user of the block runs after these kernels
    ctx->stats.kernel_launches++;
  } else {
    size_t want = (outer * inner + kBlock - 1) / kBlock;
    sum_axis_cols_kernel<<<(int)std::max<size_t>(1, std::min(want, cap)), kBlock, 0, st>>>(in, out, outer, k, inner);
  }
  DLS_CUDA(cudaGetLastError());
  ctx->stats.kernel_launches++;
  return DLS_OK;
}

// ---- N5: RNG --------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t wang_hash(uint32_t v) {  // Tensors:107-116
  v = (v ^ 61u) ^ (v >> 16);
  v *= 9u;
  v ^= v << 4;
  v *= 0x27d4eb2du;
  v ^= v >> 15;
  return v;
}
__device__ __forceinline__ uint32_t lcg_hash(uint32_t v) { return 1103515245u * v + 12345u; }  // Tensors:99-103
__device__ __forceinline__ uint32_t xorshift(uint32_t s) {                                      // Tensors:403-407
  uint32_t t1 = s ^ (s << 13);
  uint32_t t2 = t1 ^ (t1 >> 17);
  return t2 ^ (t2 << 5);
}
__device__ __forceinline__ float u01(uint32_t h) { return __fdiv_rn(__uint2float_rn(h), 4294967296.0f); }

template <int HASH>
__global__ void __launch_bounds__(kBlock) random_kernel(float *__restrict__ out, size_t n, uint32_t seed) {
  // Tensors:436-439: buffer[i] = hash(i ^ seed) / 4294967296.0f
  for (size_t i = (size_t)blockIdx.x * kBlock + threadIdx.x; i < n; i += (size_t)gridDim.x * kBlock) {
    uint32_t x = (uint32_t)i ^ seed;
    out[i] = u01(HASH == DLS_HASH_WANG ? wang_hash(x) : lcg_hash(x));
  }
}

template <int HASH>
__global__ void __launch_bounds__(kBlock) random_normal_kernel(float2 *__restrict__ out, size_t n_pairs, uint32_t seed) {
  // Tensors:409-425: Box-Muller, work-item i writes elements 2i and 2i+1
  for (size_t i = (size_t)blockIdx.x * kBlock + threadIdx.x; i < n_pairs; i += (size_t)gridDim.x * kBlock) {
    uint32_t x = (uint32_t)i ^ seed;
    uint32_t r1 = HASH == DLS_HASH_WANG ? wang_hash(x) : lcg_hash(x);
    uint32_t r2 = xorshift(r1);
    float u1 = u01(r1), u2 = u01(r2);
    float r = __fsqrt_rn(__fmul_rn(-2.0f, logf(u1)));
    float theta = __fmul_rn(6.28318530717958647692f, u2);
    out[i] = make_float2(__fmul_rn(r, cosf(theta)), __fmul_rn(r, sinf(theta)));
  }
}

static int grid_for(dls_ctx *ctx, size_t work) {
  size_t want = (work + kBlock - 1) / kBlock;
  size_t cap = (size_t)ctx->sm_count * 8;
  return (int)std::max<size_t>(1, std::min(want, cap));
}

int launch_random(dls_ctx *ctx, int stream, float *out, size_t n, uint32_t seed, int hash_kind) {
  if (n == 0) return DLS_OK;
  cudaStream_t st = ctx->streams[stream];
  if (hash_kind == DLS_HASH_LCG) random_kernel<DLS_HASH_LCG><<<grid_for(ctx, n), kBlock, 0, st>>>(out, n, seed);
  else random_kernel<DLS_HASH_WANG><<<grid_for(ctx, n), kBlock, 0, st>>>(out, n, seed);
  DLS_CUDA(cudaGetLastError());
  ctx->stats.kernel_launches++;
  return DLS_OK;
}

int launch_random_normal(dls_ctx *ctx, int stream, float *out, size_t n_padded, uint32_t seed, int hash_kind) {
  if (n_padded == 0) return DLS_OK;
  cudaStream_t st = ctx->streams[stream];
  size_t pairs = n_padded / 2;
  float2 *o2 = reinterpret_cast<float2 *>(out);
  if (hash_kind == DLS_HASH_LCG) random_normal_kernel<DLS_HASH_LCG><<<grid_for(ctx, pairs), kBlock, 0, st>>>(o2, pairs, seed);
  else random_normal_kernel<DLS_HASH_WANG><<<grid_for(ctx, pairs), kBlock, 0, st>>>(o2, pairs, seed);
  DLS_CUDA(cudaGetLastError());
  ctx->stats.kernel_launches++;
  return DLS_OK;
}

// ---- N6: fused SGD step ----------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kBlock) sgd_kernel(float *__restrict__ w, const float *__restrict__ d, size_t n, float lr) {
  const size_t n4 = n / 4;
  float4 *w4 = reinterpret_cast<float4 *>(w);
  const float4 *d4 = reinterpret_cast<const float4 *>(d);
  for (size_t i = (size_t)blockIdx.x * kBlock + threadIdx.x; i < n4; i += (size_t)gridDim.x * kBlock) {
    float4 a = w4[i], g = __ldcs(d4 + i);
    a.x = __fsub_rn(a.x, __fmul_rn(lr, g.x));
    a.y = __fsub_rn(a.y, __fmul_rn(lr, g.y));
    a.z = __fsub_rn(a.z, __fmul_rn(lr, g.z));
    a.w = __fsub_rn(a.w, __fmul_rn(lr, g.w));
    w4[i] = a;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    size_t i = n4 * 4 + threadIdx.x;
    w[i] = __fsub_rn(w[i], __fmul_rn(lr, d[i]));
  }
}

int launch_sgd(dls_ctx *ctx, int stream, float *w, const float *d, size_t n, float lr) {
  if (n == 0) return DLS_OK;
  sgd_kernel<<<grid_for(ctx, (n + 3) / 4), kBlock, 0, ctx->streams[stream]>>>(w, d, n, lr);
  DLS_CUDA(cudaGetLastError());
  ctx->stats.kernel_launches++;
  return DLS_OK;
}

// ---- N6b: stateful optimizers, one pass over the weight, its state and the gradient -----------------------------------
// The reference's extension point is `TensorOptimizerApi.delta` (TW:155-161, Weights:46-51); it ships learning rate /
// L1 / L2 only for scalars (CumulativeFloatLayersSpec:18-47) and links Adagrad & co. from the README.  These kernels
// are the fused form of the two stateful rules every model uses.  Arithmetic: one IEEE fp32 rounding per written
// operation below, evaluated exactly in the written order (no FMA contraction: the file is built with -fmad=false),
// so the CPU restatement used by the tests reproduces them bit for bit.
struct OptArgs {
  float lr, a, b, c, d, e;   // momentum: a = mu, b = weight_decay, c = nesterov (0/1)
                             // adam:     a = beta1, b = beta2, c = eps, d = weight_decay, lr = lr / (1 - beta1^t), e = 1 / (1 - beta2^t)
};

__device__ __forceinline__ void momentum_one(float &w, float &v, float g, const OptArgs &o) {
  if (o.b != 0.f) g = __fadd_rn(g, __fmul_rn(o.b, w));            // g += wd * w
  v = __fadd_rn(__fmul_rn(o.a, v), g);                             // v = mu * v + g
  const float step = o.c != 0.f ? __fadd_rn(g, __fmul_rn(o.a, v)) : v;   // nesterov: g + mu * v
  w = __fsub_rn(w, __fmul_rn(o.lr, step));                         // w -= lr * step
}

__device__ __forceinline__ void adam_one(float &w, float &m, float &v, float g, const OptArgs &o) {
  if (o.d != 0.f) g = __fadd_rn(g, __fmul_rn(o.d, w));            // g += wd * w   (L2 form, as in the original paper's code)
  m = __fadd_rn(__fmul_rn(o.a, m), __fmul_rn(__fsub_rn(1.f, o.a), g));                   // m = b1 m + (1 - b1) g
  v = __fadd_rn(__fmul_rn(o.b, v), __fmul_rn(__fmul_rn(__fsub_rn(1.f, o.b), g), g));     // v = b2 v + ((1 - b2) g) g
  const float denom = __fadd_rn(__fsqrt_rn(__fmul_rn(v, o.e)), o.c);                     // sqrt(v / (1 - b2^t)) + eps
  w = __fsub_rn(w, __fdiv_rn(__fmul_rn(o.lr, m), denom));                                // w -= (lr / (1 - b1^t)) m / denom
}

template <int KIND>  // 0: momentum SGD, 1: Adam
__global__ void __launch_bounds__(kBlock) optimizer_kernel(float *__restrict__ w, float *__restrict__ s1, float *__restrict__ s2,
                                                           const float *__restrict__ grad, size_t n,
                                                           const OptArgs *__restrict__ hyper) {
  // hyper-parameters live in device memory: Adam's bias corrections change every step, and a captured step (graph.cu)
  // would otherwise replay the values of the step it was recorded on
  const OptArgs o = *hyper;
  const size_t n4 = n / 4;
  float4 *w4 = reinterpret_cast<float4 *>(w), *a4 = reinterpret_cast<float4 *>(s1), *b4 = reinterpret_cast<float4 *>(s2);
  const float4 *g4 = reinterpret_cast<const float4 *>(grad);
  for (size_t i = (size_t)blockIdx.x * kBlock + threadIdx.x; i < n4; i += (size_t)gridDim.x * kBlock) {
    float4 x = w4[i], p = a4[i], g = __ldcs(g4 + i);
    if (KIND == 0) {
      momentum_one(x.x, p.x, g.x, o);
      momentum_one(x.y, p.y, g.y, o);
      momentum_one(x.z, p.z, g.z, o);
      momentum_one(x.w, p.w, g.w, o);
    } else {
      float4 q = b4[i];
      adam_one(x.x, p.x, q.x, g.x, o);
      adam_one(x.y, p.y, q.y, g.y, o);
      adam_one(x.z, p.z, q.z, g.z, o);
      adam_one(x.w, p.w, q.w, g.w, o);
      b4[i] = q;
    }
    w4[i] = x;
    a4[i] = p;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const size_t i = n4 * 4 + threadIdx.x;
    float x = w[i], p = s1[i];
    if (KIND == 0) {
      momentum_one(x, p, grad[i], o);
    } else {
      float q = s2[i];
      adam_one(x, p, q, grad[i], o);
      s2[i] = q;
    }
    w[i] = x;
    s1[i] = p;
  }
}

int launch_optimizer(dls_ctx *ctx, int stream, int kind, float *w, float *s1, float *s2, const float *grad, size_t n,
                     const float *hyper) {
  if (n == 0) return DLS_OK;
  const OptArgs *o = reinterpret_cast<const OptArgs *>(hyper);
  const int grid = grid_for(ctx, (n + 3) / 4);
  if (kind == 0) optimizer_kernel<0><<<grid, kBlock, 0, ctx->streams[stream]>>>(w, s1, s2, grad, n, o);
  else optimizer_kernel<1><<<grid, kBlock, 0, ctx->streams[stream]>>>(w, s1, s2, grad, n, o);
  DLS_CUDA(cudaGetLastError());
  ctx->stats.kernel_launches++;
  return DLS_OK;
}

}  // namespace dls
