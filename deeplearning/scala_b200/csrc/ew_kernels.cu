// ew_kernels.cu -- the two hand-written sm_100a kernels every closure tree lowers to (no JIT, no per-tree source):
//   N1 ew_vec_kernel    fused elementwise over contiguous / broadcast leaves, 128-bit coalesced loads and stores
//   N2 ew_gather_kernel general affine gather (translate / permute / split / scale views with padding), one
//                       launch coordinate per lane, bit-exact index arithmetic of KB:348-411
// Both run the same accumulator-machine program (program.h); control flow is uniform across the grid.
// Replaces the generated `jit_kernel` (KB:135-221) and its launch (Tensors:1203-1251).
#include "common.h"
#include "reduce.cuh"

namespace dls {

constexpr int kVecUnroll = 2;          // float4 tiles per thread per trip: 2 x 16 B x n_leaves in flight per lane
constexpr int kInlineParams = 128;

struct VLeafDev {
  const float *ptr;                    // parameter pointer + base offset
  long long ls[kMaxVecOuter];
  int32_t inner;
  int32_t streaming;
};

struct VecArgs {
  float *out;
  long long total_tiles;               // float4 tiles
  long long tiles_per_row;             // ceil(inner / 4)
  long long n_valid;                   // launch coordinates (tail masking in reduce mode)
  long long dims[kMaxVecOuter];
  int32_t v_outer;
  int32_t n_slots;
  int32_t reduce;
  int32_t stream_store;
  ReduceArgs red;
  VLeafDev leaves[kInlineVLeaves];
  Instr code[kInlineCode];
};

struct GatherArgs {
  float *out;
  const Instr *code;
  const GLeaf *leaves;
  const double *dpool;
  long long n;
  long long shape[kMaxRank];
  int32_t rank;
  int32_t n_roots;
  int32_t n_slots;
  int32_t reduce;
  ReduceArgs red;
  const float *params[kInlineParams];
};

// ---- arithmetic: one IEEE rounding per tree node, never contracted into FMA (oracle definition) -----------------
__device__ __forceinline__ float apply_unary(int op, float a) {
  switch (op) {
    case OP_NEG: return -a;
    case OP_ABS: return fabsf(a);
    case OP_SQRT: return __fsqrt_rn(a);
    case OP_TANH: return tanhf(a);
    case OP_EXP: return expf(a);
    default: return logf(a);
  }
}

__device__ __forceinline__ float apply_binary(int op, float a, float s) {
  switch (op) {
    case OP_LOAD: return s;
    case OP_ADD: return __fadd_rn(a, s);
    case OP_SUB: return __fsub_rn(a, s);
    case OP_MUL: return __fmul_rn(a, s);
    case OP_DIV: return __fdiv_rn(a, s);
    case OP_MOD: return fmodf(a, s);
    case OP_MIN: return fminf(a, s);
    case OP_MAX: return fmaxf(a, s);
    case OP_RSUB: return __fsub_rn(s, a);
    case OP_RDIV: return __fdiv_rn(s, a);
    default: return fmodf(s, a);  // OP_RMOD
  }
}

// ---- N1 ------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ float4 splat(float x) { return make_float4(x, x, x, x); }

#define DLS_V4(expr_x, expr_y, expr_z, expr_w) make_float4(expr_x, expr_y, expr_z, expr_w)

template <bool IDX32>
__global__ void __launch_bounds__(kBlock) ew_vec_kernel(const __grid_constant__ VecArgs A) {
  extern __shared__ float4 slot_mem[];  // [slot][unroll][thread]
  typedef typename std::conditional<IDX32, unsigned int, unsigned long long>::type idx_t;
  const idx_t total = (idx_t)A.total_tiles;
  const idx_t tpr = (idx_t)A.tiles_per_row;
  const idx_t nthreads = (idx_t)gridDim.x * kBlock;
  float tsum = 0.f;
  for (idx_t t0 = (idx_t)blockIdx.x * kBlock + threadIdx.x; t0 < total; t0 += nthreads * kVecUnroll) {
    bool valid[kVecUnroll];
    idx_t c4[kVecUnroll];
    long long oc[kVecUnroll][kMaxVecOuter];
    idx_t tile[kVecUnroll];
#pragma unroll
    for (int u = 0; u < kVecUnroll; ++u) {
      idx_t t = t0 + (idx_t)u * nthreads;
      tile[u] = t;
      valid[u] = t < total;
      idx_t row = 0;
      c4[u] = t;
      if (A.v_outer > 0) {
        row = t / tpr;
        c4[u] = t - row * tpr;
      }
      oc[u][0] = oc[u][1] = oc[u][2] = 0;
      if (A.v_outer == 1) {
        oc[u][0] = (long long)row;
      } else if (A.v_outer == 2) {
        idx_t q = row / (idx_t)A.dims[1];
        oc[u][1] = (long long)(row - q * (idx_t)A.dims[1]);
        oc[u][0] = (long long)q;
      } else if (A.v_outer == 3) {
        idx_t q = row / (idx_t)A.dims[2];
        oc[u][2] = (long long)(row - q * (idx_t)A.dims[2]);
        idx_t q2 = q / (idx_t)A.dims[1];
        oc[u][1] = (long long)(q - q2 * (idx_t)A.dims[1]);
        oc[u][0] = (long long)q2;
      }
    }
    float4 acc[kVecUnroll];
#pragma unroll
    for (int u = 0; u < kVecUnroll; ++u) acc[u] = splat(0.f);

    auto fetch_leaf = [&](unsigned int li, float4 *dst) {
      const VLeafDev &L = A.leaves[li];
#pragma unroll
      for (int u = 0; u < kVecUnroll; ++u) {
        dst[u] = splat(0.f);
        if (!valid[u]) continue;
        const float *p = L.ptr + oc[u][0] * L.ls[0] + oc[u][1] * L.ls[1] + oc[u][2] * L.ls[2];
        if (L.inner) {
          const float4 *q = reinterpret_cast<const float4 *>(p) + c4[u];
          dst[u] = L.streaming ? __ldcs(q) : __ldg(q);
        } else {
          dst[u] = splat(__ldg(p));
        }
      }
    };

    for (int pc = 0;; ++pc) {
      const Instr in = A.code[pc];
      const int op = in.op;
      if (op == OP_END) break;
      float4 s[kVecUnroll];
#pragma unroll
      for (int u = 0; u < kVecUnroll; ++u) s[u] = splat(0.f);
      if (in.src == SRC_LEAF) {
        fetch_leaf(in.operand, s);
      } else if (in.src == SRC_LIT) {
#pragma unroll
        for (int u = 0; u < kVecUnroll; ++u) s[u] = splat(__uint_as_float(in.operand));
      } else if (in.src == SRC_SLOT) {
#pragma unroll
        for (int u = 0; u < kVecUnroll; ++u) s[u] = slot_mem[(in.operand * kVecUnroll + u) * kBlock + threadIdx.x];
      }
#define DLS_EACH(EXPR)                                                                 \
  _Pragma("unroll") for (int u = 0; u < kVecUnroll; ++u) {                             \
    float4 r_;                                                                         \
    { const float a = acc[u].x, b = s[u].x; (void)a; (void)b; r_.x = (EXPR); }                           \
    { const float a = acc[u].y, b = s[u].y; (void)a; (void)b; r_.y = (EXPR); }                           \
    { const float a = acc[u].z, b = s[u].z; (void)a; (void)b; r_.z = (EXPR); }                           \
    { const float a = acc[u].w, b = s[u].w; (void)a; (void)b; r_.w = (EXPR); }                           \
    acc[u] = r_;                                                                       \
  }
      // the switch is on a grid-uniform value: one indirect branch per instruction, then straight-line lanes
      if (op <= OP_LOG) {
        switch (op) {
          case OP_LOAD: DLS_EACH(b) break;
          case OP_ADD: DLS_EACH(__fadd_rn(a, b)) break;
          case OP_SUB: DLS_EACH(__fsub_rn(a, b)) break;
          case OP_MUL: DLS_EACH(__fmul_rn(a, b)) break;
          case OP_DIV: DLS_EACH(__fdiv_rn(a, b)) break;
          case OP_MOD: DLS_EACH(fmodf(a, b)) break;
          case OP_MIN: DLS_EACH(fminf(a, b)) break;
          case OP_MAX: DLS_EACH(fmaxf(a, b)) break;
          case OP_RSUB: DLS_EACH(__fsub_rn(b, a)) break;
          case OP_RDIV: DLS_EACH(__fdiv_rn(b, a)) break;
          case OP_RMOD: DLS_EACH(fmodf(b, a)) break;
          case OP_NEG: DLS_EACH(-a) break;
          case OP_ABS: DLS_EACH(fabsf(a)) break;
          case OP_SQRT: DLS_EACH(__fsqrt_rn(a)) break;
          case OP_TANH: DLS_EACH(tanhf(a)) break;
          case OP_EXP: DLS_EACH(expf(a)) break;
          default: DLS_EACH(logf(a)) break;
        }
      } else if (op == OP_SAVE) {
#pragma unroll
        for (int u = 0; u < kVecUnroll; ++u) slot_mem[(in.operand * kVecUnroll + u) * kBlock + threadIdx.x] = acc[u];
      } else {  // OP_MAC / OP_MSUB
        float4 f[kVecUnroll];
        fetch_leaf(in.aux, f);
#pragma unroll
        for (int u = 0; u < kVecUnroll; ++u) {
          float4 p = DLS_V4(__fmul_rn(f[u].x, s[u].x), __fmul_rn(f[u].y, s[u].y), __fmul_rn(f[u].z, s[u].z),
                            __fmul_rn(f[u].w, s[u].w));
          acc[u] = op == OP_MAC ? DLS_V4(__fadd_rn(acc[u].x, p.x), __fadd_rn(acc[u].y, p.y), __fadd_rn(acc[u].z, p.z),
                                         __fadd_rn(acc[u].w, p.w))
                                : DLS_V4(__fsub_rn(acc[u].x, p.x), __fsub_rn(acc[u].y, p.y), __fsub_rn(acc[u].z, p.z),
                                         __fsub_rn(acc[u].w, p.w));
        }
      }
    }

#pragma unroll
    for (int u = 0; u < kVecUnroll; ++u) {
      if (!valid[u]) continue;
      if (A.reduce) {
        // only a rank-collapsed launch (v_outer == 0) can have a ragged last tile
        long long first = (long long)tile[u] * 4;
        long long left = A.v_outer == 0 ? A.n_valid - first : 4;
        float part = acc[u].x;
        if (left > 1) part += acc[u].y;
        if (left > 2) part += acc[u].z;
        if (left > 3) part += acc[u].w;
        tsum += part;
      } else {
        float4 *q = reinterpret_cast<float4 *>(A.out) + tile[u];
        if (A.stream_store) __stcs(q, acc[u]);
        else *q = acc[u];
      }
    }
  }
  if (A.reduce) grid_sum_finish(tsum, A.red);
}

// ---- N2 ------------------------------------------------------------------------------------------------------------
template <bool IDX32>
__device__ __forceinline__ float gather_leaf(const GatherArgs &A, unsigned int li, const long long *g) {
  const GLeaf *L = A.leaves + li;
  const int mode = L->mode;
  const float *p = A.params[L->param];
  if (mode == G_SCALAR) return __ldg(p + L->base);
  if (mode == G_STRIDED) {
    long long addr = L->base;
    for (int d = 0; d < A.rank; ++d) addr += g[d] * L->ls[d];
    return __ldg(p + addr);
  }
  const int rows = L->rows;
  bool ok = true;
  long long addr = 0;
  if (mode == G_MONO) {
    for (int y = 0; y < rows; ++y) {
      const GRow &R = L->row[y];
      long long idx = R.off;
      if (R.col >= 0) idx += g[R.col] * R.coef;
      if (R.checked) ok = ok && idx >= 0 && idx < R.extent;
      addr += idx * R.stride;
    }
  } else {  // G_GENERIC: `(int)(g0 * c0 + g1 * c1 + ... + offset)` evaluated left to right in double (KB:363-390)
    const double *M = A.dpool + L->mat_off;
    const int cols = A.rank + 1;
    for (int y = 0; y < rows; ++y) {
      const GRow &R = L->row[y];
      double s = 0.0;
      bool started = false;
      for (int x = 0; x < cols; ++x) {
        double c = M[y * cols + x];
        if (c == 0.0) continue;
        double term = x < A.rank ? (c == 1.0 ? (double)g[x] : __dmul_rn((double)g[x], c)) : c;
        s = started ? __dadd_rn(s, term) : term;
        started = true;
      }
      long long idx = (long long)s;  // truncation toward zero (KB:386)
      if (R.checked) ok = ok && idx >= 0 && idx < R.extent;
      addr += idx * R.stride;
    }
  }
  return ok ? __ldg(p + addr) : L->padding;
}

template <bool IDX32>
__global__ void __launch_bounds__(kBlock) ew_gather_kernel(const __grid_constant__ GatherArgs A) {
  extern __shared__ float slot_f[];  // [slot][thread]
  typedef typename std::conditional<IDX32, unsigned int, unsigned long long>::type idx_t;
  const idx_t n = (idx_t)A.n;
  const idx_t nthreads = (idx_t)gridDim.x * kBlock;
  float tsum = 0.f;
  for (idx_t e = (idx_t)blockIdx.x * kBlock + threadIdx.x; e < n; e += nthreads) {
    long long g[kMaxRank];
    {
      idx_t rem = e;
#pragma unroll
      for (int d = kMaxRank - 1; d >= 0; --d) {
        if (d < A.rank) {
          idx_t s = (idx_t)A.shape[d];
          idx_t q = rem / s;
          g[d] = (long long)(rem - q * s);
          rem = q;
        } else {
          g[d] = 0;
        }
      }
    }
    float acc = 0.f;
    for (int pc = 0;; ++pc) {
      Instr in;
      *reinterpret_cast<uint2 *>(&in) = __ldg(reinterpret_cast<const uint2 *>(A.code) + pc);
      const int op = in.op;
      if (op == OP_END) break;
      float s = 0.f;
      if (in.src == SRC_LEAF) s = gather_leaf<IDX32>(A, in.operand, g);
      else if (in.src == SRC_LIT) s = __uint_as_float(in.operand);
      else if (in.src == SRC_SLOT) s = slot_f[in.operand * kBlock + threadIdx.x];
      if (op <= OP_RMOD) acc = apply_binary(op, acc, s);
      else if (op <= OP_LOG) acc = apply_unary(op, acc);
      else if (op == OP_SAVE) slot_f[in.operand * kBlock + threadIdx.x] = acc;
      else if (op == OP_STORE) {
        if (A.reduce) tsum += acc;
        else A.out[(long long)e * A.n_roots + in.operand] = acc;
      } else {
        float f = gather_leaf<IDX32>(A, in.aux, g);
        float p = __fmul_rn(f, s);
        acc = op == OP_MAC ? __fadd_rn(acc, p) : __fsub_rn(acc, p);
      }
    }
    if (A.n_roots == 1) {
      if (A.reduce) tsum += acc;
      else A.out[e] = acc;
    }
  }
  if (A.reduce) grid_sum_finish(tsum, A.red);
}


// ---- host side -------------------------------------------------------------------------------------------------------
int query_ew_occupancy(dls_ctx *ctx) {
  int b = 0;
  DLS_CUDA(cudaFuncSetAttribute(ew_gather_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  DLS_CUDA(cudaFuncSetAttribute(ew_gather_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  DLS_CUDA(cudaFuncSetAttribute(ew_vec_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  DLS_CUDA(cudaFuncSetAttribute(ew_vec_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  DLS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, ew_vec_kernel<true>, kBlock, 0));
  ctx->vec_blocks_per_sm = b > 0 ? b : 1;
  DLS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, ew_gather_kernel<true>, kBlock, 0));
  ctx->gather_blocks_per_sm = b > 0 ? b : 1;
  return DLS_OK;
}

static int upload_tables(Plan &P) {
  if (P.d_tables || P.kind != PLAN_GATHER) return DLS_OK;
  size_t code_b = P.code.size() * sizeof(Instr);
  size_t leaf_b = P.gleaves.size() * sizeof(GLeaf);
  size_t pool_b = P.dpool.size() * sizeof(double);
  auto up = [](size_t x) { return (x + 255) & ~size_t(255); };
  P.d_code_off = 0;
  P.d_leaf_off = up(code_b);
  P.d_pool_off = P.d_leaf_off + up(leaf_b);
  P.d_tables_bytes = P.d_pool_off + up(pool_b) + 256;
  std::vector<uint8_t> host(P.d_tables_bytes, 0);
  memcpy(host.data() + P.d_code_off, P.code.data(), code_b);
  if (leaf_b) memcpy(host.data() + P.d_leaf_off, P.gleaves.data(), leaf_b);
  if (pool_b) memcpy(host.data() + P.d_pool_off, P.dpool.data(), pool_b);
  DLS_CUDA(cudaMalloc(&P.d_tables, P.d_tables_bytes));
  DLS_CUDA(cudaMemcpy(P.d_tables, host.data(), P.d_tables_bytes, cudaMemcpyHostToDevice));
  return DLS_OK;
}

int launch_plan(dls_ctx *ctx, int stream, CachedPlan &cp, dls_buf *const *params, dls_buf *out, bool reduce_sum) {
  Plan &P = cp.plan;
  cudaStream_t st = ctx->streams[stream];
  if (P.n_elems == 0) {
    if (reduce_sum) DLS_CUDA(cudaMemsetAsync(out->ptr, 0, sizeof(float), st));
    return DLS_OK;
  }
  ReduceArgs red{ctx->red_partials[stream], ctx->red_counter[stream], out->ptr};
  if (P.kind == PLAN_VEC) {
    static thread_local VecArgs A;
    A.out = out->ptr;
    A.tiles_per_row = (P.v_inner + 3) / 4;
    long long rows = 1;
    for (int j = 0; j < P.v_outer; ++j) rows *= P.v_dims[j];
    A.total_tiles = rows * A.tiles_per_row;
    A.n_valid = P.n_elems;
    for (int j = 0; j < kMaxVecOuter; ++j) A.dims[j] = P.v_dims[j];
    A.v_outer = P.v_outer;
    A.n_slots = P.n_slots;
    A.reduce = reduce_sum ? 1 : 0;
    A.stream_store = (size_t)P.n_elems * 4 > (size_t)64 << 20;
    A.red = red;
    for (size_t l = 0; l < P.vleaves.size(); ++l) {
      const VLeaf &V = P.vleaves[l];
      A.leaves[l].ptr = params[V.param]->ptr + V.base;
      for (int j = 0; j < kMaxVecOuter; ++j) A.leaves[l].ls[j] = V.ls[j];
      A.leaves[l].inner = V.inner;
      A.leaves[l].streaming = V.streaming;
    }
    memcpy(A.code, P.code.data(), P.code.size() * sizeof(Instr));
    size_t smem = (size_t)P.n_slots * kVecUnroll * kBlock * sizeof(float4);
    int per_sm = ctx->vec_blocks_per_sm;
    if (smem) {
      int b = 0;
      DLS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, ew_vec_kernel<true>, kBlock, smem));
      per_sm = b > 0 ? b : 1;
    }
    long long want = (A.total_tiles + (long long)kBlock * kVecUnroll - 1) / ((long long)kBlock * kVecUnroll);
    long long cap = (long long)ctx->sm_count * per_sm;
    if (reduce_sum && cap > kReduceMaxBlocks) cap = kReduceMaxBlocks;
    int grid = (int)std::min<long long>(want, cap);
    if (A.total_tiles < (1ll << 31) - (long long)cap * kBlock * kVecUnroll)
      ew_vec_kernel<true><<<grid, kBlock, smem, st>>>(A);
    else
      ew_vec_kernel<false><<<grid, kBlock, smem, st>>>(A);
  } else {
    DLS_CHECK(P.n_params <= kInlineParams, DLS_ERR_UNSUPPORTED, "more than 128 distinct parameters in one gather tree");
    int rc = upload_tables(P);
    if (rc) return rc;
    static thread_local GatherArgs A;
    A.out = out->ptr;
    A.code = reinterpret_cast<const Instr *>((const uint8_t *)P.d_tables + P.d_code_off);
    A.leaves = reinterpret_cast<const GLeaf *>((const uint8_t *)P.d_tables + P.d_leaf_off);
    A.dpool = reinterpret_cast<const double *>((const uint8_t *)P.d_tables + P.d_pool_off);
    A.n = P.n_elems;
    for (int d = 0; d < kMaxRank; ++d) A.shape[d] = d < P.out_rank ? P.out_shape[d] : 1;
    A.rank = P.out_rank;
    A.n_roots = P.n_roots;
    A.n_slots = P.n_slots;
    A.reduce = reduce_sum ? 1 : 0;
    A.red = red;
    for (int i = 0; i < P.n_params; ++i) A.params[i] = params[i]->ptr;
    size_t smem = (size_t)P.n_slots * kBlock * sizeof(float);
    int per_sm = ctx->gather_blocks_per_sm;
    if (smem) {
      int b = 0;
      DLS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, ew_gather_kernel<true>, kBlock, smem));
      per_sm = b > 0 ? b : 1;
    }
    long long want = (A.n + kBlock - 1) / kBlock;
    long long cap = (long long)ctx->sm_count * per_sm;
    if (reduce_sum && cap > kReduceMaxBlocks) cap = kReduceMaxBlocks;
    int grid = (int)std::min<long long>(want, cap);
    if (A.n < (1ll << 31) - (long long)cap * kBlock)
      ew_gather_kernel<true><<<grid, kBlock, smem, st>>>(A);
    else
      ew_gather_kernel<false><<<grid, kBlock, smem, st>>>(A);
  }
  DLS_CUDA(cudaGetLastError());
  ctx->stats.kernel_launches++;
  return DLS_OK;
}

}  // namespace dls
