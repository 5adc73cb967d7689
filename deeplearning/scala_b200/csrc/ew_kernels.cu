// ew_kernels.cu -- the two hand-written sm_100a kernels every closure tree lowers to (no JIT, no per-tree source):
//   N1 ew_vec_kernel    fused elementwise over contiguous / broadcast leaves, 128-bit coalesced loads and stores
//   N2 ew_gather_kernel general affine gather (translate / permute / split / scale views with padding), one
//                       launch coordinate per lane, bit-exact index arithmetic of KB:348-411
// Both run the same accumulator-machine program (program.h); control flow is uniform across the grid.
// Replaces the generated `jit_kernel` (KB:135-221) and its launch (Tensors:1203-1251).
#include "common.h"
#include "reduce.cuh"

namespace dls {

// float4 tiles per thread per trip (template parameter U of ew_vec_kernel): U x 16 B x n_leaves in flight per lane, and
// the interpreter's fetch / decode / dispatch cost per program instruction is shared by 4 * U elements
constexpr int kInlineParams = 128;

struct VLeafDev {
  const float *ptr;                    // parameter pointer + base offset
  uint32_t ls[kMaxVecOuter];           // element stride per collapsed outer dim (0 <= stride < 2^31)
  uint32_t flags;                      // bit 0: contiguous along the inner dim (128-bit loads), else broadcast along it
                                       // bit 1: read-once stream (evict-first loads)
};

// leaf reference kinds of the vector kernel, patched into Instr.src / Instr.aux by launch_plan (a plan stores SRC_LEAF)
enum VSrc : uint32_t { VSRC_DENSE = 4, VSRC_CONTIG = 5, VSRC_BCAST = 6 };

struct FastDiv {                       // n / d for n < 2^31: umulhi(n, mul) >> shr   (d == 1: mul == 0)
  uint32_t d, mul, shr;
};

struct VecArgs {
  float *out;
  uint32_t total_tiles;                // float4 tiles (< 2^31)
  uint32_t tail_valid;                 // valid elements in the last tile (1..4); only a flat launch can be ragged
  FastDiv tpr;                         // tiles per row = ceil(inner / 4)
  FastDiv dim1, dim2;                  // extents of collapsed outer dims 1 and 2 (OUTER == 3 only)
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
  const GLoop *loops;
  const long long *lpool;
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

__device__ __forceinline__ uint32_t fast_div(uint32_t n, const FastDiv &f) {
  return f.mul ? __umulhi(n, f.mul) >> f.shr : n;
}

#define DLS_EACH(EXPR)                                                                       \
  _Pragma("unroll") for (int u = 0; u < U; ++u) {                                   \
    float4 r_;                                                                               \
    { const float a = acc[u].x, b = s[u].x; (void)a; (void)b; r_.x = (EXPR); }               \
    { const float a = acc[u].y, b = s[u].y; (void)a; (void)b; r_.y = (EXPR); }               \
    { const float a = acc[u].z, b = s[u].z; (void)a; (void)b; r_.z = (EXPR); }               \
    { const float a = acc[u].w, b = s[u].w; (void)a; (void)b; r_.w = (EXPR); }               \
    acc[u] = r_;                                                                             \
  }

#define DLS_EACH1(EXPR)                                                                      \
  _Pragma("unroll") for (int u = 0; u < U; ++u) {                                   \
    float4 r_;                                                                               \
    { const float a = acc[u].x; r_.x = (EXPR); }                                             \
    { const float a = acc[u].y; r_.y = (EXPR); }                                             \
    { const float a = acc[u].z; r_.z = (EXPR); }                                             \
    { const float a = acc[u].w; r_.w = (EXPR); }                                             \
    acc[u] = r_;                                                                             \
  }

// OUTER = number of collapsed outer dims the launch has: 0 (flat), 1 (rows x inner) or 3 (up to three outer dims).
// Every lane owns U float4 tiles per trip, one grid stride apart (coalesced 128-bit accesses); tiles past the
// end are clamped to the last tile (recomputed, never stored) so that no load needs a predicate.
template <int OUTER, int U>
__global__ void __launch_bounds__(kBlock, U == 4 ? 3 : 1) ew_vec_kernel(const __grid_constant__ VecArgs A) {
  extern __shared__ float4 slot_mem[];  // [slot][unroll][thread]
  const uint32_t total = A.total_tiles;
  const uint32_t nthreads = gridDim.x * kBlock;
  float tsum = 0.f;
  for (uint32_t t0 = blockIdx.x * kBlock + threadIdx.x; t0 < total; t0 += nthreads * U) {
    uint32_t tile[U], c4[U], r0[U], r1[U], r2[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint32_t t = min(t0 + (uint32_t)u * nthreads, total - 1);
      tile[u] = t;
      r0[u] = r1[u] = r2[u] = 0;
      if (OUTER == 0) {
        c4[u] = t;
      } else {
        const uint32_t row = fast_div(t, A.tpr);
        c4[u] = t - row * A.tpr.d;
        if (OUTER == 1) {
          r0[u] = row;
        } else {
          const uint32_t q = fast_div(row, A.dim2);
          r2[u] = row - q * A.dim2.d;
          const uint32_t q2 = fast_div(q, A.dim1);
          r1[u] = q - q2 * A.dim1.d;
          r0[u] = q2;
        }
      }
    }
    float4 acc[U];

    // element offsets are 32-bit (the host only picks this kernel when every leaf spans < 2^32 floats).  The launch
    // code classifies every leaf reference (VSRC_*), so each addressing mode is its own grid-uniform branch:
    //   dense  -- same layout as the output: offset = tile * 4, one IMAD.WIDE + LDG.128 per tile
    //   contig -- contiguous along the inner dim, own outer strides
    //   bcast  -- constant along the inner dim: one scalar load per tile, splat
    auto fetch_dense = [&](uint32_t li, float4 *dst) {
      const float4 *p = reinterpret_cast<const float4 *>(A.leaves[li].ptr);
#pragma unroll
      for (int u = 0; u < U; ++u) dst[u] = __ldg(p + tile[u]);
    };
    auto fetch_contig = [&](uint32_t li, float4 *dst) {
      const VLeafDev &L = A.leaves[li];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        uint32_t off = c4[u] * 4u;
        if (OUTER >= 1) off += r0[u] * L.ls[0];
        if (OUTER == 3) off += r1[u] * L.ls[1] + r2[u] * L.ls[2];
        dst[u] = __ldg(reinterpret_cast<const float4 *>(L.ptr + off));
      }
    };
    auto fetch_bcast = [&](uint32_t li, float4 *dst) {
      const VLeafDev &L = A.leaves[li];
      __syncwarp();  // control flow is grid-uniform, so this is free; it keeps ptxas from if-converting this arm into
                     // the dense one (both arms' loads would then issue, predicated, on every fetch)
#pragma unroll
      for (int u = 0; u < U; ++u) {
        uint32_t off = 0u;
        if (OUTER >= 1) off += r0[u] * L.ls[0];
        if (OUTER == 3) off += r1[u] * L.ls[1] + r2[u] * L.ls[2];
        dst[u] = splat(__ldg(L.ptr + off));
      }
    };
    auto fetch_leaf = [&](uint32_t kind, uint32_t li, float4 *dst) {
      switch (kind) {
        case VSRC_DENSE: fetch_dense(li, dst); break;
        case VSRC_CONTIG: fetch_contig(li, dst); break;
        default: fetch_bcast(li, dst); break;
      }
    };

    for (int pc = 0;; ++pc) {
      // one 64-bit constant-bank load per instruction: {op:8, src:8, aux:16, operand:32}
      const uint2 w = reinterpret_cast<const uint2 *>(A.code)[pc];
      Instr in;
      in.op = (uint8_t)(w.x & 0xffu);
      in.src = (uint8_t)((w.x >> 8) & 0xffu);
      in.aux = (uint16_t)(w.x >> 16);
      in.operand = w.y;
      const int op = in.op;
      if (op == OP_END) break;
      if (op >= OP_NEG && op <= OP_LOG) {
        switch (op) {
          case OP_NEG: DLS_EACH1(-a) break;
          case OP_ABS: DLS_EACH1(fabsf(a)) break;
          case OP_SQRT: DLS_EACH1(__fsqrt_rn(a)) break;
          case OP_TANH: DLS_EACH1(tanhf(a)) break;
          case OP_EXP: DLS_EACH1(expf(a)) break;
          default: DLS_EACH1(logf(a)) break;
        }
        continue;
      }
      if (op == OP_SAVE) {
#pragma unroll
        for (int u = 0; u < U; ++u) slot_mem[(in.operand * U + u) * kBlock + threadIdx.x] = acc[u];
        continue;
      }
      float4 s[U];
      if (in.src >= VSRC_DENSE) {
        fetch_leaf(in.src, in.operand, s);
      } else if (in.src == SRC_LIT) {
#pragma unroll
        for (int u = 0; u < U; ++u) s[u] = splat(__uint_as_float(in.operand));
      } else {
#pragma unroll
        for (int u = 0; u < U; ++u) s[u] = slot_mem[(in.operand * U + u) * kBlock + threadIdx.x];
      }
      if (op >= OP_MAC) {  // OP_MAC / OP_MSUB: acc +- leaf[aux] * src, two roundings; aux = kind << 12 | leaf
        float4 f[U];
        fetch_leaf(in.aux >> 12, in.aux & 0xfffu, f);
#pragma unroll
        for (int u = 0; u < U; ++u)
          s[u] = make_float4(__fmul_rn(f[u].x, s[u].x), __fmul_rn(f[u].y, s[u].y), __fmul_rn(f[u].z, s[u].z),
                             __fmul_rn(f[u].w, s[u].w));
        if (op == OP_MAC) { DLS_EACH(__fadd_rn(a, b)) } else { DLS_EACH(__fsub_rn(a, b)) }
        continue;
      }
      switch (op) {  // grid-uniform: one branch per instruction, then straight-line lanes
        case OP_LOAD: DLS_EACH(__fadd_rn(b, -0.0f)) break;  // x + (-0) == x bit for bit; keeps src out of acc's registers
        case OP_ADD: DLS_EACH(__fadd_rn(a, b)) break;
        case OP_SUB: DLS_EACH(__fsub_rn(a, b)) break;
        case OP_MUL: DLS_EACH(__fmul_rn(a, b)) break;
        case OP_DIV: DLS_EACH(__fdiv_rn(a, b)) break;
        case OP_MOD: DLS_EACH(fmodf(a, b)) break;
        case OP_MIN: DLS_EACH(fminf(a, b)) break;
        case OP_MAX: DLS_EACH(fmaxf(a, b)) break;
        case OP_RSUB: DLS_EACH(__fsub_rn(b, a)) break;
        case OP_RDIV: DLS_EACH(__fdiv_rn(b, a)) break;
        default: DLS_EACH(fmodf(b, a)) break;  // OP_RMOD
      }
    }

    if (A.reduce && t0 + (uint32_t)(U - 1) * nthreads < total - 1) {
      // every tile of the trip is live and full: pairwise sum per tile
#pragma unroll
      for (int u = 0; u < U; ++u) tsum += (acc[u].x + acc[u].y) + (acc[u].z + acc[u].w);
      continue;
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const bool live = t0 + (uint32_t)u * nthreads < total;   // clamped duplicates contribute nothing
      if (A.reduce) {
        const uint32_t left = tile[u] == total - 1 ? A.tail_valid : 4u;
        const float hi = (left > 2 ? acc[u].z : 0.f) + (left > 3 ? acc[u].w : 0.f);
        const float lo = acc[u].x + (left > 1 ? acc[u].y : 0.f);
        if (live) tsum += lo + hi;
      } else if (live) {
        reinterpret_cast<float4 *>(A.out)[tile[u]] = acc[u];
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

// One leaf of a folded chain: the launch-coordinate part of its address is computed once, the per-term constant comes
// from the loop table (uniform loads: every lane is at the same term).
struct LoopLeaf {
  const float *p;
  const long long *tab;
  long long lin;                       // LL_LINEAR: sum_x g_x * ls_x
  long long gy[kLoopMonoRows];         // LL_MONO: g_col * coef per row
  long long extent[kLoopMonoRows], stride[kLoopMonoRows];
  unsigned checked;
  int rows, kind;
  float padding;
};

__device__ __forceinline__ void loop_leaf_init(LoopLeaf &S, const GatherArgs &A, const GLoop *G, int pos, const long long *g) {
  const GLeaf *L = A.leaves + G->leaf[pos];
  S.p = A.params[L->param];
  S.tab = A.lpool + G->table[pos];
  S.kind = G->leaf_kind[pos];
  S.padding = L->padding;
  S.rows = L->rows;
  S.lin = 0;
  S.checked = 0;
  if (S.kind == LL_LINEAR) {
    for (int d = 0; d < A.rank; ++d) S.lin += g[d] * L->ls[d];
  } else {
#pragma unroll
    for (int y = 0; y < kLoopMonoRows; ++y) {
      const GRow &R = L->row[y];
      const bool live = y < S.rows;
      S.gy[y] = live && R.col >= 0 ? g[R.col] * R.coef : 0;
      S.extent[y] = live ? R.extent : 1;
      S.stride[y] = live ? R.stride : 0;
      if (live && R.checked) S.checked |= 1u << y;
    }
  }
}

__device__ __forceinline__ float loop_leaf_at(const LoopLeaf &S, int t) {
  if (S.kind == LL_LINEAR) return __ldg(S.p + S.lin + __ldg(S.tab + t));
  bool ok = true;
  long long addr = 0;
  const long long *row = S.tab + (long long)t * S.rows;
#pragma unroll
  for (int y = 0; y < kLoopMonoRows; ++y) {
    if (y < S.rows) {
      const long long idx = S.gy[y] + __ldg(row + y);
      if (S.checked & (1u << y)) ok = ok && idx >= 0 && idx < S.extent[y];
      addr += idx * S.stride[y];
    }
  }
  return ok ? __ldg(S.p + addr) : S.padding;
}

__device__ __noinline__ float run_loop(const GatherArgs &A, const GLoop *G, const long long *g, float acc) {
  LoopLeaf a, b;
  loop_leaf_init(a, A, G, 0, g);
  const int n = G->count;
  if (G->kind == LOOP_MAC) {
    loop_leaf_init(b, A, G, 1, g);
    for (int t = 0; t < n; ++t) acc = __fadd_rn(acc, __fmul_rn(loop_leaf_at(a, t), loop_leaf_at(b, t)));
  } else {
    for (int t = 0; t < n; ++t) acc = __fadd_rn(acc, loop_leaf_at(a, t));
  }
  return acc;
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
      } else if (op == OP_LOOP) {
        acc = run_loop(A, A.loops + in.operand, g, acc);
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


// ---- N2 fast path: 32-bit affine gathers, V inner elements per lane -------------------------------------------------------
struct G4Args {
  float *out;
  uint32_t n_tiles;                    // lanes of work: rows * ceil(inner / V)
  int32_t rank;                        // launch rank (dims renumbered innermost-first)
  int32_t n_roots, n_slots, reduce;
  int32_t stage_roots;                 // n_roots > 1: collect a warp's outputs in shared memory, store them coalesced
  FastDiv tpr;                         // tiles per row
  FastDiv dims[kMaxRank];              // extents of dims 1 .. rank-1 (index 0 unused)
  ReduceArgs red;
  const float *params[kG4Params];
  G4Leaf leaves[kG4Leaves];
  Instr code[kG4Code];
};

// `vd`: the launch dim the V elements of a lane run along (0 = the output's inner dim; the tiled kernel passes the dim its
// leaves are contiguous along)
template <int V, int R>
__device__ __forceinline__ void g4_leaf(const G4Args &A, unsigned li, const int *g, float *v, const int vd = 0) {
  const G4Leaf &L = A.leaves[li];
  const float *p = A.params[L.param];
  int addr = L.base;
#pragma unroll
  for (int d = 0; d < R; ++d) addr += g[d] * L.ls[d];   // ls[d] == 0 and g[d] == 0 beyond the launch rank
  const int step = L.ls[vd];
  unsigned valid = (1u << V) - 1u;
  const int nc = L.n_checks;
#pragma unroll
  for (int c = 0; c < kG4Checks; ++c) {
    if (c < nc) {
      const G4Check C = L.chk[c];
      if (C.shift) {  // dyadic row: (int)(g * coef / 2^shift + off / 2^shift), truncation toward zero (KB:386)
        // along the innermost dim (col 0) the host launches V == 1, so g[0] is the element's own coordinate
        const int n = g[C.col] * C.coef + C.off;
        const int idx = n >= 0 ? n >> C.shift : -((-n) >> C.shift);
        if ((unsigned)idx >= (unsigned)C.extent) valid = 0u;
        addr += idx * C.stride;
      } else if (C.col != vd || V == 1) {
        const int idx = g[C.col] * C.coef + C.off;
        if ((unsigned)idx >= (unsigned)C.extent) valid = 0u;
      } else {
#pragma unroll
        for (int e = 0; e < V; ++e) {
          const int idx = (g[vd] + e) * C.coef + C.off;
          if ((unsigned)idx >= (unsigned)C.extent) valid &= ~(1u << e);
        }
      }
    }
  }
  if (V == 4 && step == 1 && valid == 15u && (addr & 3) == 0) {
    const float4 q = __ldg(reinterpret_cast<const float4 *>(p + addr));
    v[0] = q.x;
    v[1] = q.y;
    v[2] = q.z;
    v[3] = q.w;
    return;
  }
#pragma unroll
  for (int e = 0; e < V; ++e) v[e] = (valid >> e) & 1u ? __ldg(p + addr + e * step) : L.padding;
}

// U tiles per lane per trip, a grid stride apart: every interpreted instruction issues U independent leaf fetches (the
// interpreter cannot hoist loads across its own dispatch, so this is where memory-level parallelism comes from) and
// its fetch / decode / branch is shared by U * V elements.
// R: launch ranks this instantiation handles (4 covers NHWC tensors; 8 is the general case).
template <int V, int U, int R>
__global__ void __launch_bounds__(kBlock, 3) ew_gather4_kernel(const __grid_constant__ G4Args A) {
  extern __shared__ float g4_slots[];   // [slot][U][V][thread], then the root staging area [U][thread][V][root]
  float *stage = g4_slots + (size_t)A.n_slots * U * V * kBlock;
  float tsum = 0.f;
  const uint32_t stride = gridDim.x * kBlock;
  // with staged roots whole warps iterate together (the tail lanes of the last warp idle inside)
  const uint32_t t_end = A.stage_roots ? ((A.n_tiles + 31u) & ~31u) : A.n_tiles;
  for (uint32_t t0 = blockIdx.x * kBlock + threadIdx.x; t0 < t_end; t0 += stride * U) {
    bool live[U];
    int g[U][R];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const uint32_t t = t0 + (uint32_t)u * stride;
      live[u] = t < A.n_tiles;
      const uint32_t tc = live[u] ? t : 0u;          // dead tiles compute tile 0 (never stored): no predicated loads
      uint32_t row = fast_div(tc, A.tpr);
      g[u][0] = (int)(tc - row * A.tpr.d) * V;
#pragma unroll
      for (int d = 1; d < R; ++d) {                  // dims beyond the launch rank have extent 1: q == row, g == 0
        const uint32_t q = fast_div(row, A.dims[d]);
        g[u][d] = (int)(row - q * A.dims[d].d);
        row = q;
      }
    }
    float acc[U][V];
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int e = 0; e < V; ++e) acc[u][e] = 0.f;
    for (int pc = 0;; ++pc) {
      const Instr in = A.code[pc];
      const int op = in.op;
      if (op == OP_END) break;
      float s[U][V];
      if (in.src == SRC_LEAF) {
#pragma unroll
        for (int u = 0; u < U; ++u) g4_leaf<V, R>(A, in.operand, g[u], s[u]);
      } else if (in.src == SRC_LIT) {
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
          for (int e = 0; e < V; ++e) s[u][e] = __uint_as_float(in.operand);
      } else if (in.src == SRC_SLOT) {
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
          for (int e = 0; e < V; ++e) s[u][e] = g4_slots[((in.operand * U + u) * V + e) * kBlock + threadIdx.x];
      }
      if (op <= OP_RMOD) {
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
          for (int e = 0; e < V; ++e) acc[u][e] = apply_binary(op, acc[u][e], s[u][e]);
      } else if (op <= OP_LOG) {
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
          for (int e = 0; e < V; ++e) acc[u][e] = apply_unary(op, acc[u][e]);
      } else if (op == OP_SAVE) {
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
          for (int e = 0; e < V; ++e) g4_slots[((in.operand * U + u) * V + e) * kBlock + threadIdx.x] = acc[u][e];
      } else if (op == OP_STORE) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
          if (A.reduce) {
            if (live[u]) {
#pragma unroll
              for (int e = 0; e < V; ++e) tsum += acc[u][e];
            }
          } else if (A.stage_roots) {
#pragma unroll
            for (int e = 0; e < V; ++e) stage[((u * kBlock + threadIdx.x) * V + e) * A.n_roots + in.operand] = acc[u][e];
          } else if (live[u]) {
            const uint32_t t = t0 + (uint32_t)u * stride;
#pragma unroll
            for (int e = 0; e < V; ++e) A.out[(size_t)(t * V + e) * A.n_roots + in.operand] = acc[u][e];
          }
        }
      } else {  // OP_MAC / OP_MSUB
        float f[U][V];
#pragma unroll
        for (int u = 0; u < U; ++u) g4_leaf<V, R>(A, in.aux, g[u], f[u]);
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
          for (int e = 0; e < V; ++e) {
            const float pr = __fmul_rn(f[u][e], s[u][e]);
            acc[u][e] = op == OP_MAC ? __fadd_rn(acc[u][e], pr) : __fsub_rn(acc[u][e], pr);
          }
      }
    }
    if (A.n_roots == 1) {
#pragma unroll
      for (int u = 0; u < U; ++u) {
        if (!live[u]) continue;
        const uint32_t t = t0 + (uint32_t)u * stride;
        if (A.reduce) {
#pragma unroll
          for (int e = 0; e < V; ++e) tsum += acc[u][e];
        } else if (V == 4) {
          reinterpret_cast<float4 *>(A.out)[t] = make_float4(acc[u][0], acc[u][1], acc[u][2], acc[u][3]);
        } else {
          A.out[t] = acc[u][0];
        }
      }
    } else if (A.stage_roots) {
      // a warp's 32 * V coordinates x n_roots outputs (per unrolled tile) are one contiguous run of the output buffer
      __syncwarp();
      const uint32_t lane = threadIdx.x & 31u;
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const uint32_t tw = t0 - lane + (uint32_t)u * stride;          // first tile of this warp's run
        if (tw >= A.n_tiles) continue;
        const uint32_t lanes_live = min(32u, A.n_tiles - tw);
        const uint32_t count = lanes_live * V * (uint32_t)A.n_roots;
        const float *src = stage + (size_t)(u * kBlock + threadIdx.x - lane) * V * A.n_roots;
        float *dst = A.out + (size_t)tw * V * A.n_roots;
        if (V == 4) {
          for (uint32_t i = lane * 4u; i < count; i += 128u)
            *reinterpret_cast<float4 *>(dst + i) = *reinterpret_cast<const float4 *>(src + i);
        } else {
          for (uint32_t i = lane; i < count; i += 32u) dst[i] = src[i];
        }
      }
      __syncwarp();
    }
  }
  if (A.reduce) grid_sum_finish(tsum, A.red);
}

// ---- N2 tiled path: views whose unit-stride dim is NOT the output's inner dim (permute / transpose, Tensors:1008-1030;
// gather semantics KB:348-411 unchanged) ------------------------------------------------------------------------------
// With one lane per inner output element such a leaf is read with 4-byte gathers, a 32-byte sector each (measured: 0.26 of
// HBM peak, DRAM reads 2.2 x algorithmic).  Here a CTA owns a TI x 32 tile of output dims (0, vd), vd = the launch dim
// along which every leaf is contiguous: the program is evaluated with the LANES ALONG vd (each warp load is one 128-byte
// line of the leaf), the results cross a padded shared-memory tile, and are stored with the lanes along the output's
// inner dim (128-byte lines again).  Coordinates, bounds checks and arithmetic are the interpreter's own, per element.
struct G4TArgs {
  int32_t vd;                          // launch dim (innermost-first numbering) the lanes follow while loading
  uint32_t n_tiles;
  uint32_t dim0, dimv;                 // extents of dims 0 and vd
  FastDiv tiles_i, tiles_j;            // tile counts along dim 0 and dim vd
  uint32_t ostride[kMaxRank];          // output element stride per launch dim (ostride[0] == 1)
};

// One leaf for the U tile rows of a thread.  The rows differ only in their dim-0 coordinate (g[0] + u * istep), so the
// affine address and every bounds check that does not involve dim 0 are evaluated once; `live` masks rows / vectors that
// lie outside the launch (they read nothing).  No dyadic rows here (the host keeps those programs on the row path).
template <int V, int U, int R>
__device__ __forceinline__ void g4t_leaf(const G4Args &A, unsigned li, const int *g, int istep, unsigned live, float (*v)[V], const int vd) {
  const G4Leaf &L = A.leaves[li];
  const float *p = A.params[L.param];
  int addr = L.base;
#pragma unroll
  for (int d = 0; d < R; ++d) addr += g[d] * L.ls[d];
  const int step = L.ls[vd], stepi = L.ls[0] * istep;
  unsigned common = (1u << V) - 1u, rows = live;
  const int nc = L.n_checks;
#pragma unroll
  for (int c = 0; c < kG4Checks; ++c) {
    if (c < nc) {
      const G4Check C = L.chk[c];
      if (C.col == 0) {
#pragma unroll
        for (int u = 0; u < U; ++u) {
          const int idx = (g[0] + u * istep) * C.coef + C.off;
          if ((unsigned)idx >= (unsigned)C.extent) rows &= ~(1u << u);
        }
      } else if (C.col == vd && V > 1) {
#pragma unroll
        for (int e = 0; e < V; ++e) {
          const int idx = (g[C.col] + e) * C.coef + C.off;
          if ((unsigned)idx >= (unsigned)C.extent) common &= ~(1u << e);
        }
      } else {
        const int idx = g[C.col] * C.coef + C.off;
        if ((unsigned)idx >= (unsigned)C.extent) common = 0u;
      }
    }
  }
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const int a = addr + u * stepi;
    const unsigned valid = (rows >> u) & 1u ? common : 0u;
    if (V == 4 && step == 1 && valid == 15u && (a & 3) == 0) {
      const float4 q = __ldg(reinterpret_cast<const float4 *>(p + a));
      v[u][0] = q.x;
      v[u][V > 1 ? 1 : 0] = q.y;
      v[u][V > 2 ? 2 : 0] = q.z;
      v[u][V > 3 ? 3 : 0] = q.w;
    } else {
#pragma unroll
      for (int e = 0; e < V; ++e) v[u][e] = (valid >> e) & 1u ? __ldg(p + a + e * step) : L.padding;
    }
  }
}

// V elements per lane along vd, LJ lanes of a warp side by side along vd (tile width TJ = LJ * V; the other 32 / LJ lane
// groups take further tile rows), U tile rows per thread: tile = TI x TJ with TI = 8 * (32 / LJ) * U.
template <int V, int LJ, int U, int R>
__global__ void __launch_bounds__(kBlock, 4) ew_gather4t_kernel(const __grid_constant__ G4Args A, const __grid_constant__ G4TArgs T) {
  constexpr int RPW = 32 / LJ;                // tile rows a warp covers per unit
  constexpr int TI = 8 * RPW * U, TJ = LJ * V;
  constexpr int RS = TJ + (V == 4 ? 4 : 1);   // row stride of the tile: conflict-free both ways (V == 4 with the XOR below)
  static_assert(TI % 32 == 0 && TJ % 8 == 0, "tile shape");
  extern __shared__ float g4t_smem[];         // [slot][U][V][thread], then tile[TI][RS]
  float *tile = g4t_smem + (size_t)A.n_slots * U * V * kBlock;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int row0 = w * RPW + lane / LJ, colv = (lane % LJ) * V;   // this thread's first tile row / first tile column
  const int vd = T.vd;
  const uint32_t ostride_v = T.ostride[vd];
  for (uint32_t tid = blockIdx.x; tid < T.n_tiles; tid += gridDim.x) {
    int g[R];
    uint32_t rest = fast_div(tid, T.tiles_i);
    const uint32_t ib = tid - rest * T.tiles_i.d;
    uint32_t q = fast_div(rest, T.tiles_j);
    const uint32_t jb = rest - q * T.tiles_j.d;
    rest = q;
    uint32_t obase = 0;
    const uint32_t i0 = ib * TI + row0, j0 = jb * TJ + colv;
#pragma unroll
    for (int d = 1; d < R; ++d) {
      if (d == vd) {
        g[d] = (int)j0;
      } else {
        q = fast_div(rest, A.dims[d]);
        g[d] = (int)(rest - q * A.dims[d].d);
        rest = q;
        obase += (uint32_t)g[d] * T.ostride[d];
      }
    }
    g[0] = (int)i0;
    unsigned live = 0u;                                // dimv % V == 0: a vector is inside or outside as a whole
#pragma unroll
    for (int u = 0; u < U; ++u)
      if (i0 + 8 * RPW * u < T.dim0 && j0 < T.dimv) live |= 1u << u;
    float acc[U][V];
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
      for (int e = 0; e < V; ++e) acc[u][e] = 0.f;
    for (int pc = 0;; ++pc) {
      const Instr in = A.code[pc];
      const int op = in.op;
      if (op == OP_END) break;
      float s[U][V];
      if (in.src == SRC_LEAF) {
        g4t_leaf<V, U, R>(A, in.operand, g, 8 * RPW, live, s, vd);
      } else if (in.src == SRC_LIT) {
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
          for (int e = 0; e < V; ++e) s[u][e] = __uint_as_float(in.operand);
      } else if (in.src == SRC_SLOT) {
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
          for (int e = 0; e < V; ++e) s[u][e] = g4t_smem[((in.operand * U + u) * V + e) * kBlock + threadIdx.x];
      }
      if (op <= OP_RMOD) {
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
          for (int e = 0; e < V; ++e) acc[u][e] = apply_binary(op, acc[u][e], s[u][e]);
      } else if (op <= OP_LOG) {
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
          for (int e = 0; e < V; ++e) acc[u][e] = apply_unary(op, acc[u][e]);
      } else if (op == OP_SAVE) {
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
          for (int e = 0; e < V; ++e) g4t_smem[((in.operand * U + u) * V + e) * kBlock + threadIdx.x] = acc[u][e];
      } else if (op == OP_MAC || op == OP_MSUB) {
        float f[U][V];
        g4t_leaf<V, U, R>(A, in.aux, g, 8 * RPW, live, f, vd);
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
          for (int e = 0; e < V; ++e) {
            const float pr = __fmul_rn(f[u][e], s[u][e]);
            acc[u][e] = op == OP_MAC ? __fadd_rn(acc[u][e], pr) : __fsub_rn(acc[u][e], pr);
          }
      }                                                // OP_STORE: single root, stored below
    }
    // tile[i][j]: element (i, j) lives in column j ^ ((i >> 3) & 3) of row i (V == 4), so that both the 16-byte row writes
    // here and the column reads below touch 32 different banks
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int il = row0 + 8 * RPW * u;
      if (V == 4) {
        float a0 = acc[u][0], a1 = acc[u][V > 1 ? 1 : 0], a2 = acc[u][V > 2 ? 2 : 0], a3 = acc[u][V > 3 ? 3 : 0];
        const int sw = (il >> 3) & 3;                  // q[e ^ sw] = a[e]
        if (sw & 1) {
          float t0 = a0; a0 = a1; a1 = t0;
          float t1 = a2; a2 = a3; a3 = t1;
        }
        if (sw & 2) {
          float t0 = a0; a0 = a2; a2 = t0;
          float t1 = a1; a1 = a3; a3 = t1;
        }
        *reinterpret_cast<float4 *>(tile + il * RS + colv) = make_float4(a0, a1, a2, a3);
      } else {
        tile[il * RS + colv] = acc[u][0];
      }
    }
    __syncthreads();
    // store: row jj of the tile = TI consecutive inner elements of the output; lane (+ 32 h) = inner coordinate
#pragma unroll
    for (int h = 0; h < TI / 32; ++h) {
      const int ii = lane + 32 * h;
      const uint32_t i = ib * TI + ii;
      const int swz = V == 4 ? (ii >> 3) & 3 : 0;
      const float *trow = tile + ii * RS;
      float *optr = A.out + (size_t)obase + (size_t)(jb * TJ + w) * ostride_v + i;
      const bool full = ib * TI + TI <= T.dim0 && jb * TJ + TJ <= T.dimv;
      if (full) {
#pragma unroll 8
        for (int k = 0; k < TJ / 8; ++k) optr[(size_t)(8 * k) * ostride_v] = trow[(w + 8 * k) ^ swz];
      } else if (i < T.dim0) {
#pragma unroll 4
        for (int k = 0; k < TJ / 8; ++k)
          if (jb * TJ + w + 8 * k < T.dimv) optr[(size_t)(8 * k) * ostride_v] = trow[(w + 8 * k) ^ swz];
      }
    }
    __syncthreads();
  }
}

constexpr int kG4SmemLimit = 75 * 1024;   // three CTAs per SM: slots + root staging of the widest im2col (2 x 9 roots x 4 KB)

// ---- host side -------------------------------------------------------------------------------------------------------
int query_ew_occupancy(dls_ctx *ctx) {
  int b = 0;
  DLS_CUDA(cudaFuncSetAttribute(ew_gather_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
  DLS_CUDA(cudaFuncSetAttribute(ew_gather_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024));
#define DLS_G4_ATTR(VV, UU, RR) \
  DLS_CUDA(cudaFuncSetAttribute(ew_gather4_kernel<VV, UU, RR>, cudaFuncAttributeMaxDynamicSharedMemorySize, kG4SmemLimit));
  DLS_G4_ATTR(4, 1, 4) DLS_G4_ATTR(1, 1, 4) DLS_G4_ATTR(4, 2, 4) DLS_G4_ATTR(1, 2, 4)
  DLS_G4_ATTR(4, 1, 8) DLS_G4_ATTR(1, 1, 8) DLS_G4_ATTR(4, 2, 8) DLS_G4_ATTR(1, 2, 8)
#undef DLS_G4_ATTR
#define DLS_G4T_ATTR(VV, LL, UU, RR) \
  DLS_CUDA(cudaFuncSetAttribute(ew_gather4t_kernel<VV, LL, UU, RR>, cudaFuncAttributeMaxDynamicSharedMemorySize, kG4SmemLimit));
  DLS_G4T_ATTR(4, 32, 4, 4) DLS_G4T_ATTR(4, 32, 4, 8) DLS_G4T_ATTR(4, 16, 2, 4) DLS_G4T_ATTR(4, 16, 2, 8)
  DLS_G4T_ATTR(4, 8, 1, 4) DLS_G4T_ATTR(4, 8, 1, 8) DLS_G4T_ATTR(1, 32, 4, 4) DLS_G4T_ATTR(1, 32, 4, 8)
  DLS_G4T_ATTR(4, 16, 4, 4) DLS_G4T_ATTR(4, 16, 4, 8)
#undef DLS_G4T_ATTR
#define DLS_VEC_ATTR(O, UU) \
  DLS_CUDA(cudaFuncSetAttribute(ew_vec_kernel<O, UU>, cudaFuncAttributeMaxDynamicSharedMemorySize, 112 * 1024));
  DLS_VEC_ATTR(0, 1) DLS_VEC_ATTR(0, 2) DLS_VEC_ATTR(0, 4)
  DLS_VEC_ATTR(1, 1) DLS_VEC_ATTR(1, 2) DLS_VEC_ATTR(1, 4)
  DLS_VEC_ATTR(3, 1) DLS_VEC_ATTR(3, 2) DLS_VEC_ATTR(3, 4)
#undef DLS_VEC_ATTR
  DLS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, ew_vec_kernel<1, 2>, kBlock, 0));
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
  size_t loop_b = P.loops.size() * sizeof(GLoop);
  size_t lpool_b = P.lpool.size() * sizeof(long long);
  auto up = [](size_t x) { return (x + 255) & ~size_t(255); };
  P.d_code_off = 0;
  P.d_leaf_off = up(code_b);
  P.d_pool_off = P.d_leaf_off + up(leaf_b);
  P.d_loop_off = P.d_pool_off + up(pool_b);
  P.d_lpool_off = P.d_loop_off + up(loop_b);
  P.d_tables_bytes = P.d_lpool_off + up(lpool_b) + 256;
  std::vector<uint8_t> host(P.d_tables_bytes, 0);
  memcpy(host.data() + P.d_code_off, P.code.data(), code_b);
  if (leaf_b) memcpy(host.data() + P.d_leaf_off, P.gleaves.data(), leaf_b);
  if (pool_b) memcpy(host.data() + P.d_pool_off, P.dpool.data(), pool_b);
  if (loop_b) memcpy(host.data() + P.d_loop_off, P.loops.data(), loop_b);
  if (lpool_b) memcpy(host.data() + P.d_lpool_off, P.lpool.data(), lpool_b);
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
    auto mk_div = [](long long d) {
      FastDiv f;
      f.d = (uint32_t)d;
      if (d <= 1) {
        f.mul = 0;
        f.shr = 0;
      } else {
        int lg = 0;
        while ((1ll << lg) < d) ++lg;
        const int p = 31 + lg;
        f.mul = (uint32_t)((((unsigned long long)1 << p) + (unsigned long long)d - 1) / (unsigned long long)d);
        f.shr = (uint32_t)(p - 32);
      }
      return f;
    };
    A.out = out->ptr;
    const long long tpr = (P.v_inner + 3) / 4;
    long long rows = 1;
    for (int j = 0; j < P.v_outer; ++j) rows *= P.v_dims[j];
    const long long total_tiles = rows * tpr;
    A.total_tiles = (uint32_t)total_tiles;
    A.tail_valid = (uint32_t)(P.v_inner - (tpr - 1) * 4);
    A.tpr = mk_div(tpr);
    A.dim1 = mk_div(P.v_dims[1]);
    A.dim2 = mk_div(P.v_dims[2]);
    A.n_slots = P.n_slots;
    A.reduce = reduce_sum ? 1 : 0;
    A.stream_store = (size_t)P.n_elems * 4 > (size_t)64 << 20;
    A.red = red;
    for (size_t l = 0; l < P.vleaves.size(); ++l) {
      const VLeaf &V = P.vleaves[l];
      A.leaves[l].ptr = params[V.param]->ptr + V.base;
      for (int j = 0; j < kMaxVecOuter; ++j) A.leaves[l].ls[j] = (uint32_t)V.ls[j];
      A.leaves[l].flags = (V.inner ? 1u : 0u) | (V.streaming ? 2u : 0u);
    }
    memcpy(A.code, P.code.data(), P.code.size() * sizeof(Instr));
    // leaf references carry their addressing mode, so that the kernel branches once per instruction, grid-uniformly
    auto leaf_kind = [&](uint32_t li) -> uint32_t {
      const VLeaf &V = P.vleaves[li];
      return V.dense ? VSRC_DENSE : V.inner ? VSRC_CONTIG : VSRC_BCAST;
    };
    for (size_t i = 0; i < P.code.size(); ++i) {
      Instr &in = A.code[i];
      if (in.op == OP_MAC || in.op == OP_MSUB) in.aux = (uint16_t)(leaf_kind(in.aux) << 12 | in.aux);
      if (in.src == SRC_LEAF) in.src = (uint8_t)leaf_kind(in.operand);
    }
    // Unroll: 4 tiles per lane once the launch fills the machine several times over (the interpreter's per-instruction
    // cost is then shared by 16 elements), fewer for small launches (more CTAs) and slot-heavy programs (shared memory).
    static const int forced_unroll = [] {
      const char *e = getenv("DLS_VEC_UNROLL");
      return e ? atoi(e) : 0;
    }();
    int U = 1;
    const long long lanes = (long long)ctx->sm_count * kBlock;
    if (total_tiles >= lanes * 4) U = 2;
    if (total_tiles >= lanes * 32 && P.n_slots <= 3) U = 4;
    if (forced_unroll == 1 || forced_unroll == 2 || (forced_unroll == 4 && P.n_slots <= 6)) U = forced_unroll;
    size_t smem = (size_t)P.n_slots * U * kBlock * sizeof(float4);
    int per_sm = 0;
#define DLS_VEC_OCC(UU)                                                                                              \
  DLS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ew_vec_kernel<1, UU>, kBlock, smem))
    if (U == 4) DLS_VEC_OCC(4);
    else if (U == 2) DLS_VEC_OCC(2);
    else DLS_VEC_OCC(1);
#undef DLS_VEC_OCC
    if (per_sm < 1) per_sm = 1;
    long long want = (total_tiles + (long long)kBlock * U - 1) / ((long long)kBlock * U);
    long long cap = (long long)ctx->sm_count * per_sm;
    if (reduce_sum && cap > kReduceMaxBlocks) cap = kReduceMaxBlocks;
    int grid = (int)std::min<long long>(want, cap);
    // a 3-outer launch with leading extents of 1 was collapsed on the host: v_outer is 0, 1, 2 or 3
#define DLS_VEC_GO(O)                                                          \
  do {                                                                         \
    if (U == 4) ew_vec_kernel<O, 4><<<grid, kBlock, smem, st>>>(A);            \
    else if (U == 2) ew_vec_kernel<O, 2><<<grid, kBlock, smem, st>>>(A);       \
    else ew_vec_kernel<O, 1><<<grid, kBlock, smem, st>>>(A);                   \
  } while (0)
    if (P.v_outer == 0) DLS_VEC_GO(0);
    else if (P.v_outer == 1) DLS_VEC_GO(1);
    else DLS_VEC_GO(3);
#undef DLS_VEC_GO
  } else if (P.g4_ok) {
    static thread_local G4Args A;
    auto mk_div = [](long long d) {
      FastDiv f;
      f.d = (uint32_t)d;
      f.mul = 0;
      f.shr = 0;
      if (d > 1) {
        int lg = 0;
        while ((1ll << lg) < d) ++lg;
        const int p = 31 + lg;
        f.mul = (uint32_t)((((unsigned long long)1 << p) + (unsigned long long)d - 1) / (unsigned long long)d);
        f.shr = (uint32_t)(p - 32);
      }
      return f;
    };
    const int V = P.g4_vec, r = P.out_rank;
    const long long W = r ? P.out_shape[r - 1] : 1;
    const long long tpr = (W + V - 1) / V;          // V == 4 only when W % 4 == 0
    A.out = out->ptr;
    A.n_tiles = (uint32_t)(P.n_elems / (W ? W : 1) * tpr);
    A.rank = r;
    A.n_roots = P.n_roots;
    A.n_slots = P.n_slots;
    A.reduce = reduce_sum ? 1 : 0;
    A.tpr = mk_div(tpr);
    for (int d = 0; d < kMaxRank; ++d) A.dims[d] = mk_div(d >= 1 && d < r ? P.out_shape[r - 1 - d] : 1);
    A.red = red;
    for (int i = 0; i < P.n_params; ++i) A.params[i] = params[i]->ptr;
    memcpy(A.leaves, P.g4leaves.data(), P.g4leaves.size() * sizeof(G4Leaf));
    memcpy(A.code, P.code.data(), P.code.size() * sizeof(Instr));
    // tiled path: every leaf that depends on the inner output dim is strided along it and contiguous along ONE other launch
    // dim (a permute / transpose view, possibly under elementwise ops); single root, no fused reduction
    static const int tiled_off = [] {
      const char *e = getenv("DLS_G4_TILED");
      return e && atoi(e) == 0 ? 1 : 0;
    }();
    int vd = 0;
    if (!tiled_off && !reduce_sum && P.n_roots == 1 && r >= 2 && P.n_elems < (1ll << 31)) {
      bool ok = true, any = false;
      for (const G4Leaf &L : P.g4leaves) {
        bool scalar = true;
        for (int d = 0; d < r; ++d) scalar = scalar && L.ls[d] == 0;
        for (int c = 0; c < L.n_checks; ++c) scalar = scalar && L.chk[c].shift == 0;
        if (scalar) continue;                                   // a broadcast value: any lane order reads it equally well
        int unit = -1;
        for (int d = 1; d < r; ++d)
          if (L.ls[d] == 1) unit = d;
        for (int c = 0; c < L.n_checks; ++c) ok = ok && L.chk[c].shift == 0;   // dyadic rows carry their own address term
        if (L.ls[0] == 1 || L.ls[0] == 0 || unit < 0 || (vd && unit != vd)) ok = false;
        if (!ok) break;
        vd = unit;
        any = true;
      }
      if (!(ok && any && P.out_shape[r - 1 - vd] >= 32 && W >= 16)) vd = 0;
    }
    if (vd) {
      static thread_local G4TArgs T;
      // the two innermost launch dims fold into one when every leaf walks them as one (stride of dim 1 = extent x stride of
      // dim 0, no bounds check on either): NHWC -> NCHW is then a batched [HW, C] transpose with 1024-long output rows
      int rr = r;
      long long shape_in[kMaxRank];                    // launch extents, innermost first
      for (int d = 0; d < kMaxRank; ++d) shape_in[d] = d < r ? P.out_shape[r - 1 - d] : 1;
      while (rr >= 3 && vd >= 2) {
        bool fold = true;
        for (int l = 0; l < (int)P.g4leaves.size() && fold; ++l) {
          const G4Leaf &L = A.leaves[l];
          fold = (long long)L.ls[1] == (long long)L.ls[0] * shape_in[0];
          for (int c = 0; c < L.n_checks; ++c) fold = fold && L.chk[c].col > 1;
        }
        if (!fold || shape_in[0] * shape_in[1] >= (1ll << 30)) break;
        shape_in[0] *= shape_in[1];
        for (int d = 1; d + 1 < kMaxRank; ++d) shape_in[d] = shape_in[d + 1];
        shape_in[kMaxRank - 1] = 1;
        for (int l = 0; l < (int)P.g4leaves.size(); ++l) {
          G4Leaf &L = A.leaves[l];
          for (int d = 1; d + 1 < kMaxRank; ++d) L.ls[d] = L.ls[d + 1];
          L.ls[kMaxRank - 1] = 0;
          for (int c = 0; c < L.n_checks; ++c) L.chk[c].col -= 1;
        }
        --rr;
        --vd;
      }
      const long long W = shape_in[0];                 // shadows the row path's inner extent from here on
      const int r = rr;
      for (int d = 0; d < kMaxRank; ++d) A.dims[d] = mk_div(d >= 1 && d < r ? shape_in[d] : 1);
      T.vd = vd;
      T.dim0 = (uint32_t)W;
      T.dimv = (uint32_t)shape_in[vd];
      // 4 elements per lane along vd (128-bit loads) when that dim allows it: 32 x 128, 64 x 64 / 32 x 64 or 32 x 32 tiles
      // by its extent; else one element per lane and a 32 x 32 tile
      const bool v4 = T.dimv % 4 == 0;
      const int LJ = !v4 ? 32 : T.dimv >= 128 ? 32 : T.dimv >= 64 ? 16 : 8;
      const int Vt = v4 ? 4 : 1, Ut = !v4 ? 4 : LJ == 32 ? 4 : LJ == 16 ? (W >= 64 ? 4 : 2) : 1;
      const int TI = 8 * (32 / LJ) * Ut, TJ = LJ * Vt;
      const long long ti = (W + TI - 1) / TI, tj = (T.dimv + TJ - 1) / TJ;
      T.tiles_i = mk_div(ti);
      T.tiles_j = mk_div(tj);
      long long others = 1, os = 1;
      for (int d = 0; d < kMaxRank; ++d) {
        T.ostride[d] = d < r ? (uint32_t)os : 0u;
        if (d < r) {
          if (d != 0 && d != vd) others *= shape_in[d];
          os *= shape_in[d];
        }
      }
      T.n_tiles = (uint32_t)(ti * tj * others);
      const size_t smem = (size_t)P.n_slots * Ut * Vt * kBlock * sizeof(float) + (size_t)TI * (TJ + (Vt == 4 ? 4 : 1)) * sizeof(float);
      if (smem <= (size_t)kG4SmemLimit && ti * tj * others < (1ll << 31)) {
        int per_sm = 0;
#define DLS_G4T_DO(VV, LL, UU)                                                                                                   \
  do {                                                                                                                           \
    DLS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ew_gather4t_kernel<VV, LL, UU, 8>, kBlock, smem));           \
    if (per_sm < 1) per_sm = 1;                                                                                                  \
    const int grid = (int)std::max<long long>(1, std::min<long long>(T.n_tiles, (long long)ctx->sm_count * per_sm));             \
    if (r <= 4) ew_gather4t_kernel<VV, LL, UU, 4><<<grid, kBlock, smem, st>>>(A, T);                                              \
    else ew_gather4t_kernel<VV, LL, UU, 8><<<grid, kBlock, smem, st>>>(A, T);                                                     \
  } while (0)
        if (!v4) DLS_G4T_DO(1, 32, 4);
        else if (LJ == 32) DLS_G4T_DO(4, 32, 4);
        else if (LJ == 16 && Ut == 4) DLS_G4T_DO(4, 16, 4);
        else if (LJ == 16) DLS_G4T_DO(4, 16, 2);
        else DLS_G4T_DO(4, 8, 1);
#undef DLS_G4T_DO
        DLS_CUDA(cudaGetLastError());
        ctx->stats.kernel_launches++;
        return DLS_OK;
      }
      // not launched after all: undo the dim folding before the row path reads the arguments
      memcpy(A.leaves, P.g4leaves.data(), P.g4leaves.size() * sizeof(G4Leaf));
      for (int d = 0; d < kMaxRank; ++d) A.dims[d] = mk_div(d >= 1 && d < P.out_rank ? P.out_shape[P.out_rank - 1 - d] : 1);
    }
    static const int forced_g4_unroll = [] {
      const char *e = getenv("DLS_G4_UNROLL");
      return e ? atoi(e) : 0;
    }();
    // two tiles per lane once there is enough work for every SM to stay full; single-root programs only unless forced
    // (the root staging area of a join doubles with U and costs occupancy)
    int U = (long long)A.n_tiles >= (long long)ctx->sm_count * kBlock * 8 && P.n_roots == 1 ? 2 : 1;
    if (forced_g4_unroll == 1 || forced_g4_unroll == 2) U = forced_g4_unroll;
    size_t smem = 0, stage_bytes = 0;
    for (;; U = 1) {
      smem = (size_t)P.n_slots * U * V * kBlock * sizeof(float);
      stage_bytes = (size_t)U * kBlock * V * P.n_roots * sizeof(float);
      if (U == 1 || smem + (P.n_roots > 1 && !reduce_sum ? stage_bytes : 0) <= (size_t)kG4SmemLimit) break;
    }
    A.stage_roots = P.n_roots > 1 && !reduce_sum && smem + stage_bytes <= (size_t)kG4SmemLimit ? 1 : 0;
    if (A.stage_roots) smem += stage_bytes;
    int per_sm = 0;
#define DLS_G4_OCC(VV, UU) DLS_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ew_gather4_kernel<VV, UU, 8>, kBlock, smem))
    if (V == 4 && U == 2) DLS_G4_OCC(4, 2);
    else if (V == 4) DLS_G4_OCC(4, 1);
    else if (U == 2) DLS_G4_OCC(1, 2);
    else DLS_G4_OCC(1, 1);
#undef DLS_G4_OCC
    if (per_sm < 1) per_sm = 1;
    long long want = ((long long)A.n_tiles + (long long)kBlock * U - 1) / ((long long)kBlock * U);
    long long cap = (long long)ctx->sm_count * per_sm;
    if (reduce_sum && cap > kReduceMaxBlocks) cap = kReduceMaxBlocks;
    const int grid = (int)std::max<long long>(1, std::min<long long>(want, cap));
#define DLS_G4_GO(RR)                                                                   \
  do {                                                                                  \
    if (V == 4 && U == 2) ew_gather4_kernel<4, 2, RR><<<grid, kBlock, smem, st>>>(A);   \
    else if (V == 4) ew_gather4_kernel<4, 1, RR><<<grid, kBlock, smem, st>>>(A);        \
    else if (U == 2) ew_gather4_kernel<1, 2, RR><<<grid, kBlock, smem, st>>>(A);        \
    else ew_gather4_kernel<1, 1, RR><<<grid, kBlock, smem, st>>>(A);                    \
  } while (0)
    if (r <= 4) DLS_G4_GO(4);
    else DLS_G4_GO(8);
#undef DLS_G4_GO
  } else {
    DLS_CHECK(P.n_params <= kInlineParams, DLS_ERR_UNSUPPORTED, "more than 128 distinct parameters in one gather tree");
    int rc = upload_tables(P);
    if (rc) return rc;
    static thread_local GatherArgs A;
    A.out = out->ptr;
    A.code = reinterpret_cast<const Instr *>((const uint8_t *)P.d_tables + P.d_code_off);
    A.leaves = reinterpret_cast<const GLeaf *>((const uint8_t *)P.d_tables + P.d_leaf_off);
    A.dpool = reinterpret_cast<const double *>((const uint8_t *)P.d_tables + P.d_pool_off);
    A.loops = reinterpret_cast<const GLoop *>((const uint8_t *)P.d_tables + P.d_loop_off);
    A.lpool = reinterpret_cast<const long long *>((const uint8_t *)P.d_tables + P.d_lpool_off);
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
