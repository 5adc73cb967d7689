// program.h -- the lowered form of one Compute.scala closure tree, shared by the host-side lowering (plan.cpp) and
// the sm_100a interpreter kernels (ew_kernels.cu).
//
// The reference turns every distinct tree into OpenCL C source (KB:135-221) and compiles it.  Here a tree becomes a
// straight-line program for a uniform-control-flow accumulator machine: every lane of every warp executes the same
// instruction stream (no divergence), operands are leaves (gathers), literals or saved slots.
#pragma once
#include <stdint.h>

namespace dls {

constexpr int kMaxRank = 8;          // launch rank / parameter rank supported by the gather path
constexpr int kMaxVecOuter = 3;      // collapsed outer dims the vector kernel handles (plus one inner dim)
constexpr int kInlineCode = 192;     // instructions carried in kernel-parameter (constant) space
constexpr int kInlineVLeaves = 24;   // vector-kernel leaves carried in kernel-parameter space
constexpr int kMaxSlots = 64;        // saved sub-expression / evaluation-stack slots (live in shared memory)

enum Op : uint8_t {
  OP_END = 0,
  OP_LOAD,   // acc = src
  OP_ADD,    // acc = acc + src      (one IEEE rounding per node, never contracted: KB:516-571 semantics as pinned
  OP_SUB,    // acc = acc - src       by the oracle)
  OP_MUL,
  OP_DIV,
  OP_MOD,    // fmodf (KB:564-570 emits C `%`; see DESIGN.md "intentional divergences")
  OP_MIN,    // fminf (KB:271-280)
  OP_MAX,
  OP_RSUB,   // acc = src - acc
  OP_RDIV,
  OP_RMOD,
  OP_NEG,
  OP_ABS,
  OP_SQRT,
  OP_TANH,
  OP_EXP,
  OP_LOG,
  OP_SAVE,   // slot[operand] = acc
  OP_MAC,    // acc = acc + (leaf[aux] * src)   two roundings
  OP_MSUB,   // acc = acc - (leaf[aux] * src)
  OP_STORE,  // out[L * n_roots + operand] = acc   (Concatenate roots, KB:632-652)
  OP_LOOP,   // acc = (((acc + t_0) + t_1) + ...) over loops[operand]: a host-unrolled sum folded back into a device loop
  OP_COUNT_
};

enum Src : uint8_t { SRC_NONE = 0, SRC_LEAF = 1, SRC_LIT = 2, SRC_SLOT = 3 };

struct Instr {
  uint8_t op;
  uint8_t src;
  uint16_t aux;      // OP_MAC/OP_MSUB: index of the first factor leaf
  uint32_t operand;  // leaf index | fp32 literal bits | slot index | output column
};
static_assert(sizeof(Instr) == 8, "Instr must be 8 bytes");

// ---- vector kernel (N1) leaf: unchecked linear gather whose innermost stride is 0 or 1 ---------------------------
struct VLeaf {
  long long base;                 // element offset into the parameter buffer
  long long ls[kMaxVecOuter];     // element stride per collapsed outer dim
  int32_t inner;                  // 1: contiguous along the inner dim (128-bit loads); 0: broadcast along it
  int32_t param;                  // index into the parameter pointer table
  int32_t streaming;              // 1: footprint == output size, read once (evict-first loads)
  int32_t dense;                  // 1: same row-major layout as the (collapsed) output: element offset = 4 * tile index
};

// ---- gather kernel (N2) leaf ---------------------------------------------------------------------------------------
enum GMode : int32_t {
  G_SCALAR = 0,   // rank-0 parameter or all-zero matrix: *(p + base); never bounds checked (KB:351-356)
  G_STRIDED = 1,  // proven in-bounds: p[base + sum_x g_x * ls_x]
  G_MONO = 2,     // integer matrix, <= 1 non-zero coefficient per row, bounds checked per row (KB:382-408)
  G_GENERIC = 3   // any matrix: double arithmetic left to right, (int) truncation (KB:386), bounds checked
};

struct GRow {
  int32_t col;        // output dim feeding this row, -1 when the row is a constant
  int32_t checked;    // 0: `const int index = 0` row without bounds check (KB:383-384) or proven in range
  long long coef;
  long long off;
  long long extent;   // parameter shape along this row
  long long stride;   // parameter element stride along this row
};

struct GLeaf {
  int32_t mode;
  int32_t param;
  float padding;
  int32_t rows;
  long long base;
  long long ls[kMaxRank];
  GRow row[kMaxRank];
  int32_t mat_off;    // G_GENERIC: index of rows*(rank+1) doubles in the constant pool
  int32_t pad_;
};

// ---- N2 fast path (gather4.cu): every leaf is an affine address plus a few bounds checks, all in 32-bit ---------------
// Launch dims are renumbered innermost-first (dim 0 = the fastest-varying output dim); a lane owns V consecutive inner
// elements, so coordinates, outer-row checks and the leaf base address are computed once per V elements and contiguous
// leaves are read with one 128-bit load.
constexpr int kG4Leaves = 16;
constexpr int kG4Code = 128;
constexpr int kG4Checks = 4;
constexpr int kG4Params = 16;
struct G4Check {
  int32_t col, coef, off, extent;   // 0 <= g[col] * coef + off < extent, else the leaf reads its padding
  // shift > 0: a row with a dyadic-rational coefficient (`scale` up by 2, 4, 8: KB:363-390 evaluates
  // `(int)(g * 0.5 + off)` in double, exact for multiples of 1/8): index = trunc_toward_zero((g * coef + off) / 2^shift)
  // with coef and off pre-multiplied by 2^shift; the row's address term index * stride is added here, not through ls[]
  int32_t shift, stride;
};
struct G4Leaf {
  int32_t param, base, n_checks;
  float padding;
  int32_t ls[kMaxRank];             // element stride per (renumbered) launch dim
  G4Check chk[kG4Checks];
};

// ---- device loop over a host-unrolled chain (N2) ----------------------------------------------------------------------
// TensorLayers.sumAs (TL:124-162) and every user-level matmul / convolution (TensorsSpec:144-210, :337-393) arrive as
// `((t_0 + t_1) + t_2) + ...` with 10^2..10^6 terms that differ only in the constant column of their leaves' matrices.
// The lowering folds such a run back into ONE loop instruction: the term template (one or two leaves) plus a table of
// per-iteration offsets.  Evaluation order and roundings are exactly those of the unrolled chain (fmul, then fadd, term
// by term), so the result is bit-identical to it.
enum LoopKind : int32_t { LOOP_ADD = 0 /* acc + a_t */, LOOP_MAC = 1 /* acc + a_t * b_t */ };
enum LoopLeafKind : int32_t {
  LL_LINEAR = 0,  // scalar / strided leaves: address = sum_x g_x * ls_x + table[t]
  LL_MONO = 1     // bounds-checked mono rows (<= kLoopMonoRows rows): index_y = g_col * coef + table[t * rows + y]
};
constexpr int kLoopMonoRows = 4;
constexpr int kLoopMinTerms = 8;

struct GLoop {
  int32_t kind;          // LoopKind
  int32_t count;
  int32_t leaf[2];       // template leaves (gleaves index); leaf[1] unused for LOOP_ADD
  int32_t leaf_kind[2];  // LoopLeafKind
  int32_t pad_[2];
  long long table[2];    // offsets into the loop pool (in long long units)
};

}  // namespace dls
