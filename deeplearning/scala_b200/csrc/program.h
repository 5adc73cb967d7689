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
  int32_t pad_;
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

}  // namespace dls
