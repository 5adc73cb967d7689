// plan.h -- host-side lowering of a serialised closure tree (wire format v1) into an interpreter program.
// Replaces the reference's per-tree code generation + clBuildProgram (KB:135-221, OpenCL:917-932): there is no JIT,
// a plan only selects one of the hand-written kernels and fills in its program / leaf tables.
#pragma once
#include <stdint.h>

#include <string>
#include <vector>

#include "program.h"

namespace dls {

// ---- wire format v1 (little endian u32 words; doubles as two words) ---------------------------------------------------
//   [0] magic 'DLST' (0x54534C44)  [1] version = 1  [2] n_params  [3] n_nodes  [4] n_roots  [5] out_rank
//   out_shape[out_rank]
//   per parameter : rank, shape[rank], padding (fp32 bits)                       (ArrayParameter, Trees:763-798)
//   per node      : opcode, then
//        W_LITERAL : fp32 bits                                                    (FloatLiteral,  Trees:309-330)
//        W_IDENT   : param                                                        (Extract(ArrayParameter), Tensors:1126-1132)
//        W_XFORM   : param, rows, cols, rows*cols f64 row-major                   (Extract(Transform(..)), Tensors:1296-1298)
//        unary     : child                                                        (Trees:371-470)
//        binary    : left, right                                                  (Trees:472-640)
//   roots[n_roots]  node ids; n_roots > 1 is a Concatenate written as the innermost output dim (Trees:953-973)
// Nodes are in topological order (children before parents); sharing is expressed by node id, parameters are numbered
// by first visit, so equal blobs <=> structurally equal trees in the sense of Trees:68-178 (plus the launch shape,
// which the reference's cache key forgets -- SURVEY appendix B).
constexpr uint32_t kWireMagic = 0x54534C44u;
constexpr uint32_t kWireVersion = 1;

enum WireOp : uint32_t {
  W_LITERAL = 1,
  W_IDENT = 2,
  W_XFORM = 3,
  W_NEG = 10,
  W_ABS = 11,
  W_SQRT = 12,
  W_TANH = 13,
  W_EXP = 14,
  W_LOG = 15,
  W_ADD = 20,
  W_SUB = 21,
  W_MUL = 22,
  W_DIV = 23,
  W_MOD = 24,
  W_MIN = 25,
  W_MAX = 26
};

struct ParamInfo {
  int rank = 0;
  long long shape[kMaxRank] = {0};
  float padding = 0.f;
  long long length() const {
    long long n = 1;
    for (int i = 0; i < rank; ++i) n *= shape[i];
    return n;
  }
};

enum PlanKind { PLAN_VEC = 0, PLAN_GATHER = 1 };

struct Plan {
  PlanKind kind = PLAN_GATHER;
  int out_rank = 0;
  long long out_shape[kMaxRank] = {0};
  long long n_elems = 1;          // number of launch coordinates (work-items in the reference)
  int n_roots = 1;                // floats written per coordinate
  int n_params = 0;
  int n_slots = 0;
  std::vector<ParamInfo> params;
  std::vector<Instr> code;        // terminated by OP_END
  // PLAN_VEC: collapsed geometry [v_dims[0..v_outer) , v_inner]
  int v_outer = 0;
  long long v_dims[kMaxVecOuter] = {1, 1, 1};
  long long v_inner = 1;
  std::vector<VLeaf> vleaves;
  // PLAN_GATHER
  std::vector<GLeaf> gleaves;
  std::vector<double> dpool;
  // PLAN_GATHER fast path (gather4.cu): filled when every leaf fits the 32-bit affine + checks form
  bool g4_ok = false;
  int g4_vec = 1;                 // inner elements per lane (4 when the innermost extent is a multiple of 4)
  std::vector<G4Leaf> g4leaves;
  std::vector<GLoop> loops;       // OP_LOOP descriptors
  std::vector<long long> lpool;   // their per-iteration offset tables
  // set when the whole launch is ONE dense contraction out[M,N] = A[M,K] * B[K,N] (+ nothing else): the runtime routes
  // it to the tcgen05 GEMM when the context's gemm_mode is a tensor-core mode (TensorsSpec:337-344 / :371-383 patterns)
  struct Contraction {
    bool valid = false;
    int param_a = 0, param_b = 0;
    long long m = 0, n = 0, k = 0;
    bool transpose_a = false, transpose_b = false;
  } contraction;
  // device-resident copy of code + leaves (uploaded by the runtime on first launch; owned by the plan cache)
  void *d_tables = nullptr;
  size_t d_code_off = 0, d_leaf_off = 0, d_pool_off = 0, d_loop_off = 0, d_lpool_off = 0, d_tables_bytes = 0;

  std::string describe() const;
};

// Returns 0 on success; on failure a dls_status code with a message in *err.
int lower_tree(const uint8_t *blob, size_t len, Plan *plan, std::string *err);

// KB:14-32 -- java.text.DecimalFormat("0.###", HALF_EVEN) applied to a matrix coefficient; returns the value of the
// printed literal and whether it is an integer literal (int arithmetic in the generated C).
double matrix_literal(double value, bool *is_int);

uint64_t hash_bytes(const uint8_t *p, size_t n);

}  // namespace dls
