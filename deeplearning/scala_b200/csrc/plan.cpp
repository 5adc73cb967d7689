// plan.cpp -- wire-format parser, leaf classification, dim collapsing and code emission.  Pure host C++ (no CUDA).
#include "plan.h"

#include <math.h>
#include <stdio.h>
#include <string.h>

#include <algorithm>
#include <map>
#include <sstream>

#include "../../../include/dls_b200.h"

namespace dls {

uint64_t hash_bytes(const uint8_t *p, size_t n) {
  // FNV-1a 64 with a final avalanche; the plan cache compares full blobs on hash equality.
  uint64_t h = 1469598103934665603ull;
  for (size_t i = 0; i < n; ++i) {
    h ^= p[i];
    h *= 1099511628211ull;
  }
  h ^= h >> 33;
  h *= 0xff51afd7ed558ccdull;
  h ^= h >> 33;
  return h;
}

double matrix_literal(double value, bool *is_int) {
  // KB:14-32: DecimalFormat with at most 3 fraction digits, RoundingMode.HALF_EVEN, applied to the exact binary value.
  // |value| * 1000 is exact in x87 80-bit long double (53 + 10 mantissa bits), rintl rounds ties to even.
  if (!(fabs(value) < 4503599627370496.0)) {  // >= 2^52, inf or nan: already integral (or not printable)
    *is_int = true;
    return value;
  }
  long double scaled = (long double)value * 1000.0L;
  long double q = rintl(scaled);
  long long qi = (long long)q;
  if (qi % 1000 == 0) {
    *is_int = true;
    return (double)(qi / 1000);
  }
  *is_int = false;
  return (double)qi / 1000.0;  // nearest double to the printed decimal literal, as the C compiler would parse it
}

namespace {

struct Node {
  uint32_t op = 0;
  int a = -1, b = -1;
  int param = -1;
  uint32_t lit = 0;
  int rows = 0, cols = 0;
  size_t mat = 0;  // offset into mats
  int refs = 0;
  int height = 0;
  int leaf = -1;   // index into the leaf table for W_IDENT / W_XFORM
};

struct Reader {
  const uint8_t *p;
  size_t n, pos = 0;
  bool ok = true;
  uint32_t u32() {
    if (pos + 4 > n) {
      ok = false;
      return 0;
    }
    uint32_t v;
    memcpy(&v, p + pos, 4);
    pos += 4;
    return v;
  }
  double f64() {
    if (pos + 8 > n) {
      ok = false;
      return 0;
    }
    double v;
    memcpy(&v, p + pos, 8);
    pos += 8;
    return v;
  }
};

inline bool is_unary(uint32_t op) { return op >= W_NEG && op <= W_LOG; }
inline bool is_binary(uint32_t op) { return op >= W_ADD && op <= W_MAX; }
inline bool is_leaf(uint32_t op) { return op == W_IDENT || op == W_XFORM; }

// One classified leaf before the kernel family is chosen.
struct LeafInfo {
  int mode = G_SCALAR;  // GMode
  int param = 0;
  float padding = 0.f;
  int rows = 0;
  long long base = 0;
  long long ls[kMaxRank] = {0};  // per launch dim (G_STRIDED / G_SCALAR: zeros)
  GRow row[kMaxRank];
  std::vector<double> mat;        // G_GENERIC: literal-rounded matrix rows*(r+1)
  bool mono_ok = false;           // every row is an integer row with <= 1 launch coordinate (row[] is filled in)
};

struct Lowering {
  Plan *plan;
  std::string *err;
  std::vector<Node> nodes;
  std::vector<double> mats;
  std::vector<int> roots;
  std::vector<LeafInfo> leaves;

  int fail(int code, const std::string &m) {
    *err = m;
    return code;
  }

  int parse(const uint8_t *blob, size_t len);
  int classify_leaf(Node &nd);
  int choose_kernel();
  int emit_all();
  int compress_loops();
  void detect_contraction();
  void build_g4();

  // ---- code emission state ----
  std::vector<int> slot_of;        // node -> permanent slot (shared sub-expressions), -1 if none
  std::vector<int> remaining;      // node -> references not yet emitted
  std::vector<char> slot_busy;
  int alloc_slot() {
    for (size_t i = 0; i < slot_busy.size(); ++i)
      if (!slot_busy[i]) {
        slot_busy[i] = 1;
        return (int)i;
      }
    slot_busy.push_back(1);
    return (int)slot_busy.size() - 1;
  }
  void free_slot(int s) { slot_busy[s] = 0; }
  void put(uint8_t op, uint8_t src = SRC_NONE, uint32_t operand = 0, uint16_t aux = 0) {
    Instr in;
    in.op = op;
    in.src = src;
    in.aux = aux;
    in.operand = operand;
    plan->code.push_back(in);
  }
  bool simple_now(int n) const {
    const Node &nd = nodes[n];
    return nd.op == W_LITERAL || is_leaf(nd.op) || slot_of[n] >= 0;
  }
  void src_of(int n, uint8_t *src, uint32_t *operand) {
    const Node &nd = nodes[n];
    if (slot_of[n] >= 0) {
      *src = SRC_SLOT;
      *operand = (uint32_t)slot_of[n];
    } else if (nd.op == W_LITERAL) {
      *src = SRC_LIT;
      *operand = nd.lit;
    } else {
      *src = SRC_LEAF;
      *operand = (uint32_t)nd.leaf;
    }
  }
  void consume(int n) {  // one reference to n has been emitted
    if (--remaining[n] == 0 && slot_of[n] >= 0) {
      free_slot(slot_of[n]);
      // keep slot_of[n] (no further references exist)
    }
  }
  bool mac_candidate(int mul, int *leaf_factor, int *other) const;
  int emit(int root, int depth);
};

int Lowering::parse(const uint8_t *blob, size_t len) {
  Reader r{blob, len};
  if (r.u32() != kWireMagic) return fail(DLS_ERR_INVALID, "tree blob: bad magic");
  if (r.u32() != kWireVersion) return fail(DLS_ERR_INVALID, "tree blob: unsupported wire version");
  uint32_t n_params = r.u32(), n_nodes = r.u32(), n_roots = r.u32(), out_rank = r.u32();
  if (!r.ok) return fail(DLS_ERR_INVALID, "tree blob: truncated header");
  if (out_rank > (uint32_t)kMaxRank) return fail(DLS_ERR_UNSUPPORTED, "launch rank > 8 is not supported");
  if (n_roots == 0 || n_roots > 65535) return fail(DLS_ERR_INVALID, "tree blob: bad root count");
  if (n_nodes == 0 || n_nodes > (1u << 26)) return fail(DLS_ERR_INVALID, "tree blob: bad node count");
  plan->out_rank = (int)out_rank;
  plan->n_roots = (int)n_roots;
  plan->n_params = (int)n_params;
  plan->n_elems = 1;
  for (uint32_t i = 0; i < out_rank; ++i) {
    plan->out_shape[i] = r.u32();
    plan->n_elems *= plan->out_shape[i];
  }
  plan->params.resize(n_params);
  for (uint32_t i = 0; i < n_params; ++i) {
    uint32_t rank = r.u32();
    if (!r.ok || rank > (uint32_t)kMaxRank) return fail(DLS_ERR_UNSUPPORTED, "parameter rank > 8 is not supported");
    plan->params[i].rank = (int)rank;
    for (uint32_t d = 0; d < rank; ++d) plan->params[i].shape[d] = r.u32();
    uint32_t bits = r.u32();
    memcpy(&plan->params[i].padding, &bits, 4);
  }
  nodes.resize(n_nodes);
  for (uint32_t i = 0; i < n_nodes; ++i) {
    Node &nd = nodes[i];
    nd.op = r.u32();
    if (!r.ok) return fail(DLS_ERR_INVALID, "tree blob: truncated node table");
    if (nd.op == W_LITERAL) {
      nd.lit = r.u32();
    } else if (nd.op == W_IDENT) {
      nd.param = (int)r.u32();
      if (nd.param < 0 || nd.param >= (int)n_params) return fail(DLS_ERR_INVALID, "tree blob: parameter index out of range");
    } else if (nd.op == W_XFORM) {
      nd.param = (int)r.u32();
      nd.rows = (int)r.u32();
      nd.cols = (int)r.u32();
      if (nd.param < 0 || nd.param >= (int)n_params) return fail(DLS_ERR_INVALID, "tree blob: parameter index out of range");
      if (nd.rows < 0 || nd.rows > kMaxRank || nd.cols < 1 || nd.cols > kMaxRank + 1)
        return fail(DLS_ERR_INVALID, "tree blob: bad transform matrix size");
      nd.mat = mats.size();
      for (int k = 0; k < nd.rows * nd.cols; ++k) mats.push_back(r.f64());
    } else if (is_unary(nd.op)) {
      nd.a = (int)r.u32();
      if (nd.a < 0 || nd.a >= (int)i) return fail(DLS_ERR_INVALID, "tree blob: child index not topological");
    } else if (is_binary(nd.op)) {
      nd.a = (int)r.u32();
      nd.b = (int)r.u32();
      if (nd.a < 0 || nd.a >= (int)i || nd.b < 0 || nd.b >= (int)i)
        return fail(DLS_ERR_INVALID, "tree blob: child index not topological");
    } else {
      return fail(DLS_ERR_INVALID, "tree blob: unknown opcode");
    }
    if (!r.ok) return fail(DLS_ERR_INVALID, "tree blob: truncated node");
    int ha = nd.a >= 0 ? nodes[nd.a].height : 0, hb = nd.b >= 0 ? nodes[nd.b].height : 0;
    nd.height = 1 + std::max(ha, hb);
  }
  roots.resize(n_roots);
  for (uint32_t i = 0; i < n_roots; ++i) {
    roots[i] = (int)r.u32();
    if (!r.ok || roots[i] < 0 || roots[i] >= (int)n_nodes) return fail(DLS_ERR_INVALID, "tree blob: bad root id");
  }
  if (r.pos != len) return fail(DLS_ERR_INVALID, "tree blob: trailing bytes");
  // reference counts over the part reachable from the roots (iterative, reverse topological sweep)
  std::vector<char> live(n_nodes, 0);
  for (int rt : roots) {
    live[rt] = 1;
    nodes[rt].refs++;
  }
  for (int i = (int)n_nodes - 1; i >= 0; --i) {
    if (!live[i]) continue;
    for (int c : {nodes[i].a, nodes[i].b})
      if (c >= 0) {
        live[c] = 1;
        nodes[c].refs++;
      }
  }
  // leaf table: dedupe identical gathers (same parameter, same matrix)
  std::map<std::string, int> seen;
  for (uint32_t i = 0; i < n_nodes; ++i) {
    Node &nd = nodes[i];
    if (!live[i] || !is_leaf(nd.op)) continue;
    std::string key((const char *)&nd.op, 4);
    key.append((const char *)&nd.param, 4);
    if (nd.op == W_XFORM) {
      key.append((const char *)&nd.rows, 4);
      key.append((const char *)&nd.cols, 4);
      key.append((const char *)(mats.data() + nd.mat), sizeof(double) * nd.rows * nd.cols);
    }
    auto it = seen.find(key);
    if (it != seen.end()) {
      nd.leaf = it->second;
      continue;
    }
    if (leaves.size() >= 65535) return fail(DLS_ERR_UNSUPPORTED, "more than 65535 distinct leaves in one tree");
    int rc = classify_leaf(nd);
    if (rc) return rc;
    nd.leaf = (int)leaves.size() - 1;
    seen.emplace(std::move(key), nd.leaf);
  }
  return 0;
}


int Lowering::classify_leaf(Node &nd) {
  const ParamInfo &P = plan->params[nd.param];
  const int q = P.rank, r = plan->out_rank;
  long long pst[kMaxRank];
  {
    long long acc = 1;
    for (int y = q - 1; y >= 0; --y) {
      pst[y] = acc;
      acc *= P.shape[y];
    }
  }
  LeafInfo L;
  L.param = nd.param;
  L.padding = P.padding;
  L.rows = q;
  for (int y = 0; y < kMaxRank; ++y) {
    L.row[y].col = -1;
    L.row[y].checked = 0;
    L.row[y].coef = 0;
    L.row[y].off = 0;
    L.row[y].extent = y < q ? P.shape[y] : 1;
    L.row[y].stride = y < q ? pst[y] : 0;
  }
  if (q == 0) {  // KB:351-356 / KB:436-451 with an empty shape: `*p`, never checked
    L.mode = G_SCALAR;
    L.mono_ok = true;
    leaves.push_back(L);
    return 0;
  }
  auto span = [&](int x) -> long long { return std::max<long long>(plan->out_shape[x] - 1, 0); };
  if (nd.op == W_IDENT) {
    if (q != r) return fail(DLS_ERR_INVALID, "identity leaf rank differs from the launch rank (KB:435-455 would not compile)");
    bool any_checked = false;
    for (int y = 0; y < q; ++y) {
      L.row[y].col = y;
      L.row[y].coef = 1;
      L.row[y].checked = plan->out_shape[y] > P.shape[y];  // only the upper bound is tested (KB:440-442)
      any_checked |= L.row[y].checked != 0;
      L.ls[y] = pst[y];
    }
    L.mode = any_checked ? G_MONO : G_STRIDED;
    L.mono_ok = true;
    leaves.push_back(L);
    return 0;
  }
  // W_XFORM (KB:348-411)
  if (nd.rows != q) return fail(DLS_ERR_INVALID, "transform matrix rows != parameter rank");
  if (nd.cols != r + 1) return fail(DLS_ERR_INVALID, "transform matrix columns != launch rank + 1");
  const double *M = mats.data() + nd.mat;
  bool all_int = true, all_mono = true, any_checked = false;
  std::vector<double> lit((size_t)q * (r + 1), 0.0);
  for (int y = 0; y < q; ++y) {
    bool row_int = true, has_term = false;
    int nz = 0, col = -1;
    for (int x = 0; x <= r; ++x) {
      double raw = M[y * (r + 1) + x];
      if (raw == 0.0) continue;  // KB:366 `if scale == 0.0` is tested on the raw value, before formatting
      has_term = true;
      bool is_int;
      double v = matrix_literal(raw, &is_int);
      lit[(size_t)y * (r + 1) + x] = v;
      row_int &= is_int;
      if (x < r && v != 0.0) {
        ++nz;
        col = x;
      }
    }
    GRow &R = L.row[y];
    if (!has_term) {  // `const int index = 0;` no bounds check (KB:383-384)
      R.checked = 0;
      continue;
    }
    if (!row_int) {
      all_int = false;
      R.checked = 1;
      any_checked = true;
      continue;
    }
    long long lo = (long long)lit[(size_t)y * (r + 1) + r], hi = lo;
    for (int x = 0; x < r; ++x) {
      long long c = (long long)lit[(size_t)y * (r + 1) + x];
      if (c > 0) hi += c * span(x);
      if (c < 0) lo += c * span(x);
    }
    R.off = (long long)lit[(size_t)y * (r + 1) + r];
    R.checked = !(lo >= 0 && hi < P.shape[y]);
    any_checked |= R.checked != 0;
    if (nz <= 1) {
      R.col = nz ? col : -1;
      R.coef = nz ? (long long)lit[(size_t)y * (r + 1) + col] : 0;
    } else {
      all_mono = false;
    }
  }
  L.mono_ok = all_int && all_mono;
  if (all_int && !any_checked) {
    L.mode = G_STRIDED;
    for (int y = 0; y < q; ++y) {
      L.base += (long long)lit[(size_t)y * (r + 1) + r] * pst[y];
      for (int x = 0; x < r; ++x) L.ls[x] += (long long)lit[(size_t)y * (r + 1) + x] * pst[y];
    }
  } else if (all_int && all_mono) {
    L.mode = G_MONO;
  } else {
    L.mode = G_GENERIC;
    L.mat = lit;
  }
  leaves.push_back(L);
  return 0;
}


bool Lowering::mac_candidate(int mul, int *leaf_factor, int *other) const {
  const Node &m = nodes[mul];
  if (m.op != W_MUL || m.refs != 1) return false;
  const Node &a = nodes[m.a], &b = nodes[m.b];
  auto plain = [&](const Node &x) { return x.op == W_LITERAL || is_leaf(x.op); };
  if (is_leaf(a.op) && plain(b)) {
    *leaf_factor = m.a;
    *other = m.b;
    return true;
  }
  if (is_leaf(b.op) && plain(a)) {
    *leaf_factor = m.b;
    *other = m.a;
    return true;
  }
  return false;
}

static uint8_t forward_op(uint32_t w) {
  switch (w) {
    case W_ADD: return OP_ADD;
    case W_SUB: return OP_SUB;
    case W_MUL: return OP_MUL;
    case W_DIV: return OP_DIV;
    case W_MOD: return OP_MOD;
    case W_MIN: return OP_MIN;
    case W_MAX: return OP_MAX;
    case W_NEG: return OP_NEG;
    case W_ABS: return OP_ABS;
    case W_SQRT: return OP_SQRT;
    case W_TANH: return OP_TANH;
    case W_EXP: return OP_EXP;
    case W_LOG: return OP_LOG;
  }
  return OP_END;
}
static uint8_t reverse_op(uint32_t w) {  // acc = src OP acc
  switch (w) {
    case W_SUB: return OP_RSUB;
    case W_DIV: return OP_RDIV;
    case W_MOD: return OP_RMOD;
    default: return forward_op(w);  // + * fmin fmax commute bit-for-bit (up to NaN payloads)
  }
}

// Emits code that leaves the value of `root` in the accumulator and consumes one reference to it.
// The taller operand chain is walked iteratively (host-unrolled sums are 10^2..10^6 deep, TL:124-162); only the
// shorter side of a node whose operands are both compound recurses, so the C stack depth stays O(log nodes).
int Lowering::emit(int root, int depth) {
  if (depth > 256) return fail(DLS_ERR_UNSUPPORTED, "expression nesting too deep for the interpreter");
  enum Kind { UN, BIN_A, BIN_B, MAC, MSUB };
  struct Pending {
    int node;
    Kind kind;
    int leaf_factor, other, mul;
  };
  std::vector<Pending> st;
  int cur = root;
  for (;;) {
    const Node &nd = nodes[cur];
    if (simple_now(cur)) {
      uint8_t src;
      uint32_t operand;
      src_of(cur, &src, &operand);
      put(OP_LOAD, src, operand);
      consume(cur);
      break;
    }
    if (is_unary(nd.op)) {
      st.push_back({cur, UN, -1, -1, -1});
      cur = nd.a;
      continue;
    }
    int lf, ot;
    if (nd.op == W_ADD && mac_candidate(nd.b, &lf, &ot)) {
      st.push_back({cur, MAC, lf, ot, nd.b});
      cur = nd.a;
      continue;
    }
    if (nd.op == W_ADD && mac_candidate(nd.a, &lf, &ot)) {
      st.push_back({cur, MAC, lf, ot, nd.a});
      cur = nd.b;
      continue;
    }
    if (nd.op == W_SUB && mac_candidate(nd.b, &lf, &ot)) {
      st.push_back({cur, MSUB, lf, ot, nd.b});
      cur = nd.a;
      continue;
    }
    bool as = simple_now(nd.a), bs = simple_now(nd.b);
    bool spine_a;
    if (bs) spine_a = true;
    else if (as) spine_a = false;
    else spine_a = nodes[nd.a].height >= nodes[nd.b].height;
    st.push_back({cur, spine_a ? BIN_A : BIN_B, -1, -1, -1});
    cur = spine_a ? nd.a : nd.b;
  }
  while (!st.empty()) {
    Pending P = st.back();
    st.pop_back();
    const Node &nd = nodes[P.node];
    uint8_t src;
    uint32_t operand;
    switch (P.kind) {
      case UN:
        put(forward_op(nd.op));
        break;
      case MAC:
      case MSUB:
        src_of(P.other, &src, &operand);
        put(P.kind == MAC ? OP_MAC : OP_MSUB, src, operand, (uint16_t)nodes[P.leaf_factor].leaf);
        consume(P.other);
        consume(P.leaf_factor);
        consume(P.mul);
        break;
      case BIN_A:
      case BIN_B: {
        int side = P.kind == BIN_A ? nd.b : nd.a;
        if (simple_now(side)) {
          src_of(side, &src, &operand);
          put(P.kind == BIN_A ? forward_op(nd.op) : reverse_op(nd.op), src, operand);
          consume(side);
        } else {
          int s = alloc_slot();
          put(OP_SAVE, SRC_NONE, (uint32_t)s);
          int rc = emit(side, depth + 1);
          if (rc) return rc;
          // acc = side, slot = spine value
          put(P.kind == BIN_A ? reverse_op(nd.op) : forward_op(nd.op), SRC_SLOT, (uint32_t)s);
          free_slot(s);
        }
        break;
      }
    }
    // the value of P.node is now in the accumulator
    if (nd.refs > 1 && slot_of[P.node] < 0 && remaining[P.node] > 1) {
      int s = alloc_slot();
      put(OP_SAVE, SRC_NONE, (uint32_t)s);
      slot_of[P.node] = s;
    }
    consume(P.node);
  }
  return 0;
}

int Lowering::emit_all() {
  slot_of.assign(nodes.size(), -1);
  remaining.resize(nodes.size());
  for (size_t i = 0; i < nodes.size(); ++i) remaining[i] = nodes[i].refs;
  for (size_t k = 0; k < roots.size(); ++k) {
    int rc = emit(roots[k], 0);
    if (rc) return rc;
    if (roots.size() > 1) put(OP_STORE, SRC_NONE, (uint32_t)k);
  }
  put(OP_END);
  plan->n_slots = (int)slot_busy.size();
  if (plan->n_slots > kMaxSlots) return fail(DLS_ERR_UNSUPPORTED, "expression needs more than 64 live temporaries");
  return 0;
}


// ---- folding host-unrolled chains back into device loops ------------------------------------------------------------------
// After emission a chain `((t0 + t1) + t2) + ...` is a run of OP_MAC (t = leaf * leaf) or OP_ADD (t = leaf) instructions whose
// leaves differ only in constant offsets.  A run of >= kLoopMinTerms compatible terms becomes one OP_LOOP.
namespace {

bool linear_mode(const LeafInfo &L) { return L.mode == G_SCALAR || L.mode == G_STRIDED; }

bool linear_compatible(const LeafInfo &A, const LeafInfo &B) {
  if (A.param != B.param || !linear_mode(A) || !linear_mode(B)) return false;
  for (int d = 0; d < kMaxRank; ++d)
    if (A.ls[d] != B.ls[d]) return false;
  return true;
}

bool mono_compatible(const LeafInfo &A, const LeafInfo &B) {
  if (A.param != B.param || !A.mono_ok || !B.mono_ok || A.rows != B.rows || A.rows > kLoopMonoRows || A.rows == 0) return false;
  if (A.mode == G_GENERIC || B.mode == G_GENERIC) return false;
  for (int y = 0; y < A.rows; ++y) {
    const GRow &a = A.row[y], &b = B.row[y];
    const long long ca = a.col >= 0 ? a.coef : 0, cb = b.col >= 0 ? b.coef : 0;
    if (ca != cb || (ca != 0 && a.col != b.col) || a.extent != b.extent || a.stride != b.stride) return false;
  }
  return true;
}

}  // namespace

int Lowering::compress_loops() {
  std::vector<Instr> &code = plan->code;
  std::vector<Instr> out;
  out.reserve(code.size());
  size_t i = 0;
  while (i < code.size()) {
    const Instr in = code[i];
    const bool mac = in.op == OP_MAC && in.src == SRC_LEAF;
    const bool add = in.op == OP_ADD && in.src == SRC_LEAF;
    if (!mac && !add) {
      out.push_back(in);
      ++i;
      continue;
    }
    size_t j = i;
    while (j < code.size() && code[j].op == in.op && code[j].src == SRC_LEAF) ++j;
    const int n_pos = mac ? 2 : 1;
    auto leaf_id = [&](size_t t, int pos) -> int { return mac && pos == 0 ? (int)code[t].aux : (int)code[t].operand; };
    size_t best = j - i;
    int kind_of[2] = {LL_LINEAR, LL_LINEAR};
    for (int pos = 0; pos < n_pos && best >= (size_t)kLoopMinTerms; ++pos) {
      const LeafInfo &L0 = leaves[leaf_id(i, pos)];
      size_t lin = 0, mono = 0;
      while (lin < best && linear_compatible(L0, leaves[leaf_id(i + lin, pos)])) ++lin;
      while (mono < best && mono_compatible(L0, leaves[leaf_id(i + mono, pos)])) ++mono;
      kind_of[pos] = lin >= mono ? LL_LINEAR : LL_MONO;
      best = std::min(best, std::max(lin, mono));
    }
    if (best < (size_t)kLoopMinTerms || best > 0x7fffffffu) {
      out.push_back(in);
      ++i;
      continue;
    }
    // a leaf position may have been classed LINEAR on a longer prefix than the run finally taken: re-check the class
    for (int pos = 0; pos < n_pos; ++pos) {
      const LeafInfo &L0 = leaves[leaf_id(i, pos)];
      bool lin = true;
      for (size_t t = 0; t < best && lin; ++t) lin = linear_compatible(L0, leaves[leaf_id(i + t, pos)]);
      kind_of[pos] = lin ? LL_LINEAR : LL_MONO;
    }
    GLoop G;
    memset(&G, 0, sizeof(G));
    G.kind = mac ? LOOP_MAC : LOOP_ADD;
    G.count = (int32_t)best;
    for (int pos = 0; pos < n_pos; ++pos) {
      LeafInfo T = leaves[leaf_id(i, pos)];   // template = the first term's leaf
      G.leaf_kind[pos] = kind_of[pos];
      G.table[pos] = (long long)plan->lpool.size();
      if (kind_of[pos] == LL_LINEAR) {
        for (size_t t = 0; t < best; ++t) plan->lpool.push_back(leaves[leaf_id(i + t, pos)].base);
        T.base = 0;
        if (T.mode == G_SCALAR) T.mode = G_STRIDED;   // a rank-0 parameter: address = table[t] = 0
      } else {
        T.mode = G_MONO;
        for (size_t t = 0; t < best; ++t) {
          const LeafInfo &Lt = leaves[leaf_id(i + t, pos)];
          for (int y = 0; y < T.rows; ++y) {
            plan->lpool.push_back(Lt.row[y].off);
            if (Lt.row[y].checked) T.row[y].checked = 1;   // checking a row that cannot fail is harmless
          }
        }
        for (int y = 0; y < T.rows; ++y) {
          T.row[y].off = 0;
          if (T.row[y].col < 0) T.row[y].coef = 0;
        }
      }
      leaves.push_back(T);
      G.leaf[pos] = (int32_t)leaves.size() - 1;
    }
    Instr L;
    L.op = OP_LOOP;
    L.src = SRC_NONE;
    L.aux = 0;
    L.operand = (uint32_t)plan->loops.size();
    plan->loops.push_back(G);
    out.push_back(L);
    i += best;
  }
  if (plan->loops.empty()) return 0;
  code.swap(out);
  // drop the leaves only the folded terms referenced, renumber the rest
  std::vector<int> remap(leaves.size(), -1);
  std::vector<LeafInfo> kept;
  auto use = [&](int l) -> int {
    if (remap[l] < 0) {
      remap[l] = (int)kept.size();
      kept.push_back(leaves[l]);
    }
    return remap[l];
  };
  for (Instr &in : code) {
    if (in.op == OP_MAC || in.op == OP_MSUB) in.aux = (uint16_t)use(in.aux);
    if (in.src == SRC_LEAF) in.operand = (uint32_t)use((int)in.operand);
  }
  for (GLoop &G : plan->loops)
    for (int pos = 0; pos < (G.kind == LOOP_MAC ? 2 : 1); ++pos) G.leaf[pos] = use(G.leaf[pos]);
  leaves.swap(kept);
  return 0;
}

// A launch that is nothing but out[M, N] = sum_t A[.., t] * B[t, ..] over dense row-major parameters: the two matmul
// formulations of the reference (TensorsSpec:337-344 single root over [i, k]; :371-383 one root per output column).
void Lowering::detect_contraction() {
  Plan &P = *plan;
  const int R = P.n_roots;
  if (P.loops.size() != (size_t)R) return;
  const size_t per_root = R == 1 ? 3 : 4;             // LOAD, MUL, LOOP [, STORE]
  if (P.code.size() != per_root * R + 1) return;
  long long K = 0, a0 = 0, sa = 0, b0 = 0, sb = 0, nb = 0;
  int la = -1, lb = -1;
  for (int c = 0; c < R; ++c) {
    const Instr *I = P.code.data() + per_root * c;
    if (I[0].op != OP_LOAD || I[0].src != SRC_LEAF || I[1].op != OP_MUL || I[1].src != SRC_LEAF || I[2].op != OP_LOOP) return;
    if (R > 1 && (I[3].op != OP_STORE || (int)I[3].operand != c)) return;
    const GLoop &G = P.loops[I[2].operand];
    if (G.kind != LOOP_MAC || G.leaf_kind[0] != LL_LINEAR || G.leaf_kind[1] != LL_LINEAR) return;
    // the first term sits outside the loop as LOAD x; MUL y -- either order of the factors
    const LeafInfo *F0 = &leaves[I[0].operand], *F1 = &leaves[I[1].operand];
    const LeafInfo &TA = leaves[G.leaf[0]], &TB = leaves[G.leaf[1]];
    if (!linear_compatible(*F0, TA)) std::swap(F0, F1);
    if (!linear_compatible(*F0, TA) || !linear_compatible(*F1, TB)) return;
    const long long *ta = P.lpool.data() + G.table[0], *tb = P.lpool.data() + G.table[1];
    const long long k = (long long)G.count + 1;
    const long long sa_c = ta[0] - F0->base, sb_c = tb[0] - F1->base;
    for (long long t = 1; t < G.count; ++t)
      if (ta[t] - ta[t - 1] != sa_c || tb[t] - tb[t - 1] != sb_c) return;
    if (c == 0) {
      K = k;
      a0 = F0->base;
      sa = sa_c;
      b0 = F1->base;
      sb = sb_c;
      la = G.leaf[0];
      lb = G.leaf[1];
    } else {
      if (k != K || sa_c != sa || sb_c != sb || F0->base != a0) return;
      if (!linear_compatible(TA, leaves[la]) || !linear_compatible(TB, leaves[lb])) return;
      if (c == 1) nb = F1->base - b0;
      if (F1->base != b0 + nb * c) return;
    }
  }
  const LeafInfo &A = leaves[la], &B = leaves[lb];
  // launch dims: a prefix walks A (rows of the output), the rest walks B (columns); with several roots the root index is
  // the column and every launch dim is a row dim
  const int r = P.out_rank;
  int split = r;
  if (R == 1) {
    split = 0;
    while (split < r && (B.ls[split] == 0 || P.out_shape[split] == 1)) ++split;
    for (int d = split; d < r; ++d)
      if (A.ls[d] != 0 && P.out_shape[d] != 1) return;
  } else {
    for (int d = 0; d < r; ++d)
      if (B.ls[d] != 0 && P.out_shape[d] != 1) return;
  }
  auto collapse = [&](const LeafInfo &L, int d0, int d1, long long *extent, long long *stride) -> bool {
    long long ext = 1, st = 0;
    bool have = false;
    for (int d = d1 - 1; d >= d0; --d) {
      if (P.out_shape[d] == 1) continue;
      if (!have) {
        st = L.ls[d];
        have = true;
      } else if (L.ls[d] != st * ext) {
        return false;
      }
      ext *= P.out_shape[d];
    }
    *extent = ext;
    *stride = have ? st : 0;
    return true;
  };
  long long M, N, rs_a, cs_b;
  if (!collapse(A, 0, split, &M, &rs_a)) return;
  if (R == 1) {
    if (!collapse(B, split, r, &N, &cs_b)) return;
  } else {
    N = R;
    cs_b = nb;
  }
  if (M <= 0 || N <= 0 || a0 != 0 || b0 != 0) return;
  const ParamInfo &PA = P.params[A.param], &PB = P.params[B.param];
  if (PA.length() != M * K || PB.length() != K * N) return;
  bool ta, tb;
  if ((sa == 1 || K == 1) && (rs_a == K || M == 1)) ta = false;
  else if ((rs_a == 1 || M == 1) && (sa == M || K == 1)) ta = true;
  else return;
  if ((cs_b == 1 || N == 1) && (sb == N || K == 1)) tb = false;
  else if ((sb == 1 || K == 1) && (cs_b == K || N == 1)) tb = true;
  else return;
  P.contraction.valid = true;
  P.contraction.param_a = A.param;
  P.contraction.param_b = B.param;
  P.contraction.m = M;
  P.contraction.n = N;
  P.contraction.k = K;
  P.contraction.transpose_a = ta;
  P.contraction.transpose_b = tb;
}

int Lowering::choose_kernel() {
  Plan &P = *plan;
  const int r = P.out_rank;
  bool vec_ok = P.n_roots == 1 && P.n_elems > 0 && (int)leaves.size() <= kInlineVLeaves &&
                (int)P.code.size() <= kInlineCode && P.n_slots <= 6 && P.loops.empty();
  for (const LeafInfo &L : leaves)
    if (L.mode != G_SCALAR && L.mode != G_STRIDED) vec_ok = false;
  if (vec_ok) {
    // collapse the launch shape: drop extent-1 dims, merge neighbours that every leaf walks contiguously
    std::vector<long long> ext;
    std::vector<std::vector<long long>> st(leaves.size());
    for (int d = 0; d < r; ++d) {
      if (P.out_shape[d] == 1) continue;
      ext.push_back(P.out_shape[d]);
      for (size_t l = 0; l < leaves.size(); ++l) st[l].push_back(leaves[l].mode == G_STRIDED ? leaves[l].ls[d] : 0);
    }
    for (int j = (int)ext.size() - 2; j >= 0; --j) {
      bool ok = true;
      for (size_t l = 0; l < leaves.size(); ++l) ok &= st[l][j] == st[l][j + 1] * ext[j + 1];
      if (!ok) continue;
      ext[j] *= ext[j + 1];
      ext.erase(ext.begin() + j + 1);
      for (size_t l = 0; l < leaves.size(); ++l) {
        st[l][j] = st[l][j + 1];
        st[l].erase(st[l].begin() + j + 1);
      }
    }
    const int k = (int)ext.size();
    if (k > kMaxVecOuter + 1) vec_ok = false;
    const long long W = k ? ext[k - 1] : 1;
    if (vec_ok && k > 1 && (W % 4) != 0) vec_ok = false;
    if (vec_ok && (P.n_elems / (W ? W : 1)) * ((W + 3) / 4) >= (1ll << 31) - (1ll << 24)) vec_ok = false;  // 32-bit tile ids
    if (vec_ok) {
      P.vleaves.clear();
      for (size_t l = 0; l < leaves.size() && vec_ok; ++l) {
        VLeaf V;
        memset(&V, 0, sizeof(V));
        V.base = leaves[l].base;
        V.param = leaves[l].param;
        long long inner = k ? st[l][k - 1] : 0;
        if (inner != 0 && inner != 1) {
          vec_ok = false;
          break;
        }
        V.inner = (int32_t)inner;
        V.streaming = V.inner;
        for (int j = 0; j + 1 < k; ++j) {
          V.ls[j] = st[l][j];
          if (V.ls[j] == 0) V.streaming = 0;
          if (V.inner && (V.ls[j] % 4) != 0) vec_ok = false;
          if (V.ls[j] < 0 || V.ls[j] >= (1ll << 31)) vec_ok = false;  // the vector kernel uses 32-bit strides
        }
        if (V.inner && (V.base % 4) != 0) vec_ok = false;
        {  // 32-bit element offsets inside the kernel: the farthest element a leaf touches must stay below 2^32
          long long far = V.inner ? W : 0;
          for (int j = 0; j + 1 < k; ++j) far += (ext[j] - 1) * V.ls[j];
          if (far >= (1ll << 32)) vec_ok = false;
        }
        P.vleaves.push_back(V);
      }
    }
    if (vec_ok) {
      P.kind = PLAN_VEC;
      P.v_outer = k > 0 ? k - 1 : 0;
      for (int j = 0; j < kMaxVecOuter; ++j) P.v_dims[j] = j < P.v_outer ? ext[j] : 1;
      P.v_inner = W;
      for (VLeaf &V : P.vleaves) {  // dense: laid out exactly like the output, so a tile's offset is tile * 4
        bool dense = V.inner == 1;
        long long expect = W;
        for (int j = P.v_outer - 1; j >= 0 && dense; --j) {
          dense = V.ls[j] == expect;
          expect *= P.v_dims[j];
        }
        V.dense = dense ? 1 : 0;
      }
      return 0;
    }
    P.vleaves.clear();
  }
  P.kind = PLAN_GATHER;
  for (const LeafInfo &L : leaves) {
    GLeaf G;
    memset(&G, 0, sizeof(G));
    G.mode = L.mode;
    G.param = L.param;
    G.padding = L.padding;
    G.rows = L.rows;
    G.base = L.base;
    for (int d = 0; d < kMaxRank; ++d) {
      G.ls[d] = L.ls[d];
      G.row[d] = L.row[d];
    }
    G.mat_off = -1;
    if (L.mode == G_GENERIC) {
      G.mat_off = (int32_t)P.dpool.size();
      P.dpool.insert(P.dpool.end(), L.mat.begin(), L.mat.end());
    }
    P.gleaves.push_back(G);
  }
  return 0;
}

// The 32-bit affine form of a gather plan for the fast N2 kernel: launch dims renumbered innermost-first.
void Lowering::build_g4() {
  Plan &P = *plan;
  P.g4_ok = false;
  if (P.kind != PLAN_GATHER || !P.loops.empty() || leaves.empty() || (int)leaves.size() > kG4Leaves ||
      (int)P.code.size() > kG4Code || P.n_params > kG4Params || P.n_elems <= 0 ||
      P.n_elems * (long long)P.n_roots >= (1ll << 31) - (1ll << 24))
    return;
  for (const ParamInfo &pi : P.params)
    if (pi.length() >= (1ll << 31)) return;
  const int r = P.out_rank;
  auto renum = [&](int d) { return r - 1 - d; };
  const long long lim = (1ll << 31) - 1;
  std::vector<G4Leaf> out;
  bool inner_scaled = false;           // a dyadic row along the innermost launch dim: one element per lane
  for (const LeafInfo &L : leaves) {
    G4Leaf G;
    memset(&G, 0, sizeof(G));
    G.param = L.param;
    G.padding = L.padding;
    long long scaled_reach = 0;
    long long base = 0, ls[kMaxRank] = {0};
    if (L.mode == G_GENERIC) {
      // rows are integer single-term (as in G_MONO) or single-term with coefficient and offset multiples of 1/8:
      // the double expression of KB:363-390 is then exact, and so is the integer form used by the kernel
      const int cols = r + 1;
      for (int y = 0; y < L.rows; ++y) {
        const double *row = L.mat.data() + (size_t)y * cols;
        int nz = 0, col = -1;
        for (int x = 0; x < r; ++x)
          if (row[x] != 0.0) {
            ++nz;
            col = x;
          }
        if (nz > 1) return;
        const double c8 = nz ? row[col] * 8.0 : 0.0, o8 = row[r] * 8.0;
        if (c8 != std::floor(c8) || o8 != std::floor(o8) || std::fabs(c8) >= 1e6 || std::fabs(o8) >= 1e9) return;
        const GRow &R = L.row[y];
        const long long span = nz ? std::max<long long>(P.out_shape[col] - 1, 0) : 0;
        const bool integral = std::fmod(c8, 8.0) == 0.0 && std::fmod(o8, 8.0) == 0.0;
        if (integral) {
          const long long coef = (long long)(c8 / 8.0), off = (long long)(o8 / 8.0);
          base += off * R.stride;
          if (nz) ls[renum(col)] += coef * R.stride;
          if (R.checked) {
            if (G.n_checks == kG4Checks) return;
            if (llabs(coef) * span + llabs(off) >= lim || R.extent >= lim) return;
            G4Check &C = G.chk[G.n_checks++];
            C.col = nz ? renum(col) : 0;
            C.coef = (int32_t)coef;
            C.off = (int32_t)off;
            C.extent = (int32_t)R.extent;
          }
        } else {
          if (G.n_checks == kG4Checks) return;
          if (std::fabs(c8) * (double)span + std::fabs(o8) >= (double)lim || R.extent >= lim || R.stride >= lim) return;
          G4Check &C = G.chk[G.n_checks++];
          C.col = nz ? renum(col) : 0;
          C.coef = (int32_t)c8;
          C.off = (int32_t)o8;
          C.extent = (int32_t)R.extent;      // always bounds checked (classify_leaf marks non-integer rows checked)
          C.shift = 3;
          C.stride = (int32_t)R.stride;
          if (nz && renum(col) == 0) inner_scaled = true;
          // farthest address this row can add, for the reach guard below
          scaled_reach += (R.extent - 1) * R.stride;
        }
      }
    } else if (L.mode == G_SCALAR || L.mode == G_STRIDED) {
      base = L.base;
      for (int d = 0; d < r; ++d) ls[renum(d)] = L.ls[d];
    } else {  // G_MONO
      for (int y = 0; y < L.rows; ++y) {
        const GRow &R = L.row[y];
        base += R.off * R.stride;
        if (R.col >= 0) ls[renum(R.col)] += R.coef * R.stride;
        if (R.checked) {
          if (G.n_checks == kG4Checks) return;
          const long long span = R.col >= 0 ? std::max<long long>(P.out_shape[R.col] - 1, 0) : 0;
          const long long coef = R.col >= 0 ? R.coef : 0;
          if (llabs(coef) * span + llabs(R.off) >= lim || R.extent >= lim) return;
          G4Check &C = G.chk[G.n_checks++];
          C.col = R.col >= 0 ? renum(R.col) : 0;
          C.coef = (int32_t)coef;
          C.off = (int32_t)R.off;
          C.extent = (int32_t)R.extent;
        }
      }
    }
    // every address a VALID element can produce is inside the parameter (< 2^31); guard the intermediate sums too
    long long reach = llabs(base) + scaled_reach;
    for (int d = 0; d < r; ++d) reach += llabs(ls[renum(d)]) * std::max<long long>(P.out_shape[d] - 1, 0);
    if (reach >= lim) return;
    G.base = (int32_t)base;
    for (int d = 0; d < kMaxRank; ++d) G.ls[d] = (int32_t)ls[d];
    out.push_back(G);
  }
  P.g4leaves.swap(out);
  const long long W = r ? P.out_shape[r - 1] : 1;
  P.g4_vec = (r && W % 4 == 0 && !inner_scaled) ? 4 : 1;
  P.g4_ok = true;
}

}  // namespace

int lower_tree(const uint8_t *blob, size_t len, Plan *plan, std::string *err) {
  Lowering L;
  L.plan = plan;
  L.err = err;
  int rc = L.parse(blob, len);
  if (rc) return rc;
  rc = L.emit_all();
  if (rc) return rc;
  rc = L.choose_kernel();
  if (rc || plan->kind == PLAN_VEC) return rc;   // small contiguous programs stay on the 128-bit vector kernel
  rc = L.compress_loops();
  if (rc) return rc;
  if (plan->loops.empty()) {
    L.build_g4();
    return 0;
  }
  plan->gleaves.clear();
  plan->dpool.clear();
  L.detect_contraction();
  return L.choose_kernel();
}

std::string Plan::describe() const {
  static const char *mode_name[] = {"scalar", "strided", "mono", "generic"};
  std::ostringstream o;
  o << "kind=" << (kind == PLAN_VEC ? "vec" : g4_ok ? (g4_vec == 4 ? "gather4" : "gather1") : "gather") << " elems=" << n_elems << " roots=" << n_roots
    << " code=" << code.size() << " slots=" << n_slots;
  if (kind == PLAN_VEC) {
    o << " outer=" << v_outer << " dims=[";
    for (int j = 0; j < v_outer; ++j) o << v_dims[j] << ",";
    o << v_inner << "] leaves=[";
    for (const VLeaf &v : vleaves)
      o << (v.inner ? (v.streaming ? "stream" : "vec") : "bcast") << ":p" << v.param << ",";
    o << "]";
  } else {
    o << " leaves=[";
    if (gleaves.size() <= 16) {
      for (const GLeaf &g : gleaves) o << mode_name[g.mode] << ":p" << g.param << ",";
    } else {
      int cnt[4] = {0, 0, 0, 0};
      for (const GLeaf &g : gleaves) cnt[g.mode]++;
      for (int m = 0; m < 4; ++m)
        if (cnt[m]) o << cnt[m] << "x" << mode_name[m] << ",";
    }
    o << "]";
  }
  int mac = 0;
  for (const Instr &i : code) mac += i.op == OP_MAC || i.op == OP_MSUB;
  o << " mac=" << mac;
  if (!loops.empty()) {
    o << " loops=[";
    if (loops.size() > 8) o << loops.size() << " x ";
    for (size_t li = 0; li < loops.size() && li < 8; ++li) {
      const GLoop &g = loops[li];
      o << (g.kind == LOOP_MAC ? "mac" : "add") << "x" << g.count << ":" << (g.leaf_kind[0] == LL_LINEAR ? "lin" : "mono")
        << (g.kind == LOOP_MAC ? (g.leaf_kind[1] == LL_LINEAR ? "*lin" : "*mono") : "") << ",";
    }
    o << "]";
  }
  if (contraction.valid)
    o << " contraction=" << contraction.m << "x" << contraction.n << "x" << contraction.k << (contraction.transpose_a ? "T" : "N")
      << (contraction.transpose_b ? "T" : "N");
  return o.str();
}

}  // namespace dls
