// common.h -- internal runtime structures behind the opaque handles of include/dls_b200.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <map>
#include <memory>
#include <mutex>
#include <string>
#include <unordered_map>
#include <vector>

#include "../../../include/dls_b200.h"
#include "plan.h"

namespace dls {

void set_error(const std::string &msg);
int cuda_fail(cudaError_t e, const char *what, const char *file, int line);

#define DLS_CUDA(expr)                                                       \
  do {                                                                       \
    cudaError_t e__ = (expr);                                                \
    if (e__ != cudaSuccess) return ::dls::cuda_fail(e__, #expr, __FILE__, __LINE__); \
  } while (0)

#define DLS_CHECK(cond, code, msg) \
  do {                             \
    if (!(cond)) {                 \
      ::dls::set_error(msg);       \
      return (code);               \
    }                              \
  } while (0)

constexpr int kMaxStreams = 32;
constexpr int kShadows = 5;               // staged operand copies cached per buffer (dls_buf::shadow)
constexpr int kReduceMaxBlocks = 4096;   // partial sums per stream-private scratch

struct Block {
  void *ptr = nullptr;
  size_t bytes = 0;
  uint32_t stream_mask = 0;                 // pool streams that touched the block before it was freed
  int pool = 0;                             // 0: the context's allocator; > 0: private arena of a captured step (graph.cu)
  cudaEvent_t free_events[kMaxStreams] = {};  // only for streams in stream_mask when more than one stream exists
};

// Private arena of one captured step: every block allocated between dls_graph_begin and dls_graph_end comes from here
// and returns here when freed, so the addresses baked into the graph's kernel nodes are never handed to eager work.
struct GraphPool {
  std::multimap<size_t, Block> free_blocks;
  uint64_t bytes = 0;
};

struct CachedPlan {
  std::vector<uint8_t> blob;
  Plan plan;
};

}  // namespace dls

namespace dls { struct PeerState; }   // peer.cu

struct dls_event {
  std::atomic<int> rc{1};
  dls_ctx *ctx = nullptr;
  cudaEvent_t ev = nullptr;
};

struct dls_buf {
  std::atomic<int> rc{1};
  dls_ctx *ctx = nullptr;
  float *ptr = nullptr;
  size_t n = 0;            // floats
  dls::Block block;
  int last_stream = 0;     // stream of the most recent writer
  cudaEvent_t ready = nullptr;  // recorded after the most recent writer (multi-stream contexts only)
  cudaEvent_t comm_ready = nullptr;  // an all-reduce of this buffer is in flight on the exchange stream (comm.cpp)
  // tensor-core operand shadows (matmul.cu): a bf16 copy and a transposed fp32 copy of this buffer for a given
  // [rows, cols] interpretation, kept until the buffer is written again or released, so a tensor that feeds several
  // contractions (an activation in forward and in the weight-gradient GEMM) is staged once.
  // slots 2 and 3 belong to the convolution path (conv.cu): the staged NHWC operand copy of an activation / delta, and
  // the packed per-tap weight matrices of a [3,3,Cin,Cout] weight; slot 4: a second conv copy of the same buffer in another
  // form (a sampled delta: compact for the weight gradient, zero-filled for backward-data).
  dls::Block shadow[dls::kShadows];
  long long shadow_rows[dls::kShadows] = {0, 0, 0, 0, 0}, shadow_cols[dls::kShadows] = {0, 0, 0, 0, 0};
};

struct dls_graph {
  std::atomic<int> rc{1};
  dls_ctx *ctx = nullptr;
  cudaGraph_t graph = nullptr;
  cudaGraphExec_t exec = nullptr;
  int pool = 0;
  uint64_t kernel_nodes = 0, nodes = 0;
};

struct dls_ctx {
  int device = 0;
  int sm_count = 148;
  int n_streams = 1;
  dls_ctx_opts opts{};
  std::vector<cudaStream_t> streams;
  uint32_t rr = 0;
  std::recursive_mutex mu;

  // caching allocator
  std::multimap<size_t, dls::Block> free_blocks;
  dls_stats stats{};

  // plan cache (replaces Tensors.kernelCache)
  std::unordered_multimap<uint64_t, std::shared_ptr<dls::CachedPlan>> plans;

  // per-stream reduction scratch: kReduceMaxBlocks partial sums + one arrival counter
  std::vector<float *> red_partials;
  std::vector<unsigned int *> red_counter;

  // occupancy-derived grid sizes, filled at create time
  int vec_blocks_per_sm = 4;
  int gather_blocks_per_sm = 4;

  // captured steps (graph.cu)
  int capture_pool = 0;                       // arena id while a capture is open, else 0
  int next_pool = 1;
  std::map<int, dls::GraphPool> pools;
  std::vector<dls_graph *> graphs;            // captured steps still alive (their CUDA objects go with the context)

  // data-parallel exchange: collectives run on their own stream, forked from / joined to the compute stream with
  // events, so that an all-reduce overlaps the backward kernels launched after it (comm.cpp)
  cudaStream_t copy_stream = nullptr;         // dls_buf_upload_async: host -> device copies that overlap the compute stream
  cudaStream_t comm_stream = nullptr;
  int exchange_sms = 0;                       // SMs a collective occupies (its channel count); one-CTA-per-SM kernels launched
                                              // while an exchange is in flight leave that many SMs to it
  std::vector<dls_buf *> comm_pending;        // buffers whose comm_ready has not been waited for yet
  void *nccl_comm = nullptr;
  int comm_rank = 0, comm_world = 1;
  dls::PeerState *peer = nullptr;             // NVLink peer segment of the one-launch end-of-step exchange (peer.cu)
};

namespace dls {

// runtime.cu
int pick_stream(dls_ctx *ctx);
int alloc_block(dls_ctx *ctx, size_t bytes, int stream, Block *out);
void free_block(dls_ctx *ctx, Block &blk);
int new_buf(dls_ctx *ctx, size_t n_floats, int stream, dls_buf **out);
int join_exchange(dls_ctx *ctx, int stream, dls_buf *buf);   // make `stream` wait for buf's pending all-reduce
int join_all_exchanges(dls_ctx *ctx, int stream);
int wait_inputs(dls_ctx *ctx, int stream, dls_buf *const *bufs, int n_bufs, dls_event *const *deps, int n_deps);
int finish_output(dls_ctx *ctx, int stream, dls_buf *const *written, int n_written, dls_buf *const *read, int n_read,
                  dls_event **done);
int wait_streams(dls_ctx *ctx, int stream, uint32_t mask);   // multi-stream contexts: `stream` waits for the tail of every stream in `mask`
int wait_readers(dls_ctx *ctx, int stream, dls_buf *buf);    // ... of every stream that touched `buf` (before an in-place write)

// graph.cu
void destroy_graphs(dls_ctx *ctx);

// peer.cu
void peer_destroy(dls_ctx *ctx);
int peer_check(dls_ctx *ctx);   // DLS_ERR_COMM once a peer exchange has timed out

// ew_kernels.cu
int launch_plan(dls_ctx *ctx, int stream, CachedPlan &cp, dls_buf *const *params, dls_buf *out, bool reduce_sum);
int query_ew_occupancy(dls_ctx *ctx);

// aux_kernels.cu
int launch_sum(dls_ctx *ctx, int stream, const float *in, size_t n, float *out);
int launch_sum_axis(dls_ctx *ctx, int stream, const float *in, size_t outer, size_t k, size_t inner, float *out);
int launch_random(dls_ctx *ctx, int stream, float *out, size_t n, uint32_t seed, int hash_kind);
int launch_random_normal(dls_ctx *ctx, int stream, float *out, size_t n_padded, uint32_t seed, int hash_kind);
int launch_sgd(dls_ctx *ctx, int stream, float *w, const float *d, size_t n, float lr);
int launch_optimizer(dls_ctx *ctx, int stream, int kind, float *w, float *s1, float *s2, const float *grad, size_t n,
                     const float *hyper);

}  // namespace dls
