// graph.cu -- captured steps: a whole training / inference step recorded once into a CUDA graph and replayed with ONE
// driver call.
//
// The reference pays one clEnqueueNDRangeKernel (+ clSetKernelArg per argument, + clCreateBuffer for the output) per
// tensor materialisation (Tensors:1203-1251, OpenCL:826); a small model such as BASELINE config C1 (MLP 784-500-10,
// batch 128) is then bound by host dispatch, not by the device: 19 launches of 3-17 us each.  Here the host code runs
// the step once between dls_graph_begin and dls_graph_end; every launch of libdlsb200 in between is recorded instead
// of executed, and dls_graph_launch replays the recording.
//
// Rules of a capture (checked where they can be):
//   * single-stream contexts only; one capture at a time per context;
//   * every plan the step needs was built before (run the step eagerly once first) -- nothing is compiled or uploaded
//     while recording;
//   * no host <-> device transfers inside (dls_buf_from_host / dls_buf_read / dls_buf_write fail): inputs live in
//     buffers allocated BEFORE the capture and are refreshed with dls_buf_write between replays, results live in buffers
//     the caller keeps and are read after a replay;
//   * buffers allocated while recording come from a private arena (common.h GraphPool) that belongs to the graph, so
//     replays never touch memory the eager allocator may hand out;
//   * a recorded step that swaps a weight for a fresh buffer (TW:63-72) copies the new value back into the weight's
//     home buffer with dls_buf_copy before the capture ends, so that the next replay reads what this one wrote.
#include "common.h"

using namespace dls;

namespace dls {

// Context teardown: the CUDA side of every captured step still alive goes before the communicator does -- NCCL counts
// the graphs that captured its kernels and ncclCommDestroy waits for them, so a step that outlives its context would
// otherwise block the teardown forever.  The handles stay valid for dls_graph_release, which then only frees the shell.
void destroy_graphs(dls_ctx *ctx) {
  for (dls_graph *g : ctx->graphs) {
    if (g->exec) cudaGraphExecDestroy(g->exec);
    if (g->graph) cudaGraphDestroy(g->graph);
    g->exec = nullptr;
    g->graph = nullptr;
    g->ctx = nullptr;
  }
  ctx->graphs.clear();
}

}  // namespace dls

extern "C" {

int dls_graph_begin(dls_ctx *ctx) {
  DLS_CHECK(ctx, DLS_ERR_INVALID, "dls_graph_begin: ctx is NULL");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CHECK(ctx->n_streams == 1, DLS_ERR_UNSUPPORTED, "dls_graph_begin: captured steps need a single-stream context");
  DLS_CHECK(ctx->capture_pool == 0, DLS_ERR_INVALID, "dls_graph_begin: a capture is already open on this context");
  DLS_CUDA(cudaSetDevice(ctx->device));
  // exchanges issued eagerly rejoin the compute stream now: their events may not be waited for inside the recording
  int rc = join_all_exchanges(ctx, 0);
  if (rc) return rc;
  // relaxed mode: the allocator may call cudaMalloc for the arena while recording
  DLS_CUDA(cudaStreamBeginCapture(ctx->streams[0], cudaStreamCaptureModeRelaxed));
  ctx->capture_pool = ctx->next_pool++;
  ctx->pools[ctx->capture_pool];
  return DLS_OK;
}

static void drop_pool(dls_ctx *ctx, int id) {
  auto it = ctx->pools.find(id);
  if (it == ctx->pools.end()) return;
  for (auto &kv : it->second.free_blocks) cudaFree(kv.second.ptr);
  ctx->pools.erase(it);  // blocks still held by live buffers go back to the driver when those buffers are released
}

int dls_graph_end(dls_ctx *ctx, dls_graph **out) {
  DLS_CHECK(ctx && out, DLS_ERR_INVALID, "dls_graph_end: NULL argument");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CHECK(ctx->capture_pool != 0, DLS_ERR_INVALID, "dls_graph_end: no capture is open on this context");
  DLS_CUDA(cudaSetDevice(ctx->device));
  const int pool = ctx->capture_pool;
  join_all_exchanges(ctx, 0);   // work forked onto the exchange stream must rejoin the origin stream before the end
  ctx->capture_pool = 0;
  cudaGraph_t graph = nullptr;
  cudaError_t e = cudaStreamEndCapture(ctx->streams[0], &graph);
  if (e != cudaSuccess || !graph) {
    drop_pool(ctx, pool);
    if (e == cudaSuccess) {
      set_error("dls_graph_end: the capture was invalidated (a synchronising call was made while recording)");
      return DLS_ERR_CUDA;
    }
    return cuda_fail(e, "cudaStreamEndCapture", __FILE__, __LINE__);
  }
  dls_graph *g = new dls_graph();
  g->ctx = ctx;
  g->graph = graph;
  g->pool = pool;
  e = cudaGraphInstantiate(&g->exec, graph, 0);
  if (e != cudaSuccess) {
    cudaGraphDestroy(graph);
    delete g;
    drop_pool(ctx, pool);
    return cuda_fail(e, "cudaGraphInstantiate", __FILE__, __LINE__);
  }
  size_t n = 0;
  cudaGraphGetNodes(graph, nullptr, &n);
  std::vector<cudaGraphNode_t> nodes(n);
  if (n) cudaGraphGetNodes(graph, nodes.data(), &n);
  g->nodes = n;
  for (size_t i = 0; i < n; ++i) {
    cudaGraphNodeType t;
    if (cudaGraphNodeGetType(nodes[i], &t) == cudaSuccess && t == cudaGraphNodeTypeKernel) g->kernel_nodes++;
  }
  ctx->graphs.push_back(g);
  *out = g;
  return DLS_OK;
}

int dls_graph_abort(dls_ctx *ctx) {
  DLS_CHECK(ctx, DLS_ERR_INVALID, "dls_graph_abort: ctx is NULL");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  if (ctx->capture_pool == 0) return DLS_OK;
  const int pool = ctx->capture_pool;
  ctx->capture_pool = 0;
  cudaGraph_t graph = nullptr;
  cudaStreamEndCapture(ctx->streams[0], &graph);
  if (graph) cudaGraphDestroy(graph);
  cudaGetLastError();
  drop_pool(ctx, pool);
  return DLS_OK;
}

int dls_graph_launch(dls_graph *g) {
  DLS_CHECK(g && g->exec && g->ctx, DLS_ERR_INVALID, "dls_graph_launch: graph is NULL, or its context has been destroyed");
  dls_ctx *ctx = g->ctx;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CHECK(ctx->capture_pool == 0, DLS_ERR_INVALID, "dls_graph_launch: a capture is open on this context");
  DLS_CUDA(cudaSetDevice(ctx->device));
  DLS_CUDA(cudaGraphLaunch(g->exec, ctx->streams[0]));
  ctx->stats.kernel_launches += g->kernel_nodes;
  return DLS_OK;
}

int dls_graph_info(const dls_graph *g, uint64_t *kernel_nodes, uint64_t *nodes, uint64_t *arena_bytes) {
  DLS_CHECK(g && g->ctx, DLS_ERR_INVALID, "dls_graph_info: graph is NULL, or its context has been destroyed");
  std::lock_guard<std::recursive_mutex> lk(g->ctx->mu);
  if (kernel_nodes) *kernel_nodes = g->kernel_nodes;
  if (nodes) *nodes = g->nodes;
  if (arena_bytes) {
    auto it = g->ctx->pools.find(g->pool);
    *arena_bytes = it == g->ctx->pools.end() ? 0 : it->second.bytes;
  }
  return DLS_OK;
}

void dls_graph_retain(dls_graph *g) {
  if (g) g->rc.fetch_add(1, std::memory_order_relaxed);
}

void dls_graph_release(dls_graph *g) {
  if (!g) return;
  if (g->rc.fetch_sub(1, std::memory_order_acq_rel) != 1) return;
  dls_ctx *ctx = g->ctx;
  if (ctx) {   // else the context went first and took the CUDA objects with it (destroy_graphs)
    std::lock_guard<std::recursive_mutex> lk(ctx->mu);
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->streams[0]);   // a replay may still be running out of the arena
    if (g->exec) cudaGraphExecDestroy(g->exec);
    if (g->graph) cudaGraphDestroy(g->graph);
    drop_pool(ctx, g->pool);
    for (size_t i = 0; i < ctx->graphs.size(); ++i)
      if (ctx->graphs[i] == g) {
        ctx->graphs.erase(ctx->graphs.begin() + i);
        break;
      }
  }
  delete g;
}

int dls_buf_copy(dls_ctx *ctx, dls_buf *dst, dls_buf *src, size_t n_floats) {
  DLS_CHECK(ctx && dst && src, DLS_ERR_INVALID, "dls_buf_copy: NULL argument");
  DLS_CHECK(n_floats <= dst->n && n_floats <= src->n, DLS_ERR_INVALID, "dls_buf_copy: range exceeds a buffer");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CUDA(cudaSetDevice(ctx->device));
  const int stream = ctx->n_streams == 1 ? 0 : dst->last_stream;
  dls_buf *both[2] = {src, dst};
  int rc = wait_inputs(ctx, stream, both, 2, nullptr, 0);
  if (!rc) rc = wait_readers(ctx, stream, dst);
  if (rc) return rc;
  if (n_floats && dst->ptr != src->ptr)
    DLS_CUDA(cudaMemcpyAsync(dst->ptr, src->ptr, n_floats * sizeof(float), cudaMemcpyDeviceToDevice, ctx->streams[stream]));
  return finish_output(ctx, stream, &dst, 1, &src, 1, nullptr);
}

}  // extern "C"
