// runtime.cu -- context, stream pool, caching allocator, refcounted buffers and events: the C ABI of
// include/dls_b200.h that replaces Compute.scala's OpenCL trait (OpenCL:1139-1433) and Memory (Memory:13-35).
#include <string.h>

#include <algorithm>

#include "common.h"

namespace dls {

static thread_local std::string g_last_error;

void set_error(const std::string &msg) { g_last_error = msg; }

int cuda_fail(cudaError_t e, const char *what, const char *file, int line) {
  char buf[512];
  snprintf(buf, sizeof buf, "%s failed: %s (%s) at %s:%d", what, cudaGetErrorString(e), cudaGetErrorName(e), file, line);
  g_last_error = buf;
  cudaGetLastError();  // clear the sticky-free error state
  if (e == cudaErrorMemoryAllocation) return DLS_ERR_OOM;
  if (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver || e == cudaErrorNoKernelImageForDevice)
    return DLS_ERR_NO_DEVICE;
  return DLS_ERR_CUDA;
}

int pick_stream(dls_ctx *ctx) { return ctx->n_streams == 1 ? 0 : (int)(ctx->rr++ % (uint32_t)ctx->n_streams); }

// ---- caching allocator ------------------------------------------------------------------------------------------------
// The reference does one clCreateBuffer / clReleaseMemObject pair per kernel launch (Tensors:1217, OpenCL:1399-1413).
// Here blocks are rounded to size classes, parked in a size-ordered free list when released and handed back to the
// next request of the same class; cross-stream reuse is ordered with the events recorded at free time.
static size_t size_class(size_t bytes) {
  if (bytes < 512) return 512;
  if (bytes < (1u << 20)) return (bytes + 511) & ~size_t(511);
  return (bytes + ((2u << 20) - 1)) & ~size_t((2u << 20) - 1);
}

int alloc_block(dls_ctx *ctx, size_t bytes, int stream, Block *out) {
  const size_t want = size_class(bytes + 16);  // 16 B slack: ragged 128-bit tail accesses stay inside the block
  if (ctx->capture_pool) {  // inside dls_graph_begin .. dls_graph_end: private arena, single stream, no events
    GraphPool &gp = ctx->pools[ctx->capture_pool];
    auto it = gp.free_blocks.lower_bound(want);
    if (it != gp.free_blocks.end() && it->first <= want + want / 8) {
      *out = it->second;
      gp.free_blocks.erase(it);
      ctx->stats.bytes_in_use += out->bytes;
      return DLS_OK;
    }
    Block b;
    b.bytes = want;
    b.pool = ctx->capture_pool;
    b.stream_mask = 1u << stream;
    cudaError_t e = cudaMalloc(&b.ptr, want);   // legal while capturing in relaxed mode; not a graph node
    if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc (captured step arena)", __FILE__, __LINE__);
    ctx->stats.driver_allocs++;
    ctx->stats.bytes_in_use += want;
    gp.bytes += want;
    *out = b;
    return DLS_OK;
  }
  auto it = ctx->free_blocks.lower_bound(want);
  if (it != ctx->free_blocks.end() && it->first <= want + want / 8) {
    Block b = it->second;
    ctx->free_blocks.erase(it);
    ctx->stats.bytes_cached -= b.bytes;
    for (int s = 0; s < ctx->n_streams; ++s) {
      if (!(b.stream_mask & (1u << s)) || !b.free_events[s]) continue;
      if (s != stream) DLS_CUDA(cudaStreamWaitEvent(ctx->streams[stream], b.free_events[s], 0));
      cudaEventDestroy(b.free_events[s]);
      b.free_events[s] = nullptr;
    }
    b.stream_mask = 1u << stream;
    ctx->stats.bytes_in_use += b.bytes;
    *out = b;
    return DLS_OK;
  }
  Block b;
  b.bytes = want;
  cudaError_t e = cudaMalloc(&b.ptr, want);
  if (e == cudaErrorMemoryAllocation) {  // give cached blocks back to the driver and retry once
    cudaGetLastError();
    dls_ctx_trim(ctx);
    e = cudaMalloc(&b.ptr, want);
  }
  if (e != cudaSuccess) return cuda_fail(e, "cudaMalloc", __FILE__, __LINE__);
  ctx->stats.driver_allocs++;
  ctx->stats.bytes_in_use += want;
  b.stream_mask = 1u << stream;
  *out = b;
  return DLS_OK;
}

void free_block(dls_ctx *ctx, Block &blk) {
  if (!blk.ptr) return;
  ctx->stats.bytes_in_use -= blk.bytes;
  if (blk.pool) {  // arena block: parked for the graph that owns the arena; returned to the driver once that graph is gone
    auto it = ctx->pools.find(blk.pool);
    if (it != ctx->pools.end()) it->second.free_blocks.emplace(blk.bytes, blk);
    else cudaFree(blk.ptr);
    blk.ptr = nullptr;
    return;
  }
  if (ctx->n_streams > 1) {
    for (int s = 0; s < ctx->n_streams; ++s) {
      if (!(blk.stream_mask & (1u << s))) continue;
      cudaEvent_t ev;
      if (cudaEventCreateWithFlags(&ev, cudaEventDisableTiming) == cudaSuccess) {
        cudaEventRecord(ev, ctx->streams[s]);
        blk.free_events[s] = ev;
      }
    }
  }
  const uint64_t cap = ctx->opts.allocator_cap_bytes;
  if (cap && ctx->stats.bytes_cached + blk.bytes > cap) {
    if (ctx->n_streams > 1) cudaDeviceSynchronize();
    for (int s = 0; s < kMaxStreams; ++s)
      if (blk.free_events[s]) cudaEventDestroy(blk.free_events[s]);
    cudaFree(blk.ptr);
  } else {
    ctx->stats.bytes_cached += blk.bytes;
    ctx->free_blocks.emplace(blk.bytes, blk);
  }
  blk.ptr = nullptr;
}

int new_buf(dls_ctx *ctx, size_t n_floats, int stream, dls_buf **out) {
  dls_buf *b = new dls_buf();
  b->ctx = ctx;
  b->n = n_floats;
  b->last_stream = stream;
  int rc = alloc_block(ctx, n_floats * sizeof(float), stream, &b->block);
  if (rc) {
    delete b;
    return rc;
  }
  b->ptr = static_cast<float *>(b->block.ptr);
  *out = b;
  return DLS_OK;
}

// Make `stream` wait for the producers of every input buffer and for the explicit dependencies
// (replaces the `waitingEvents` list handed to clEnqueueNDRangeKernel, Tensors:1235, OpenCL:826).
int join_exchange(dls_ctx *ctx, int stream, dls_buf *b) {
  if (!b || !b->comm_ready) return DLS_OK;
  cudaEvent_t ev = b->comm_ready;
  b->comm_ready = nullptr;
  for (size_t i = 0; i < ctx->comm_pending.size(); ++i)
    if (ctx->comm_pending[i] == b) {
      ctx->comm_pending.erase(ctx->comm_pending.begin() + i);
      break;
    }
  cudaError_t e = cudaStreamWaitEvent(ctx->streams[stream], ev, 0);
  cudaEventDestroy(ev);
  if (e != cudaSuccess) return cuda_fail(e, "cudaStreamWaitEvent(exchange)", __FILE__, __LINE__);
  return DLS_OK;
}

int join_all_exchanges(dls_ctx *ctx, int stream) {
  while (!ctx->comm_pending.empty()) {
    int rc = join_exchange(ctx, stream, ctx->comm_pending.back());
    if (rc) return rc;
  }
  return DLS_OK;
}

int wait_inputs(dls_ctx *ctx, int stream, dls_buf *const *bufs, int n_bufs, dls_event *const *deps, int n_deps) {
  cudaStream_t st = ctx->streams[stream];
  for (int i = 0; i < n_bufs; ++i)
    if (bufs[i] && bufs[i]->comm_ready) {
      int rc = join_exchange(ctx, stream, bufs[i]);
      if (rc) return rc;
    }
  if (ctx->n_streams > 1) {
    for (int i = 0; i < n_bufs; ++i) {
      dls_buf *b = bufs[i];
      if (b && b->ready && b->last_stream != stream) DLS_CUDA(cudaStreamWaitEvent(st, b->ready, 0));
    }
  }
  for (int i = 0; i < n_deps; ++i)
    if (deps && deps[i] && deps[i]->ev) DLS_CUDA(cudaStreamWaitEvent(st, deps[i]->ev, 0));
  return DLS_OK;
}

// Multi-stream contexts order work by the last WRITER's event only (`dls_buf::ready`).  The reference never writes a
// buffer in place, so that is all it needs; the in-place entry points of this library (dls_sgd_update,
// dls_optimizer_step, dls_buf_write, dls_buf_copy) must in addition wait for every stream that may still be READING the
// buffer.  `stream_mask` records every pool stream that touched the block; the writer's stream waits for the current
// tail of each of them (conservative: a marker event per stream, not per reader).
int wait_streams(dls_ctx *ctx, int stream, uint32_t mask) {
  if (ctx->n_streams <= 1) return DLS_OK;
  for (int s = 0; s < ctx->n_streams; ++s) {
    if (s == stream || !(mask & (1u << s))) continue;
    cudaEvent_t ev;
    DLS_CUDA(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    cudaError_t e = cudaEventRecord(ev, ctx->streams[s]);
    if (e == cudaSuccess) e = cudaStreamWaitEvent(ctx->streams[stream], ev, 0);
    cudaEventDestroy(ev);   // released by the runtime once the wait has been satisfied
    if (e != cudaSuccess) return cuda_fail(e, "wait_streams", __FILE__, __LINE__);
  }
  return DLS_OK;
}

int wait_readers(dls_ctx *ctx, int stream, dls_buf *buf) {
  if (!buf) return DLS_OK;
  return wait_streams(ctx, stream, buf->block.stream_mask);
}

int finish_output(dls_ctx *ctx, int stream, dls_buf *const *written, int n_written, dls_buf *const *read, int n_read,
                  dls_event **done) {
  cudaStream_t st = ctx->streams[stream];
  for (int i = 0; i < n_read; ++i)
    if (read[i]) read[i]->block.stream_mask |= 1u << stream;
  for (int i = 0; i < n_written; ++i) {
    dls_buf *b = written[i];
    if (!b) continue;
    b->block.stream_mask |= 1u << stream;
    b->last_stream = stream;
    for (int k = 0; k < kShadows; ++k)   // the contents changed: operand shadows are stale
      if (b->shadow[k].ptr) free_block(ctx, b->shadow[k]);
    if (ctx->n_streams > 1) {
      if (!b->ready) DLS_CUDA(cudaEventCreateWithFlags(&b->ready, cudaEventDisableTiming));
      DLS_CUDA(cudaEventRecord(b->ready, st));
    }
  }
  if (done) {
    DLS_CHECK(ctx->capture_pool == 0, DLS_ERR_INVALID, "completion events cannot be requested inside a captured step");
    dls_event *e = new dls_event();
    e->ctx = ctx;
    cudaError_t err = cudaEventCreate(&e->ev);
    if (err == cudaSuccess) err = cudaEventRecord(e->ev, st);
    if (err != cudaSuccess) {
      delete e;
      return cuda_fail(err, "cudaEventRecord", __FILE__, __LINE__);
    }
    *done = e;
  }
  return DLS_OK;
}

}  // namespace dls

using namespace dls;

extern "C" {

int dls_abi_version(void) { return DLS_ABI_VERSION; }

const char *dls_last_error(void) { return g_last_error.c_str(); }

int dls_ctx_create(const dls_ctx_opts *opts, dls_ctx **out) {
  DLS_CHECK(out != nullptr, DLS_ERR_INVALID, "dls_ctx_create: out is NULL");
  dls_ctx_opts o{};
  if (opts) o = *opts;
  if (o.streams_per_device <= 0) o.streams_per_device = 1;
  DLS_CHECK(o.streams_per_device <= kMaxStreams, DLS_ERR_INVALID, "dls_ctx_create: at most 32 streams per device");
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    cudaGetLastError();
    set_error("no CUDA device is visible: libdlsb200 has no CPU fallback");
    return DLS_ERR_NO_DEVICE;
  }
  DLS_CHECK(o.device >= 0 && o.device < count, DLS_ERR_INVALID, "dls_ctx_create: device ordinal out of range");
  cudaDeviceProp prop;
  DLS_CUDA(cudaGetDeviceProperties(&prop, o.device));
  if (prop.major != 10) {
    set_error(std::string("device is sm_") + std::to_string(prop.major * 10 + prop.minor) +
              ": libdlsb200 is built for sm_100a only");
    return DLS_ERR_NO_DEVICE;
  }
  DLS_CUDA(cudaSetDevice(o.device));
  dls_ctx *ctx = new dls_ctx();
  ctx->device = o.device;
  ctx->opts = o;
  ctx->sm_count = prop.multiProcessorCount;
  ctx->n_streams = o.streams_per_device;
  ctx->streams.resize(ctx->n_streams);
  ctx->red_partials.resize(ctx->n_streams);
  ctx->red_counter.resize(ctx->n_streams);
  for (int s = 0; s < ctx->n_streams; ++s) {
    DLS_CUDA(cudaStreamCreateWithFlags(&ctx->streams[s], cudaStreamNonBlocking));
    DLS_CUDA(cudaMalloc(&ctx->red_partials[s], kReduceMaxBlocks * sizeof(float)));
    DLS_CUDA(cudaMalloc(&ctx->red_counter[s], 256));
    DLS_CUDA(cudaMemset(ctx->red_counter[s], 0, 256));
  }
  int rc = query_ew_occupancy(ctx);
  if (rc) {
    delete ctx;
    return rc;
  }
  *out = ctx;
  return DLS_OK;
}

int dls_ctx_sync(dls_ctx *ctx) {
  DLS_CHECK(ctx, DLS_ERR_INVALID, "dls_ctx_sync: ctx is NULL");
  DLS_CHECK(ctx->capture_pool == 0, DLS_ERR_INVALID, "dls_ctx_sync: a capture is open on this context");
  DLS_CUDA(cudaSetDevice(ctx->device));
  if (ctx->comm_stream) DLS_CUDA(cudaStreamSynchronize(ctx->comm_stream));
  if (ctx->copy_stream) DLS_CUDA(cudaStreamSynchronize(ctx->copy_stream));
  for (cudaStream_t s : ctx->streams) DLS_CUDA(cudaStreamSynchronize(s));
  {
    // Forget the joins owed to exchanges that have finished: an event recorded by eager work must not be waited for
    // inside a capture opened later (cudaErrorStreamCaptureIsolation).  An exchange another thread issued after the
    // synchronisation above is not complete yet and stays owed.
    std::lock_guard<std::recursive_mutex> lk(ctx->mu);
    size_t kept = 0;
    for (dls_buf *b : ctx->comm_pending) {
      if (!b || !b->comm_ready) continue;
      if (cudaEventQuery(b->comm_ready) == cudaSuccess) {
        cudaEventDestroy(b->comm_ready);
        b->comm_ready = nullptr;
      } else {
        cudaGetLastError();   // cudaErrorNotReady is not an error here
        ctx->comm_pending[kept++] = b;
      }
    }
    ctx->comm_pending.resize(kept);
  }
  return peer_check(ctx);
}

int dls_ctx_trim(dls_ctx *ctx) {
  DLS_CHECK(ctx, DLS_ERR_INVALID, "dls_ctx_trim: ctx is NULL");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  cudaDeviceSynchronize();
  for (auto &kv : ctx->free_blocks) {
    for (int s = 0; s < kMaxStreams; ++s)
      if (kv.second.free_events[s]) cudaEventDestroy(kv.second.free_events[s]);
    cudaFree(kv.second.ptr);
  }
  ctx->free_blocks.clear();
  ctx->stats.bytes_cached = 0;
  return DLS_OK;
}

void dls_ctx_destroy(dls_ctx *ctx) {
  if (!ctx) return;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  dls_graph_abort(ctx);
  destroy_graphs(ctx);
  dls_comm_destroy(ctx);
  dls_ctx_trim(ctx);
  for (auto &kv : ctx->pools)
    for (auto &fb : kv.second.free_blocks) cudaFree(fb.second.ptr);
  ctx->pools.clear();
  for (auto &kv : ctx->plans)
    if (kv.second->plan.d_tables) cudaFree(kv.second->plan.d_tables);
  ctx->plans.clear();
  if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
  for (int s = 0; s < ctx->n_streams; ++s) {
    cudaFree(ctx->red_partials[s]);
    cudaFree(ctx->red_counter[s]);
    cudaStreamDestroy(ctx->streams[s]);
  }
  delete ctx;
}

int dls_ctx_stats(dls_ctx *ctx, dls_stats *out) {
  DLS_CHECK(ctx && out, DLS_ERR_INVALID, "dls_ctx_stats: NULL argument");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  *out = ctx->stats;
  return DLS_OK;
}

void *dls_ctx_stream(dls_ctx *ctx, int index) {
  if (!ctx || index < 0 || index >= ctx->n_streams) return nullptr;
  return ctx->streams[index];
}

// ---- buffers --------------------------------------------------------------------------------------------------------
int dls_buf_alloc(dls_ctx *ctx, size_t n_floats, dls_buf **out) {
  DLS_CHECK(ctx && out, DLS_ERR_INVALID, "dls_buf_alloc: NULL argument");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CUDA(cudaSetDevice(ctx->device));
  return new_buf(ctx, n_floats, 0, out);
}

int dls_buf_from_host(dls_ctx *ctx, const float *src, size_t n_floats, dls_buf **out, dls_event **done) {
  DLS_CHECK(ctx && out && (src || n_floats == 0), DLS_ERR_INVALID, "dls_buf_from_host: NULL argument");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CHECK(ctx->capture_pool == 0, DLS_ERR_INVALID, "dls_buf_from_host: host transfers cannot be recorded into a captured step");
  DLS_CUDA(cudaSetDevice(ctx->device));
  int stream = pick_stream(ctx);
  dls_buf *b = nullptr;
  int rc = new_buf(ctx, n_floats, stream, &b);
  if (rc) return rc;
  if (n_floats) {
    cudaError_t e = cudaMemcpyAsync(b->ptr, src, n_floats * sizeof(float), cudaMemcpyHostToDevice, ctx->streams[stream]);
    if (e != cudaSuccess) {
      dls_buf_release(b);
      return cuda_fail(e, "cudaMemcpyAsync(H2D)", __FILE__, __LINE__);
    }
    ctx->stats.h2d_bytes += n_floats * sizeof(float);
  }
  rc = finish_output(ctx, stream, &b, 1, nullptr, 0, done);
  if (rc) {
    dls_buf_release(b);
    return rc;
  }
  *out = b;
  return DLS_OK;
}

void dls_buf_retain(dls_buf *buf) {
  if (buf) buf->rc.fetch_add(1, std::memory_order_relaxed);
}

void dls_buf_release(dls_buf *buf) {
  if (!buf) return;
  if (buf->rc.fetch_sub(1, std::memory_order_acq_rel) != 1) return;
  dls_ctx *ctx = buf->ctx;
  {
    std::lock_guard<std::recursive_mutex> lk(ctx->mu);
    if (buf->comm_ready) join_exchange(ctx, buf->last_stream, buf);   // the block may be reused right away
    if (buf->ready) cudaEventDestroy(buf->ready);
    for (int k = 0; k < kShadows; ++k) free_block(ctx, buf->shadow[k]);
    free_block(ctx, buf->block);
  }
  delete buf;
}

size_t dls_buf_length(const dls_buf *buf) { return buf ? buf->n : 0; }

void *dls_buf_device_ptr(const dls_buf *buf) { return buf ? buf->ptr : nullptr; }

int dls_buf_read(dls_ctx *ctx, dls_buf *buf, size_t offset, float *dst, size_t n_floats, dls_event *const *deps, int n_deps,
                 dls_event **done) {
  DLS_CHECK(ctx && buf && (dst || n_floats == 0), DLS_ERR_INVALID, "dls_buf_read: NULL argument");
  DLS_CHECK(offset + n_floats <= buf->n, DLS_ERR_INVALID, "dls_buf_read: range exceeds the buffer");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CHECK(ctx->capture_pool == 0, DLS_ERR_INVALID, "dls_buf_read: host transfers cannot be recorded into a captured step");
  DLS_CUDA(cudaSetDevice(ctx->device));
  int stream = buf->last_stream;
  int rc = wait_inputs(ctx, stream, &buf, 1, deps, n_deps);
  if (rc) return rc;
  if (n_floats) {
    DLS_CUDA(cudaMemcpyAsync(dst, buf->ptr + offset, n_floats * sizeof(float), cudaMemcpyDeviceToHost, ctx->streams[stream]));
    ctx->stats.d2h_bytes += n_floats * sizeof(float);
  }
  if (done) return finish_output(ctx, stream, nullptr, 0, &buf, 1, done);
  DLS_CUDA(cudaStreamSynchronize(ctx->streams[stream]));
  return peer_check(ctx);   // a timed-out peer exchange invalidates whatever was read
}

int dls_buf_write(dls_ctx *ctx, dls_buf *buf, size_t offset, const float *src, size_t n_floats, dls_event **done) {
  DLS_CHECK(ctx && buf && (src || n_floats == 0), DLS_ERR_INVALID, "dls_buf_write: NULL argument");
  DLS_CHECK(offset + n_floats <= buf->n, DLS_ERR_INVALID, "dls_buf_write: range exceeds the buffer");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CHECK(ctx->capture_pool == 0, DLS_ERR_INVALID, "dls_buf_write: host transfers cannot be recorded into a captured step");
  DLS_CUDA(cudaSetDevice(ctx->device));
  int stream = buf->last_stream;
  int rc = join_exchange(ctx, stream, buf);
  if (!rc) rc = wait_readers(ctx, stream, buf);
  if (rc) return rc;
  if (n_floats) {
    DLS_CUDA(cudaMemcpyAsync(buf->ptr + offset, src, n_floats * sizeof(float), cudaMemcpyHostToDevice, ctx->streams[stream]));
    ctx->stats.h2d_bytes += n_floats * sizeof(float);
  }
  return finish_output(ctx, stream, &buf, 1, nullptr, 0, done);
}

int dls_buf_upload_async(dls_ctx *ctx, dls_buf *buf, size_t offset, const float *src, size_t n_floats, dls_event *const *deps,
                         int n_deps, dls_event **done) {
  DLS_CHECK(ctx && buf && done && (src || n_floats == 0), DLS_ERR_INVALID, "dls_buf_upload_async: NULL argument");
  DLS_CHECK(offset + n_floats <= buf->n, DLS_ERR_INVALID, "dls_buf_upload_async: range exceeds the buffer");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CHECK(ctx->capture_pool == 0, DLS_ERR_INVALID, "dls_buf_upload_async: host transfers cannot be recorded into a captured step");
  DLS_CUDA(cudaSetDevice(ctx->device));
  if (!ctx->copy_stream) DLS_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
  for (int i = 0; i < n_deps; ++i)
    if (deps && deps[i] && deps[i]->ev) DLS_CUDA(cudaStreamWaitEvent(ctx->copy_stream, deps[i]->ev, 0));
  if (n_floats) {
    DLS_CUDA(cudaMemcpyAsync(buf->ptr + offset, src, n_floats * sizeof(float), cudaMemcpyHostToDevice, ctx->copy_stream));
    ctx->stats.h2d_bytes += n_floats * sizeof(float);
  }
  for (int k = 0; k < kShadows; ++k)   // the contents change: operand shadows are stale
    if (buf->shadow[k].ptr) free_block(ctx, buf->shadow[k]);
  dls_event *e = new dls_event();
  e->ctx = ctx;
  cudaError_t err = cudaEventCreate(&e->ev);
  if (err == cudaSuccess) err = cudaEventRecord(e->ev, ctx->copy_stream);
  if (err != cudaSuccess) {
    delete e;
    return cuda_fail(err, "cudaEventRecord(copy stream)", __FILE__, __LINE__);
  }
  *done = e;
  return DLS_OK;
}

int dls_stream_wait_event(dls_ctx *ctx, int stream_index, dls_event *ev) {
  DLS_CHECK(ctx && ev && ev->ev && stream_index >= 0 && stream_index < ctx->n_streams, DLS_ERR_INVALID,
            "dls_stream_wait_event: bad argument");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CHECK(ctx->capture_pool == 0, DLS_ERR_INVALID, "dls_stream_wait_event: not inside a captured step");
  DLS_CUDA(cudaSetDevice(ctx->device));
  DLS_CUDA(cudaStreamWaitEvent(ctx->streams[stream_index], ev->ev, 0));
  return DLS_OK;
}

int dls_host_alloc(size_t n_floats, float **out) {
  DLS_CHECK(out, DLS_ERR_INVALID, "dls_host_alloc: NULL argument");
  void *p = nullptr;
  DLS_CUDA(cudaHostAlloc(&p, n_floats ? n_floats * sizeof(float) : 4, cudaHostAllocDefault));
  *out = static_cast<float *>(p);
  return DLS_OK;
}

void dls_host_free(float *ptr) {
  if (ptr) cudaFreeHost(ptr);
}

// ---- events ---------------------------------------------------------------------------------------------------------
void dls_event_retain(dls_event *ev) {
  if (ev) ev->rc.fetch_add(1, std::memory_order_relaxed);
}

void dls_event_release(dls_event *ev) {
  if (!ev) return;
  if (ev->rc.fetch_sub(1, std::memory_order_acq_rel) != 1) return;
  if (ev->ev) cudaEventDestroy(ev->ev);
  delete ev;
}

int dls_event_wait(dls_event *ev) {
  DLS_CHECK(ev, DLS_ERR_INVALID, "dls_event_wait: NULL event");
  DLS_CUDA(cudaEventSynchronize(ev->ev));
  return DLS_OK;
}

int dls_event_query(dls_event *ev) {
  DLS_CHECK(ev, DLS_ERR_INVALID, "dls_event_query: NULL event");
  cudaError_t e = cudaEventQuery(ev->ev);
  if (e == cudaSuccess) return DLS_OK;
  if (e == cudaErrorNotReady) return 1;
  return cuda_fail(e, "cudaEventQuery", __FILE__, __LINE__);
}

struct CallbackBox {
  dls_event_cb cb;
  void *user;
};

static void CUDART_CB host_trampoline(void *p) {
  CallbackBox *box = static_cast<CallbackBox *>(p);
  box->cb(box->user, DLS_OK);
  delete box;
}

int dls_event_on_complete(dls_event *ev, dls_event_cb cb, void *user) {
  DLS_CHECK(ev && cb, DLS_ERR_INVALID, "dls_event_on_complete: NULL argument");
  dls_ctx *ctx = ev->ctx;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  // a side stream waits on the event and then runs the host function: never blocks a compute stream
  // (the reference bounces its clSetEventCallback to an ExecutionContext for the same reason, OpenCL:404-411)
  static thread_local cudaStream_t cb_stream = nullptr;
  if (!cb_stream) DLS_CUDA(cudaStreamCreateWithFlags(&cb_stream, cudaStreamNonBlocking));
  DLS_CUDA(cudaStreamWaitEvent(cb_stream, ev->ev, 0));
  DLS_CUDA(cudaLaunchHostFunc(cb_stream, host_trampoline, new CallbackBox{cb, user}));
  return DLS_OK;
}

int dls_event_elapsed_ms(dls_event *start, dls_event *end, float *ms) {
  DLS_CHECK(start && end && ms, DLS_ERR_INVALID, "dls_event_elapsed_ms: NULL argument");
  DLS_CUDA(cudaEventElapsedTime(ms, start->ev, end->ev));
  return DLS_OK;
}

int dls_event_record(dls_ctx *ctx, int stream_index, dls_event **out) {
  DLS_CHECK(ctx && out && stream_index >= 0 && stream_index < ctx->n_streams, DLS_ERR_INVALID,
            "dls_event_record: bad argument");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  return finish_output(ctx, stream_index, nullptr, 0, nullptr, 0, out);
}

}  // extern "C"
