// comm.cpp -- data-parallel exchange (N7): one NCCL all-reduce of the packed fp32 gradient arena over NVLink 5 /
// NVSwitch.  New functionality: the reference has no collective (its only multi-device mechanism is the shared
// command-queue pool, OpenCL:423-448).  libnccl is bound at run time with dlopen so that the library loads on
// machines without NCCL; DLS_NCCL_LIB overrides the library path.
#include <climits>
#include <dlfcn.h>
#include <stdlib.h>
#include <string.h>

#include "common.h"

namespace {

typedef struct { char internal[128]; } nccl_unique_id;
typedef void *nccl_comm_t;
typedef int (*fn_get_unique_id)(nccl_unique_id *);
typedef int (*fn_comm_init_rank)(nccl_comm_t *, int, nccl_unique_id, int);
typedef int (*fn_all_reduce)(const void *, void *, size_t, int, int, nccl_comm_t, cudaStream_t);
typedef int (*fn_comm_destroy)(nccl_comm_t);
typedef const char *(*fn_get_error_string)(int);
typedef int (*fn_group)(void);
typedef int (*fn_all_gather)(const void *, void *, size_t, int, nccl_comm_t, cudaStream_t);
// ncclConfig_t as NCCL 2.17 laid it out (nccl.h, `ncclConfig_v21700`); later releases append fields and accept this prefix,
// told apart by `size` / `version`.  Only maxCTAs is set: the cap on the CTAs (= SMs) a collective of THIS communicator
// occupies.  An environment variable cannot do this reliably: NCCL caches NCCL_MAX_NCHANNELS on first use, and the
// process may have created other communicators (torch.distributed) before this one.
struct nccl_config_v21700 {
  size_t size;
  unsigned int magic, version;
  int blocking, cga_cluster_size, min_ctas, max_ctas;
  const char *net_name;
  int split_share;
};
typedef int (*fn_comm_init_rank_config)(nccl_comm_t *, int, nccl_unique_id, int, nccl_config_v21700 *);

struct NcclApi {
  fn_comm_init_rank_config comm_init_rank_config = nullptr;
  void *handle = nullptr;
  fn_get_unique_id get_unique_id = nullptr;
  fn_comm_init_rank comm_init_rank = nullptr;
  fn_all_reduce all_reduce = nullptr;
  fn_comm_destroy comm_destroy = nullptr;
  fn_get_error_string get_error_string = nullptr;
  fn_group group_start = nullptr, group_end = nullptr;
  fn_all_gather all_gather = nullptr;
};

NcclApi g_nccl;
std::mutex g_nccl_mu;

int load_nccl() {
  std::lock_guard<std::mutex> lk(g_nccl_mu);
  if (g_nccl.handle) return DLS_OK;
  const char *env = getenv("DLS_NCCL_LIB");
  const char *names[] = {env, "libnccl.so.2", "libnccl.so"};
  void *h = nullptr;
  for (const char *n : names) {
    if (!n || !*n) continue;
    h = dlopen(n, RTLD_NOW | RTLD_GLOBAL);
    if (h) break;
  }
  if (!h) {
    dls::set_error(std::string("cannot load libnccl (set DLS_NCCL_LIB): ") + (dlerror() ? dlerror() : "not found"));
    return DLS_ERR_COMM;
  }
  g_nccl.get_unique_id = (fn_get_unique_id)dlsym(h, "ncclGetUniqueId");
  g_nccl.comm_init_rank = (fn_comm_init_rank)dlsym(h, "ncclCommInitRank");
  g_nccl.all_reduce = (fn_all_reduce)dlsym(h, "ncclAllReduce");
  g_nccl.comm_destroy = (fn_comm_destroy)dlsym(h, "ncclCommDestroy");
  g_nccl.get_error_string = (fn_get_error_string)dlsym(h, "ncclGetErrorString");
  g_nccl.group_start = (fn_group)dlsym(h, "ncclGroupStart");
  g_nccl.group_end = (fn_group)dlsym(h, "ncclGroupEnd");
  g_nccl.all_gather = (fn_all_gather)dlsym(h, "ncclAllGather");
  g_nccl.comm_init_rank_config = (fn_comm_init_rank_config)dlsym(h, "ncclCommInitRankConfig");
  if (!g_nccl.get_unique_id || !g_nccl.comm_init_rank || !g_nccl.all_reduce || !g_nccl.comm_destroy) {
    dls::set_error("libnccl is missing a required symbol");
    dlclose(h);
    return DLS_ERR_COMM;
  }
  g_nccl.handle = h;
  return DLS_OK;
}

int nccl_fail(int code, const char *what) {
  std::string m = std::string(what) + " failed: ";
  m += g_nccl.get_error_string ? g_nccl.get_error_string(code) : "nccl error";
  dls::set_error(m);
  return DLS_ERR_COMM;
}

}  // namespace

namespace dls {

// Small host-side all-gather over the context's communicator (setup paths only: it synchronises the stream).
int comm_allgather_bytes(dls_ctx *ctx, const void *mine, void *all_host, size_t bytes) {
  DLS_CHECK(ctx->nccl_comm && g_nccl.all_gather, DLS_ERR_COMM, "all-gather: no communicator, or libnccl lacks ncclAllGather");
  char *dev = nullptr;
  const size_t world = (size_t)ctx->comm_world;
  DLS_CUDA(cudaMalloc((void **)&dev, bytes * (world + 1)));
  cudaError_t ce = cudaMemcpyAsync(dev, mine, bytes, cudaMemcpyHostToDevice, ctx->streams[0]);
  int e = 0;
  if (ce == cudaSuccess) e = g_nccl.all_gather(dev, dev + bytes, bytes, /*ncclInt8*/ 0, ctx->nccl_comm, ctx->streams[0]);
  if (ce == cudaSuccess && !e) ce = cudaMemcpyAsync(all_host, dev + bytes, bytes * world, cudaMemcpyDeviceToHost, ctx->streams[0]);
  if (ce == cudaSuccess && !e) ce = cudaStreamSynchronize(ctx->streams[0]);
  cudaFree(dev);
  if (e) return nccl_fail(e, "ncclAllGather");
  if (ce != cudaSuccess) return cuda_fail(ce, "all-gather staging", __FILE__, __LINE__);
  return DLS_OK;
}

}  // namespace dls

extern "C" {

int dls_comm_unique_id(uint8_t *id128) {
  DLS_CHECK(id128, DLS_ERR_INVALID, "dls_comm_unique_id: NULL argument");
  int rc = load_nccl();
  if (rc) return rc;
  nccl_unique_id id;
  int e = g_nccl.get_unique_id(&id);
  if (e) return nccl_fail(e, "ncclGetUniqueId");
  memcpy(id128, id.internal, 128);
  return DLS_OK;
}

int dls_comm_init(dls_ctx *ctx, const uint8_t *id128, int rank, int world_size) {
  DLS_CHECK(ctx && id128 && world_size >= 1 && rank >= 0 && rank < world_size, DLS_ERR_INVALID, "dls_comm_init: bad argument");
  int rc = load_nccl();
  if (rc) return rc;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CUDA(cudaSetDevice(ctx->device));
  if (ctx->nccl_comm) dls_comm_destroy(ctx);
  nccl_unique_id id;
  memcpy(id.internal, id128, 128);
  nccl_comm_t comm = nullptr;
  // a collective kernel runs one CTA per channel; a persistent one-CTA-per-SM GEMM launched next to it would otherwise
  // queue the CTAs that find their SM taken behind the whole first wave (measured: the overlap hid almost nothing).  The
  // gradient buckets are a few MB: eight CTAs move them under the rest of the backward pass and leave 140 SMs to its kernels
  // (C5 at 8 GPUs, ms per step: 4 CTAs 0.331, 8 CTAs 0.317, 16 CTAs 0.320).
  int ctas = 8;
  if (const char *e = getenv("DLS_EXCHANGE_CTAS")) ctas = atoi(e);
  if (ctas < 1 || ctas > 64) ctas = 8;
  int e = -1;
  if (g_nccl.comm_init_rank_config) {
    nccl_config_v21700 cfg;
    cfg.size = sizeof(cfg);
    cfg.magic = 0xcafebeef;
    cfg.version = 21700;
    cfg.blocking = cfg.cga_cluster_size = cfg.min_ctas = cfg.split_share = INT_MIN;   // NCCL_CONFIG_UNDEF_INT
    cfg.net_name = nullptr;
    cfg.max_ctas = ctas;
    e = g_nccl.comm_init_rank_config(&comm, world_size, id, rank, &cfg);
    if (e) comm = nullptr;
  }
  bool capped = e == 0;
  if (e) e = g_nccl.comm_init_rank(&comm, world_size, id, rank);
  if (e) return nccl_fail(e, "ncclCommInitRank");
  ctx->nccl_comm = comm;
  if (capped) {
    ctx->exchange_sms = ctas;
  } else {
    // no per-communicator cap in this libnccl: fall back to what the environment promises
    const char *env = getenv("NCCL_MAX_NCHANNELS");
    const int ch = env ? atoi(env) : 0;
    ctx->exchange_sms = ch > 0 && ch <= 64 ? ch : 16;
  }
  ctx->comm_rank = rank;
  ctx->comm_world = world_size;
  return DLS_OK;
}

int dls_allreduce_sum(dls_ctx *ctx, dls_buf *buf, size_t n_floats, dls_event *const *deps, int n_deps, dls_event **done) {
  DLS_CHECK(ctx && buf, DLS_ERR_INVALID, "dls_allreduce_sum: NULL argument");
  DLS_CHECK(n_floats <= buf->n, DLS_ERR_INVALID, "dls_allreduce_sum: length exceeds the buffer");
  DLS_CHECK(ctx->nccl_comm, DLS_ERR_COMM, "dls_allreduce_sum: dls_comm_init has not been called on this context");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CUDA(cudaSetDevice(ctx->device));
  if (ctx->n_streams == 1 && !done && n_deps == 0) return dls_allreduce_sum_many(ctx, &buf, &n_floats, 1);
  // multi-stream contexts order everything with per-buffer events already: run in line on the producing stream
  const int stream = buf->last_stream;
  int rc = dls::wait_inputs(ctx, stream, &buf, 1, deps, n_deps);
  if (rc) return rc;
  int e = g_nccl.all_reduce(buf->ptr, buf->ptr, n_floats, /*ncclFloat32*/ 7, /*ncclSum*/ 0, ctx->nccl_comm, ctx->streams[stream]);
  if (e) return nccl_fail(e, "ncclAllReduce");
  return dls::finish_output(ctx, stream, &buf, 1, nullptr, 0, done);
}

// One bucket of gradients: every buffer is SUM-all-reduced in place by ONE grouped NCCL launch on the exchange stream.
// The exchange stream is forked from the compute stream with an event recorded here; the compute stream carries on
// with the rest of the backward pass and only waits (join_exchange, via wait_inputs) when a launch touches one of the
// buffers again -- normally the weight update at the end of the step.  Every rank issues its buckets in program order
// on one stream, so the order matches across ranks.  All of it is recordable: inside a captured step the fork and the
// joins become graph edges.
int dls_allreduce_sum_many(dls_ctx *ctx, dls_buf *const *bufs, const size_t *n_floats, int n_bufs) {
  DLS_CHECK(ctx && (n_bufs == 0 || (bufs && n_floats)) && n_bufs >= 0, DLS_ERR_INVALID, "dls_allreduce_sum_many: bad argument");
  DLS_CHECK(ctx->nccl_comm, DLS_ERR_COMM, "dls_allreduce_sum_many: dls_comm_init has not been called on this context");
  DLS_CHECK(ctx->n_streams == 1, DLS_ERR_UNSUPPORTED, "dls_allreduce_sum_many: single-stream contexts only");
  for (int i = 0; i < n_bufs; ++i)
    DLS_CHECK(bufs[i] && n_floats[i] <= bufs[i]->n, DLS_ERR_INVALID, "dls_allreduce_sum_many: NULL buffer or length exceeds it");
  if (n_bufs == 0) return DLS_OK;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CUDA(cudaSetDevice(ctx->device));
  int rc = dls::wait_inputs(ctx, 0, bufs, n_bufs, nullptr, 0);   // joins an exchange still pending on any of them
  if (rc) return rc;
  if (!ctx->comm_stream) DLS_CUDA(cudaStreamCreateWithFlags(&ctx->comm_stream, cudaStreamNonBlocking));
  cudaEvent_t fork = nullptr;
  DLS_CUDA(cudaEventCreateWithFlags(&fork, cudaEventDisableTiming));
  cudaError_t ce = cudaEventRecord(fork, ctx->streams[0]);
  if (ce == cudaSuccess) ce = cudaStreamWaitEvent(ctx->comm_stream, fork, 0);
  cudaEventDestroy(fork);
  if (ce != cudaSuccess) return dls::cuda_fail(ce, "fork onto the exchange stream", __FILE__, __LINE__);
  const bool group = n_bufs > 1 && g_nccl.group_start && g_nccl.group_end;
  int e = group ? g_nccl.group_start() : 0;
  if (e) return nccl_fail(e, "ncclGroupStart");
  for (int i = 0; i < n_bufs && !e; ++i)
    e = g_nccl.all_reduce(bufs[i]->ptr, bufs[i]->ptr, n_floats[i], /*ncclFloat32*/ 7, /*ncclSum*/ 0, ctx->nccl_comm, ctx->comm_stream);
  if (group) {
    int e2 = g_nccl.group_end();
    if (!e) e = e2;
  }
  if (e) return nccl_fail(e, "ncclAllReduce");
  for (int i = 0; i < n_bufs; ++i) {
    dls_buf *b = bufs[i];
    if (b->comm_ready) continue;   // the same buffer twice in one bucket
    cudaEvent_t joined = nullptr;
    DLS_CUDA(cudaEventCreateWithFlags(&joined, cudaEventDisableTiming));
    ce = cudaEventRecord(joined, ctx->comm_stream);
    if (ce != cudaSuccess) {
      cudaEventDestroy(joined);
      return dls::cuda_fail(ce, "cudaEventRecord(exchange)", __FILE__, __LINE__);
    }
    b->comm_ready = joined;
    ctx->comm_pending.push_back(b);
  }
  return dls::finish_output(ctx, 0, bufs, n_bufs, nullptr, 0, nullptr);
}

int dls_comm_destroy(dls_ctx *ctx) {
  if (!ctx || !ctx->nccl_comm) return DLS_OK;
  cudaSetDevice(ctx->device);
  cudaDeviceSynchronize();
  dls::peer_destroy(ctx);
  if (ctx->comm_stream) {
    cudaStreamSynchronize(ctx->comm_stream);
    cudaStreamDestroy(ctx->comm_stream);
    ctx->comm_stream = nullptr;
  }
  if (g_nccl.comm_destroy) g_nccl.comm_destroy(ctx->nccl_comm);
  ctx->nccl_comm = nullptr;
  ctx->comm_world = 1;
  ctx->comm_rank = 0;
  return DLS_OK;
}

}  // extern "C"
