// comm.cpp -- data-parallel exchange (N7): one NCCL all-reduce of the packed fp32 gradient arena over NVLink 5 /
// NVSwitch.  New functionality: the reference has no collective (its only multi-device mechanism is the shared
// command-queue pool, OpenCL:423-448).  libnccl is bound at run time with dlopen so that the library loads on
// machines without NCCL; DLS_NCCL_LIB overrides the library path.
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

struct NcclApi {
  void *handle = nullptr;
  fn_get_unique_id get_unique_id = nullptr;
  fn_comm_init_rank comm_init_rank = nullptr;
  fn_all_reduce all_reduce = nullptr;
  fn_comm_destroy comm_destroy = nullptr;
  fn_get_error_string get_error_string = nullptr;
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
  int e = g_nccl.comm_init_rank(&comm, world_size, id, rank);
  if (e) return nccl_fail(e, "ncclCommInitRank");
  ctx->nccl_comm = comm;
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
  const int stream = buf->last_stream;  // in order behind the kernels that produced the gradients
  int rc = dls::wait_inputs(ctx, stream, &buf, 1, deps, n_deps);
  if (rc) return rc;
  int e = g_nccl.all_reduce(buf->ptr, buf->ptr, n_floats, /*ncclFloat32*/ 7, /*ncclSum*/ 0, ctx->nccl_comm, ctx->streams[stream]);
  if (e) return nccl_fail(e, "ncclAllReduce");
  return dls::finish_output(ctx, stream, &buf, 1, nullptr, 0, done);
}

int dls_comm_destroy(dls_ctx *ctx) {
  if (!ctx || !ctx->nccl_comm) return DLS_OK;
  if (g_nccl.comm_destroy) g_nccl.comm_destroy(ctx->nccl_comm);
  ctx->nccl_comm = nullptr;
  ctx->comm_world = 1;
  ctx->comm_rank = 0;
  return DLS_OK;
}

}  // extern "C"
