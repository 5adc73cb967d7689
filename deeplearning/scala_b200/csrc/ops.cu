// ops.cu -- compute entry points of the C ABI: tree launches through the plan cache, reductions, RNG, SGD.
#include <string.h>

#include "common.h"
#include "gemm_tc.h"

using namespace dls;

namespace dls {

// Plan cache lookup keyed by the serialised tree (replaces kernelCache.getIfPresent / get(loader), Tensors:1165, :1258).
static int get_plan(dls_ctx *ctx, const uint8_t *tree, size_t len, std::shared_ptr<CachedPlan> *out, bool insert) {
  const uint64_t h = hash_bytes(tree, len);
  auto range = ctx->plans.equal_range(h);
  for (auto it = range.first; it != range.second; ++it) {
    CachedPlan &cp = *it->second;
    if (cp.blob.size() == len && memcmp(cp.blob.data(), tree, len) == 0) {
      if (insert) ctx->stats.plan_cache_hits++;
      *out = it->second;
      return DLS_OK;
    }
  }
  if (!insert) {
    out->reset();
    return DLS_OK;
  }
  auto cp = std::make_shared<CachedPlan>();
  std::string err;
  int rc = lower_tree(tree, len, &cp->plan, &err);
  if (rc) {
    set_error(err);
    return rc;
  }
  cp->blob.assign(tree, tree + len);
  ctx->stats.plan_cache_misses++;
  const size_t cap = ctx->opts.plan_cache_capacity > 0 ? (size_t)ctx->opts.plan_cache_capacity : 4096;
  // crude bound: drop everything (plans are cheap to rebuild, no compilation involved).  Captured steps have the
  // device tables of the plans they recorded baked into their kernel nodes, and a device-wide synchronise would
  // invalidate an open capture: while a capture is open or a captured step is alive the cache is allowed to grow.
  if (ctx->plans.size() >= cap && ctx->capture_pool == 0 && ctx->graphs.empty()) {
    cudaDeviceSynchronize();
    for (auto &kv : ctx->plans)
      if (kv.second->plan.d_tables) cudaFree(kv.second->plan.d_tables);
    ctx->plans.clear();
  }
  ctx->plans.emplace(h, cp);
  *out = cp;
  return DLS_OK;
}

}  // namespace dls

extern "C" {

int dls_launch_tree(dls_ctx *ctx, const uint8_t *tree, size_t tree_len, dls_buf *const *params, int n_params,
                    dls_event *const *deps, int n_deps, uint32_t flags, dls_buf **out, dls_event **done) {
  DLS_CHECK(ctx && tree && out, DLS_ERR_INVALID, "dls_launch_tree: NULL argument");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CUDA(cudaSetDevice(ctx->device));
  std::shared_ptr<CachedPlan> cp;
  int rc = get_plan(ctx, tree, tree_len, &cp, true);
  if (rc) return rc;
  Plan &P = cp->plan;
  DLS_CHECK(n_params == P.n_params, DLS_ERR_INVALID, "dls_launch_tree: parameter count does not match the tree");
  for (int i = 0; i < n_params; ++i) {
    DLS_CHECK(params && params[i], DLS_ERR_INVALID, "dls_launch_tree: NULL parameter buffer");
    DLS_CHECK((long long)params[i]->n >= P.params[i].length(), DLS_ERR_INVALID,
              "dls_launch_tree: parameter buffer shorter than its declared shape");
  }
  const bool reduce = (flags & DLS_LAUNCH_REDUCE_SUM) != 0;
  const int stream = pick_stream(ctx);
  rc = wait_inputs(ctx, stream, params, n_params, deps, n_deps);
  if (rc) return rc;
  dls_buf *o = nullptr;
  rc = new_buf(ctx, reduce ? 1 : (size_t)P.n_elems * P.n_roots, stream, &o);
  if (rc) return rc;
  // a tree that is nothing but one dense contraction (the unrolled matmuls of TensorsSpec:337-344 / :371-383) runs on the
  // tcgen05 GEMM when the context asks for tensor-core contractions; in fp32 mode the folded device loop keeps the
  // chain's exact left-fold rounding
  const Plan::Contraction &ct = P.contraction;
  if (ct.valid && !reduce && ctx->opts.gemm_mode != DLS_GEMM_FP32)
    rc = run_matmul(ctx, stream, ctx->opts.gemm_mode, params[ct.param_a], params[ct.param_b], nullptr, (size_t)ct.m, (size_t)ct.n,
                    (size_t)ct.k, ct.transpose_a, ct.transpose_b, 0, 0, o);
  else
    rc = launch_plan(ctx, stream, *cp, params, o, reduce);
  if (!rc) rc = finish_output(ctx, stream, &o, 1, params, n_params, done);
  if (rc) {
    dls_buf_release(o);
    return rc;
  }
  *out = o;
  return DLS_OK;
}

int dls_plan_cache_contains(dls_ctx *ctx, const uint8_t *tree, size_t tree_len, int *found) {
  DLS_CHECK(ctx && tree && found, DLS_ERR_INVALID, "dls_plan_cache_contains: NULL argument");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  std::shared_ptr<CachedPlan> cp;
  get_plan(ctx, tree, tree_len, &cp, false);
  *found = cp ? 1 : 0;
  return DLS_OK;
}

int dls_plan_describe(dls_ctx *ctx, const uint8_t *tree, size_t tree_len, char *text, size_t cap) {
  (void)ctx;  // lowering is device independent: callable without a context (and without a GPU)
  DLS_CHECK(tree && text && cap > 0, DLS_ERR_INVALID, "dls_plan_describe: NULL argument");
  Plan P;
  std::string err;
  int rc = lower_tree(tree, tree_len, &P, &err);
  if (rc) {
    set_error(err);
    return rc;
  }
  std::string d = P.describe();
  snprintf(text, cap, "%s", d.c_str());
  return DLS_OK;
}

int dls_sum(dls_ctx *ctx, dls_buf *in, size_t n_floats, dls_event *const *deps, int n_deps, dls_buf **out, dls_event **done) {
  DLS_CHECK(ctx && in && out, DLS_ERR_INVALID, "dls_sum: NULL argument");
  DLS_CHECK(n_floats <= in->n, DLS_ERR_INVALID, "dls_sum: length exceeds the buffer");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CUDA(cudaSetDevice(ctx->device));
  const int stream = pick_stream(ctx);
  int rc = wait_inputs(ctx, stream, &in, 1, deps, n_deps);
  if (rc) return rc;
  dls_buf *o = nullptr;
  rc = new_buf(ctx, 1, stream, &o);
  if (rc) return rc;
  rc = launch_sum(ctx, stream, in->ptr, n_floats, o->ptr);
  if (!rc) rc = finish_output(ctx, stream, &o, 1, &in, 1, done);
  if (rc) {
    dls_buf_release(o);
    return rc;
  }
  *out = o;
  return DLS_OK;
}

int dls_sum_axis(dls_ctx *ctx, dls_buf *in, size_t outer, size_t k, size_t inner, dls_event *const *deps, int n_deps,
                 dls_buf **out, dls_event **done) {
  DLS_CHECK(ctx && in && out, DLS_ERR_INVALID, "dls_sum_axis: NULL argument");
  DLS_CHECK(outer * k * inner <= in->n, DLS_ERR_INVALID, "dls_sum_axis: view exceeds the buffer");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CUDA(cudaSetDevice(ctx->device));
  const int stream = pick_stream(ctx);
  int rc = wait_inputs(ctx, stream, &in, 1, deps, n_deps);
  if (rc) return rc;
  dls_buf *o = nullptr;
  rc = new_buf(ctx, outer * inner, stream, &o);
  if (rc) return rc;
  rc = launch_sum_axis(ctx, stream, in->ptr, outer, k, inner, o->ptr);
  if (!rc) rc = finish_output(ctx, stream, &o, 1, &in, 1, done);
  if (rc) {
    dls_buf_release(o);
    return rc;
  }
  *out = o;
  return DLS_OK;
}

static int rng_common(dls_ctx *ctx, size_t n_floats, uint32_t seed, bool normal, dls_buf **out, dls_event **done) {
  DLS_CHECK(ctx && out, DLS_ERR_INVALID, "dls_random: NULL argument");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CUDA(cudaSetDevice(ctx->device));
  const int stream = pick_stream(ctx);
  // Tensors:508-518: the normal generator writes pairs, odd sizes allocate one extra float
  const size_t padded = normal ? n_floats + (n_floats & 1) : n_floats;
  dls_buf *o = nullptr;
  int rc = new_buf(ctx, padded, stream, &o);
  if (rc) return rc;
  o->n = n_floats;
  rc = normal ? launch_random_normal(ctx, stream, o->ptr, padded, seed, ctx->opts.rng_hash)
              : launch_random(ctx, stream, o->ptr, n_floats, seed, ctx->opts.rng_hash);
  if (!rc) rc = finish_output(ctx, stream, &o, 1, nullptr, 0, done);
  if (rc) {
    dls_buf_release(o);
    return rc;
  }
  *out = o;
  return DLS_OK;
}

int dls_random(dls_ctx *ctx, size_t n_floats, uint32_t seed, dls_buf **out, dls_event **done) {
  return rng_common(ctx, n_floats, seed, false, out, done);
}

int dls_random_normal(dls_ctx *ctx, size_t n_floats, uint32_t seed, dls_buf **out, dls_event **done) {
  return rng_common(ctx, n_floats, seed, true, out, done);
}

int dls_sgd_update(dls_ctx *ctx, dls_buf *weight, dls_buf *delta, size_t n_floats, float learning_rate,
                   dls_event *const *deps, int n_deps, dls_event **done) {
  DLS_CHECK(ctx && weight && delta, DLS_ERR_INVALID, "dls_sgd_update: NULL argument");
  DLS_CHECK(n_floats <= weight->n && n_floats <= delta->n, DLS_ERR_INVALID, "dls_sgd_update: length exceeds a buffer");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CUDA(cudaSetDevice(ctx->device));
  const int stream = pick_stream(ctx);
  dls_buf *ins[2] = {weight, delta};
  int rc = wait_inputs(ctx, stream, ins, 2, deps, n_deps);
  if (!rc) rc = wait_readers(ctx, stream, weight);
  if (rc) return rc;
  rc = launch_sgd(ctx, stream, weight->ptr, delta->ptr, n_floats, learning_rate);
  if (rc) return rc;
  return finish_output(ctx, stream, &weight, 1, &delta, 1, done);
}

int dls_optimizer_step(dls_ctx *ctx, int kind, dls_buf *weight, dls_buf *state1, dls_buf *state2, dls_buf *grad, size_t n_floats,
                       dls_buf *hyper) {
  DLS_CHECK(ctx && weight && state1 && grad && hyper, DLS_ERR_INVALID, "dls_optimizer_step: NULL argument");
  DLS_CHECK(kind == DLS_OPT_MOMENTUM || kind == DLS_OPT_ADAM, DLS_ERR_INVALID, "dls_optimizer_step: unknown optimizer kind");
  DLS_CHECK(kind != DLS_OPT_ADAM || state2, DLS_ERR_INVALID, "dls_optimizer_step: Adam needs two state buffers");
  DLS_CHECK(hyper->n >= 6, DLS_ERR_INVALID, "dls_optimizer_step: the hyper-parameter buffer holds six floats");
  DLS_CHECK(n_floats <= weight->n && n_floats <= state1->n && n_floats <= grad->n && (!state2 || n_floats <= state2->n),
            DLS_ERR_INVALID, "dls_optimizer_step: length exceeds a buffer");
  DLS_CHECK(weight != grad && state1 != grad && state1 != weight && state2 != weight && state2 != state1 && state2 != grad,
            DLS_ERR_INVALID, "dls_optimizer_step: buffers must be distinct");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CUDA(cudaSetDevice(ctx->device));
  const int stream = pick_stream(ctx);
  dls_buf *ins[5] = {weight, state1, grad, hyper, state2};
  int rc = wait_inputs(ctx, stream, ins, state2 ? 5 : 4, nullptr, 0);
  if (!rc) rc = wait_readers(ctx, stream, weight);
  if (!rc) rc = wait_readers(ctx, stream, state1);
  if (!rc) rc = wait_readers(ctx, stream, state2);
  if (rc) return rc;
  rc = launch_optimizer(ctx, stream, kind, weight->ptr, state1->ptr, state2 ? state2->ptr : nullptr, grad->ptr, n_floats, hyper->ptr);
  if (rc) return rc;
  dls_buf *outs[3] = {weight, state1, state2};
  dls_buf *reads[2] = {grad, hyper};
  return finish_output(ctx, stream, outs, state2 ? 3 : 2, reads, 2, nullptr);
}

}  // extern "C"
