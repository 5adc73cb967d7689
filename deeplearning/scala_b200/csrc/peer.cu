// peer.cu -- N8: the end-of-step exchange as ONE kernel over NVLink peer memory.
//
// New functionality (the reference has no collective, see comm.cpp).  The tail of a data-parallel step is latency
// bound: the gradients of the first layers appear last, are small (C5: 0.37 MB), and the step cannot finish before they
// are summed over the ranks and every weight is updated.  Over NCCL that tail is one collective launch (25-40 us at 8
// GPUs for any size) followed by one update launch per weight.  Here it is a single launch per rank:
//
//   phase 1  every CTA copies its chunks of this rank's gradients into this rank's staging area (peer visible)
//            and raises one flag per peer (st.release.sys into the peer's flag array);
//   phase 2  the updates that need no exchange -- segments whose gradient was already reduced (the large bucket that NCCL
//            moved under the backward pass), i.e. a multi-tensor SGD -- run on kPeerGrid CTAs of their own in the same
//            launch (on the exchange CTAs themselves when the launch has nothing to exchange);
//   phase 3  it waits for the same CTA of every peer (ld.acquire.sys on its own flag array);
//   phase 4  it reads the chunk from every rank's staging area (remote ranks over NVLink, 16-byte loads, all in
//            flight together), adds them in rank order 0..W-1 -- the same order on every rank, so replicas stay
//            bit-identical -- and applies  w = w - lr * sum  (or writes the sum back for a stateful optimizer).
//
// One-shot: every rank reads W-1 remote copies, (W-1) x bytes over its NVLink ports; that beats a ring or tree below
// a few MB, which is where this path is used (harness.Exchange: `peer_bytes`).  There is no grid-wide barrier: CTA b
// only ever exchanges with CTA b of the peers, and the chunk -> CTA map is a fixed function of the staging offset, so
// the synchronisation is per CTA.  Staging is double buffered on the CTA's epoch parity: a rank that starts epoch e + 2
// has seen every peer's flag for epoch e + 1, which a peer raises only after it finished reading epoch e.  The epoch
// counters live in device memory and advance inside the kernel, so the launch is recordable into a captured step.
//
// The segment is one cudaMalloc per rank, exported with cudaIpcGetMemHandle; the 64-byte handles are all-gathered over
// the NCCL communicator the context already has (comm.cpp), so no extra rendezvous channel is needed.
#include <string.h>

#include "common.h"

namespace dls {

constexpr int kPeerGrid = 128;        // CTAs per launch; the same on every rank (part of the protocol)
constexpr int kPeerBlock = 256;
constexpr int kPeerChunk = kPeerBlock * 4;   // floats per chunk: one float4 per thread
constexpr int kPeerMaxSeg = 12;
constexpr int kPeerMaxWorld = 8;

struct PeerSeg {
  float *w;          // weight to update in place, or nullptr: write the sum back into g
  float *g;          // this rank's gradient
  uint32_t n;        // floats
  uint32_t chunk0;   // first chunk id: staging offset / kPeerChunk for exchanged segments, a running count otherwise
  uint32_t exchange; // 1: sum over the ranks first
  uint32_t pad_;
};

struct PeerArgs {
  PeerSeg seg[kPeerMaxSeg];
  float *stage[kPeerMaxWorld];      // every rank's staging area as mapped into this process: [2][stage_floats]
  uint32_t *flags[kPeerMaxWorld];   // every rank's flag array: [world][kPeerGrid]
  uint32_t *epochs;                 // this rank's [kPeerGrid]
  int *error;                       // host-mapped: a peer did not arrive in time
  uint32_t stage_floats;
  int n_seg, rank, world;
  float lr;
  int split_local;                  // 1: CTAs kPeerGrid .. 2 * kPeerGrid - 1 do phase 2, the first kPeerGrid skip it
};

struct PeerState {
  void *segment = nullptr;            // flags, then the two staging halves
  uint32_t *epochs = nullptr;
  int *error_host = nullptr, *error_dev = nullptr;
  void *peer_base[kPeerMaxWorld] = {};
  size_t stage_floats = 0;
  int world = 0, rank = 0;
};

namespace {

constexpr size_t kFlagBytes = 4096;   // head of the segment: the arrival flags [world][kPeerGrid]
static_assert(kPeerMaxWorld * kPeerGrid * 4 <= kFlagBytes, "flag array does not fit its page");
constexpr unsigned long long kPeerTimeoutNs = 20ull * 1000 * 1000 * 1000;

__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_relaxed_sys_u32(const uint32_t *p) {
  uint32_t v;
  asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ float4 ld_peer(const float *p) {   // never served from this SM's L1
  float4 v;
  asm volatile("ld.relaxed.sys.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}

// four consecutive floats of a segment; the last vector of a segment whose length is not a multiple of four is ragged
__device__ __forceinline__ float4 load4(const float *p, uint32_t i, uint32_t n) {
  if (i + 4 <= n) return *reinterpret_cast<const float4 *>(p + i);
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
  if (i < n) v.x = p[i];
  if (i + 1 < n) v.y = p[i + 1];
  if (i + 2 < n) v.z = p[i + 2];
  return v;
}
__device__ __forceinline__ void store4(float *p, uint32_t i, uint32_t n, float4 v) {
  if (i + 4 <= n) {
    *reinterpret_cast<float4 *>(p + i) = v;
    return;
  }
  if (i < n) p[i] = v.x;
  if (i + 1 < n) p[i + 1] = v.y;
  if (i + 2 < n) p[i + 2] = v.z;
}
__device__ __forceinline__ float4 sgd4(float4 w, float4 s, float lr) {   // data - delta * lr, one rounding per operation
  w.x = __fsub_rn(w.x, __fmul_rn(s.x, lr));
  w.y = __fsub_rn(w.y, __fmul_rn(s.y, lr));
  w.z = __fsub_rn(w.z, __fmul_rn(s.z, lr));
  w.w = __fsub_rn(w.w, __fmul_rn(s.w, lr));
  return w;
}

__global__ void __launch_bounds__(kPeerBlock) peer_step_kernel(const __grid_constant__ PeerArgs A) {
  // A launch with exchanged segments runs 2 * kPeerGrid CTAs: the first kPeerGrid run the exchange (phases 1, 3, 4), the
  // other kPeerGrid only the updates that need no exchange (phase 2) -- the large weight whose gradient NCCL already summed is
  // 25 MB of traffic, and a CTA that did its share of it before looking at its flags put ~6 us in front of the exchange's
  // own latency chain.  Without exchanged segments (single GPU) kPeerGrid CTAs do phase 2 and nothing else.
  const bool does_local = blockIdx.x >= kPeerGrid || A.split_local == 0;
  const bool does_exchange = blockIdx.x < kPeerGrid;
  const uint32_t b = blockIdx.x % kPeerGrid, tid = threadIdx.x;

  // does this CTA own a chunk of an exchanged segment?  (uniform over the CTA, identical on every rank)
  bool mine = false;
  for (int s = 0; does_exchange && s < A.n_seg; ++s) {
    const PeerSeg &S = A.seg[s];
    if (!S.exchange) continue;
    const uint32_t chunks = (S.n + kPeerChunk - 1) / kPeerChunk;
    const uint32_t first = (b + kPeerGrid - S.chunk0 % kPeerGrid) % kPeerGrid;   // smallest c with (chunk0 + c) % grid == b
    mine |= first < chunks;
  }
  uint32_t epoch = 0;
  if (mine) epoch = A.epochs[b] + 1;
  const size_t half = (size_t)(epoch & 1) * A.stage_floats;

  // phase 1: publish this rank's chunks
  if (mine) {
    float *own = A.stage[A.rank] + half;
    for (int s = 0; s < A.n_seg; ++s) {
      const PeerSeg &S = A.seg[s];
      if (!S.exchange) continue;
      const uint32_t chunks = (S.n + kPeerChunk - 1) / kPeerChunk;
      for (uint32_t c = (b + kPeerGrid - S.chunk0 % kPeerGrid) % kPeerGrid; c < chunks; c += kPeerGrid) {
        const uint32_t i = c * kPeerChunk + tid * 4;
        if (i < S.n) *reinterpret_cast<float4 *>(own + (size_t)S.chunk0 * kPeerChunk + i) = load4(S.g, i, S.n);
      }
    }
    // release pattern of a CTA: the barrier orders every thread's staging writes before the signalling threads, whose
    // system-scope release (st.release.sys = fence + store) then publishes them cumulatively -- one fence per peer instead
    // of one per thread (256 system fences per CTA measured ~10 us of the tail)
    __syncthreads();
    if (tid < (uint32_t)A.world && tid != (uint32_t)A.rank) st_release_sys(A.flags[tid] + A.rank * kPeerGrid + b, epoch);
  }

  // phase 2: updates that need no exchange
  for (int s = 0; does_local && s < A.n_seg; ++s) {
    const PeerSeg &S = A.seg[s];
    if (S.exchange || !S.w) continue;
    const uint32_t chunks = (S.n + kPeerChunk - 1) / kPeerChunk;
    uint32_t c = (b + kPeerGrid - S.chunk0 % kPeerGrid) % kPeerGrid;
    for (; c + 3 * kPeerGrid < chunks; c += 4 * kPeerGrid) {   // four chunks in flight; all but the last chunk are full
      float4 g[4], w[4];
      bool full = (c + 3 * kPeerGrid + 1) * (size_t)kPeerChunk <= S.n;
      if (!full) break;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t i = (c + u * kPeerGrid) * kPeerChunk + tid * 4;
        g[u] = *reinterpret_cast<const float4 *>(S.g + i);
        w[u] = *reinterpret_cast<const float4 *>(S.w + i);
      }
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const uint32_t i = (c + u * kPeerGrid) * kPeerChunk + tid * 4;
        *reinterpret_cast<float4 *>(S.w + i) = sgd4(w[u], g[u], A.lr);
      }
    }
    for (; c < chunks; c += kPeerGrid) {
      const uint32_t i = c * kPeerChunk + tid * 4;
      if (i < S.n) store4(S.w, i, S.n, sgd4(load4(S.w, i, S.n), load4(S.g, i, S.n), A.lr));
    }
  }
  if (!mine) return;

  // phase 3: the same CTA of every peer has published its chunks
  if (tid < (uint32_t)A.world && tid != (uint32_t)A.rank) {
    const uint32_t *flag = A.flags[A.rank] + tid * kPeerGrid + b;
    const unsigned long long t0 = global_ns();
    // poll with relaxed loads, acquire once at the end (an acquire per poll makes every iteration a fence)
    while ((int32_t)(ld_relaxed_sys_u32(flag) - epoch) < 0) {
      if (global_ns() - t0 > kPeerTimeoutNs) {   // a peer died or never launched: give up loudly instead of hanging the GPU
        *A.error = 1 + (int)tid;
        break;
      }
    }
    (void)ld_acquire_sys(flag);
  }
  __syncthreads();

  // phase 4: sum in rank order, update
  for (int s = 0; s < A.n_seg; ++s) {
    const PeerSeg &S = A.seg[s];
    if (!S.exchange) continue;
    const uint32_t chunks = (S.n + kPeerChunk - 1) / kPeerChunk;
    for (uint32_t c = (b + kPeerGrid - S.chunk0 % kPeerGrid) % kPeerGrid; c < chunks; c += kPeerGrid) {
      const uint32_t i = c * kPeerChunk + tid * 4;
      if (i >= S.n) continue;
      const size_t off = half + (size_t)S.chunk0 * kPeerChunk + i;
      float4 part[kPeerMaxWorld];
#pragma unroll
      for (int r = 0; r < kPeerMaxWorld; ++r)
        if (r < A.world) part[r] = ld_peer(A.stage[r] + off);
      float4 sum = part[0];
#pragma unroll
      for (int r = 1; r < kPeerMaxWorld; ++r)
        if (r < A.world) {
          sum.x = __fadd_rn(sum.x, part[r].x);
          sum.y = __fadd_rn(sum.y, part[r].y);
          sum.z = __fadd_rn(sum.z, part[r].z);
          sum.w = __fadd_rn(sum.w, part[r].w);
        }
      if (S.w) store4(S.w, i, S.n, sgd4(load4(S.w, i, S.n), sum, A.lr));
      else store4(S.g, i, S.n, sum);
    }
  }
  if (tid == 0) A.epochs[b] = epoch;
}

}  // namespace

int comm_allgather_bytes(dls_ctx *ctx, const void *mine, void *all_host, size_t bytes);   // comm.cpp

// Teardown: the cudaIpc contract makes it undefined behaviour for an exporter to free a segment that an importer still
// has open, and ranks leave at different times with no channel left to agree on an order (the communicator may already
// be gone on the other side).  So a rank closes its imports -- always legal -- and does NOT free the segment it exported:
// the driver reclaims it when the process exits, which is also what happens to every mapping after a crash.  The cost
// is one staging area (16 MB by default) per dls_peer_init for the life of the process.
void peer_destroy(dls_ctx *ctx) {
  PeerState *P = ctx->peer;
  if (!P) return;
  cudaDeviceSynchronize();
  bool imported = false;
  for (int r = 0; r < P->world; ++r) {
    if (r == P->rank || !P->peer_base[r]) continue;
    imported = true;
    cudaIpcCloseMemHandle(P->peer_base[r]);
  }
  if (P->segment && !imported && P->world == 1) cudaFree(P->segment);
  cudaGetLastError();
  if (P->epochs) cudaFree(P->epochs);
  if (P->error_host) cudaFreeHost(P->error_host);
  delete P;
  ctx->peer = nullptr;
}

int peer_check(dls_ctx *ctx) {
  PeerState *P = ctx->peer;
  if (!P || !P->error_host || *P->error_host == 0) return DLS_OK;
  set_error("peer exchange: rank " + std::to_string(*P->error_host - 1) + " did not arrive within 20 s; the step's results are invalid");
  return DLS_ERR_COMM;
}

}  // namespace dls

using namespace dls;

extern "C" {

int dls_peer_init(dls_ctx *ctx, size_t staging_floats) {
  DLS_CHECK(ctx && staging_floats > 0, DLS_ERR_INVALID, "dls_peer_init: bad argument");
  DLS_CHECK(ctx->nccl_comm, DLS_ERR_COMM, "dls_peer_init: dls_comm_init has not been called on this context");
  DLS_CHECK(ctx->comm_world <= kPeerMaxWorld, DLS_ERR_UNSUPPORTED, "dls_peer_init: at most 8 ranks (one NVSwitch domain)");
  DLS_CHECK(ctx->capture_pool == 0, DLS_ERR_INVALID, "dls_peer_init: not inside a captured step");
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  DLS_CUDA(cudaSetDevice(ctx->device));
  peer_destroy(ctx);
  PeerState *P = new PeerState();
  ctx->peer = P;
  P->world = ctx->comm_world;
  P->rank = ctx->comm_rank;
  P->stage_floats = (staging_floats + kPeerChunk - 1) / kPeerChunk * kPeerChunk;
  const size_t bytes = kFlagBytes + 2 * P->stage_floats * sizeof(float);
  cudaError_t e = cudaMalloc(&P->segment, bytes);
  if (e == cudaSuccess) e = cudaMemset(P->segment, 0, bytes);
  if (e == cudaSuccess) e = cudaMalloc((void **)&P->epochs, kPeerGrid * sizeof(uint32_t));
  if (e == cudaSuccess) e = cudaMemset(P->epochs, 0, kPeerGrid * sizeof(uint32_t));
  if (e == cudaSuccess) e = cudaHostAlloc((void **)&P->error_host, sizeof(int), cudaHostAllocMapped);
  if (e == cudaSuccess) {
    *P->error_host = 0;
    e = cudaHostGetDevicePointer((void **)&P->error_dev, P->error_host, 0);
  }
  cudaIpcMemHandle_t mine;
  memset(&mine, 0, sizeof(mine));
  if (e == cudaSuccess) e = cudaIpcGetMemHandle(&mine, P->segment);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();   // the flags are zero before any peer can learn the handle
  // Both all-gathers below are collective: a rank whose setup failed still takes part and reports it in the second one,
  // so that every rank gives up together instead of waiting for it.
  std::vector<cudaIpcMemHandle_t> all(P->world);
  int rc = comm_allgather_bytes(ctx, &mine, all.data(), sizeof(mine));
  if (rc) {
    peer_destroy(ctx);
    return rc;
  }
  for (int r = 0; r < P->world && e == cudaSuccess; ++r) {
    if (r == P->rank) {
      P->peer_base[r] = P->segment;
      continue;
    }
    e = cudaIpcOpenMemHandle(&P->peer_base[r], all[r], cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) P->peer_base[r] = nullptr;
  }
  // second round: nobody launches an exchange before every rank has mapped every segment
  char ok = e == cudaSuccess ? 1 : 0;
  std::vector<char> oks(P->world);
  rc = comm_allgather_bytes(ctx, &ok, oks.data(), 1);
  if (e != cudaSuccess) {
    cudaGetLastError();
    peer_destroy(ctx);
    return cuda_fail(e, "peer segment setup (cudaIpc needs the ranks on peer-accessible GPUs of one node)", __FILE__, __LINE__);
  }
  if (rc) {
    peer_destroy(ctx);
    return rc;
  }
  for (int r = 0; r < P->world; ++r)
    if (!oks[r]) {
      peer_destroy(ctx);
      set_error("dls_peer_init: rank " + std::to_string(r) + " could not map the peer segments");
      return DLS_ERR_COMM;
    }
  return DLS_OK;
}

int dls_peer_ready(dls_ctx *ctx) { return ctx && ctx->peer ? 1 : 0; }

// One launch: for every i, grads[i] is summed over the ranks when exchange[i] != 0, then
// weights[i] = weights[i] - lr * grads[i] in place; with weights == NULL the sums are written back into grads[i].
int dls_peer_sgd_step(dls_ctx *ctx, dls_buf *const *weights, dls_buf *const *grads, const size_t *n_floats, const int *exchange,
                      int n, float lr) {
  DLS_CHECK(ctx && grads && n_floats && n >= 0, DLS_ERR_INVALID, "dls_peer_sgd_step: bad argument");
  if (n == 0) return DLS_OK;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  PeerState *P = ctx->peer;
  bool any_exchange = false;
  for (int i = 0; i < n; ++i) {
    DLS_CHECK(grads[i] && n_floats[i] <= grads[i]->n && n_floats[i] < (1ull << 31), DLS_ERR_INVALID,
              "dls_peer_sgd_step: NULL gradient or length exceeds it");
    DLS_CHECK(!weights || (weights[i] && n_floats[i] <= weights[i]->n && weights[i] != grads[i]), DLS_ERR_INVALID,
              "dls_peer_sgd_step: NULL weight, length exceeds it, or weight aliases its gradient");
    DLS_CHECK(weights || (exchange && exchange[i]), DLS_ERR_INVALID, "dls_peer_sgd_step: nothing to do for a segment");
    any_exchange |= exchange && exchange[i];
  }
  DLS_CHECK(!any_exchange || P, DLS_ERR_COMM, "dls_peer_sgd_step: dls_peer_init has not been called on this context");
  DLS_CHECK(ctx->n_streams == 1, DLS_ERR_UNSUPPORTED, "dls_peer_sgd_step: single-stream contexts only");
  DLS_CUDA(cudaSetDevice(ctx->device));
  int rc = peer_check(ctx);
  if (rc) return rc;
  std::vector<dls_buf *> touched;
  for (int i = 0; i < n; ++i) {
    touched.push_back(grads[i]);
    if (weights) touched.push_back(weights[i]);
  }
  rc = wait_inputs(ctx, 0, touched.data(), (int)touched.size(), nullptr, 0);
  if (rc) return rc;

  PeerArgs A;
  memset(&A, 0, sizeof(A));
  A.lr = lr;
  A.rank = P ? P->rank : 0;
  A.world = P ? P->world : 1;
  if (P) {
    for (int r = 0; r < P->world; ++r) {
      A.flags[r] = reinterpret_cast<uint32_t *>(P->peer_base[r]);
      A.stage[r] = reinterpret_cast<float *>(static_cast<char *>(P->peer_base[r]) + kFlagBytes);
    }
    A.epochs = P->epochs;
    A.error = P->error_dev;
    A.stage_floats = (uint32_t)P->stage_floats;
  }
  auto launch = [&]() -> int {
    if (A.n_seg == 0) return DLS_OK;
    bool any_ex = false, any_local = false;
    for (int i = 0; i < A.n_seg; ++i) {
      any_ex |= A.seg[i].exchange != 0;
      any_local |= A.seg[i].exchange == 0 && A.seg[i].w != nullptr;
    }
    A.split_local = any_ex && any_local ? 1 : 0;
    peer_step_kernel<<<A.split_local ? 2 * kPeerGrid : kPeerGrid, kPeerBlock, 0, ctx->streams[0]>>>(A);
    DLS_CUDA(cudaGetLastError());
    ctx->stats.kernel_launches++;
    A.n_seg = 0;
    return DLS_OK;
  };
  // a launch takes up to kPeerMaxSeg segments and one staging half; every rank cuts at the same places because the
  // cut depends on the sizes only
  uint32_t stage_chunk = 0, local_chunk = 0;
  for (int i = 0; i < n; ++i) {
    const bool ex = exchange && exchange[i];
    size_t done = 0;
    do {   // an exchanged segment larger than the staging half goes out in pieces
      size_t piece = n_floats[i] - done;
      if (ex) {
        const size_t room = P->stage_floats - (size_t)stage_chunk * kPeerChunk;
        if (room == 0 || A.n_seg == kPeerMaxSeg) {
          rc = launch();
          if (rc) return rc;
          stage_chunk = 0;
          continue;
        }
        if (piece > room) piece = room;
      } else if (A.n_seg == kPeerMaxSeg) {
        rc = launch();
        if (rc) return rc;
        stage_chunk = 0;
      }
      PeerSeg &S = A.seg[A.n_seg++];
      S.w = weights ? weights[i]->ptr + done : nullptr;
      S.g = grads[i]->ptr + done;
      S.n = (uint32_t)piece;
      S.exchange = ex ? 1u : 0u;
      const uint32_t chunks = (uint32_t)((piece + kPeerChunk - 1) / kPeerChunk);
      if (ex) {
        S.chunk0 = stage_chunk;
        stage_chunk += chunks;
      } else {
        S.chunk0 = local_chunk;
        local_chunk += chunks;
      }
      done += piece;
    } while (done < n_floats[i]);
  }
  rc = launch();
  if (rc) return rc;
  if (weights) return finish_output(ctx, 0, weights, n, grads, n, nullptr);
  return finish_output(ctx, 0, grads, n, nullptr, 0, nullptr);
}

}  // extern "C"
