// gemm_tc.cu -- N4: dense contraction on the 5th-generation tensor cores (tcgen05.mma, accumulators in TMEM, operands
// staged by TMA), hand-written for sm_100a.  D[M,N] (fp32, row-major) = A * B with BF16 or TF32 operands and fp32
// accumulation, plus an optional fused epilogue (bias along rows or columns, tanh / relu).
//
// This is the lowering target for the contractions the reference can only express as host-unrolled elementwise trees:
// the two matmul formulations of TensorsSpec:337-344 / :371-383, the im2col convolution of TensorsSpec:144-210 and the
// weight-gradient sums of TensorLayers.sumAs (TL:124-162).  There is nothing comparable in the reference's device layer.
//
// Kernel anatomy (one persistent CTA per SM, 192 threads):
//   warp 0, lane 0 : TMA producer   -- cp.async.bulk.tensor.2d into a 4-stage ring of 128B-swizzled smem tiles
//   warp 1, lane 0 : MMA issuer     -- tcgen05.mma.cta_group::1 (M=128, N=256, K=32 bytes per instruction), commit -> mbarrier
//   warps 2..5     : epilogue       -- tcgen05.ld (32 lanes x 32 columns) -> registers -> swizzled smem -> TMA store
// Two 256-column accumulators (all 512 TMEM columns) are double buffered, so the epilogue of tile i overlaps the main
// loop of tile i+1.  Operands may be K-major ([rows][K], K contiguous) or MN-major ([K][rows], rows contiguous): the
// major-ness only changes the TMA boxes and the UMMA shared-memory descriptors, never the data in HBM.
#include <cuda.h>
#include <cuda_bf16.h>

#include "common.h"
#include "gemm_tc.h"

namespace dls {

namespace tc {

constexpr int BM = 128;              // UMMA M (one TMEM lane per output row)
constexpr int BN = 256;              // UMMA N (one TMEM column per output column)
constexpr int STAGES = 4;
constexpr int A_STAGE_BYTES = BM * 128;
constexpr int B_STAGE_BYTES = BN * 128;
constexpr int STAGE_BYTES = A_STAGE_BYTES + B_STAGE_BYTES;   // 48 KiB
constexpr int EPI_WARPS = 4;
constexpr int EPI_CHUNK = 32;                                // columns per tcgen05.ld / TMA-store box
constexpr int EPI_BUF_BYTES = 32 * EPI_CHUNK * 4;            // one warp's 32 rows x 32 fp32 columns
constexpr int EPI_BYTES = EPI_WARPS * 2 * EPI_BUF_BYTES;     // double buffered per warp: 32 KiB
constexpr int NUM_THREADS = 32 * (2 + EPI_WARPS);
constexpr int BAR_BYTES = 256;
constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + BAR_BYTES + 1024;  // + slack for 1 KiB alignment
constexpr int TMEM_COLS = 512;

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
// Spin on the phase parity; a wait that lasts seconds can only be a protocol bug, so trap instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok = 0;
  long long t0 = 0;
  for (uint32_t spins = 0;; ++spins) {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(bar), "r"(parity)
        : "memory");
    if (ok) return;
    if ((spins & 0xFFF) == 0xFFF) {
      const long long now = clock64();
      if (t0 == 0) t0 = now;
      else if (now - t0 > 8000000000LL) __trap();
    }
  }
}

__device__ __forceinline__ void tma_load_2d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
      "l"(map), "r"(bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tma_load_3d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];" ::"r"(dst),
      "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tma_load_4d(uint32_t dst, const CUtensorMap *map, uint32_t bar, int c0, int c1, int c2, int c3) {
  asm volatile(
      "cp.async.bulk.tensor.4d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5, %6}], [%2];" ::"r"(dst),
      "l"(map), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3)
      : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap *map, uint32_t src, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];" ::"l"(map), "r"(src), "r"(c0),
               "r"(c1)
               : "memory");
}

__device__ __forceinline__ void umma(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate,
                                     bool tf32) {
  if (tf32) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
  }
}
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
}

// UMMA shared-memory matrix descriptor (cute::UMMA::SmemDescriptor bit layout): start address, leading / stride byte
// offsets in 16-byte units, version 1 (Blackwell), layout type 2 = SWIZZLE_128B.
// layout: 2 = SWIZZLE_128B (16-byte chunks XORed with the row of an 8-row atom), 1 = SWIZZLE_128B_BASE32B (32-byte chunks XORed
// with the row of a 4-row atom: the form tcgen05 reads MN-major 32-bit operands in; TMA: CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B)
__device__ __forceinline__ uint64_t smem_desc(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes, uint32_t layout = 2) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)layout << 61;
  return d;
}

struct Tile {
  int m0, n0;
};
// Grouped raster: 8 tile-rows share their B tiles through L2 while the group sweeps N.
__device__ __forceinline__ Tile tile_coord(int t, int tiles_m, int tiles_n, int bn = BN) {
  constexpr int GROUP = 8;
  const int per_group = GROUP * tiles_n;
  const int g = t / per_group;
  const int first_m = g * GROUP;
  const int gsz = min(GROUP, tiles_m - first_m);
  const int r = t - g * per_group;
  Tile c;
  c.m0 = (first_m + r % gsz) * BM;
  c.n0 = (r / gsz) * bn;
  return c;
}

// Implicit-GEMM convolution: pixel tile `t` of an NHWC tensor -> (first image, first row).  A tile is `bb` whole images
// x `bh` rows x the full width, `bb * bh * W` = 128 (CONV 1: the rows of one output tile) or = BLOCK_K (CONV 2: one k block).
__device__ __forceinline__ void pixel_tile(const ConvGeom &g, int t, int &b0, int &y0) {
  if (g.bb == 1) {
    b0 = t / g.tiles_per_img;
    y0 = (t - b0 * g.tiles_per_img) * g.bh;
  } else {
    b0 = t * g.bb;
    y0 = 0;
  }
}

#ifdef DLS_GEMM_TRACE
// Debug build only (make NVFLAGS_EXTRA=-DDLS_GEMM_TRACE): CTA 0 records (event id, globaltimer ns) pairs of its pipeline,
// 160 fire-and-forget slots per warp (no atomics: a returning atomic costs about a microsecond and would distort the trace).
__device__ unsigned long long g_gemm_trace[1024];
__device__ unsigned int g_gemm_trace_n;
#define DLS_TRACE(id)                                                                                  \
  do {                                                                                                 \
    if (blockIdx.x == 0 && tr_n < 160) {                                                               \
      unsigned long long t_;                                                                           \
      asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_));                                           \
      g_gemm_trace[(threadIdx.x >> 5) * 160 + tr_n++] = ((unsigned long long)(id) << 48) | (t_ & 0xFFFFFFFFFFFFull); \
    }                                                                                                  \
  } while (0)
#define DLS_TRACE_DECL unsigned tr_n = 0;
#else
#define DLS_TRACE(id)
#define DLS_TRACE_DECL
#endif

// 32 lanes x 32 consecutive fp32 columns of the accumulator -> 32 registers per lane (asynchronous until tcgen05.wait::ld)
__device__ __forceinline__ void tmem_ld_32x32(uint32_t (&v)[32], uint32_t taddr) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}

// Epilogue arithmetic of one 32-column chunk.  The (bias, activation) combination is dispatched ONCE per chunk and each
// combination is a straight-line loop: an epilogue warp is alone on its scheduler, so a per-element branch costs its
// full pipeline latency with nothing to hide it (measured: 1.25 us per chunk with the branches inside the loop).
template <int BIAS, int ACT>
__device__ __forceinline__ void epi_math(const uint32_t (&v)[32], float (&f)[32], float row_bias, const float *__restrict__ bias,
                                         int col0, int N) {
#pragma unroll
  for (int j = 0; j < 32; ++j) {
    float x = __uint_as_float(v[j]);
    if (BIAS == 1) x = __fadd_rn(x, row_bias);
    if (BIAS == 2) x = __fadd_rn(x, (col0 + j < N) ? __ldg(bias + col0 + j) : 0.f);
    if (ACT == 1) x = tanhf(x);
    if (ACT == 2) x = fmaxf(x, 0.f);
    f[j] = x;
  }
}

__device__ __forceinline__ void epi_apply(const uint32_t (&v)[32], float (&f)[32], int bias_mode, int act, float row_bias,
                                          const float *__restrict__ bias, int col0, int N) {
  switch (bias_mode * 3 + act) {
    case 0: epi_math<0, 0>(v, f, row_bias, bias, col0, N); break;
    case 1: epi_math<0, 1>(v, f, row_bias, bias, col0, N); break;
    case 2: epi_math<0, 2>(v, f, row_bias, bias, col0, N); break;
    case 3: epi_math<1, 0>(v, f, row_bias, bias, col0, N); break;
    case 4: epi_math<1, 1>(v, f, row_bias, bias, col0, N); break;
    case 5: epi_math<1, 2>(v, f, row_bias, bias, col0, N); break;
    case 6: epi_math<2, 0>(v, f, row_bias, bias, col0, N); break;
    case 7: epi_math<2, 1>(v, f, row_bias, bias, col0, N); break;
    default: epi_math<2, 2>(v, f, row_bias, bias, col0, N); break;
  }
}

// x *= act'(g) for the 32 columns of a chunk (`left` = columns that exist from col0 on): the delta of the layer below in the
// same kernel.  One rounding per node of `delta * (1 - y * y)` / `delta * min(y * FLT_MAX, 1)`, like the elementwise kernels.
__device__ __forceinline__ void epi_gate(float (&f)[32], const float *g, int left, int mode, bool vec, float scale) {
  float h[32];
  if (vec) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      float4 q = make_float4(0.f, 0.f, 0.f, 0.f);
      if (4 * j < left) q = *(reinterpret_cast<const float4 *>(g) + j);   // plain loads: in mode 3 the gate is the output buffer
      h[4 * j] = q.x;
      h[4 * j + 1] = q.y;
      h[4 * j + 2] = q.z;
      h[4 * j + 3] = q.w;
    }
  } else {
#pragma unroll
    for (int j = 0; j < 32; ++j) h[j] = j < left ? g[j] : 0.f;
  }
  if (mode == 1) {
#pragma unroll
    for (int j = 0; j < 32; ++j) f[j] = __fmul_rn(f[j], __fsub_rn(1.0f, __fmul_rn(h[j], h[j])));
  } else if (mode == 3) {
#pragma unroll
    for (int j = 0; j < 32; ++j) f[j] = __fsub_rn(h[j], __fmul_rn(f[j], scale));
  } else {
#pragma unroll
    for (int j = 0; j < 32; ++j) f[j] = __fmul_rn(f[j], fminf(__fmul_rn(h[j], 3.4028234663852886e38f), 1.0f));
  }
}

// the chunk once more as bf16 (round to nearest even), 16 bytes per store; `left` = elements to the end of the copy's row
__device__ __forceinline__ void epi_copy_bf16(const float (&f)[32], __nv_bfloat16 *dst, long long left) {
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    if (8 * q < left) {
      __nv_bfloat162 p0 = __floats2bfloat162_rn(f[8 * q], f[8 * q + 1]), p1 = __floats2bfloat162_rn(f[8 * q + 2], f[8 * q + 3]);
      __nv_bfloat162 p2 = __floats2bfloat162_rn(f[8 * q + 4], f[8 * q + 5]), p3 = __floats2bfloat162_rn(f[8 * q + 6], f[8 * q + 7]);
      uint4 o;
      o.x = *reinterpret_cast<uint32_t *>(&p0);
      o.y = *reinterpret_cast<uint32_t *>(&p1);
      o.z = *reinterpret_cast<uint32_t *>(&p2);
      o.w = *reinterpret_cast<uint32_t *>(&p3);
      reinterpret_cast<uint4 *>(dst)[q] = o;
    }
  }
}

// CONV selects where the operand tiles come from:
//   0  plain GEMM: 2-D boxes of A and B.
//   1  implicit-GEMM 3x3 'same' convolution, forward or backward-data (TensorsSpec:144-210 without the im2col matrix):
//      M = pixels of an NHWC tensor, K = 9 taps x channels.  The A tile of tap (oy, ox) is ONE 4-D TMA box
//      (channels, W, rows, images) whose start is shifted by the tap; whatever falls outside the image is zero-filled by
//      the TMA unit -- that fill IS the 'same' padding, and images never bleed into each other because the batch is its
//      own tensor dimension.  B = the weights of that tap, a 3-D box of [9][N][K] (K-major).
//   2  convolution weight gradient: M = 9 taps x channels, N = output channels, K = pixels; both operands MN-major, A again
//      through tap-shifted 4-D boxes.
template <bool TF32, bool A_MN, bool B_MN, int CONV>
__global__ void __launch_bounds__(NUM_THREADS, 1)
    gemm_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                   const __grid_constant__ CUtensorMap map_c, float *C, int M, int N, int K, long long ldc,
                   const float *__restrict__ bias, int bias_mode, int act, int direct_store, int splits, int kb_per_split,
                   int c_rows, const ConvGeom cg, const float *gate, int gate_mode, void *__restrict__ copy, long long copy_ld,
                   float gate_scale) {
  // splits > 1 (split-K): work item w = split * n_tiles + tile covers k blocks [split * kb_per_split, ...) and writes its
  // partial tile at row offset split * tiles_m * BM of a [splits * tiles_m * BM, N] scratch matrix (c_rows = its row count);
  // the epilogue (bias / activation) then runs in splitk_reduce_kernel.
  constexpr int ELEM = TF32 ? 4 : 2;
  constexpr int BLOCK_K = 128 / ELEM;       // one 128-byte swizzle row of K per stage
  constexpr int UMMA_K = 32 / ELEM;         // K per tcgen05.mma
  constexpr int MN_ATOM = 128 / ELEM;       // rows covered by one 128-byte line of an MN-major operand

  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t *epi_smem = smem + STAGES * STAGE_BYTES;
  uint64_t *bars = (uint64_t *)(epi_smem + EPI_BYTES);
  // bars[0..S) full, [S..2S) empty, [2S..2S+2) tmem_full, [2S+2..2S+4) tmem_empty, then the TMEM base address
  const uint32_t bar0 = smem_u32(bars);
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (STAGES + s); };
  auto tfull_bar = [&](int s) { return bar0 + 8u * (2 * STAGES + s); };
  auto tempty_bar = [&](int s) { return bar0 + 8u * (2 * STAGES + 2 + s); };
  uint32_t *tmem_slot = (uint32_t *)(bars + 2 * STAGES + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // plain GEMM: tile width bn = cg.n_mma (256, or 128 for problems with too few 256-wide tiles to fill the machine);
  // convolution modes: one tile column of cg.n_mma (<= 256) output channels
  const int bn = CONV ? BN : cg.n_mma;
  const int tiles_m = (M + BM - 1) / BM, tiles_n = (N + bn - 1) / bn;
  const int n_tiles = tiles_m * tiles_n;
  const int num_kb = CONV == 1 ? 9 * cg.kc_per_tap : (K + BLOCK_K - 1) / BLOCK_K;
  const int n_work = n_tiles * splits;

  DLS_TRACE_DECL
  if (threadIdx.x == 0) DLS_TRACE(1);
  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar(s), 1);
      mbar_init(tempty_bar(s), EPI_WARPS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_c) : "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===== TMA producer =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int w = blockIdx.x; w < n_work; w += gridDim.x) {
        const Tile tc = tile_coord(w % n_tiles, tiles_m, tiles_n, bn);
        const int kb0 = (w / n_tiles) * kb_per_split, kb1 = min(num_kb, kb0 + kb_per_split);
        int img0 = 0, row0 = 0;
        if (CONV == 1) pixel_tile(cg, tc.m0 / BM, img0, row0);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          DLS_TRACE(10);
          const uint32_t a_dst = smem_u32(smem + stage * STAGE_BYTES);
          const uint32_t b_dst = a_dst + A_STAGE_BYTES;
          const int k0 = kb * BLOCK_K;
          if (CONV == 1) {
            // k block = (tap, channel block): the activation box shifted by the tap, and that tap's weights
            const int tap = kb / cg.kc_per_tap, kc = kb - tap * cg.kc_per_tap;
            const int oy = tap / 3 - 1, ox = tap - (tap / 3) * 3 - 1;
            mbar_expect_tx(full_bar(stage), A_STAGE_BYTES + cg.b_bytes);
            tma_load_4d(a_dst, &map_a, full_bar(stage), kc * BLOCK_K, -cg.tap_sign * ox, row0 - cg.tap_sign * oy, img0);
            tma_load_3d(b_dst, &map_b, full_bar(stage), kc * BLOCK_K, 0, tap);
            if (++stage == STAGES) {
              stage = 0;
              phase ^= 1;
            }
            continue;
          }
          if (CONV == 2) {
            // k block = 64 pixels; A atom j = (tap, 64-channel block) `tc.m0 / 64 + j`: the pixel box shifted by that tap;
            // B atoms = 64-column blocks of the delta at the same (unshifted) pixels
            int pi, py;
            pixel_tile(cg, kb, pi, py);
            const int rb0 = tc.m0 / MN_ATOM, n_rb = 9 * cg.cblks;
            const int a_atoms = min(BM / MN_ATOM, n_rb - rb0), b_atoms = (cg.n_mma + MN_ATOM - 1) / MN_ATOM;
            mbar_expect_tx(full_bar(stage), (a_atoms + b_atoms) * BLOCK_K * 128);
            for (int j = 0; j < a_atoms; ++j) {
              const int rb = rb0 + j, tap = rb / cg.cblks, cb = rb - tap * cg.cblks;
              const int oy = tap / 3 - 1, ox = tap - (tap / 3) * 3 - 1;
              tma_load_4d(a_dst + j * BLOCK_K * 128, &map_a, full_bar(stage), cb * MN_ATOM, -ox, py * cg.px_stride - oy, pi);
            }
            for (int j = 0; j < b_atoms; ++j) tma_load_2d(b_dst + j * BLOCK_K * 128, &map_b, full_bar(stage), j * MN_ATOM, k0);
            if (++stage == STAGES) {
              stage = 0;
              phase ^= 1;
            }
            continue;
          }
          mbar_expect_tx(full_bar(stage), A_STAGE_BYTES + bn * 128);
          if (A_MN) {
#pragma unroll
            for (int j = 0; j < BM / MN_ATOM; ++j)
              tma_load_2d(a_dst + j * BLOCK_K * 128, &map_a, full_bar(stage), tc.m0 + j * MN_ATOM, k0);
          } else {
            tma_load_2d(a_dst, &map_a, full_bar(stage), k0, tc.m0);
          }
          if (B_MN) {
            for (int j = 0; j < bn / MN_ATOM; ++j)
              tma_load_2d(b_dst + j * BLOCK_K * 128, &map_b, full_bar(stage), tc.n0 + j * MN_ATOM, k0);
          } else {
            tma_load_2d(b_dst, &map_b, full_bar(stage), k0, tc.n0);
          }
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer =====
    if (lane == 0) {
      // instruction descriptor (cute::UMMA::InstrDescriptor): D = fp32, A/B format, major-ness, N >> 3, M >> 4
      const uint32_t fmt = TF32 ? 2u : 1u;
      const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((A_MN ? 1u : 0u) << 15) | ((B_MN ? 1u : 0u) << 16) |
                             ((uint32_t)(cg.n_mma >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
      // K-major: 8-row groups are 1024 B apart (SBO), one swizzle atom along K (LBO unused).
      // MN-major: 8 K-rows of 128 B form an atom (SBO = 1024 B), atoms along MN are BLOCK_K * 128 B apart (LBO).
      const uint32_t a_lbo = A_MN ? BLOCK_K * 128 : 0, b_lbo = B_MN ? BLOCK_K * 128 : 0;
      const uint32_t a_kstep = A_MN ? UMMA_K * 128 : 32, b_kstep = B_MN ? UMMA_K * 128 : 32;
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int w = blockIdx.x; w < n_work; w += gridDim.x, ++it) {
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(tempty_bar(as), aphase ^ 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t d_tmem = tmem_base + as * BN;
        const int kb0 = (w / n_tiles) * kb_per_split, kb1 = min(num_kb, kb0 + kb_per_split);
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(full_bar(stage), phase);
          DLS_TRACE(20);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t a_addr = smem_u32(smem + stage * STAGE_BYTES);
          const uint32_t b_addr = a_addr + A_STAGE_BYTES;
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
            // MN-major tf32: 4-row atoms of 512 B (SBO), 32-byte-chunk swizzle; everything else: 8-row atoms of 1024 B
            const uint64_t ad = smem_desc(a_addr + k * a_kstep, a_lbo, (TF32 && A_MN) ? 512 : 1024, (TF32 && A_MN) ? 1u : 2u);
            const uint64_t bd = smem_desc(b_addr + k * b_kstep, b_lbo, (TF32 && B_MN) ? 512 : 1024, (TF32 && B_MN) ? 1u : 2u);
            umma(d_tmem, ad, bd, idesc, ((kb - kb0) | k) != 0 ? 1u : 0u, TF32);
          }
          umma_commit(empty_bar(stage));   // frees the smem slot once these MMAs have read it
          if (++stage == STAGES) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit(tfull_bar(as));        // accumulator complete -> epilogue
        DLS_TRACE(21);
      }
    }
  } else {
    // ===== epilogue warps: TMEM lane quarter = warp % 4 =====
    const int quarter = warp & 3;
    uint8_t *my_smem = epi_smem + (warp - 2) * 2 * EPI_BUF_BYTES;
    int it = 0;
    int buf = 0;
    for (int w = blockIdx.x; w < n_work; w += gridDim.x, ++it) {
      Tile tc = tile_coord(w % n_tiles, tiles_m, tiles_n, bn);
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      mbar_wait(tfull_bar(as), aphase);
      if (warp == 2 && lane == 0) DLS_TRACE(30);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      float row_bias = 0.f;
      if (bias_mode == 1 && tc.m0 + quarter * 32 + lane < M) row_bias = __ldg(bias + tc.m0 + quarter * 32 + lane);
      tc.m0 += (w / n_tiles) * tiles_m * BM;                      // split-K: rows of this split's partial matrix
      const int row = tc.m0 + quarter * 32 + lane;
      long long gate_row = row;                                   // row of the gate tensor that belongs to output row `row`
      long long down_row = -1, up_row = -1;                       // rows of the sampled / zero-filled operand copies
      if (CONV == 1 && (cg.gate_stride > 1 || cg.copy_down || cg.copy_up)) {
        const int hw = cg.H * cg.W, b = row / hw, rem = row - b * hw, y = rem / cg.W, x = rem - y * cg.W;
        gate_row = ((long long)(b * cg.H + y) * cg.gate_stride * cg.W + x) * cg.gate_stride;
        const int cs = cg.copy_s;
        if (cg.copy_down && y % cs == 0 && x % cs == 0) down_row = ((long long)b * (cg.H / cs) + y / cs) * (cg.W / cs) + x / cs;
        if (cg.copy_up) up_row = ((long long)(b * cg.H + y) * cs * cg.W + x) * cs;
      }
#pragma unroll 1
      for (int c = 0; c < bn / EPI_CHUNK; ++c) {
        const int col0 = tc.n0 + c * EPI_CHUNK;
        if (col0 >= N) break;                                     // warp uniform
        uint32_t v[32];
        tmem_ld_32x32(v, tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(as * BN + c * EPI_CHUNK));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        float f[32];
        epi_apply(v, f, bias_mode, act, row_bias, bias, col0, N);
        if (gate_mode && row < M) epi_gate(f, gate + gate_row * N + col0, N - col0, gate_mode, (N & 3) == 0, gate_scale);
        if (copy && row < M) epi_copy_bf16(f, reinterpret_cast<__nv_bfloat16 *>(copy) + (long long)row * copy_ld + col0, copy_ld - col0);
        if (CONV == 1 && row < M) {
          if (cg.copy_plain) epi_copy_bf16(f, reinterpret_cast<__nv_bfloat16 *>(cg.copy_plain) + (long long)row * N + col0, N - col0);
          if (down_row >= 0) epi_copy_bf16(f, reinterpret_cast<__nv_bfloat16 *>(cg.copy_down) + down_row * N + col0, N - col0);
          if (cg.copy_up) {
            __nv_bfloat16 *u0 = reinterpret_cast<__nv_bfloat16 *>(cg.copy_up) + up_row * N + col0;
            epi_copy_bf16(f, u0, N - col0);
            float z[32];
#pragma unroll
            for (int j = 0; j < 32; ++j) z[j] = 0.f;
            for (int dy = 0; dy < cg.copy_s; ++dy)
              for (int dx = 0; dx < cg.copy_s; ++dx)
                if (dy | dx) epi_copy_bf16(z, u0 + ((long long)dy * cg.W * cg.copy_s + dx) * N, N - col0);
          }
        }
        if (direct_store) {
          if (row < c_rows) {
            float *dst = C + (long long)row * ldc + col0;
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (col0 + j < N) dst[j] = f[j];
          }
        } else {
          // the TMA store that last read this staging buffer must have finished reading it
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          __syncwarp();
          uint8_t *sb = my_smem + buf * EPI_BUF_BYTES + lane * 128;   // 128B-swizzled: 16 B chunk j -> j ^ (row % 8)
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float4 q = make_float4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
            *reinterpret_cast<float4 *>(sb + ((j ^ (lane & 7)) << 4)) = q;
          }
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          __syncwarp();
          if (lane == 0) {
            tma_store_2d(&map_c, smem_u32(my_smem + buf * EPI_BUF_BYTES), col0, tc.m0 + quarter * 32);
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
          buf ^= 1;
        }
      }
      // all of this warp's TMEM reads of accumulator `as` are complete: hand it back to the MMA issuer
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive(tempty_bar(as));
      if (warp == 2 && lane == 0) DLS_TRACE(31);
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }

  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) DLS_TRACE(2);
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
  }
}

// ---- CTA-pair variant (cta_group::2) for the large plain GEMMs ---------------------------------------------------------
// A cluster of two CTAs (two SMs of one TPC) owns a 256 x 256 output tile.  CTA r of the pair loads ITS 128 rows of the A
// tile and ITS 128 rows (of N) of the B tile -- 32 KB per k block instead of the 48 KB a lone CTA needs for half the
// flops, which is what the single-CTA kernel runs out of (tensor pipe 75 % active at ~96 B/clk/SM from L2) -- and one thread
// of the leader CTA issues tcgen05.mma.cta_group::2 (M = 256, N = 256), which reads both CTAs' shared memory and writes
// each CTA's 128 accumulator rows into that CTA's TMEM.  Protocol:
//   full[s]   lives in the leader: its producer arrives with expect_tx = the pair's 64 KB; both CTAs' TMA loads signal it
//             (cp.async.bulk.tensor ... .cta_group::2 with the barrier address of CTA 0);
//   empty[s]  in both CTAs, arrived by the leader's tcgen05.commit multicast, so each producer knows its slot is free;
//   tfull[a]  in both CTAs (commit multicast): each CTA's four epilogue warps drain their own 128 TMEM lanes;
//   tempty[a] in the leader, 8 arrivals: the epilogue warps of BOTH CTAs (remote mbarrier.arrive through mapa).
constexpr int STAGES2 = 6;
constexpr int STAGE2_BYTES = 2 * BM * 128;                   // A 128 x 128 B + B 128 x 128 B per CTA and k block
constexpr int SMEM2_BYTES = STAGES2 * STAGE2_BYTES + EPI_BYTES + BAR_BYTES + 1024;
constexpr int BN2 = 256;

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// the same barrier / smem offset in CTA `rank` of the cluster
__device__ __forceinline__ uint32_t map_to_cta(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t cluster_addr) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_addr) : "memory");
}
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap *map, uint32_t leader_bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(dst),
      "l"(map), "r"(leader_bar), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void umma_pair(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate,
                                          bool tf32) {
  if (tf32) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
  }
}
__device__ __forceinline__ void umma_commit_pair(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
               "h"((uint16_t)3)
               : "memory");
}

template <bool TF32, bool A_MN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
    gemm2_tc_kernel(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
                    const __grid_constant__ CUtensorMap map_c, float *C, int M, int N, int K, long long ldc,
                    const float *__restrict__ bias, int bias_mode, int act, int direct_store, const float *gate, int gate_mode,
                    void *__restrict__ copy, long long copy_ld, float gate_scale) {
  constexpr int ELEM = TF32 ? 4 : 2;
  constexpr int BLOCK_K = 128 / ELEM;
  constexpr int UMMA_K = 32 / ELEM;
  constexpr int MN_ATOM = 128 / ELEM;

  extern __shared__ uint8_t smem_raw[];
  uint8_t *smem = (uint8_t *)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint8_t *epi_smem = smem + STAGES2 * STAGE2_BYTES;
  uint64_t *bars = (uint64_t *)(epi_smem + EPI_BYTES);
  const uint32_t bar0 = smem_u32(bars);
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (STAGES2 + s); };
  auto tfull_bar = [&](int s) { return bar0 + 8u * (2 * STAGES2 + s); };
  auto tempty_bar = [&](int s) { return bar0 + 8u * (2 * STAGES2 + 2 + s); };
  uint32_t *tmem_slot = (uint32_t *)(bars + 2 * STAGES2 + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();          // 0 = leader
  const int pair = blockIdx.x >> 1, n_pairs = gridDim.x >> 1;
  const int tiles_m = (M + 2 * BM - 1) / (2 * BM), tiles_n = (N + BN2 - 1) / BN2;
  const int n_tiles = tiles_m * tiles_n;
  const int num_kb = (K + BLOCK_K - 1) / BLOCK_K;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES2; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int s = 0; s < 2; ++s) {
      mbar_init(tfull_bar(s), 1);
      mbar_init(tempty_bar(s), 2 * EPI_WARPS);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_b) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&map_c) : "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  cluster_sync_all();                                // both CTAs' barriers are initialised before anything signals them
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;

  // tile t of the pair: 256 rows x 256 columns; this CTA's rows start at m0 + 128 * rank, its B rows at n0 + 128 * rank
  auto pair_tile = [&](int t, int &m0, int &n0) {
    constexpr int GROUP = 4;                         // 4 pair-rows (1024 rows) share their B tiles through L2
    const int per_group = GROUP * tiles_n;
    const int g = t / per_group;
    const int first_m = g * GROUP;
    const int gsz = min(GROUP, tiles_m - first_m);
    const int r = t - g * per_group;
    m0 = (first_m + r % gsz) * 2 * BM;
    n0 = (r / gsz) * BN2;
  };

  if (warp == 0) {
    // ===== TMA producer (both CTAs) =====
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int t = pair; t < n_tiles; t += n_pairs) {
        int m0, n0;
        pair_tile(t, m0, n0);
        const int my_m = m0 + BM * (int)rank, my_n = n0 + (BN2 / 2) * (int)rank;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          const uint32_t a_dst = smem_u32(smem + stage * STAGE2_BYTES);
          const uint32_t b_dst = a_dst + BM * 128;
          const uint32_t lbar = map_to_cta(full_bar(stage), 0);
          const int k0 = kb * BLOCK_K;
          if (rank == 0) mbar_expect_tx(full_bar(stage), 2 * STAGE2_BYTES);
          if (A_MN) {
#pragma unroll
            for (int j = 0; j < BM / MN_ATOM; ++j) tma_load_2d_pair(a_dst + j * BLOCK_K * 128, &map_a, lbar, my_m + j * MN_ATOM, k0);
          } else {
            tma_load_2d_pair(a_dst, &map_a, lbar, k0, my_m);
          }
          if (B_MN) {
#pragma unroll
            for (int j = 0; j < (BN2 / 2) / MN_ATOM; ++j)
              tma_load_2d_pair(b_dst + j * BLOCK_K * 128, &map_b, lbar, my_n + j * MN_ATOM, k0);
          } else {
            tma_load_2d_pair(b_dst, &map_b, lbar, k0, my_n);
          }
          if (++stage == STAGES2) {
            stage = 0;
            phase ^= 1;
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===== MMA issuer (leader CTA only) =====
    if (lane == 0 && rank == 0) {
      const uint32_t fmt = TF32 ? 2u : 1u;
      const uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((A_MN ? 1u : 0u) << 15) | ((B_MN ? 1u : 0u) << 16) |
                             ((uint32_t)(BN2 >> 3) << 17) | ((uint32_t)((2 * BM) >> 4) << 24);
      const uint32_t a_lbo = A_MN ? BLOCK_K * 128 : 0, b_lbo = B_MN ? BLOCK_K * 128 : 0;
      const uint32_t a_kstep = A_MN ? UMMA_K * 128 : 32, b_kstep = B_MN ? UMMA_K * 128 : 32;
      int stage = 0;
      uint32_t phase = 0;
      int it = 0;
      for (int t = pair; t < n_tiles; t += n_pairs, ++it) {
        const int as = it & 1;
        const uint32_t aphase = (it >> 1) & 1;
        mbar_wait(tempty_bar(as), aphase ^ 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t d_tmem = tmem_base + as * BN2;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(full_bar(stage), phase);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t a_addr = smem_u32(smem + stage * STAGE2_BYTES);
          const uint32_t b_addr = a_addr + BM * 128;
#pragma unroll
          for (int k = 0; k < BLOCK_K / UMMA_K; ++k) {
            // MN-major tf32: 4-row atoms of 512 B (SBO), 32-byte-chunk swizzle; everything else: 8-row atoms of 1024 B
            const uint64_t ad = smem_desc(a_addr + k * a_kstep, a_lbo, (TF32 && A_MN) ? 512 : 1024, (TF32 && A_MN) ? 1u : 2u);
            const uint64_t bd = smem_desc(b_addr + k * b_kstep, b_lbo, (TF32 && B_MN) ? 512 : 1024, (TF32 && B_MN) ? 1u : 2u);
            umma_pair(d_tmem, ad, bd, idesc, (kb | k) != 0 ? 1u : 0u, TF32);
          }
          umma_commit_pair(empty_bar(stage));        // frees the slot in BOTH CTAs once these MMAs have read it
          if (++stage == STAGES2) {
            stage = 0;
            phase ^= 1;
          }
        }
        umma_commit_pair(tfull_bar(as));             // accumulator complete -> the epilogue warps of both CTAs
      }
    }
  } else {
    // ===== epilogue warps (both CTAs): this CTA's 128 rows of the pair tile =====
    const int quarter = warp & 3;
    uint8_t *my_smem = epi_smem + (warp - 2) * 2 * EPI_BUF_BYTES;
    const uint32_t leader_tempty0 = map_to_cta(tempty_bar(0), 0);
    int it = 0;
    int buf = 0;
    for (int t = pair; t < n_tiles; t += n_pairs, ++it) {
      int m0, n0;
      pair_tile(t, m0, n0);
      m0 += BM * (int)rank;
      const int as = it & 1;
      const uint32_t aphase = (it >> 1) & 1;
      mbar_wait(tfull_bar(as), aphase);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const int row = m0 + quarter * 32 + lane;
      float row_bias = 0.f;
      if (bias_mode == 1 && row < M) row_bias = __ldg(bias + row);
#pragma unroll 1
      for (int c = 0; c < BN2 / EPI_CHUNK; ++c) {
        const int col0 = n0 + c * EPI_CHUNK;
        if (col0 >= N) break;
        uint32_t v[32];
        tmem_ld_32x32(v, tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(as * BN2 + c * EPI_CHUNK));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        float f[32];
        epi_apply(v, f, bias_mode, act, row_bias, bias, col0, N);
        if (gate_mode && row < M) epi_gate(f, gate + (long long)row * N + col0, N - col0, gate_mode, (N & 3) == 0, gate_scale);
        if (copy && row < M) epi_copy_bf16(f, reinterpret_cast<__nv_bfloat16 *>(copy) + (long long)row * copy_ld + col0, copy_ld - col0);
        if (direct_store) {
          if (row < M) {
            float *dst = C + (long long)row * ldc + col0;
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (col0 + j < N) dst[j] = f[j];
          }
        } else {
          if (lane == 0) asm volatile("cp.async.bulk.wait_group.read 1;" ::: "memory");
          __syncwarp();
          uint8_t *sb = my_smem + buf * EPI_BUF_BYTES + lane * 128;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            float4 q = make_float4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
            *reinterpret_cast<float4 *>(sb + ((j ^ (lane & 7)) << 4)) = q;
          }
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          __syncwarp();
          if (lane == 0 && m0 + quarter * 32 < M) {
            tma_store_2d(&map_c, smem_u32(my_smem + buf * EPI_BUF_BYTES), col0, m0 + quarter * 32);
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
          buf ^= 1;
        }
      }
      asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(leader_tempty0 + 8u * as);
    }
    if (lane == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }

  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  cluster_sync_all();                                // neither CTA leaves while the other may still read its smem / signal it
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
  }
}

// ---- fp32 -> bf16 operand staging ------------------------------------------------------------------------------------
// [rows, cols] fp32 (dense) -> [rows, pitch] bf16 with pitch a multiple of 8 elements so every row starts 16-byte
// aligned (a TMA requirement).  Round to nearest even.
__global__ void __launch_bounds__(256) to_bf16_flat_kernel(const float4 *__restrict__ in, uint4 *__restrict__ out, size_t n8) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n8; i += (size_t)gridDim.x * blockDim.x) {
    const float4 a = __ldcs(in + 2 * i), b = __ldcs(in + 2 * i + 1);
    __nv_bfloat162 p0 = __floats2bfloat162_rn(a.x, a.y), p1 = __floats2bfloat162_rn(a.z, a.w);
    __nv_bfloat162 p2 = __floats2bfloat162_rn(b.x, b.y), p3 = __floats2bfloat162_rn(b.z, b.w);
    uint4 o;
    o.x = *reinterpret_cast<uint32_t *>(&p0);
    o.y = *reinterpret_cast<uint32_t *>(&p1);
    o.z = *reinterpret_cast<uint32_t *>(&p2);
    o.w = *reinterpret_cast<uint32_t *>(&p3);
    out[i] = o;
  }
}
__global__ void __launch_bounds__(256) to_bf16_rows_kernel(const float *__restrict__ in, __nv_bfloat16 *__restrict__ out,
                                                           long long rows, long long cols, long long pitch) {
  const long long total = rows * pitch;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / pitch, c = i - r * pitch;
    out[i] = __float2bfloat16_rn(c < cols ? in[r * cols + c] : 0.f);
  }
}

// fp32 -> tf32 (10-bit mantissa), round to nearest, ties away: tcgen05 kind::tf32 ignores the low 13 mantissa bits of its
// 32-bit operands, i.e. truncates; truncation is biased (every product shrinks by ~2^-11 twice) and misses the stated
// 1e-3, so operands are rounded once when they are staged.
__device__ __forceinline__ float round_tf32(float x) {
  uint32_t r;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(x));
  return __uint_as_float(r);
}

__global__ void __launch_bounds__(256) to_tf32_rows_kernel(const float *__restrict__ in, float *__restrict__ out, long long rows,
                                                           long long cols, long long pitch) {
  const long long total = rows * pitch;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / pitch, c = i - r * pitch;
    out[i] = c < cols ? round_tf32(in[r * cols + c]) : 0.f;
  }
}
__global__ void __launch_bounds__(256) to_tf32_flat_kernel(const float4 *__restrict__ in, float4 *__restrict__ out, size_t n4) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    float4 v = __ldcs(in + i);
    v.x = round_tf32(v.x);
    v.y = round_tf32(v.y);
    v.z = round_tf32(v.z);
    v.w = round_tf32(v.w);
    out[i] = v;
  }
}

// fp32 [rows, cols] dense -> tf32-rounded fp32 [cols, pitch] (its transpose, pitch % 4 == 0): TF32 operands that arrive MN-major are
// made K-major once, because tcgen05 reads MN-major 32-bit operands only through the 32-byte-atom swizzle.
__global__ void __launch_bounds__(256) transpose_f32_kernel(const float *__restrict__ in, float *__restrict__ out, int rows,
                                                            int cols, long long pitch) {
  __shared__ float tile[32][33];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;   // 32 x 8
  const int tiles_c = (cols + 31) / 32, tiles_r = (rows + 31) / 32;
  for (int t = blockIdx.x; t < tiles_c * tiles_r; t += gridDim.x) {
    const int r0 = (t / tiles_c) * 32, c0 = (t % tiles_c) * 32;
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
      const int r = r0 + ty + j, c = c0 + tx;
      tile[ty + j][tx] = (r < rows && c < cols) ? in[(long long)r * cols + c] : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int j = 0; j < 32; j += 8) {
      const int c = c0 + ty + j, r = r0 + tx;
      if (c < cols && r < rows) out[(long long)c * pitch + r] = round_tf32(tile[tx][ty + j]);
    }
    __syncthreads();
  }
}

typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                    const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                    CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static encode_tiled_fn get_encoder() {
  static encode_tiled_fn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void *p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (encode_tiled_fn)p;
    else
      cudaGetLastError();
  });
  return fn;
}

// 2-D row-major tensor [outer][inner] with a 128B-swizzled box.
static int make_map(CUtensorMap *map, CUtensorMapDataType dt, int elem_bytes, const void *base, uint64_t inner, uint64_t outer,
                    uint64_t pitch_elems, uint32_t box_inner, uint32_t box_outer, bool atom32 = false) {
  encode_tiled_fn enc = get_encoder();
  DLS_CHECK(enc != nullptr, DLS_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
  cuuint64_t dims[2] = {inner, outer};
  cuuint64_t strides[1] = {pitch_elems * (uint64_t)elem_bytes};
  cuuint32_t box[2] = {box_inner, box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, dt, 2, const_cast<void *>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult " + std::to_string((int)r) + " (inner " + std::to_string(inner) +
              ", outer " + std::to_string(outer) + ", pitch " + std::to_string(pitch_elems) + ")");
    return DLS_ERR_CUDA;
  }
  return DLS_OK;
}

// rank-N row-major tensor (dims[0] innermost, strides in bytes for dims 1..rank-1) with a 128B-swizzled box; elements
// outside the tensor read as zero.
static int make_map_nd(CUtensorMap *map, CUtensorMapDataType dt, const void *base, int rank, const uint64_t *dims,
                       const uint64_t *strides_bytes, const uint32_t *box, const uint32_t *elem_strides = nullptr, bool atom32 = false) {
  encode_tiled_fn enc = get_encoder();
  DLS_CHECK(enc != nullptr, DLS_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
  cuuint64_t d[5], s[4];
  cuuint32_t b[5], e[5];
  for (int i = 0; i < rank; ++i) {
    d[i] = dims[i];
    b[i] = box[i];
    e[i] = elem_strides ? elem_strides[i] : 1;   // traversal stride: box[i] / e[i] elements are loaded along dim i
    if (i + 1 < rank) s[i] = strides_bytes[i];
  }
  CUresult r = enc(map, dt, (cuuint32_t)rank, const_cast<void *>(base), d, s, b, e, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    std::string t = "cuTensorMapEncodeTiled (rank " + std::to_string(rank) + ") failed with CUresult " + std::to_string((int)r) + ": dims";
    for (int i = 0; i < rank; ++i) t += " " + std::to_string(dims[i]) + "/" + std::to_string(box[i]);
    set_error(t);
    return DLS_ERR_CUDA;
  }
  return DLS_OK;
}

// out[i] = act(sum_s partial[s * stride + i] + bias) (* act'(gate[i])), i < m * n: the second half of a split-K contraction;
// optionally also the bf16 operand copy of the result.
__global__ void __launch_bounds__(256) splitk_reduce_kernel(const float *__restrict__ partial, float *out, size_t m,
                                                            size_t n, size_t stride, int splits, const float *__restrict__ bias,
                                                            int bias_mode, int act, const float *gate, int gate_mode,
                                                            __nv_bfloat16 *__restrict__ copy, long long copy_ld, float gate_scale) {
  const size_t total = m * n;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    float x = __ldg(partial + i);
    for (int s = 1; s < splits; ++s) x = __fadd_rn(x, __ldg(partial + (size_t)s * stride + i));
    if (bias_mode == 1) x = __fadd_rn(x, __ldg(bias + i / n));
    else if (bias_mode == 2) x = __fadd_rn(x, __ldg(bias + i % n));
    if (act == 1) x = tanhf(x);
    else if (act == 2) x = fmaxf(x, 0.f);
    if (gate_mode == 1) {
      const float h = gate[i];
      x = __fmul_rn(x, __fsub_rn(1.0f, __fmul_rn(h, h)));
    } else if (gate_mode == 2) {
      x = __fmul_rn(x, fminf(__fmul_rn(gate[i], 3.4028234663852886e38f), 1.0f));
    } else if (gate_mode == 3) {
      x = __fsub_rn(gate[i], __fmul_rn(x, gate_scale));   // plain load: the gate is the output buffer
    }
    out[i] = x;
    if (copy) copy[(i / n) * copy_ld + i % n] = __float2bfloat16_rn(x);
  }
}

struct SplitPlan {
  int splits = 1, kb_per_split = 0, c_rows = 0;
};

// SMs a persistent launch may occupy now: all of them, minus the ones an in-flight collective of the data-parallel exchange
// pins (one CTA per NCCL channel).  Split-K plans are sized with the same number -- a plan cut for 148 SMs but launched on
// 144 runs its last four work items as a second wave and takes twice as long.
static int usable_sms(const dls_ctx *ctx) {
  return ctx->comm_pending.empty() ? ctx->sm_count : std::max(1, ctx->sm_count - ctx->exchange_sms);
}

template <bool TF32, bool A_MN, bool B_MN, int CONV = 0>
static int launch_variant(dls_ctx *ctx, cudaStream_t st, const CUtensorMap &ma, const CUtensorMap &mb, const CUtensorMap &mc,
                          float *C, int M, int N, int K, long long ldc, const GemmEpilogue &epi, int direct, const SplitPlan &sp,
                          const ConvGeom &cg = ConvGeom()) {
  auto kern = gemm_tc_kernel<TF32, A_MN, B_MN, CONV>;
  // the attribute is per DEVICE (one process may hold a context on every GPU of the node), guarded by ctx->mu per context
  static std::atomic<bool> configured[64];
  const int dev = ctx->device & 63;
  if (!configured[dev].load(std::memory_order_acquire)) {
    DLS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
    configured[dev].store(true, std::memory_order_release);
  }
  const int bn = CONV ? BN : cg.n_mma;
  const int tiles = ((M + BM - 1) / BM) * ((N + bn - 1) / bn) * sp.splits;
  // one persistent CTA per SM; SMs hosting an in-flight collective (data-parallel exchange) are left to it
  const int sms = usable_sms(ctx);
  const int grid = tiles < sms ? tiles : sms;
  kern<<<grid, NUM_THREADS, SMEM_BYTES, st>>>(ma, mb, mc, C, M, N, K, ldc, epi.bias, epi.bias_mode, epi.act, direct, sp.splits,
                                              sp.kb_per_split, sp.c_rows, cg, epi.gate, epi.gate_mode, epi.copy, epi.copy_ld,
                                              epi.gate_scale);
  DLS_CUDA(cudaGetLastError());
  ctx->stats.kernel_launches++;
  return DLS_OK;
}

template <bool TF32, bool A_MN, bool B_MN>
static int launch_variant2(dls_ctx *ctx, cudaStream_t st, const CUtensorMap &ma, const CUtensorMap &mb, const CUtensorMap &mc, float *C,
                           int M, int N, int K, long long ldc, const GemmEpilogue &epi, int direct) {
  auto kern = gemm2_tc_kernel<TF32, A_MN, B_MN>;
  static std::atomic<bool> configured[64];
  const int dev = ctx->device & 63;
  if (!configured[dev].load(std::memory_order_acquire)) {
    DLS_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM2_BYTES));
    configured[dev].store(true, std::memory_order_release);
  }
  const int tiles = ((M + 2 * BM - 1) / (2 * BM)) * ((N + BN2 - 1) / BN2);
  const int pairs = std::max(1, std::min(tiles, usable_sms(ctx) / 2));
  kern<<<2 * pairs, NUM_THREADS, SMEM2_BYTES, st>>>(ma, mb, mc, C, M, N, K, ldc, epi.bias, epi.bias_mode, epi.act, direct, epi.gate,
                                                    epi.gate_mode, epi.copy, epi.copy_ld, epi.gate_scale);
  DLS_CUDA(cudaGetLastError());
  ctx->stats.kernel_launches++;
  return DLS_OK;
}

}  // namespace tc

int launch_to_bf16(dls_ctx *ctx, int stream, const float *in, void *out, long long rows, long long cols, long long pitch) {
  cudaStream_t st = ctx->streams[stream];
  if (rows * cols == 0) return DLS_OK;
  if (pitch == cols && ((uintptr_t)in % 16) == 0) {
    const size_t n8 = (size_t)(rows * cols) / 8;       // cols % 8 == 0 here
    const int grid = (int)std::min<size_t>((n8 + 255) / 256, (size_t)ctx->sm_count * 8);
    tc::to_bf16_flat_kernel<<<grid, 256, 0, st>>>((const float4 *)in, (uint4 *)out, n8);
  } else {
    const long long total = rows * pitch;
    const int grid = (int)std::min<long long>((total + 255) / 256, (long long)ctx->sm_count * 8);
    tc::to_bf16_rows_kernel<<<grid, 256, 0, st>>>(in, (__nv_bfloat16 *)out, rows, cols, pitch);
  }
  DLS_CUDA(cudaGetLastError());
  ctx->stats.kernel_launches++;
  return DLS_OK;
}

int launch_to_tf32(dls_ctx *ctx, int stream, const float *in, float *out, long long rows, long long cols, long long pitch) {
  cudaStream_t st = ctx->streams[stream];
  if (rows * cols == 0) return DLS_OK;
  if (pitch == cols && ((uintptr_t)in % 16) == 0) {
    const size_t n4 = (size_t)(rows * cols) / 4;
    const int grid = (int)std::min<size_t>((n4 + 255) / 256, (size_t)ctx->sm_count * 8);
    tc::to_tf32_flat_kernel<<<grid, 256, 0, st>>>((const float4 *)in, (float4 *)out, n4);
  } else {
    const long long total = rows * pitch;
    const int grid = (int)std::min<long long>((total + 255) / 256, (long long)ctx->sm_count * 8);
    tc::to_tf32_rows_kernel<<<grid, 256, 0, st>>>(in, out, rows, cols, pitch);
  }
  DLS_CUDA(cudaGetLastError());
  ctx->stats.kernel_launches++;
  return DLS_OK;
}

int launch_transpose_f32(dls_ctx *ctx, int stream, const float *in, float *out, int rows, int cols, long long pitch) {
  if ((long long)rows * cols == 0) return DLS_OK;
  const long long tiles = (long long)((rows + 31) / 32) * ((cols + 31) / 32);
  const int grid = (int)std::min<long long>(tiles, (long long)ctx->sm_count * 16);
  tc::transpose_f32_kernel<<<grid, 256, 0, ctx->streams[stream]>>>(in, out, rows, cols, pitch);
  DLS_CUDA(cudaGetLastError());
  ctx->stats.kernel_launches++;
  return DLS_OK;
}

int launch_gemm_tc(dls_ctx *ctx, int stream, int kind, const void *A, bool a_mn, long long lda, const void *B, bool b_mn,
                   long long ldb, float *C, long long ldc, int M, int N, int K, const GemmEpilogue &epi) {
  using namespace tc;
  DLS_CHECK(kind == DLS_GEMM_BF16 || kind == DLS_GEMM_TF32, DLS_ERR_INVALID, "launch_gemm_tc: kind must be BF16 or TF32");
  const bool tf32 = kind == DLS_GEMM_TF32;
  const int elem = tf32 ? 4 : 2;
  const int block_k = 128 / elem, mn_atom = 128 / elem;
  DLS_CHECK((lda * elem) % 16 == 0 && (ldb * elem) % 16 == 0 && ((uintptr_t)A % 16) == 0 && ((uintptr_t)B % 16) == 0,
            DLS_ERR_INVALID, "launch_gemm_tc: operand rows must be 16-byte aligned");
  const CUtensorMapDataType dt = tf32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  CUtensorMap ma, mb, mc;
  int rc;
  // K-major [rows][K]: box = BLOCK_K x tile rows.  MN-major [K][rows]: box = 128 bytes of rows x BLOCK_K.
  rc = a_mn ? make_map(&ma, dt, elem, A, (uint64_t)M, (uint64_t)K, (uint64_t)lda, mn_atom, block_k, tf32)
            : make_map(&ma, dt, elem, A, (uint64_t)K, (uint64_t)M, (uint64_t)lda, block_k, BM);
  if (rc) return rc;
  // CTA-pair kernel (cta_group::2, 256 x 256 tiles) for contractions that fill the machine at least twice with such tiles;
  // DLS_GEMM_2CTA=0 keeps everything on the single-CTA kernel
  static const bool pair_off = [] {
    const char *e = getenv("DLS_GEMM_2CTA");
    return e && atoi(e) == 0;
  }();
  const long long tiles2 = (long long)((M + 2 * BM - 1) / (2 * BM)) * ((N + BN2 - 1) / BN2);
  const bool pair_mode = !pair_off && tiles2 >= ctx->sm_count && K >= 4 * block_k;
  // too few 256-wide tiles to occupy half the machine and a K too short to split: 128-wide tiles double the CTAs at work
  // (the weight gradient / backward-data contractions of a dense layer over a small batch are output bound)
  const int tiles256 = ((M + BM - 1) / BM) * ((N + BN - 1) / BN);
  const int num_kb_all = (K + block_k - 1) / block_k;
  const bool will_split = tiles256 * 2 <= tc::usable_sms(ctx) && num_kb_all >= 16 && N % 4 == 0;
  const int bn = (!pair_mode && !will_split && N > 128 && tiles256 * 2 <= ctx->sm_count) ? 128 : BN;
  rc = b_mn ? make_map(&mb, dt, elem, B, (uint64_t)N, (uint64_t)K, (uint64_t)ldb, mn_atom, block_k, tf32)
            : make_map(&mb, dt, elem, B, (uint64_t)K, (uint64_t)N, (uint64_t)ldb, block_k, pair_mode ? BN2 / 2 : bn);
  if (rc) return rc;
  ConvGeom plain;
  plain.n_mma = bn;
  // split-K: a contraction with few output tiles and a long K (the weight gradients of a convolution: M x N = Cin*9 x Cout,
  // K = batch * H * W) would leave most SMs idle; cut K into `splits` ranges, one work item each, then reduce the partials.
  const int tiles = ((M + BM - 1) / BM) * ((N + bn - 1) / bn);
  const int num_kb = (K + block_k - 1) / block_k;
  SplitPlan sp;
  sp.kb_per_split = num_kb;
  sp.c_rows = M;
  const int sms = tc::usable_sms(ctx);
  if (will_split) {
    int want = std::min(sms / tiles, num_kb / 8);
    sp.kb_per_split = (num_kb + want - 1) / want;
    sp.splits = (num_kb + sp.kb_per_split - 1) / sp.kb_per_split;
  }
  cudaStream_t st = ctx->streams[stream];
  Block scratch;
  float *dst = C;
  long long ld_dst = ldc;
  GemmEpilogue kepi = epi;
  const int m_pad = ((M + BM - 1) / BM) * BM;
  if (sp.splits > 1) {
    const long long ldp = (N + 3) & ~3LL;
    rc = alloc_block(ctx, (size_t)sp.splits * m_pad * ldp * sizeof(float), stream, &scratch);
    if (rc) return rc;
    dst = (float *)scratch.ptr;
    ld_dst = ldp;
    sp.c_rows = sp.splits * m_pad;
    kepi = GemmEpilogue();
  }
  const int direct = ((ld_dst % 4) != 0 || ((uintptr_t)dst % 16) != 0) ? 1 : 0;   // unaligned output: scalar stores
  if (direct == 0) {
    rc = make_map(&mc, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, dst, (uint64_t)N, (uint64_t)sp.c_rows, (uint64_t)ld_dst, EPI_CHUNK, 32);
  } else {
    mc = ma;   // unused by the kernel in direct-store mode
  }
  if (!rc && pair_mode && sp.splits == 1) {
    if (tf32 && !a_mn && !b_mn) rc = launch_variant2<true, false, false>(ctx, st, ma, mb, mc, dst, M, N, K, ld_dst, kepi, direct);
    else if (tf32 && !a_mn && b_mn) rc = launch_variant2<true, false, true>(ctx, st, ma, mb, mc, dst, M, N, K, ld_dst, kepi, direct);
    else if (tf32 && a_mn && !b_mn) rc = launch_variant2<true, true, false>(ctx, st, ma, mb, mc, dst, M, N, K, ld_dst, kepi, direct);
    else if (tf32) rc = launch_variant2<true, true, true>(ctx, st, ma, mb, mc, dst, M, N, K, ld_dst, kepi, direct);
    else if (!a_mn && !b_mn) rc = launch_variant2<false, false, false>(ctx, st, ma, mb, mc, dst, M, N, K, ld_dst, kepi, direct);
    else if (!a_mn && b_mn) rc = launch_variant2<false, false, true>(ctx, st, ma, mb, mc, dst, M, N, K, ld_dst, kepi, direct);
    else if (a_mn && !b_mn) rc = launch_variant2<false, true, false>(ctx, st, ma, mb, mc, dst, M, N, K, ld_dst, kepi, direct);
    else rc = launch_variant2<false, true, true>(ctx, st, ma, mb, mc, dst, M, N, K, ld_dst, kepi, direct);
  } else if (!rc) {
    if (tf32 && !a_mn && !b_mn) rc = launch_variant<true, false, false>(ctx, st, ma, mb, mc, dst, M, N, K, ld_dst, kepi, direct, sp, plain);
    else if (tf32 && !a_mn && b_mn) rc = launch_variant<true, false, true>(ctx, st, ma, mb, mc, dst, M, N, K, ld_dst, kepi, direct, sp, plain);
    else if (tf32 && a_mn && !b_mn) rc = launch_variant<true, true, false>(ctx, st, ma, mb, mc, dst, M, N, K, ld_dst, kepi, direct, sp, plain);
    else if (tf32) rc = launch_variant<true, true, true>(ctx, st, ma, mb, mc, dst, M, N, K, ld_dst, kepi, direct, sp, plain);
    else if (!a_mn && !b_mn) rc = launch_variant<false, false, false>(ctx, st, ma, mb, mc, dst, M, N, K, ld_dst, kepi, direct, sp, plain);
    else if (!a_mn && b_mn) rc = launch_variant<false, false, true>(ctx, st, ma, mb, mc, dst, M, N, K, ld_dst, kepi, direct, sp, plain);
    else if (a_mn && !b_mn) rc = launch_variant<false, true, false>(ctx, st, ma, mb, mc, dst, M, N, K, ld_dst, kepi, direct, sp, plain);
    else rc = launch_variant<false, true, true>(ctx, st, ma, mb, mc, dst, M, N, K, ld_dst, kepi, direct, sp, plain);
  }
  if (!rc && sp.splits > 1) {
    // partial s lives at rows [s * m_pad, s * m_pad + M) of the scratch matrix; when N is not a multiple of 4 the scratch
    // pitch differs from N, so reduce row by row through the (m, n) index
    if (ld_dst == N) {
      const size_t total = (size_t)M * N;
      const int grid = (int)std::min<size_t>((total + 255) / 256, (size_t)ctx->sm_count * 8);
      tc::splitk_reduce_kernel<<<grid, 256, 0, st>>>(dst, C, (size_t)M, (size_t)N, (size_t)m_pad * ld_dst, sp.splits, epi.bias,
                                                     epi.bias_mode, epi.act, epi.gate, epi.gate_mode, (__nv_bfloat16 *)epi.copy,
                                                     epi.copy_ld, epi.gate_scale);
      cudaError_t e = cudaGetLastError();
      if (e != cudaSuccess) rc = cuda_fail(e, "splitk_reduce_kernel", __FILE__, __LINE__);
      ctx->stats.kernel_launches++;
    } else {
      set_error("launch_gemm_tc: split-K needs N % 4 == 0");
      rc = DLS_ERR_UNSUPPORTED;
    }
  }
  if (scratch.ptr) free_block(ctx, scratch);
  return rc;
}

// Pixel tiling of an NHWC tensor for the implicit-GEMM kernels: `pixels` (128 or BLOCK_K) consecutive pixels must be `bb` whole
// images x `bh` rows x the full width.
bool conv_tiling(int H, int W, int pixels, ConvGeom *g) {
  if (W <= 0 || H <= 0 || W > 256 || pixels % W) return false;
  g->H = H;
  g->W = W;
  g->bw = W;
  if ((long long)W * H >= pixels) {
    g->bh = pixels / W;
    if (H % g->bh) return false;
    g->bb = 1;
    g->tiles_per_img = H / g->bh;
  } else {
    if (pixels % (W * H)) return false;
    g->bh = H;
    g->bb = pixels / (W * H);
    g->tiles_per_img = 0;
  }
  return g->bh <= 256 && g->bb <= 256;
}

bool conv_tc_supported(int kind, int H, int W, int k_channels, int n) {
  ConvGeom g;
  const int block_k = kind == DLS_GEMM_TF32 ? 32 : 64;
  return (kind == DLS_GEMM_TF32 || kind == DLS_GEMM_BF16) && conv_tiling(H, W, tc::BM, &g) && k_channels > 0 &&
         k_channels % block_k == 0 && n >= 1 && n <= tc::BN;
}

int launch_conv_tc(dls_ctx *ctx, int stream, int kind, const void *act, int batch, int H, int W, int kp, const void *wts, int N,
                   float *C, int tap_sign, const GemmEpilogue &epi, int gate_stride, void *copy_plain, void *copy_down, void *copy_up,
                   int copy_s) {
  using namespace tc;
  DLS_CHECK(conv_tc_supported(kind, H, W, kp, N), DLS_ERR_UNSUPPORTED, "launch_conv_tc: shape not supported by the implicit-GEMM kernel");
  const bool tf32 = kind == DLS_GEMM_TF32;
  const int elem = tf32 ? 4 : 2, block_k = 128 / elem;
  ConvGeom cg;
  conv_tiling(H, W, BM, &cg);
  cg.kc_per_tap = kp / block_k;
  cg.tap_sign = tap_sign;
  cg.gate_stride = gate_stride;
  cg.copy_plain = copy_plain;
  cg.copy_down = copy_down;
  cg.copy_up = copy_up;
  cg.copy_s = copy_s;
  cg.n_mma = (N + 15) & ~15;
  cg.b_bytes = cg.n_mma * 128;
  const CUtensorMapDataType dt = tf32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  CUtensorMap ma, mb, mc;
  {
    const uint64_t dims[4] = {(uint64_t)kp, (uint64_t)W, (uint64_t)H, (uint64_t)batch};
    const uint64_t str[3] = {(uint64_t)kp * elem, (uint64_t)kp * elem * W, (uint64_t)kp * elem * W * H};
    const uint32_t box[4] = {(uint32_t)block_k, (uint32_t)cg.bw, (uint32_t)cg.bh, (uint32_t)cg.bb};
    int rc = make_map_nd(&ma, dt, act, 4, dims, str, box);
    if (rc) return rc;
  }
  {
    const uint64_t dims[3] = {(uint64_t)kp, (uint64_t)N, 9};
    const uint64_t str[2] = {(uint64_t)kp * elem, (uint64_t)kp * elem * N};
    const uint32_t box[3] = {(uint32_t)block_k, (uint32_t)cg.n_mma, 1};
    int rc = make_map_nd(&mb, dt, wts, 3, dims, str, box);
    if (rc) return rc;
  }
  const long long P = (long long)batch * H * W;
  DLS_CHECK(P < (1LL << 31), DLS_ERR_UNSUPPORTED, "launch_conv_tc: too many pixels");
  const int direct = ((N % 4) != 0 || ((uintptr_t)C % 16) != 0) ? 1 : 0;
  int rc = DLS_OK;
  if (direct == 0) rc = make_map(&mc, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, C, (uint64_t)N, (uint64_t)P, (uint64_t)N, EPI_CHUNK, 32);
  else mc = ma;
  if (rc) return rc;
  SplitPlan sp;
  sp.kb_per_split = 9 * cg.kc_per_tap;
  sp.c_rows = (int)P;
  cudaStream_t st = ctx->streams[stream];
  if (tf32) return launch_variant<true, false, false, 1>(ctx, st, ma, mb, mc, C, (int)P, N, 9 * kp, N, epi, direct, sp, cg);
  return launch_variant<false, false, false, 1>(ctx, st, ma, mb, mc, C, (int)P, N, 9 * kp, N, epi, direct, sp, cg);
}

// dw[(tap * cin + c) * cout + f] = sum_s partial[s * stride + (tap * cxp + c) * ldp + f]: the second half of the split
// weight-gradient contraction; drops the zero rows of the padded channels.
__global__ void __launch_bounds__(256) conv_dw_reduce_kernel(const float *__restrict__ partial, float *__restrict__ dw, int splits,
                                                             size_t stride, int cxp, int cin, int cout, int ldp) {
  const int total = 9 * cin * cout;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
    const int f = i % cout, c = (i / cout) % cin, tap = i / (cout * cin);
    const float *src = partial + (size_t)(tap * cxp + c) * ldp + f;
    float x = __ldg(src);
    for (int sp = 1; sp < splits; ++sp) x = __fadd_rn(x, __ldg(src + (size_t)sp * stride));
    dw[i] = x;
  }
}

bool conv_dw_tc_supported(int kind, int H, int W, int cin, int cout) {
  ConvGeom g;
  const int atom = kind == DLS_GEMM_TF32 ? 32 : 64;   // pixels per k block = channels per 128-byte operand row
  return (kind == DLS_GEMM_BF16 || kind == DLS_GEMM_TF32) && conv_tiling(H, W, atom, &g) && cin >= 1 && cout >= 1 && cout <= tc::BN;
}

int launch_conv_dw_tc(dls_ctx *ctx, int stream, int kind, const void *act, int cxp, const void *dz, int cdp, int batch, int H, int W,
                      int cin, int cout, float *dw, int dzs) {
  using namespace tc;
  const bool tf32 = kind == DLS_GEMM_TF32;
  const int elem = tf32 ? 4 : 2, atom = 128 / elem;
  const CUtensorMapDataType dt = tf32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  DLS_CHECK(conv_dw_tc_supported(kind, H, W, cin, cout) && cxp % atom == 0 && cdp % atom == 0, DLS_ERR_UNSUPPORTED,
            "launch_conv_dw_tc: shape not supported by the implicit-GEMM weight-gradient kernel");
  ConvGeom cg;
  // the k pixels: all of them, or (sampled delta) the H/dzs x W/dzs grid of delta-carrying ones
  DLS_CHECK(conv_tiling(H / dzs, W / dzs, atom, &cg), DLS_ERR_UNSUPPORTED, "launch_conv_dw_tc: the sampled pixel grid does not tile");
  cg.cblks = cxp / atom;
  cg.n_mma = (cout + 15) & ~15;
  cg.px_stride = dzs;
  const long long P = (long long)batch * (H / dzs) * (W / dzs);
  DLS_CHECK(P < (1LL << 31), DLS_ERR_UNSUPPORTED, "launch_conv_dw_tc: too many pixels");
  const int M = 9 * cxp, N = cout;
  CUtensorMap ma, mb, mc;
  {
    const uint64_t dims[4] = {(uint64_t)cxp, (uint64_t)W, (uint64_t)H, (uint64_t)batch};
    const uint64_t str[3] = {(uint64_t)cxp * elem, (uint64_t)cxp * elem * W, (uint64_t)cxp * elem * W * H};
    const uint32_t box[4] = {(uint32_t)atom, (uint32_t)(cg.bw * dzs), (uint32_t)(cg.bh * dzs), (uint32_t)cg.bb};
    const uint32_t est[4] = {1, (uint32_t)dzs, (uint32_t)dzs, 1};
    int rc = make_map_nd(&ma, dt, act, 4, dims, str, box, est, tf32);
    if (rc) return rc;
  }
  int rc = make_map(&mb, dt, elem, dz, (uint64_t)cdp, (uint64_t)P, (uint64_t)cdp, atom, atom, tf32);
  if (rc) return rc;
  const int tiles = (M + BM - 1) / BM;
  const int num_kb = (int)((P + atom - 1) / atom);
  SplitPlan sp;
  int want = std::max(1, std::min(usable_sms(ctx) / tiles, num_kb / 8));
  sp.kb_per_split = (num_kb + want - 1) / want;
  sp.splits = (num_kb + sp.kb_per_split - 1) / sp.kb_per_split;
  const int m_pad = tiles * BM;
  const int ldp = (N + 3) & ~3;
  sp.c_rows = sp.splits * m_pad;
  Block scratch;
  rc = alloc_block(ctx, (size_t)sp.c_rows * ldp * sizeof(float), stream, &scratch);
  if (rc) return rc;
  float *part = (float *)scratch.ptr;
  rc = make_map(&mc, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, part, (uint64_t)N, (uint64_t)sp.c_rows, (uint64_t)ldp, EPI_CHUNK, 32);
  cudaStream_t st = ctx->streams[stream];
  if (!rc && tf32) rc = launch_variant<true, true, true, 2>(ctx, st, ma, mb, mc, part, M, N, (int)P, ldp, GemmEpilogue(), 0, sp, cg);
  else if (!rc) rc = launch_variant<false, true, true, 2>(ctx, st, ma, mb, mc, part, M, N, (int)P, ldp, GemmEpilogue(), 0, sp, cg);
  if (!rc) {
    const int total = 9 * cin * cout;
    conv_dw_reduce_kernel<<<std::min((total + 255) / 256, ctx->sm_count * 8), 256, 0, st>>>(part, dw, sp.splits, (size_t)m_pad * ldp, cxp,
                                                                                            cin, cout, ldp);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) rc = cuda_fail(e, "conv_dw_reduce_kernel", __FILE__, __LINE__);
    ctx->stats.kernel_launches++;
  }
  free_block(ctx, scratch);
  return rc;
}

}  // namespace dls

#ifdef DLS_GEMM_TRACE
extern "C" int dls_debug_gemm_trace(unsigned long long *out, int cap) {
  static unsigned long long host[1024];
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(host, dls::tc::g_gemm_trace, sizeof(host));
  int n = 0;
  for (int i = 0; i < 1024 && n < cap; ++i)
    if (host[i]) out[n++] = host[i];
  memset(host, 0, sizeof(host));
  cudaMemcpyToSymbol(dls::tc::g_gemm_trace, host, sizeof(host));
  return n;
}
#endif
