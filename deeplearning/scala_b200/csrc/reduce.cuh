// reduce.cuh -- deterministic block / grid sums shared by the interpreter kernels and the reduction kernels.
#pragma once
#include <cuda_runtime.h>

namespace dls {

constexpr int kBlock = 256;

struct ReduceArgs {
  float *partials;
  unsigned int *counter;
  float *out;
};

// ---- deterministic block sum + last-block finish (one launch; replaces the two-stage K3 of Tensors:358-392, :697-754)
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ float block_sum(float v) {
  __shared__ float warp_part[kBlock / 32];
  v = warp_sum(v);
  __syncthreads();
  if ((threadIdx.x & 31) == 0) warp_part[threadIdx.x >> 5] = v;
  __syncthreads();
  float r = 0.f;
  if (threadIdx.x < 32) {
    r = threadIdx.x < kBlock / 32 ? warp_part[threadIdx.x] : 0.f;
    r = warp_sum(r);
  }
  return r;  // valid in warp 0
}

__device__ __forceinline__ void grid_sum_finish(float thread_sum, const ReduceArgs &R) {
  __shared__ bool is_last;
  float b = block_sum(thread_sum);
  if (threadIdx.x == 0) {
    R.partials[blockIdx.x] = b;
    __threadfence();
    unsigned int ticket = atomicAdd(R.counter, 1u);
    is_last = ticket == gridDim.x - 1;
  }
  __syncthreads();
  if (!is_last) return;
  __threadfence();
  float s = 0.f;
  for (unsigned int i = threadIdx.x; i < gridDim.x; i += kBlock) s += __ldcg(R.partials + i);
  s = block_sum(s);
  if (threadIdx.x == 0) {
    R.out[0] = s;
    *R.counter = 0u;  // ready for the next launch on this stream
  }
}


}  // namespace dls
