// scan.cuh -- block-level scan primitives and the tile-aggregate scan used by every
// device-wide scan in this library (reduce per tile -> scan of aggregates -> apply).
#pragma once
#include "common.cuh"

namespace psc {

constexpr int SC_THREADS = 256;

// Exclusive sum over the block's threads; returns the exclusive prefix, *total = block sum.
// `smem` must hold SC_THREADS/32 + 1 elements.  All threads must call.
template <typename T>
__device__ __forceinline__ T block_excl_sum(T v, T *smem, T *total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  T incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    T t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  __syncthreads();  // protect smem reuse across consecutive calls
  if (lane == 31) smem[warp] = incl;
  __syncthreads();
  T woff = 0, tot = 0;
#pragma unroll
  for (int w = 0; w < SC_THREADS / 32; ++w) {
    T s = smem[w];
    if (w < warp) woff += s;
    tot += s;
  }
  *total = tot;
  return woff + incl - v;
}

// Single-block exclusive scan over `n` tile aggregates (sum), in place; total to *total_d.
void scan_aggregates_sum(psc_context *ctx, uint64_t *agg_d, long n, uint64_t *total_d);

}  // namespace psc
