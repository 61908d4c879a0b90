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

// Exclusive max over the block's threads: max of v over strictly earlier threads (identity 0);
// *total = max over the whole block.  Used for "index of the last group head so far".
template <typename T>
__device__ __forceinline__ T block_excl_max(T v, T *smem, T *total) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  T incl = v;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    T t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl = max(incl, t);
  }
  __syncthreads();
  if (lane == 31) smem[warp] = incl;
  __syncthreads();
  T wmax = 0, tot = 0;
#pragma unroll
  for (int w = 0; w < SC_THREADS / 32; ++w) {
    T s = smem[w];
    if (w < warp) wmax = max(wmax, s);
    tot = max(tot, s);
  }
  *total = tot;
  T prev = __shfl_up_sync(0xffffffffu, incl, 1);
  if (lane == 0) prev = 0;
  return max(wmax, prev);
}

// Single-block exclusive scan over `n` tile aggregates (sum), in place; total to *total_d.
void scan_aggregates_sum(psc_context *ctx, uint64_t *agg_d, long n, uint64_t *total_d);
// Single-block exclusive running max over `n` tile aggregates, in place (identity 0).
void scan_aggregates_max(psc_context *ctx, uint64_t *agg_d, long n);
void scan_aggregates_sum_max(psc_context *ctx, uint64_t *asum_d, uint64_t *amax_d, long n, uint64_t *total_d);

}  // namespace psc
