// scan.cu -- device-wide exclusive scan (uint32 -> uint64), the "device exclusive scan of the
// gap array" of the merge stage (replaces the sequential gap decoding of
// /root/reference/include/merge.hpp:123-156 and gap_array.hpp:116-124).
#include "scan.cuh"

namespace psc {

constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SC_THREADS * SCAN_ITEMS;

__global__ void scan_aggregates_sum_kernel(uint64_t *agg, long n, uint64_t *total_out) {
  __shared__ uint64_t sm[SC_THREADS / 32 + 1];
  uint64_t carry = 0;
  for (long base = 0; base < n; base += SC_THREADS) {
    long i = base + threadIdx.x;
    uint64_t v = i < n ? agg[i] : 0, tot;
    uint64_t ex = block_excl_sum<uint64_t>(v, sm, &tot);
    if (i < n) agg[i] = carry + ex;
    carry += tot;
  }
  if (threadIdx.x == 0 && total_out) *total_out = carry;
}

__global__ void scan_aggregates_max_kernel(uint64_t *agg, long n) {
  __shared__ uint64_t sm[SC_THREADS / 32 + 1];
  uint64_t carry = 0;
  for (long base = 0; base < n; base += SC_THREADS) {
    long i = base + threadIdx.x;
    uint64_t v = i < n ? agg[i] : 0, tot;
    uint64_t ex = block_excl_max<uint64_t>(v, sm, &tot);
    if (i < n) agg[i] = max(carry, ex);
    carry = max(carry, tot);
  }
}

__global__ void __launch_bounds__(SC_THREADS)
    scan_reduce_u32_kernel(const uint32_t *__restrict__ in, long n, uint64_t *__restrict__ agg) {
  __shared__ uint64_t sm[SC_THREADS / 32 + 1];
  long base = (long)blockIdx.x * SCAN_TILE + (long)threadIdx.x * SCAN_ITEMS;
  uint64_t s = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) if (base + k < n) s += in[base + k];
  uint64_t tot;
  block_excl_sum<uint64_t>(s, sm, &tot);
  if (threadIdx.x == 0) agg[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(SC_THREADS)
    scan_apply_u32_kernel(const uint32_t *__restrict__ in, long n, const uint64_t *__restrict__ agg,
                          uint64_t *__restrict__ out) {
  __shared__ uint64_t sm[SC_THREADS / 32 + 1];
  long base = (long)blockIdx.x * SCAN_TILE + (long)threadIdx.x * SCAN_ITEMS;
  uint32_t v[SCAN_ITEMS];
  uint64_t s = 0;
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) { v[k] = base + k < n ? in[base + k] : 0; s += v[k]; }
  uint64_t tot;
  uint64_t ex = block_excl_sum<uint64_t>(s, sm, &tot) + agg[blockIdx.x];
#pragma unroll
  for (int k = 0; k < SCAN_ITEMS; ++k) {
    if (base + k < n) out[base + k] = ex;
    ex += v[k];
  }
}

// one launch for both aggregate arrays of the prefix-doubling grouping: exclusive sum of `asum`
// (total to *total_out) and exclusive running max of `amax`, 1024 threads per step
__global__ void __launch_bounds__(1024)
    scan_aggregates_sum_max_kernel(uint64_t *asum, uint64_t *amax, long n, uint64_t *total_out) {
  __shared__ uint64_t ws[32], wm[32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint64_t carry_s = 0, carry_m = 0;
  for (long base = 0; base < n; base += 1024) {
    long i = base + threadIdx.x;
    uint64_t vs = i < n ? asum[i] : 0, vm = i < n ? amax[i] : 0;
    uint64_t is = vs, im = vm;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint64_t ts = __shfl_up_sync(0xffffffffu, is, o), tm = __shfl_up_sync(0xffffffffu, im, o);
      if (lane >= o) { is += ts; im = max(im, tm); }
    }
    if (lane == 31) { ws[warp] = is; wm[warp] = im; }
    __syncthreads();
    uint64_t os = 0, om = 0, tots = 0, totm = 0;
    for (int w = 0; w < 32; ++w) {
      uint64_t a = ws[w], b = wm[w];
      if (w < warp) { os += a; om = max(om, b); }
      tots += a; totm = max(totm, b);
    }
    uint64_t pm = __shfl_up_sync(0xffffffffu, im, 1);
    if (lane == 0) pm = 0;
    if (i < n) { asum[i] = carry_s + os + is - vs; amax[i] = max(carry_m, max(om, pm)); }
    carry_s += tots;
    carry_m = max(carry_m, totm);
    __syncthreads();
  }
  if (threadIdx.x == 0 && total_out) *total_out = carry_s;
}
void scan_aggregates_sum_max(psc_context *ctx, uint64_t *asum_d, uint64_t *amax_d, long n, uint64_t *total_d) {
  scan_aggregates_sum_max_kernel<<<1, 1024, 0, ctx->stream>>>(asum_d, amax_d, n, total_d);
  PSC_LAUNCH_CHECK();
  ctx->launches++;
}

void scan_aggregates_sum(psc_context *ctx, uint64_t *agg_d, long n, uint64_t *total_d) {
  scan_aggregates_sum_kernel<<<1, SC_THREADS, 0, ctx->stream>>>(agg_d, n, total_d);
  PSC_LAUNCH_CHECK();
  ctx->launches++;
}
void scan_aggregates_max(psc_context *ctx, uint64_t *agg_d, long n) {
  scan_aggregates_max_kernel<<<1, SC_THREADS, 0, ctx->stream>>>(agg_d, n);
  PSC_LAUNCH_CHECK();
  ctx->launches++;
}

void exclusive_scan_u32_to_u64(psc_context *ctx, const uint32_t *in, uint64_t *out, long n,
                               uint64_t *total_d) {
  if (n <= 0) {
    if (total_d) PSC_CUDA(cudaMemsetAsync(total_d, 0, 8, ctx->stream));
    return;
  }
  long tiles = div_up(n, SCAN_TILE);
  DevBuf<uint64_t> agg(ctx, tiles);
  scan_reduce_u32_kernel<<<(unsigned)tiles, SC_THREADS, 0, ctx->stream>>>(in, n, agg.p);
  PSC_LAUNCH_CHECK();
  scan_aggregates_sum(ctx, agg.p, tiles, total_d);
  scan_apply_u32_kernel<<<(unsigned)tiles, SC_THREADS, 0, ctx->stream>>>(in, n, agg.p, out);
  PSC_LAUNCH_CHECK();
  ctx->launches += 2;
}

}  // namespace psc
