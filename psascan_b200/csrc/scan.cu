// scan.cu -- device-wide exclusive scan (uint32 -> uint64), the "device exclusive scan of the
// gap array" of the merge stage (replaces the sequential gap decoding of
// /root/reference/include/merge.hpp:123-156 and gap_array.hpp:116-124).
#include "scan.cuh"

namespace psc {

constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SC_THREADS * SCAN_ITEMS;
static_assert(SCAN_TILE == COARSE_TILE, "coarse carries are the scan's tile sums");

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
                          uint64_t *__restrict__ out, uint64_t sub_after_first) {
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
    if (base + k < n) out[base + k] = (base + k > 0) ? ex - sub_after_first : ex;
    ex += v[k];
  }
}

// Tile aggregates -> exclusive prefixes, in place, by ONE block: 1024 threads x 8 consecutive
// aggregates per step.  The tile kernels before and after it carry the real traffic.
constexpr int AG_THREADS = 1024;
constexpr int AG_ITEMS = 8;

template <bool SUM, bool MAX>
__global__ void __launch_bounds__(AG_THREADS)
    scan_aggregates_kernel(uint64_t *asum, uint64_t *amax, long n, uint64_t *total_out) {
  __shared__ uint64_t ws[32], wm[32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint64_t carry_s = 0, carry_m = 0;
  for (long base = 0; base < n; base += AG_THREADS * AG_ITEMS) {
    const long i0 = base + (long)threadIdx.x * AG_ITEMS;
    uint64_t vs[AG_ITEMS], vm[AG_ITEMS];
    uint64_t ts = 0, tm = 0;
#pragma unroll
    for (int k = 0; k < AG_ITEMS; ++k) {
      vs[k] = (SUM && i0 + k < n) ? asum[i0 + k] : 0;
      vm[k] = (MAX && i0 + k < n) ? amax[i0 + k] : 0;
      ts += vs[k];
      tm = max(tm, vm[k]);
    }
    uint64_t is = ts, im = tm;  // inclusive over the threads of the warp
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      uint64_t a = __shfl_up_sync(0xffffffffu, is, o), b = __shfl_up_sync(0xffffffffu, im, o);
      if (lane >= o) { is += a; im = max(im, b); }
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
    uint64_t ex_s = carry_s + os + is - ts, ex_m = max(carry_m, max(om, pm));
#pragma unroll
    for (int k = 0; k < AG_ITEMS; ++k) {
      if (i0 + k < n) {
        if (SUM) asum[i0 + k] = ex_s;
        if (MAX) amax[i0 + k] = ex_m;
      }
      ex_s += vs[k];
      ex_m = max(ex_m, vm[k]);
    }
    carry_s += tots;
    carry_m = max(carry_m, totm);
    __syncthreads();
  }
  if (SUM && threadIdx.x == 0 && total_out) *total_out = carry_s;
}

void scan_aggregates_sum(psc_context *ctx, uint64_t *agg_d, long n, uint64_t *total_d) {
  scan_aggregates_kernel<true, false><<<1, AG_THREADS, 0, ctx->stream>>>(agg_d, nullptr, n, total_d);
  PSC_LAUNCH_CHECK();
  ctx->launches++;
}
// carry[j] = sum of in[0 .. j * SCAN_TILE) for j in [0, tiles), carry[tiles] = sum of everything
// (tiles = ceil(n / SCAN_TILE)): the coarse level of a two-level search in the scanned array
long coarse_carry_tiles(long n) { return div_up(n, (long)SCAN_TILE); }
void coarse_carries_u32(psc_context *ctx, const uint32_t *in, long n, uint64_t *carry_d) {
  const long tiles = coarse_carry_tiles(n);
  if (tiles <= 0) { PSC_CUDA(cudaMemsetAsync(carry_d, 0, 8, ctx->stream)); return; }
  scan_reduce_u32_kernel<<<(unsigned)tiles, SC_THREADS, 0, ctx->stream>>>(in, n, carry_d);
  PSC_LAUNCH_CHECK();
  ctx->launches++;
  scan_aggregates_sum(ctx, carry_d, tiles, carry_d + tiles);
}

void exclusive_scan_u32_to_u64(psc_context *ctx, const uint32_t *in, uint64_t *out, long n,
                               uint64_t *total_d, uint64_t sub_after_first) {
  if (n <= 0) {
    if (total_d) PSC_CUDA(cudaMemsetAsync(total_d, 0, 8, ctx->stream));
    return;
  }
  long tiles = div_up(n, SCAN_TILE);
  DevBuf<uint64_t> agg(ctx, tiles);
  scan_reduce_u32_kernel<<<(unsigned)tiles, SC_THREADS, 0, ctx->stream>>>(in, n, agg.p);
  PSC_LAUNCH_CHECK();
  scan_aggregates_sum(ctx, agg.p, tiles, total_d);
  scan_apply_u32_kernel<<<(unsigned)tiles, SC_THREADS, 0, ctx->stream>>>(in, n, agg.p, out, sub_after_first);
  PSC_LAUNCH_CHECK();
  ctx->launches += 2;
}

}  // namespace psc
