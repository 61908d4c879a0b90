// radix.cu -- our own onesweep LSD radix sort of (key64, val32) pairs (no CUB / Thrust).
//
// Replaces the vendored suffix sorter call of the reference
// (inmem_psascan_src/initial_partial_sufsort.hpp:272-289 -> libsais()/divsufsort()) as the
// sorting engine of the prefix-doubling block sorter in sufsort.cu.
//
// One histogram kernel reads the keys once and counts every digit of the digit list; each
// digit pass is then ONE kernel: a tile of TILE keys is ranked with warp match-any, the tile's
// digit counts are published to a decoupled look-back chain (status word = flag<<62 | count),
// keys and values are staged through shared memory in digit order and written out in
// coalesced runs.  HBM traffic per pass: 12 B read + 12 B written per pair.
#include "common.cuh"

namespace psc {

constexpr int RADIX = 256;
constexpr int RS_THREADS = 256;
constexpr int RS_WARPS = RS_THREADS / 32;
constexpr int RS_ITEMS = 16;
constexpr int RS_TILE = RS_THREADS * RS_ITEMS;  // 4096 pairs = 48 KB of input per tile
constexpr int MAX_PASSES = 8;

struct ShiftList { int s[MAX_PASSES]; int n; };

constexpr uint64_t ST_MASK = (1ull << 62) - 1;
constexpr uint64_t ST_LOCAL = 1ull << 62;
constexpr uint64_t ST_INCL = 2ull << 62;

__device__ __forceinline__ uint64_t ld_relaxed_u64(const uint64_t *p) {
  uint64_t v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u64(uint64_t *p, uint64_t v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

// ---- histogram of all digits in one read of the keys -----------------------------------
__global__ void __launch_bounds__(512) radix_hist_kernel(const uint64_t *__restrict__ keys, long n,
                                                        ShiftList sl,
                                                        unsigned long long *__restrict__ ghist) {
  __shared__ uint32_t sh[MAX_PASSES * RADIX];
  for (int i = threadIdx.x; i < MAX_PASSES * RADIX; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  long stride = (long)gridDim.x * blockDim.x;
  for (long i = (long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
    uint64_t k = keys[i];
#pragma unroll
    for (int p = 0; p < MAX_PASSES; ++p)
      if (p < sl.n) atomicAdd(&sh[p * RADIX + (int)((k >> sl.s[p]) & 255)], 1u);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < sl.n * RADIX; i += blockDim.x)
    if (sh[i]) atomicAdd(&ghist[i], (unsigned long long)sh[i]);
}

// ---- one digit pass -----------------------------------------------------------------------
__global__ void __launch_bounds__(RS_THREADS)
    radix_pass_kernel(const uint64_t *__restrict__ kin, const uint32_t *__restrict__ vin,
                      uint64_t *__restrict__ kout, uint32_t *__restrict__ vout, long n, int shift,
                      const uint64_t *__restrict__ gscan,  // [256] exclusive global digit offsets
                      uint64_t *status,                    // [tiles][256] look-back chain
                      uint32_t *tile_counter) {
  __shared__ uint32_t warp_hist[RS_WARPS][RADIX];
  __shared__ uint32_t digit_off[RADIX];
  __shared__ long long gbase_adj[RADIX];
  __shared__ uint32_t wsum[RS_WARPS];
  __shared__ uint32_t s_tile;
  __shared__ __align__(16) uint64_t stage[RS_TILE];

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) s_tile = atomicAdd(tile_counter, 1u);
  for (int i = tid; i < RS_WARPS * RADIX; i += RS_THREADS) (&warp_hist[0][0])[i] = 0;
  __syncthreads();
  const long tile = s_tile;
  const long tile_beg = tile * RS_TILE;
  const int valid = (int)min((long)RS_TILE, n - tile_beg);

  // load: element order inside the tile is (warp, item, lane)
  uint64_t key[RS_ITEMS];
  uint32_t val[RS_ITEMS];
  uint16_t rnk[RS_ITEMS];
  const int wbase = warp * (RS_ITEMS * 32) + lane;
#pragma unroll
  for (int k = 0; k < RS_ITEMS; ++k) {
    int e = wbase + k * 32;
    if (e < valid) { key[k] = kin[tile_beg + e]; val[k] = vin[tile_beg + e]; }
    else { key[k] = ~0ull; val[k] = 0; }
  }
  // rank inside the warp
  const uint32_t lane_lt = (1u << lane) - 1;
  uint32_t *wh = warp_hist[warp];
#pragma unroll
  for (int k = 0; k < RS_ITEMS; ++k) {
    int d = (int)((key[k] >> shift) & 255);
    uint32_t peers = __match_any_sync(0xffffffffu, d);
    int leader = 31 - __clz(peers);
    uint32_t base = 0;
    if (lane == leader) { base = wh[d]; wh[d] = base + __popc(peers); }
    base = __shfl_sync(0xffffffffu, base, leader);
    rnk[k] = (uint16_t)(base + __popc(peers & lane_lt));
    __syncwarp();
  }
  __syncthreads();
  // per digit (thread = digit): prefix over warps, tile count
  uint32_t cnt = 0;
  {
#pragma unroll
    for (int w = 0; w < RS_WARPS; ++w) { uint32_t t = warp_hist[w][tid]; warp_hist[w][tid] = cnt; cnt += t; }
  }
  // exclusive scan of the 256 tile counts (padding keys included: they sit at the tile's end)
  uint32_t incl = cnt;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
  if (lane == 31) wsum[warp] = incl;
  __syncthreads();
  uint32_t woff = 0;
#pragma unroll
  for (int w = 0; w < RS_WARPS; ++w) if (w < warp) woff += wsum[w];
  const uint32_t doff = woff + incl - cnt;
  digit_off[tid] = doff;
  // publish + look back
  uint32_t cnt_valid = cnt;
  if (tid == 255) cnt_valid -= (uint32_t)(RS_TILE - valid);
  uint64_t excl = 0;
  uint64_t *my = status + (size_t)tile * RADIX + tid;
  if (tile == 0) {
    st_relaxed_u64(my, ST_INCL | cnt_valid);
  } else {
    st_relaxed_u64(my, ST_LOCAL | cnt_valid);
    long t = tile - 1;
    while (true) {
      uint64_t v = ld_relaxed_u64(status + (size_t)t * RADIX + tid);
      uint64_t f = v & ~ST_MASK;
      if (f == 0) continue;
      excl += v & ST_MASK;
      if (f == ST_INCL) break;
      --t;
    }
    st_relaxed_u64(my, ST_INCL | (excl + cnt_valid));
  }
  gbase_adj[tid] = (long long)(gscan[tid] + excl) - (long long)doff;
  __syncthreads();
  // keys -> shared memory in digit order
#pragma unroll
  for (int k = 0; k < RS_ITEMS; ++k) {
    int d = (int)((key[k] >> shift) & 255);
    uint32_t pos = digit_off[d] + wh[d] + rnk[k];
    rnk[k] = (uint16_t)pos;
    stage[pos] = key[k];
  }
  __syncthreads();
  uint8_t dg[RS_ITEMS];
#pragma unroll
  for (int k = 0; k < RS_ITEMS; ++k) {
    int i = tid + k * RS_THREADS;
    if (i < valid) {
      uint64_t kk = stage[i];
      int d = (int)((kk >> shift) & 255);
      dg[k] = (uint8_t)d;
      kout[gbase_adj[d] + i] = kk;
    }
  }
  __syncthreads();
  uint32_t *vstage = reinterpret_cast<uint32_t *>(stage);
#pragma unroll
  for (int k = 0; k < RS_ITEMS; ++k) vstage[rnk[k]] = val[k];
  __syncthreads();
#pragma unroll
  for (int k = 0; k < RS_ITEMS; ++k) {
    int i = tid + k * RS_THREADS;
    if (i < valid) vout[gbase_adj[dg[k]] + i] = vstage[i];
  }
}

void radix_sort_pairs(psc_context *ctx, uint64_t *keys_a, uint32_t *vals_a, uint64_t *keys_b,
                      uint32_t *vals_b, long n, const int *shifts, int n_shifts,
                      uint64_t **keys_out, uint32_t **vals_out) {
  *keys_out = keys_a;
  *vals_out = vals_a;
  if (n <= 1 || n_shifts <= 0) return;
  if (n_shifts > MAX_PASSES) { set_error("radix_sort_pairs: too many digit passes"); throw CudaError{PSC_EARG}; }
  PhaseTimer timer(ctx, &ctx->stats.ms_radix);
  ShiftList sl;
  sl.n = n_shifts;
  for (int p = 0; p < MAX_PASSES; ++p) sl.s[p] = p < n_shifts ? shifts[p] : 0;

  DevBuf<unsigned long long> ghist(ctx, MAX_PASSES * RADIX);
  ghist.zero();
  int hist_blocks = (int)std::min<long>(div_up(n, 512 * 8), (long)ctx->n_sm * 4);
  radix_hist_kernel<<<hist_blocks, 512, 0, ctx->stream>>>(keys_a, n, sl, ghist.p);
  PSC_LAUNCH_CHECK();
  ctx->launches++;
  std::vector<unsigned long long> h(MAX_PASSES * RADIX);
  d2h(ctx, h.data(), ghist.p, h.size());
  sync(ctx);

  std::vector<uint64_t> scans(MAX_PASSES * RADIX, 0);
  bool skip[MAX_PASSES];
  for (int p = 0; p < n_shifts; ++p) {
    skip[p] = false;
    uint64_t run = 0;
    for (int d = 0; d < RADIX; ++d) {
      if (h[p * RADIX + d] == (unsigned long long)n) skip[p] = true;
      scans[p * RADIX + d] = run;
      run += h[p * RADIX + d];
    }
  }
  DevBuf<uint64_t> gscan(ctx, MAX_PASSES * RADIX);
  h2d(ctx, gscan.p, scans.data(), scans.size());

  long tiles = div_up(n, RS_TILE);
  DevBuf<uint64_t> status(ctx, (size_t)tiles * RADIX);
  DevBuf<uint32_t> counter(ctx, 1);
  uint64_t *kin = keys_a, *kout = keys_b;
  uint32_t *vin = vals_a, *vout = vals_b;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> evs;
  for (int p = 0; p < n_shifts; ++p) {
    if (skip[p]) continue;
    status.zero();
    counter.zero();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0, ctx->stream);
    radix_pass_kernel<<<(unsigned)tiles, RS_THREADS, 0, ctx->stream>>>(
        kin, vin, kout, vout, n, sl.s[p], gscan.p + p * RADIX, status.p, counter.p);
    cudaEventRecord(e1, ctx->stream);
    evs.push_back({e0, e1});
    PSC_LAUNCH_CHECK();
    ctx->launches++;
    ctx->stats.radix_passes++;
    ctx->stats.radix_elems += n;
    std::swap(kin, kout);
    std::swap(vin, vout);
  }
  // the host vector `scans` must outlive the async copy
  sync(ctx);
  for (auto &e : evs) {
    float ms = 0;
    if (cudaEventElapsedTime(&ms, e.first, e.second) == cudaSuccess) ctx->stats.ms_radix_pass += ms;
    cudaEventDestroy(e.first); cudaEventDestroy(e.second);
  }
  *keys_out = kin;
  *vals_out = vin;
}

}  // namespace psc
