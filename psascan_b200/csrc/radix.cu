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
// Shared memory (dynamic): stage_k[TILE] u64 | stage_v[TILE] u32 | gbase_adj[256] u32 |
// warp_hist[WARPS][256] u32 | digit_off[256] u32 | wsum[8] u32 + tile id
template <int THREADS, int ITEMS>
struct PassCfg {
  static constexpr int WARPS = THREADS / 32;
  static constexpr int TILE = THREADS * ITEMS;
  static constexpr size_t SMEM = (size_t)TILE * 12 + RADIX * 4 + (size_t)WARPS * RADIX * 4 + RADIX * 4 + 64;
};

// lanes of the warp holding the same digit.  USE_MATCH: match.any (one ADU instruction, measured
// ~60 cycles of the SM's ADU pipe per warp) else eight ballots (ALU/vote pipes; 1.55x faster per pass).
// The ballot form is written in PTX: four instructions per digit bit (test, vote, select, and-xnor);
// the compiler's own translation of the C++ took seven.
template <bool USE_MATCH>
__device__ __forceinline__ uint32_t digit_peers(int d) {
  if (USE_MATCH) return __match_any_sync(0xffffffffu, d);
  uint32_t peers;
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      ".reg .b32 t, bal, m;\n\t"
      "mov.b32 %0, 0xffffffff;\n\t"
#define PSC_PEER_BIT(B)                                        \
      "and.b32 t, %1, " #B ";\n\t"                             \
      "setp.ne.u32 p, t, 0;\n\t"                               \
      "vote.sync.ballot.b32 bal, p, 0xffffffff;\n\t"           \
      "selp.b32 m, 0xffffffff, 0, p;\n\t"                      \
      "lop3.b32 %0, %0, bal, m, 0x90;\n\t" /* a & ~(b ^ c) */
      PSC_PEER_BIT(1) PSC_PEER_BIT(2) PSC_PEER_BIT(4) PSC_PEER_BIT(8)
      PSC_PEER_BIT(16) PSC_PEER_BIT(32) PSC_PEER_BIT(64) PSC_PEER_BIT(128)
#undef PSC_PEER_BIT
      "}"
      : "=r"(peers)
      : "r"(d));
  return peers;
}

// Order of work in a tile (ncu stall samples of the previous version: 15 % of the time was spent
// waiting in the look-back, 25 % on the initial loads):
//   load -> early per-warp counts -> tile counts, digit offsets, PUBLISH local counts
//   -> ranks (chunks of 4 items) and staging of keys AND values in shared memory
//   -> only now the look-back (predecessors have long published) -> coalesced run writes.
template <int THREADS, int ITEMS, int MINB, int MATCH>
__global__ void __launch_bounds__(THREADS, MINB)
    radix_pass_kernel(const uint64_t *__restrict__ kin, const uint32_t *__restrict__ vin,
                      uint64_t *__restrict__ kout, uint32_t *__restrict__ vout, long n, int shift,
                      const uint64_t *__restrict__ gscan,  // [256] exclusive global digit offsets
                      uint64_t *status,                    // [tiles][256] look-back chain
                      uint32_t *tile_counter,
                      int pf_dist) {  // > 0: pull the tile `pf_dist` ahead into L2 (bulk prefetch)
  constexpr int WARPS = THREADS / 32;
  constexpr int TILE = THREADS * ITEMS;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint64_t *stage_k = reinterpret_cast<uint64_t *>(smem_raw);
  uint32_t *stage_v = reinterpret_cast<uint32_t *>(stage_k + TILE);
  uint32_t *gbase_adj = stage_v + TILE;          // [RADIX]
  uint32_t *warp_hist = gbase_adj + RADIX;       // [WARPS][RADIX]
  uint32_t *digit_off = warp_hist + WARPS * RADIX;
  uint32_t *wsum = digit_off + RADIX;            // [8] + tile id at [8]

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  if (tid == 0) wsum[8] = atomicAdd(tile_counter, 1u);
  for (int i = tid; i < WARPS * RADIX; i += THREADS) warp_hist[i] = 0;
  __syncthreads();
  const long tile = wsum[8];
  const long tile_beg = tile * TILE;
  const int valid = (int)min((long)TILE, n - tile_beg);
  if (pf_dist > 0 && tid == THREADS - 32) {
    // tiles are handed out in order, so tile + (number of resident CTAs) is what this SM slot
    // will most likely load next: have the copy engine fetch it into L2 while this tile is processed
    const long pb = (tile + pf_dist) * TILE;
    if (pb < n) {
      const long cnt = min((long)TILE, n - pb);
      const uint32_t kb = (uint32_t)((cnt * 8) & ~15L), vb = (uint32_t)((cnt * 4) & ~15L);
      if (kb) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(kin + pb), "r"(kb) : "memory");
      if (vin && vb) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(vin + pb), "r"(vb) : "memory");
    }
  }

  // load: element order inside the tile is (warp, item, lane)
  uint64_t key[ITEMS];
  uint32_t val[ITEMS];
  const int wbase = warp * (ITEMS * 32) + lane;
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    int e = wbase + k * 32;
    key[k] = (e < valid) ? kin[tile_beg + e] : ~0ull;
  }
  if (vin) {
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) { int e = wbase + k * 32; val[k] = (e < valid) ? vin[tile_beg + e] : 0u; }
  } else {  // implicit values 0..n-1 (first pass of round 0)
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) val[k] = (uint32_t)(tile_beg + wbase + k * 32);
  }
  // early counts: per-warp digit histogram (fire-and-forget shared atomics, no dependent chain)
  uint32_t *wh = warp_hist + warp * RADIX;
  uint8_t dgt[ITEMS];
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) { dgt[k] = (uint8_t)((key[k] >> shift) & 255); atomicAdd(&wh[dgt[k]], 1u); }
  __syncthreads();
  // per digit (thread = digit): tile count, exclusive scan over digits, per-warp start offsets
  uint32_t cnt = 0, incl = 0;
  if (tid < RADIX) {
#pragma unroll
    for (int w = 0; w < WARPS; ++w) cnt += warp_hist[w * RADIX + tid];
    // publish the local count at once: successors can add it while this tile is still ranking
    uint32_t cnt_valid = cnt - ((tid == 255) ? (uint32_t)(TILE - valid) : 0u);
    st_relaxed_u64(status + (size_t)tile * RADIX + tid, (tile == 0 ? ST_INCL : ST_LOCAL) | cnt_valid);
    incl = cnt;  // padding keys included: they sit at the tile's end
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) wsum[warp] = incl;
  }
  __syncthreads();
  uint32_t doff = 0;
  if (tid < RADIX) {
    uint32_t woff = 0;
#pragma unroll
    for (int w = 0; w < 8; ++w) if (w < warp) woff += wsum[w];
    doff = woff + incl - cnt;
    digit_off[tid] = doff;
    uint32_t run = doff;
#pragma unroll
    for (int w = 0; w < WARPS; ++w) { uint32_t t = warp_hist[w * RADIX + tid]; warp_hist[w * RADIX + tid] = run; run += t; }
  }
  __syncthreads();
  // ranks: peers of each item, then one shared atomic per distinct digit (issued back to back;
  // same-address atomics of one warp retire in issue order, which keeps the sort stable)
  const uint32_t lane_lt = (1u << lane) - 1;
#pragma unroll
  for (int c0 = 0; c0 < ITEMS; c0 += 4) {
    uint32_t peers[4], base[4];
    int lead[4];
#pragma unroll
    for (int j = 0; j < 4; ++j)
      peers[j] = (c0 + j < MATCH) ? digit_peers<true>(dgt[c0 + j]) : digit_peers<false>(dgt[c0 + j]);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      lead[j] = 31 - __clz(peers[j]);
      base[j] = 0;
      if (lane == lead[j]) base[j] = atomicAdd(&wh[dgt[c0 + j]], (uint32_t)__popc(peers[j]));
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      uint32_t pos = __shfl_sync(0xffffffffu, base[j], lead[j]) + __popc(peers[j] & lane_lt);
      stage_k[pos] = key[c0 + j];
      stage_v[pos] = val[c0 + j];
    }
  }
  // look back (thread = digit) while the other warps finish staging
  if (tid < RADIX) {
    uint64_t excl = 0;
    if (tile > 0) {
      // predecessors are read LB at a time (independent loads, one L2 round trip): measured
      // depth of the walk was 14 tiles = 14 dependent round trips with one load per step
      constexpr int LB = 8;
      long t = tile - 1;
      bool done = false;
      while (!done) {
        uint64_t v[LB];
#pragma unroll
        for (int j = 0; j < LB; ++j) v[j] = (t - j >= 0) ? ld_relaxed_u64(status + (size_t)(t - j) * RADIX + tid) : ST_INCL;
        int used = 0;
#pragma unroll
        for (int j = 0; j < LB; ++j) {
          uint64_t f = v[j] & ~ST_MASK;
          if (!done && used == j && f != 0) {
            excl += v[j] & ST_MASK;
            used = j + 1;
            if (f == ST_INCL) done = true;
          }
        }
        t -= used;  // an unpublished status stops the batch; it is read again
      }
      uint32_t cnt_valid = cnt - ((tid == 255) ? (uint32_t)(TILE - valid) : 0u);
      st_relaxed_u64(status + (size_t)tile * RADIX + tid, ST_INCL | (excl + cnt_valid));
    }
    gbase_adj[tid] = (uint32_t)(gscan[tid] + excl) - doff;  // n < 2^32: index arithmetic mod 2^32
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) {
    int i = tid + k * THREADS;
    if (i < valid) {
      uint64_t kk = stage_k[i];
      uint32_t o = gbase_adj[(int)((kk >> shift) & 255)] + (uint32_t)i;
      kout[o] = kk;
      vout[o] = stage_v[i];
    }
  }
}

typedef void (*PassFn)(const uint64_t *, const uint32_t *, uint64_t *, uint32_t *, long, int, const uint64_t *,
                       uint64_t *, uint32_t *, int);
struct PassVariant { const char *name; PassFn fn; int threads, tile; size_t smem; int minb; };
#define PASS_VARIANT(T, I, B, M) \
  { #T "x" #I "b" #B "m" #M, radix_pass_kernel<T, I, B, M>, T, PassCfg<T, I>::TILE, PassCfg<T, I>::SMEM, B }
// first entry = default (measured best on B200: 2.17 ms per pass of 2^28 pairs = 2.97 TB/s);
// PSC_RADIX_VARIANT selects another one for tuning runs (tools/tune_radix.py)
static const PassVariant g_variants[] = {
    PASS_VARIANT(512, 12, 2, 0), PASS_VARIANT(384, 12, 3, 0), PASS_VARIANT(512, 8, 3, 0), PASS_VARIANT(512, 16, 2, 0),
    PASS_VARIANT(256, 12, 4, 0), PASS_VARIANT(1024, 8, 1, 0), PASS_VARIANT(512, 12, 2, 12),
    PASS_VARIANT(256, 12, 5, 0), PASS_VARIANT(256, 8, 6, 0), PASS_VARIANT(384, 8, 4, 0), PASS_VARIANT(256, 16, 3, 0),
    PASS_VARIANT(128, 12, 8, 0), PASS_VARIANT(256, 8, 8, 0),
};
static const PassVariant &pass_variant() {
  static int idx = -1;
  if (idx < 0) {
    idx = 0;
    if (const char *e = getenv("PSC_RADIX_VARIANT")) {
      for (size_t i = 0; i < sizeof(g_variants) / sizeof(g_variants[0]); ++i)
        if (!strcmp(e, g_variants[i].name)) idx = (int)i;
    }
    for (auto &v : g_variants)
      cudaFuncSetAttribute((const void *)v.fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)v.smem);
  }
  return g_variants[idx];
}

__global__ void iota_kernel(uint32_t *v, long n) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) v[i] = (uint32_t)i;
}

void radix_sort_pairs(psc_context *ctx, uint64_t *keys_a, uint32_t *vals_a, uint64_t *keys_b,
                      uint32_t *vals_b, long n, const int *shifts, int n_shifts,
                      uint64_t **keys_out, uint32_t **vals_out, const unsigned long long *pre_hist_d,
                      bool vals_iota) {
  *keys_out = keys_a;
  *vals_out = vals_a;
  if (n <= 1 || n_shifts <= 0) {
    if (vals_iota && n > 0) { iota_kernel<<<(unsigned)div_up(n, 256), 256, 0, ctx->stream>>>(vals_a, n); PSC_LAUNCH_CHECK(); }
    return;
  }
  if (n_shifts > MAX_PASSES) { set_error("radix_sort_pairs: too many digit passes"); throw CudaError{PSC_EARG}; }
  PhaseTimer timer(ctx, &ctx->stats.ms_radix);
  ShiftList sl;
  sl.n = n_shifts;
  for (int p = 0; p < MAX_PASSES; ++p) sl.s[p] = p < n_shifts ? shifts[p] : 0;

  DevBuf<unsigned long long> ghist;
  if (!pre_hist_d) {
    ghist.alloc(ctx, MAX_PASSES * RADIX);
    ghist.zero();
    int hist_blocks = (int)std::min<long>(div_up(n, 512 * 8), (long)ctx->n_sm * 4);
    radix_hist_kernel<<<hist_blocks, 512, 0, ctx->stream>>>(keys_a, n, sl, ghist.p);
    PSC_LAUNCH_CHECK();
    ctx->launches++;
    pre_hist_d = ghist.p;
  }
  std::vector<unsigned long long> h(MAX_PASSES * RADIX);
  d2h(ctx, h.data(), pre_hist_d, h.size());
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

  const PassVariant &pv = pass_variant();
  long tiles = div_up(n, pv.tile);
  static int pf_mult = -1;  // prefetch distance in units of resident CTAs (PSC_RADIX_PF, 0 = off)
  if (pf_mult < 0) { const char *e = getenv("PSC_RADIX_PF"); pf_mult = e ? atoi(e) : 1; }
  const int pf_dist = pf_mult * ctx->n_sm * pv.minb;
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
    pv.fn<<<(unsigned)tiles, pv.threads, pv.smem, ctx->stream>>>(
        kin, vals_iota ? nullptr : vin, kout, vout, n, sl.s[p], gscan.p + p * RADIX, status.p, counter.p, pf_dist);
    vals_iota = false;
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
  if (vals_iota) {  // every pass was skipped (all keys equal): materialise the implicit values
    iota_kernel<<<(unsigned)div_up(n, 256), 256, 0, ctx->stream>>>(vin, n);
    PSC_LAUNCH_CHECK();
  }
  *keys_out = kin;
  *vals_out = vin;
}

}  // namespace psc
