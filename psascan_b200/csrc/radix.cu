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
// Shared memory (dynamic): stage_k[TILE] u64 | stage_v[TILE] u32 (PAIRS) | stage_d[TILE] u8 (KEYGEN) |
// gbase_adj[256] u32 | warp_hist[WARPS][256] u32 | digit_off[256] u32 | wsum[8] u32 + tile id
template <int THREADS, int ITEMS, bool PAIRS, bool KEYGEN>
struct PassCfg {
  static constexpr int WARPS = THREADS / 32;
  static constexpr int TILE = THREADS * ITEMS;
  static constexpr size_t SMEM = (size_t)TILE * (8 + (PAIRS ? 4 : 0) + (KEYGEN ? 1 : 0)) + RADIX * 4 +
                                 (size_t)WARPS * RADIX * 4 + RADIX * 4 + 64;
};

// lanes of the warp holding the same digit: eight ballots (ALU/vote pipes) instead of match.any (one
// ADU instruction, measured ~60 cycles of the SM's ADU pipe per warp; the ballots are 1.55x faster per
// pass).  Written in PTX: four instructions per digit bit (test, vote, select, and-xnor); the
// compiler's own translation of the C++ took seven.
__device__ __forceinline__ uint32_t digit_peers(int d) {
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

// Decoupled look-back of one digit: sum of the counts of the tiles before `tile`.  Predecessors are
// read LB at a time (independent loads, one L2 round trip per batch).  Measured per pass of 2^28 pairs:
// LB = 1: 1.78 ms, 2: 1.67, 4: 1.58, 8: 1.73, 16: 2.03, 32: 2.26 -- most walks end within a few tiles
// and every status load that is not needed competes with the tile data for the L2.
template <int LB>
__device__ __forceinline__ void look_back_consume(const uint64_t (&v)[LB], uint64_t &excl, long &t, bool &done) {
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
template <int LB>
__device__ __forceinline__ void look_back_load(const uint64_t *status, long t, int tid, uint64_t (&v)[LB]) {
#pragma unroll
  for (int j = 0; j < LB; ++j) v[j] = (t - j >= 0) ? ld_relaxed_u64(status + (size_t)(t - j) * RADIX + tid) : ST_INCL;
}
// continues the walk at tile t (the caller may already have consumed a prefetched batch)
template <int LB>
__device__ __forceinline__ uint64_t look_back(const uint64_t *status, long t, int tid, int nap, uint64_t excl, bool done) {
  while (!done) {
    uint64_t v[LB];
    look_back_load<LB>(status, t, tid, v);
    const long t0 = t;
    look_back_consume<LB>(v, excl, t, done);
    if (t == t0 && nap > 0) __nanosleep(nap);
  }
  return excl;
}

// Order of work in a tile (ncu stall samples of the previous version: 15 % of the time was spent
// waiting in the look-back, 25 % on the initial loads):
//   load -> early per-warp counts -> tile counts, digit offsets, PUBLISH local counts
//   -> ranks (chunks of 4 items) and staging of keys AND values in shared memory
//   -> only now the look-back (predecessors have long published) -> coalesced run writes.
// PAIRS = false: keys only (the block sorter's packed mode keeps the text position in the low key bits).
// KEYGEN: first pass of round 0 for 8-bit renamed codes: the keys are built from the text tile (staged
// and renamed in shared memory) instead of being read, values are the positions themselves, and the
// digit is any byte of the 8 symbols at the position, also one that the stored key does not keep.
template <int THREADS, int ITEMS, int MINB, bool PAIRS, bool KEYGEN, int CH = 4>
__global__ void __launch_bounds__(THREADS, MINB)
    radix_pass_kernel(const uint64_t *__restrict__ kin, const uint32_t *__restrict__ vin,
                      uint64_t *__restrict__ kout, uint32_t *__restrict__ vout, long n, int shift,
                      const uint64_t *__restrict__ gscan,  // [256] exclusive global digit offsets
                      uint64_t *status,                    // [tiles][256] look-back chain
                      uint32_t *tile_counter,
                      int pf_dist,  // > 0: pull the tile `pf_dist` ahead into L2 (bulk prefetch)
                      int lb_mode,  // look-back tuning (PSC_RADIX_LB): batch size and back-off, 0 = default
                      const __grid_constant__ KeyGen kg) {
  constexpr int WARPS = THREADS / 32;
  constexpr int TILE = THREADS * ITEMS;
  extern __shared__ __align__(16) unsigned char smem_raw[];
  uint64_t *stage_k = reinterpret_cast<uint64_t *>(smem_raw);
  uint32_t *stage_v = reinterpret_cast<uint32_t *>(stage_k + TILE);
  uint8_t *stage_d = reinterpret_cast<uint8_t *>(stage_v + (PAIRS ? TILE : 0));
  uint32_t *gbase_adj = reinterpret_cast<uint32_t *>(stage_d + (KEYGEN ? TILE : 0));  // [RADIX]
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
  if (!KEYGEN && pf_dist > 0 && tid == THREADS - 32) {
    // tiles are handed out in order, so tile + (number of resident CTAs) is what this SM slot
    // will most likely load next: have the copy engine fetch it into L2 while this tile is processed
    const long pb = (tile + pf_dist) * TILE;
    if (pb < n) {
      const long cnt = min((long)TILE, n - pb);
      const uint32_t kb = (uint32_t)((cnt * 8) & ~15L), vb = (uint32_t)((cnt * 4) & ~15L);
      if (kb) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(kin + pb), "r"(kb) : "memory");
      if (PAIRS && vin && vb) asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(vin + pb), "r"(vb) : "memory");
    }
  }

  // load: element order inside the tile is (warp, item, lane)
  uint64_t key[ITEMS];
  uint32_t val[PAIRS ? ITEMS : 1];
  uint8_t dgt[ITEMS];
  const int wbase = warp * (ITEMS * 32) + lane;
  if constexpr (KEYGEN) {
    // the tile's symbols (+8 for the last keys), renamed (initial_partial_sufsort.hpp:61-80), as bytes
    // in the not yet used staging area
    __shared__ uint8_t s_code[256];
    uint32_t *cw = reinterpret_cast<uint32_t *>(stage_k);
    for (int i = tid; i < 256; i += THREADS) s_code[i] = kg.code[i];
    __syncthreads();
    const uint32_t last4 = (uint32_t)kg.last * 0x01010101u;
    auto rename4 = [&](uint32_t t4, long p, int nb) -> uint32_t {
      uint32_t packed = (uint32_t)s_code[t4 & 255] | ((uint32_t)s_code[(t4 >> 8) & 255] << 8) |
                        ((uint32_t)s_code[(t4 >> 16) & 255] << 16) | ((uint32_t)s_code[t4 >> 24] << 24);
      if (__vcmpeq4(t4, last4)) {  // an occurrence of the block's last symbol: the gt_end bit decides its code
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const long i = p + j;
          if (((t4 >> (8 * j)) & 255) == (uint32_t)kg.last &&
              (i == kg.m - 1 || kg.gt_end == nullptr || bit_get(kg.gt_end, i + 1)))
            packed = (packed & ~(255u << (8 * j))) | ((uint32_t)kg.code_last_high << (8 * j));
        }
      }
      if (nb < 4) packed &= (nb <= 0) ? 0u : (0xffffffffu >> (8 * (4 - nb)));
      return packed;
    };
    if ((((uintptr_t)kg.text + (uintptr_t)tile_beg) & 15) == 0 && tile_beg + TILE + 8 <= kg.m) {
      // whole tile inside the text and 16-byte aligned: one 128-bit load per thread
      for (int q = tid; q < TILE / 16; q += THREADS) {
        const uint4 t = __ldg(reinterpret_cast<const uint4 *>(kg.text + tile_beg) + q);
        const long p = tile_beg + 16 * q;
        *reinterpret_cast<uint4 *>(cw + 4 * q) =
            make_uint4(rename4(t.x, p, 4), rename4(t.y, p + 4, 4), rename4(t.z, p + 8, 4), rename4(t.w, p + 12, 4));
      }
      if (tid < 2) {
        const long p = tile_beg + TILE + 4 * tid;
        cw[TILE / 4 + tid] = rename4(*reinterpret_cast<const uint32_t *>(kg.text + p), p, 4);
      }
    } else {
      for (int w = tid; w < TILE / 4 + 2; w += THREADS) {
        const long p = tile_beg + 4 * w;
        uint32_t t4 = 0;
        const int nb = (int)max(0L, min(4L, kg.m - p));
        if (p + 8 <= kg.m) t4 = load_u32_any(kg.text + p);
        else for (int j = 0; j < nb; ++j) t4 |= (uint32_t)kg.text[p + j] << (8 * j);
        cw[w] = rename4(t4, p, nb);
      }
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) {
      const int e = wbase + k * 32;
      const int w = e >> 2, sh = (e & 3) * 8;
      const uint32_t w0 = cw[w], w1 = cw[w + 1], w2 = cw[w + 2];
      const uint32_t lo = __funnelshift_r(w0, w1, sh), hi = __funnelshift_r(w1, w2, sh);
      const uint64_t be = ((uint64_t)__byte_perm(lo, 0, 0x0123) << 32) | (uint64_t)__byte_perm(hi, 0, 0x0123);
      uint64_t kk = be >> kg.kshift;
      if (kg.vbits) kk = (kk << kg.vbits) | (uint64_t)(tile_beg + e);
      key[k] = (e < valid) ? kk : ~0ull;
      dgt[k] = (e < valid) ? (uint8_t)((be >> kg.dshift) & 255) : (uint8_t)255;
      if (PAIRS) val[k] = (uint32_t)(tile_beg + e);
    }
    __syncthreads();  // the codes alias the key staging area
  } else {
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) {
      int e = wbase + k * 32;
      key[k] = (e < valid) ? kin[tile_beg + e] : ~0ull;
    }
    if constexpr (PAIRS) {
      if (vin) {
#pragma unroll
        for (int k = 0; k < ITEMS; ++k) { int e = wbase + k * 32; val[k] = (e < valid) ? vin[tile_beg + e] : 0u; }
      } else {  // implicit values 0..n-1 (first pass of round 0)
#pragma unroll
        for (int k = 0; k < ITEMS; ++k) val[k] = (uint32_t)(tile_beg + wbase + k * 32);
      }
    }
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) dgt[k] = (uint8_t)((key[k] >> shift) & 255);
  }
  // early counts: per-warp digit histogram (fire-and-forget shared atomics, no dependent chain)
  uint32_t *wh = warp_hist + warp * RADIX;
#pragma unroll
  for (int k = 0; k < ITEMS; ++k) atomicAdd(&wh[dgt[k]], 1u);
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
  // (lb_mode & 64): the first look-back batch is loaded half way through the ranking, so its L2 round trip
  // overlaps with work instead of being waited for at the end
  constexpr int PF_AT = (ITEMS / 2) / CH * CH;
  uint64_t pre[4];
  const bool pre_on = (lb_mode & 64) && tid < RADIX && tile > 0;
#pragma unroll
  for (int c0 = 0; c0 < ITEMS; c0 += CH) {
    if (c0 == PF_AT && pre_on) look_back_load<4>(status, tile - 1, tid, pre);
    uint32_t peers[CH], base[CH];
    int lead[CH];
#pragma unroll
    for (int j = 0; j < CH; ++j) peers[j] = digit_peers(dgt[c0 + j]);
#pragma unroll
    for (int j = 0; j < CH; ++j) {
      lead[j] = 31 - __clz(peers[j]);
      base[j] = 0;
      if (lane == lead[j]) base[j] = atomicAdd(&wh[dgt[c0 + j]], (uint32_t)__popc(peers[j]));
      // items of one warp with the same digit must take their offsets in item order (stability of the LSD
      // sort): the leader lanes differ from item to item, so the warp reconverges before the next atomic
      __syncwarp();
    }
#pragma unroll
    for (int j = 0; j < CH; ++j) {
      uint32_t pos = __shfl_sync(0xffffffffu, base[j], lead[j]) + __popc(peers[j] & lane_lt);
      stage_k[pos] = key[c0 + j];
      if (PAIRS) stage_v[pos] = val[c0 + j];
      if (KEYGEN) stage_d[pos] = dgt[c0 + j];
    }
  }
  // look back (thread = digit) while the other warps finish staging
  if (tid < RADIX) {
    uint64_t excl = 0;
    if (tile > 0) {
      const int nap = ((lb_mode >> 3) & 7) * 64;  // ns to sleep when a batch made no progress
      long t = tile - 1;
      bool done = false;
      if (pre_on) look_back_consume<4>(pre, excl, t, done);
      switch (lb_mode & 7) {
        case 1: excl = look_back<8>(status, t, tid, nap, excl, done); break;
        case 2: excl = look_back<16>(status, t, tid, nap, excl, done); break;
        case 3: excl = look_back<2>(status, t, tid, nap, excl, done); break;
        case 4: excl = look_back<3>(status, t, tid, nap, excl, done); break;
        case 5: excl = look_back<5>(status, t, tid, nap, excl, done); break;
        case 6: excl = look_back<6>(status, t, tid, nap, excl, done); break;
        default: excl = look_back<4>(status, t, tid, nap, excl, done); break;
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
      const int d = KEYGEN ? (int)stage_d[i] : (int)((kk >> shift) & 255);
      uint32_t o = gbase_adj[d] + (uint32_t)i;
      kout[o] = kk;
      if (PAIRS) vout[o] = stage_v[i];
    }
  }
}

typedef void (*PassFn)(const uint64_t *, const uint32_t *, uint64_t *, uint32_t *, long, int, const uint64_t *,
                       uint64_t *, uint32_t *, int, int, const KeyGen);
// fn[pairs][keygen]
struct PassVariant { const char *name; PassFn fn[2][2]; int threads, items, minb_keys; size_t smem[2][2]; };
#define PASS_VARIANT(T, I, B, BK) PASS_VARIANT_C(T, I, B, BK, 4)
#define PASS_VARIANT_C(T, I, B, BK, CH)                                                                        \
  { #T "x" #I "b" #B "k" #BK "c" #CH, {{radix_pass_kernel<T, I, BK, false, false, CH>, radix_pass_kernel<T, I, BK, false, true, CH>},   \
                       {radix_pass_kernel<T, I, B, true, false, CH>, radix_pass_kernel<T, I, B, true, true, CH>}},      \
    T, I, BK, {{PassCfg<T, I, false, false>::SMEM, PassCfg<T, I, false, true>::SMEM},                   \
               {PassCfg<T, I, true, false>::SMEM, PassCfg<T, I, true, true>::SMEM}} }
// PSC_RADIX_VARIANT / PSC_RADIX_VARIANT_K (keys only) select a variant for tuning runs (tools/tune_radix.py);
// the defaults are named in pass_variant().  B = CTAs per SM the pair kernel is compiled for, BK = the
// keys-only kernel, C = items ranked per step of the ranking loop (1: 1.46 ms, 2: 1.47, 4: 1.51, 6: 1.72,
// 12: 1.93 ms per 2^28 pairs -- more items in flight only cost registers).
static const PassVariant g_variants[] = {
    PASS_VARIANT(512, 12, 2, 2), PASS_VARIANT(512, 12, 2, 3), PASS_VARIANT(384, 12, 3, 3), PASS_VARIANT(384, 12, 3, 4),
    PASS_VARIANT(512, 8, 3, 3), PASS_VARIANT(512, 16, 2, 2), PASS_VARIANT(256, 12, 4, 5), PASS_VARIANT(384, 16, 2, 3),
    PASS_VARIANT(256, 16, 3, 4), PASS_VARIANT(384, 8, 4, 4), PASS_VARIANT(256, 16, 3, 5), PASS_VARIANT(512, 16, 2, 3),
    PASS_VARIANT_C(512, 12, 2, 2, 2), PASS_VARIANT_C(512, 12, 2, 2, 3), PASS_VARIANT_C(512, 12, 2, 2, 6),
    PASS_VARIANT_C(512, 12, 2, 2, 1), PASS_VARIANT_C(512, 16, 2, 2, 2),
};
static const PassVariant &pass_variant(bool pairs) {
  static int idx[2] = {-1, -1};
  static std::mutex mu;
  static bool attr_set[64] = {false};  // the shared-memory opt-in is per DEVICE: every GPU of the process needs it
  std::lock_guard<std::mutex> lk(mu);
  if (idx[0] < 0) {
    // measured best on B200 (2^28 elements per pass): pairs 1.46 ms = 4.40 TB/s, keys only 1.14 ms = 3.75 TB/s
    const char *defaults[2] = {"512x16b2k2c4", "512x12b2k2c1"};
    const char *names[2] = {getenv("PSC_RADIX_VARIANT_K"), getenv("PSC_RADIX_VARIANT")};
    for (int m = 0; m < 2; ++m) {
      idx[m] = 0;
      for (size_t i = 0; i < sizeof(g_variants) / sizeof(g_variants[0]); ++i)
        if (!strcmp(defaults[m], g_variants[i].name)) idx[m] = (int)i;
    }
    for (int m = 0; m < 2; ++m)
      for (size_t i = 0; names[m] && i < sizeof(g_variants) / sizeof(g_variants[0]); ++i)
        if (!strcmp(names[m], g_variants[i].name)) idx[m] = (int)i;
  }
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !attr_set[dev]) {
    for (auto &v : g_variants)
      for (int a = 0; a < 2; ++a)
        for (int b = 0; b < 2; ++b)
          cudaFuncSetAttribute((const void *)v.fn[a][b], cudaFuncAttributeMaxDynamicSharedMemorySize, (int)v.smem[a][b]);
    attr_set[dev] = true;
  }
  return g_variants[idx[pairs ? 1 : 0]];
}

__global__ void iota_kernel(uint32_t *v, long n) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) v[i] = (uint32_t)i;
}

void radix_sort_pairs(psc_context *ctx, uint64_t *keys_a, uint32_t *vals_a, uint64_t *keys_b,
                      uint32_t *vals_b, long n, const int *shifts, int n_shifts,
                      uint64_t **keys_out, uint32_t **vals_out, const unsigned long long *pre_hist_d,
                      bool vals_iota, const KeyGen *kg) {
  const bool pairs = (vals_a != nullptr);  // keys only: the caller keeps the payload in the low key bits
  *keys_out = keys_a;
  if (vals_out) *vals_out = vals_a;
  if (n <= 1 || n_shifts <= 0) {
    if (kg) { set_error("radix_sort_pairs: key generation needs a digit pass"); throw CudaError{PSC_EARG}; }
    if (pairs && vals_iota && n > 0) { iota_kernel<<<(unsigned)div_up(n, 256), 256, 0, ctx->stream>>>(vals_a, n); PSC_LAUNCH_CHECK(); }
    return;
  }
  if (n_shifts > MAX_PASSES) { set_error("radix_sort_pairs: too many digit passes"); throw CudaError{PSC_EARG}; }
  if (kg && !pre_hist_d) { set_error("radix_sort_pairs: key generation needs the digit counts"); throw CudaError{PSC_EARG}; }
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
  if (kg) skip[0] = false;  // the first pass also creates the keys
  DevBuf<uint64_t> gscan(ctx, MAX_PASSES * RADIX);
  h2d(ctx, gscan.p, scans.data(), scans.size());

  const PassVariant &pv = pass_variant(pairs);
  const int tile = pv.threads * pv.items;
  long tiles = div_up(n, tile);
  static int pf_mult = -1, lb_mode = 0;  // prefetch distance in units of resident CTAs (PSC_RADIX_PF, 0 = off)
  if (pf_mult < 0) {
    const char *e = getenv("PSC_RADIX_PF"); pf_mult = e ? atoi(e) : 1;
    e = getenv("PSC_RADIX_LB"); lb_mode = e ? atoi(e) : 64;
  }
  DevBuf<uint64_t> status(ctx, (size_t)tiles * RADIX);
  DevBuf<uint32_t> counter(ctx, 1);
  uint64_t *kin = keys_a, *kout = keys_b;
  uint32_t *vin = vals_a, *vout = vals_b;
  KeyGen nokg;
  memset(&nokg, 0, sizeof(nokg));
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> evs;
  std::vector<bool> ev_gen;
  for (int p = 0; p < n_shifts; ++p) {
    if (skip[p]) continue;
    status.zero();
    counter.zero();
    const bool gen = (kg != nullptr && p == 0);
    // CTAs per SM this launch gets: bounded by its shared memory and by what it was compiled for
    const int resident = std::max(1, std::min((int)((227 * 1024) / (pv.smem[pairs][gen] + 1024)), pairs ? 8 : pv.minb_keys));
    const int pf_dist = pf_mult * ctx->n_sm * resident;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    cudaEventRecord(e0, ctx->stream);
    const bool first_iota = vals_iota;
    pv.fn[pairs][gen]<<<(unsigned)tiles, pv.threads, pv.smem[pairs][gen], ctx->stream>>>(
        kin, vals_iota ? nullptr : vin, kout, vout, n, sl.s[p], gscan.p + p * RADIX, status.p, counter.p, pf_dist,
        lb_mode, gen ? *kg : nokg);
    vals_iota = false;
    cudaEventRecord(e1, ctx->stream);
    evs.push_back({e0, e1});
    ev_gen.push_back(gen);
    PSC_LAUNCH_CHECK();
    ctx->launches++;
    ctx->stats.radix_passes++;
    ctx->stats.radix_elems += n;
    const long pass_bytes = n * (long)((gen ? 1 : 8 + ((pairs && !first_iota) ? 4 : 0)) + 8 + (pairs ? 4 : 0));
    ctx->stats.radix_bytes += pass_bytes;
    if (gen) ctx->stats.radix_keygen_bytes += pass_bytes;
    std::swap(kin, kout);
    std::swap(vin, vout);
  }
  // the host vector `scans` must outlive the async copy
  sync(ctx);
  for (size_t q = 0; q < evs.size(); ++q) {
    float ms = 0;
    if (cudaEventElapsedTime(&ms, evs[q].first, evs[q].second) == cudaSuccess) {
      ctx->stats.ms_radix_pass += ms;
      if (ev_gen[q]) ctx->stats.ms_radix_keygen += ms;
    }
    cudaEventDestroy(evs[q].first); cudaEventDestroy(evs[q].second);
  }
  if (pairs && vals_iota) {  // every pass was skipped (all keys equal): materialise the implicit values
    iota_kernel<<<(unsigned)div_up(n, 256), 256, 0, ctx->stream>>>(vin, n);
    PSC_LAUNCH_CHECK();
  }
  *keys_out = kin;
  if (vals_out) *vals_out = vin;
}

}  // namespace psc
