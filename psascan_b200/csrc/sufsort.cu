// sufsort.cu -- stage 1: block suffix sort by prefix doubling over packed 64-bit keys.
//
// Replaces inmem_psascan<saidx_t,12>() (inmem_psascan_src/inmem_psascan.hpp:64-304) and the
// vendored sorter behind it (initial_partial_sufsort.hpp:272-289).  What is kept from the
// reference is the *renaming* that makes a plain suffix sort of the block equal to whole-text
// order (initial_partial_sufsort.hpp:61-80): the block's last symbol, every symbol greater
// than it, and every occurrence of it that is followed by a suffix greater than suffix(end)
// (the gt_end bit) move up by one.  Everything else is new:
//   round 0   key = the first k renamed symbols packed at b bits each (k*b <= 64), LSD radix
//             sort of (key, position)
//   round r   only suffixes still sharing their key stay "active"; key = (group head << rank bits |
//             ISA[pos + h] + 1), sorted again, h doubles.
#include <algorithm>
#include <cmath>
#include "scan.cuh"
#include "texthash.cuh"

namespace psc {

constexpr int PD_ITEMS = 8;
constexpr int PD_TILE = SC_THREADS * PD_ITEMS;


// ---- symbol statistics ---------------------------------------------------------------------
// hist[c] for c in 0..255, hist[256] = number of positions with text[i] == last that are NOT
// incremented by the renaming (i < m-1 and gt_end[i+1] == 0).
__global__ void __launch_bounds__(256)
    symbol_hist_kernel(const uint8_t *__restrict__ text, long m, const uint32_t *__restrict__ gt_end,
                       int last, unsigned long long *__restrict__ hist) {
  __shared__ uint32_t sh[257];
  for (int i = threadIdx.x; i < 257; i += blockDim.x) sh[i] = 0;
  __syncthreads();
  auto count = [&](int c, long i) {
    atomicAdd(&sh[c], 1u);
    if (c == last && i < m - 1 && gt_end != nullptr && !bit_get(gt_end, i + 1)) atomicAdd(&sh[256], 1u);
  };
  // 16 symbols per load where the address allows it, single bytes at both ends
  const long head = min(m, (long)((16 - ((uintptr_t)text & 15)) & 15));
  const long nvec = (m - head) / 16;
  const long gtid = (long)blockIdx.x * blockDim.x + threadIdx.x, stride = (long)gridDim.x * blockDim.x;
  const uint4 *v = reinterpret_cast<const uint4 *>(text + head);
  for (long q = gtid; q < nvec; q += stride) {
    const uint4 x = __ldg(v + q);
    const uint32_t w[4] = {x.x, x.y, x.z, x.w};
    const long i0 = head + q * 16;
#pragma unroll
    for (int j = 0; j < 16; ++j) count((int)((w[j >> 2] >> (8 * (j & 3))) & 255), i0 + j);
  }
  for (long i = gtid; i < head; i += stride) count(text[i], i);
  for (long i = head + nvec * 16 + gtid; i < m; i += stride) count(text[i], i);
  __syncthreads();
  for (int i = threadIdx.x; i < 257; i += blockDim.x)
    if (sh[i]) atomicAdd(&hist[i], (unsigned long long)sh[i]);
}

struct RenameLut {
  uint16_t code[256];  // dense code (1..sigma') of the renamed symbol; for c == last: the LOW code
  uint16_t code_last_high;
  int last;
  int bits;  // bits per code (1-based codes, 0 = past the block end)
  int k;     // symbols per round-0 key
  int kbits; // bits per symbol in the round-0 key: `bits`, or one less in dense mode
  int dense; // 1: the round-0 key packs code - 1 (sigma is a power of two, see sort_block)
};

// ---- round 0 keys --------------------------------------------------------------------------
// keys[i] = the k renamed symbols at i, packed MSB first; also counts every 8-bit digit of the
// keys (the radix sort's histogram), so the keys are not read again before the first pass.
// The renamed codes of a tile are staged in shared memory once (the 256-entry code table too: indexed
// per thread it must not sit in the constant bank, where 32 different indices serialise).
__device__ __forceinline__ int code_slot(int e) { return e + ((e >> 6) << 1); }

__global__ void __launch_bounds__(SC_THREADS)
    init_keys_kernel(const uint8_t *__restrict__ text, long m, const uint32_t *__restrict__ gt_end,
                     RenameLut lut, int passes, uint64_t *__restrict__ keys,
                     unsigned long long *__restrict__ ghist) {
  extern __shared__ uint16_t codes[];  // code_slot(PD_TILE + k) + 2, then 8 * 256 uint32 histogram
  uint32_t *sh = reinterpret_cast<uint32_t *>(codes + ((code_slot(PD_TILE + lut.k) + 3) & ~1));
  __shared__ uint16_t scode[256];  // the table indexed per thread: shared memory, not the constant bank
  for (int e = threadIdx.x; e < 8 * 256; e += SC_THREADS) sh[e] = 0;
  scode[threadIdx.x] = lut.code[threadIdx.x];
  __syncthreads();
  const long t0 = (long)blockIdx.x * PD_TILE;
  const int span = PD_TILE + lut.k;
  for (int e = threadIdx.x; e < span; e += SC_THREADS) {
    long i = t0 + e;
    uint16_t cd = 0;
    if (i < m) {
      int c = text[i];
      cd = scode[c];
      if (c == lut.last && (i == m - 1 || gt_end == nullptr || bit_get(gt_end, i + 1)))
        cd = lut.code_last_high;
      cd -= (uint16_t)lut.dense;
    }
    codes[code_slot(e)] = cd;
  }
  __syncthreads();
  const int k = lut.k, bits = lut.kbits;
  const uint64_t mask = (k * bits >= 64) ? ~0ull : ((1ull << (k * bits)) - 1ull);
  // two consecutive positions per thread and step (rolling update for the second one): one 16-byte
  // store per lane, a warp writes 512 contiguous bytes
#pragma unroll
  for (int j = 0; j < PD_ITEMS / 2; ++j) {
    const int e0 = 2 * (threadIdx.x + j * SC_THREADS);
    const long i0 = t0 + e0;
    if (i0 >= m) break;
    uint64_t key = 0;
    for (int q = 0; q < k; ++q) key = (key << bits) | codes[code_slot(e0 + q)];
    const uint64_t k0 = key;
    const uint64_t k1 = ((key << bits) | codes[code_slot(e0 + k)]) & mask;
    if (i0 + 1 < m) *reinterpret_cast<ulonglong2 *>(keys + i0) = make_ulonglong2(k0, k1);
    else keys[i0] = k0;
    for (int p = 0; p < passes; ++p) {
      atomicAdd(&sh[p * 256 + (int)((k0 >> (8 * p)) & 255)], 1u);
      if (i0 + 1 < m) atomicAdd(&sh[p * 256 + (int)((k1 >> (8 * p)) & 255)], 1u);
    }
  }
  __syncthreads();
  for (int e = threadIdx.x; e < passes * 256; e += SC_THREADS)
    if (sh[e]) atomicAdd(&ghist[e], (unsigned long long)sh[e]);
}

// ---- grouping after a sort ------------------------------------------------------------------
// K[0..n) sorted keys.  head(i) = key(i) != key(i-1); an element is "active" unless it is alone in
// its group.  ONE independent pass per tile (no scan across tiles): writes SA (and ISA when it exists)
// for the sorted elements and compacts the still-active ones into (keyY = SA slot of the group head,
// valY = text position).  Nothing but K is needed to place an element:
//   round 0 (hshift < 0): element i sits in SA slot i.
//   later rounds: K[i] = (SA slot H of the head of the element's previous group) << hshift | new key
//     bits, and the members of a previous group are contiguous in K, so slot(i) = H + (i - first
//     index with the same H).
// What a tile needs from before its first element (start of the group / previous group that element
// belongs to) is found in K itself: 32 predecessors in one round trip, else a binary search (K is
// sorted).  The active elements of a tile go to a range reserved with one atomicAdd: their order
// across tiles is arbitrary, which the following sort does not care about.
//   vbits == 0: values come from V; sa == nullptr means the sorted value array itself is the SA
//               (round 0), only active elements are touched.
//   vbits  > 0: "packed" round 0, K[i] = key << vbits | text position; every sa[i] is written.
//   key0: key of text position 0 (round 0, vbits == 0): its SA slot goes to *i0_out.
__device__ __forceinline__ long first_index_ge(const uint64_t *__restrict__ K, long hi, int shift, uint64_t want) {
  long lo = 0;  // first idx in [0, hi] with (K[idx] >> shift) >= want, given (K[hi] >> shift) >= want
  while (lo < hi) { long mid = (lo + hi) >> 1; if ((K[mid] >> shift) < want) lo = mid + 1; else hi = mid; }
  return lo;
}

// Every warp works on its own GP_ITEMS * 32 consecutive elements (item k of lane l = element k * 32 + l,
// so loads are coalesced, neighbours come from shuffles, and "the last head at or before me" is a bit
// scan of a ballot) and needs nothing from the other warps: no shared memory, no block barrier.  (The
// first versions -- 8 consecutive elements per thread, block-wide scans -- were bound by the load/store
// unit, 32 sectors per warp load, and by the barrier + atomic round trips of a 2048-element tile.)
constexpr int GP_THREADS = 256, GP_ITEMS = 8, GP_WARPS = GP_THREADS / 32;
constexpr int GP_WCHUNK = GP_ITEMS * 32, GP_TILE = GP_THREADS * GP_ITEMS;

// index+1 of the first element of the run of equal (K & mask) that ends at wbeg - 1, given that
// K[wbeg - 1] & mask == want; all lanes of the warp must call
__device__ __forceinline__ uint32_t run_start_before(const uint64_t *__restrict__ K, uint32_t wbeg, int shift,
                                                     uint64_t mask, uint64_t want, int lane) {
  const long idx = (long)wbeg - 1 - lane;
  const uint64_t kv = idx >= 0 ? K[idx] : 0;
  const uint32_t eq = __ballot_sync(0xffffffffu, idx >= 0 && (kv & mask) == want);
  const int nn = (eq == 0xffffffffu) ? 32 : (__ffs(~eq) - 1);
  uint32_t start = wbeg - nn;
  if (nn == 32) {
    if (lane == 0) start = (uint32_t)first_index_ge(K, (long)wbeg - 32, shift, want >> shift);
    start = __shfl_sync(0xffffffffu, start, 0);
  }
  return start + 1;
}

// MODE 0: round 0, (key, position) pairs; 1: round 0, packed keys; 2: later rounds.  Indices fit 32 bits
// (blocks are < 2^31 symbols); keys are compared under a mask instead of being shifted.
template <int MODE>
__global__ void __launch_bounds__(GP_THREADS)
    group_pass_kernel(const uint64_t *__restrict__ K, const uint32_t *__restrict__ V, uint32_t n, int vbits, int hshift,
                      uint32_t *__restrict__ sa, uint32_t *__restrict__ isa,
                      uint64_t *__restrict__ keyY, uint32_t *__restrict__ valY,
                      const uint64_t *__restrict__ key0, uint32_t *__restrict__ i0_out,
                      unsigned long long *__restrict__ total) {
  constexpr bool later = (MODE == 2), packed = (MODE == 1);
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const uint32_t wbeg = (blockIdx.x * GP_WARPS + warp) * GP_WCHUNK;
  if (wbeg >= n) return;
  const uint64_t vmask = packed ? ((1ull << vbits) - 1ull) : 0ull;
  const uint64_t kmask = ~vmask;                                   // the bits that make a group
  const uint64_t omask = later ? (~0ull << hshift) : 0ull;         // the bits that name the previous group
  uint64_t kw[GP_ITEMS];
#pragma unroll
  for (int k = 0; k < GP_ITEMS; ++k) {
    const uint32_t i = wbeg + k * 32 + lane;
    kw[k] = i < n ? K[i] : 0;
  }
  const uint64_t before = wbeg > 0 ? K[wbeg - 1] : 0;
  const uint64_t after = (wbeg + GP_WCHUNK < n) ? K[wbeg + GP_WCHUNK] : 0;
  const uint64_t k0 = (MODE == 0 && key0) ? *key0 : 0;
  // flags as ballots (warp-uniform): group heads, active elements, heads of the previous groups
  uint32_t hb[GP_ITEMS], ab[GP_ITEMS], ob[later ? GP_ITEMS : 1];
  uint32_t w_act = 0;
#pragma unroll
  for (int k = 0; k < GP_ITEMS; ++k) {
    const uint32_t i = wbeg + k * 32 + lane;
    // neighbours: shuffles inside the item; lanes 0 and 31 re-read theirs (an L1 hit: the warp has just
    // loaded those lines), which is cheaper than two more 64-bit shuffles per item
    uint64_t prevw = __shfl_up_sync(0xffffffffu, kw[k], 1);
    uint64_t nextw = __shfl_down_sync(0xffffffffu, kw[k], 1);
    if (lane == 0) prevw = (k == 0) ? before : ((i - 1 < n) ? K[i - 1] : 0);
    if (lane == 31) nextw = (k == GP_ITEMS - 1) ? after : ((i + 1 < n) ? K[i + 1] : 0);
    const bool valid = i < n;
    bool dpk, dnk;  // differs from the previous / next element in the group bits
    if (MODE == 0) { dpk = kw[k] != prevw; dnk = kw[k] != nextw; }
    else { dpk = ((kw[k] ^ prevw) & kmask) != 0; dnk = ((kw[k] ^ nextw) & kmask) != 0; }
    const uint64_t dp = kw[k] ^ prevw;
    const bool head = valid && ((i == 0) || dpk);
    const bool head_next = (i + 1 >= n) || dnk;
    const bool act = valid && !(head && head_next);
    hb[k] = __ballot_sync(0xffffffffu, head);
    ab[k] = __ballot_sync(0xffffffffu, act);
    if (later) ob[k] = __ballot_sync(0xffffffffu, valid && ((i == 0) || (dp & omask) != 0));
    w_act += __popc(ab[k]);
  }
  // what the warp's first element needs from before it: start of its group / of its previous group
  const uint64_t first = __shfl_sync(0xffffffffu, kw[0], 0);
  uint32_t hd = 0, os = 1;
  if (!(hb[0] & 1u)) hd = run_start_before(K, wbeg, packed ? vbits : 0, kmask, first & kmask, lane);
  if (later && !(ob[0] & 1u)) os = run_start_before(K, wbeg, hshift, omask, first & omask, lane);
  // room for the warp's active elements (their order across warps is arbitrary)
  unsigned long long act_off = 0;
  if (w_act) {
    if (lane == 0) act_off = atomicAdd(total, (unsigned long long)w_act);
    act_off = __shfl_sync(0xffffffffu, act_off, 0);
  }
  const bool touch_all = (MODE != 0);
  const uint32_t lane_le = 0xffffffffu >> (31 - lane), lane_lt = lane_le >> 1;
#pragma unroll
  for (int k = 0; k < GP_ITEMS; ++k) {
    const uint32_t ibase = wbeg + k * 32;
    const uint32_t i = ibase + lane;
    // last (previous-)group head at or before this element
    if (MODE == 0 && ab[k] == 0 && !__any_sync(0xffffffffu, key0 && kw[k] == k0)) {
      if (hb[k]) hd = ibase + (31 - __clz(hb[k])) + 1;  // nothing to write for these 32 elements
      continue;
    }
    const uint32_t hm = hb[k] & lane_le;
    const uint32_t my_hd = hm ? ibase + (31 - __clz(hm)) + 1 : hd;
    if (hb[k]) hd = ibase + (31 - __clz(hb[k])) + 1;
    uint32_t my_os = 1;
    if (later) {
      const uint32_t om = ob[k] & lane_le;
      my_os = om ? ibase + (31 - __clz(om)) + 1 : os;
      if (ob[k]) os = ibase + (31 - __clz(ob[k])) + 1;
    }
    const bool act = (ab[k] >> lane) & 1u;
    const unsigned long long my_off = act_off + __popc(ab[k] & lane_lt);
    act_off += __popc(ab[k]);
    if (i >= n) continue;
    const bool is_k0 = (MODE == 0) && key0 && kw[k] == k0;
    if (!act && !touch_all && !is_k0) continue;  // round 0: singletons stay where the sort put them
    const uint32_t v = packed ? (uint32_t)(kw[k] & vmask) : V[i];
    // SA slots of this element and of its group's head: slot(j) = sbase + j inside the previous group
    const uint32_t sbase = later ? (uint32_t)(kw[k] >> hshift) - (my_os - 1) : 0u;
    const uint32_t slot = sbase + i, hp = sbase + my_hd - 1;
    if (v == 0 && i0_out) *i0_out = slot;
    if (!act && !touch_all) continue;
    if (MODE != 0) sa[slot] = v;
    // ISA[v] = SA slot of v's group head.  It already holds the head of the previous group (H, in the
    // key): only members of groups that split off a new head need the random 4-byte write (on
    // repetitive texts most groups survive the early rounds unsplit)
    if (isa && (!later || hp != (uint32_t)(kw[k] >> hshift))) isa[v] = hp;
    if (act) {
      keyY[my_off] = (uint64_t)hp;
      valY[my_off] = v;
    }
  }
}

// next key from the ISA: (group head << rbits) | ISA[position + h] + 1, 0 when past the block end
__global__ void __launch_bounds__(256)
    second_key_kernel(uint64_t *__restrict__ keyY, const uint32_t *__restrict__ valY, long a,
                      const uint32_t *__restrict__ isa, long m, long h, int rbits) {
  long j = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= a) return;
  long p = (long)valY[j] + h;
  uint32_t r = p < m ? isa[p] + 1u : 0u;  // <= m: fits rbits = bits_for(m)
  keyY[j] = (keyY[j] << rbits) | (uint64_t)r;
}

__device__ __forceinline__ uint32_t renamed_code(const uint8_t *__restrict__ text, long m,
                                                 const uint32_t *__restrict__ gt_end, const uint16_t *code,
                                                 int last, uint32_t code_last_high, long i) {
  if (i >= m) return 0;
  int c = text[i];
  if (c == last && (i == m - 1 || gt_end == nullptr || bit_get(gt_end, i + 1))) return code_last_high;
  return code[c];
}

// next key straight from the text ("text round"): (group head << k1*bits) | the k1 renamed symbols
// that follow the h already compared.  Needs no ISA, so texts whose ties die out within
// h + k1 symbols never pay for the m-element ISA scatter.
__global__ void __launch_bounds__(256)
    second_key_text_kernel(uint64_t *__restrict__ keyY, const uint32_t *__restrict__ valY, long a,
                           const uint8_t *__restrict__ text, long m, const uint32_t *__restrict__ gt_end,
                           RenameLut lut, long h, int k1) {
  __shared__ uint16_t scode[256];
  scode[threadIdx.x] = lut.code[threadIdx.x];
  __syncthreads();
  long j = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= a) return;
  const long v = (long)valY[j], p = v + h;
  uint64_t key = keyY[j];
  if (p >= m) {
    // a suffix that ended inside the prefix compared so far: in dense mode its padding looked like the
    // smallest symbol, so several of them (and longer suffixes) can share a group.  It sorts before every
    // suffix that goes on (their first code below is >= 1), the shorter one first.
    key = (key << (k1 * lut.bits)) | (uint64_t)(m - 1 - v);
  } else {
    for (int s = 0; s < k1; ++s)
      key = (key << lut.bits) | renamed_code(text, m, gt_end, scode, lut.last, lut.code_last_high, p + s);
  }
  keyY[j] = key;
}

// isa[sa[i]] = i for the whole block (only when doubling rounds are needed), then the group heads
// of the active elements on top
__global__ void __launch_bounds__(256)
    isa_scatter_kernel(const uint32_t *__restrict__ sa, long m, uint32_t *__restrict__ isa) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < m) isa[sa[i]] = (uint32_t)i;
}
// The same scatter for blocks whose ISA does not fit in L2 (measured: the direct scatter moves a
// 32-byte sector in and out of HBM per element, 5.1 ms for 2^27 elements): the pairs (sa[i], i) are
// first grouped by ISA range of 2^22 entries (16 MB), then scattered range after range with the
// target lines resident in L2.  The SA is a permutation, so range b receives exactly 2^22 pairs and
// its slots in the pair array are known without a counting pass.
constexpr int IP_THREADS = 512;
constexpr int IP_ITEMS = 8;
constexpr int IP_TILE = IP_THREADS * IP_ITEMS;  // 4096 pairs, 32 KB of staging
constexpr int IP_SHIFT = 22;
constexpr int IP_MAX_BUCKETS = 512;  // m < 2^31

__global__ void isa_cursor_init_kernel(unsigned long long *cursor, int nb) {
  int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b < nb) cursor[b] = (unsigned long long)b << IP_SHIFT;
}

__global__ void __launch_bounds__(IP_THREADS)
    isa_partition_kernel(const uint32_t *__restrict__ sa, long m, int nb, unsigned long long *__restrict__ cursor,
                         uint64_t *__restrict__ pairs) {
  __shared__ uint32_t cnt[IP_MAX_BUCKETS];   // count, then running slot inside the tile
  __shared__ uint32_t toff[IP_MAX_BUCKETS];  // first slot of the range inside the tile
  __shared__ unsigned long long gofs[IP_MAX_BUCKETS];
  __shared__ uint64_t stage[IP_TILE];
  __shared__ uint32_t wsum[IP_THREADS / 32];
  static_assert(IP_MAX_BUCKETS <= IP_THREADS, "one range per thread in the scan");
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long base = (long)blockIdx.x * IP_TILE;
  const int valid = (int)min((long)IP_TILE, m - base);
  for (int b = tid; b < nb; b += IP_THREADS) cnt[b] = 0;
  __syncthreads();
  uint32_t v[IP_ITEMS];
#pragma unroll
  for (int k = 0; k < IP_ITEMS; ++k) {
    int e = tid + k * IP_THREADS;
    v[k] = e < valid ? sa[base + e] : 0u;
    if (e < valid) atomicAdd(&cnt[v[k] >> IP_SHIFT], 1u);
  }
  __syncthreads();
  {
    uint32_t c = tid < nb ? cnt[tid] : 0u, incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    uint32_t woff = 0;
    for (int w = 0; w < warp; ++w) woff += wsum[w];
    uint32_t ex = woff + incl - c;
    if (tid < nb) {
      toff[tid] = ex; cnt[tid] = ex;
      if (c) gofs[tid] = atomicAdd(&cursor[tid], (unsigned long long)c);
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < IP_ITEMS; ++k) {
    int e = tid + k * IP_THREADS;
    if (e < valid) stage[atomicAdd(&cnt[v[k] >> IP_SHIFT], 1u)] = ((uint64_t)v[k] << 32) | (uint32_t)(base + e);
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < IP_ITEMS; ++k) {
    int e = tid + k * IP_THREADS;
    if (e < valid) {
      uint64_t x = stage[e];
      int b = (int)(x >> (32 + IP_SHIFT));
      pairs[gofs[b] + (unsigned long long)(e - toff[b])] = x;
    }
  }
}

__global__ void __launch_bounds__(256)
    isa_from_pairs_kernel(const uint64_t *__restrict__ pairs, long m, uint32_t *__restrict__ isa) {
  long i = ((long)blockIdx.x * blockDim.x + threadIdx.x) * 2;
  if (i + 2 <= m) {
    ulonglong2 x = *reinterpret_cast<const ulonglong2 *>(pairs + i);
    isa[x.x >> 32] = (uint32_t)x.x;
    isa[x.y >> 32] = (uint32_t)x.y;
  } else if (i < m) {
    uint64_t x = pairs[i];
    isa[x >> 32] = (uint32_t)x;
  }
}

__global__ void __launch_bounds__(256)
    isa_fix_kernel(const uint64_t *__restrict__ keyY, const uint32_t *__restrict__ valY, long a,
                   uint32_t *__restrict__ isa) {
  long j = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (j < a) isa[valY[j]] = (uint32_t)keyY[j];
}

static int bits_for(uint64_t v) { int b = 0; while (v) { ++b; v >>= 1; } return b; }

// ---- host driver ------------------------------------------------------------------------------
void sort_block(psc_context *ctx, const uint8_t *text_d, long m, const uint32_t *gt_end_bits_d,
                BlockSortResult *res) {
  PhaseTimer timer(ctx, &ctx->stats.ms_sort);
  res->isa.release();
  if (m == 0) { res->sa.alloc(ctx, 0); return; }
  uint8_t last = 0;
  PSC_CUDA(cudaMemcpyAsync(&last, text_d + m - 1, 1, cudaMemcpyDeviceToHost, ctx->stream));
  sync(ctx);

  // symbol statistics -> dense codes of the renamed alphabet
  DevBuf<unsigned long long> hist_d(ctx, 257);
  hist_d.zero();
  int hb = (int)std::min<long>(div_up(m, 256 * 16), (long)ctx->n_sm * 8);
  { KTrace kt(ctx, "symbol_hist"); symbol_hist_kernel<<<hb, 256, 0, ctx->stream>>>(text_d, m, gt_end_bits_d, last, hist_d.p); }
  PSC_LAUNCH_CHECK();
  ctx->launches++;
  unsigned long long hist[257];
  d2h(ctx, hist, hist_d.p, 257);
  sync(ctx);
  if (hist[255] != 0) { set_error("byte 255 in the text"); throw CudaError{PSC_ESENTINEL}; }
  RenameLut lut;
  memset(&lut, 0, sizeof(lut));
  lut.last = last;
  double p2 = 0;
  int sigma = 0;
  for (int c = 0; c < 256; ++c) {
    if (c == last) {
      unsigned long long low = hist[256], high = hist[c] - low;
      if (low) { lut.code[c] = (uint16_t)++sigma; p2 += ((double)low / m) * ((double)low / m); }
      lut.code_last_high = (uint16_t)++sigma;  // position m-1 always takes it
      p2 += ((double)high / m) * ((double)high / m);
    } else if (hist[c]) {
      lut.code[c] = (uint16_t)++sigma;
      p2 += ((double)hist[c] / m) * ((double)hist[c] / m);
    }
  }
  lut.bits = bits_for((uint64_t)sigma);
  int max_text_rounds = 6;
  if (const char *e = getenv("PSC_TEXT_ROUNDS")) max_text_rounds = atoi(e);
  // dense mode: when the renamed alphabet has exactly 2^b symbols the round-0 key packs code - 1 at b bits
  // instead of code at b + 1 (200 MB of DNA in one block: 16 symbols in 32 bits, 4 digit passes instead of 6).
  // The padding of suffixes shorter than k then equals the smallest symbol: such a suffix may share a group
  // with longer ones, and the text round (always the next round) puts it first (second_key_text_kernel).
  const int k1_text = (64 - bits_for((uint64_t)(m > 1 ? m - 1 : 1))) / lut.bits;  // symbols a text round adds
  lut.dense = (sigma >= 2 && (sigma & (sigma - 1)) == 0 && lut.bits <= 7 && max_text_rounds >= 1 &&
               k1_text >= 2 && (k1_text - 1) * lut.bits >= 7 && !getenv("PSC_NO_DENSE")) ? 1 : 0;
  // the same for exactly 256 renamed symbols (255 byte values plus the split of the block's last symbol, i.e.
  // every block with a tail of a text that uses all byte values): 8-bit codes keep the key-building pass
  const bool dense8 = (sigma == 256 && m >= 64 && max_text_rounds >= 1 && k1_text >= 2 && !getenv("PSC_NO_DENSE"));
  if (dense8) lut.dense = 1;
  lut.kbits = lut.bits - lut.dense;
  int kmax = 64 / lut.kbits;
  // smallest k with m * p2^k <= 1/16 (expected share of suffixes still tied after round 0)
  int k = 1;
  {
    double v = (double)m * 16.0;
    while (k < kmax && v * p2 > 1.0) { v *= p2; ++k; }
    if (v * p2 > 1.0) k = kmax;
  }
  int passes = (k * lut.kbits + 7) / 8;
  k = std::min(kmax, (passes * 8) / lut.kbits);  // fill the last digit
  if (k > m) k = (int)std::max<long>(1, m);
  passes = (k * lut.kbits + 7) / 8;
  lut.k = k;

  DevBuf<uint64_t> keyA(ctx, m), keyB(ctx, m);
  DevBuf<uint32_t> valA(ctx, m), valB(ctx, m);
  long tiles = div_up(m, PD_TILE);
  DevBuf<unsigned long long> ghist(ctx, 8 * 256);
  DevBuf<uint64_t> key_of_zero(ctx, 1);
  DevBuf<uint32_t> i0_d(ctx, 1);
  int shifts[8];
  uint64_t *K; uint32_t *V;
  int vbits = 0;  // > 0: packed round 0, sorted keys = key << vbits | text position, no value arrays
  // 8-bit codes: digit p of key(i) IS the code at i + k-1-p, so every digit histogram is the symbol
  // histogram shifted by a few positions (computed here instead of `passes` shared atomics per suffix
  // in a kernel), and the first digit pass builds its keys straight from the text (radix.cu, KEYGEN):
  // no key array is written and read back before the sort.
  const bool codes8 = (lut.kbits == 8 && m >= 64);
  const int dn = codes8 ? lut.dense : 0;  // round-0 code = code - dn
  if (codes8) {
    uint8_t head[8];
    uint32_t gw = 0xffffffffu;
    d2h(ctx, head, text_d, 8);
    if (gt_end_bits_d) d2h(ctx, &gw, gt_end_bits_d, 1);
    sync(ctx);
    unsigned long long ch[256] = {0};
    for (int c = 0; c < 256; ++c) {
      if (c == last) {
        if (hist[256]) ch[lut.code[c] - dn] += hist[256];
        ch[lut.code_last_high - dn] += hist[c] - hist[256];
      } else if (hist[c]) ch[lut.code[c] - dn] += hist[c];
    }
    int hc[8];
    for (int j = 0; j < 8; ++j) {
      int c = head[j];
      hc[j] = lut.code[c];
      if (c == last && (gt_end_bits_d == nullptr || ((gw >> (j + 1)) & 1u))) hc[j] = lut.code_last_high;
      hc[j] -= dn;
    }
    // packed mode: the position rides in the low bits of the key (8 B per element and pass instead of
    // 12) when all k symbols still fit above it.  (With fewer key bits -- the last symbol kept only in
    // part, which the code supports -- the digit passes of 2^30 random bytes take 29 ms instead of 34,
    // but 6 % of the suffixes stay tied instead of 0.1 % and the next round costs more than that.)
    const int pbits = bits_for((uint64_t)(m - 1));
    int kbits = std::min(64 - pbits, 8 * k);
    // PSC_PACKED=0/1 forces the mode, PSC_PACKED_MAXBITS caps the key bits (tests: partly kept symbols on small texts)
    const char *pe = getenv("PSC_PACKED"), *pm = getenv("PSC_PACKED_MAXBITS");
    const int packed_env = pe ? atoi(pe) : 2;
    if (pm && packed_env != 0) kbits = std::max(1, std::min(kbits, atoi(pm)));
    bool packed = packed_env == 1 || (packed_env == 2 && kbits == 8 * k);
    if (kbits < 8) packed = false;
    if (!packed) kbits = 8 * k;
    const int S = (kbits + 7) / 8;  // symbols looked at = digit passes
    std::vector<unsigned long long> gh(8 * 256, 0);
    for (int p = 0; p < S; ++p) {
      const int off = S - 1 - p;  // digit p = symbol `off`: positions [off, m + off); codes past the end are 0
      for (int d = 0; d < 256; ++d) gh[p * 256 + d] = ch[d];
      for (int j = 0; j < off; ++j) gh[p * 256 + hc[j]]--;
      gh[p * 256 + 0] += (unsigned long long)off;
      shifts[p] = (packed ? pbits : 0) + kbits - 8 * (S - p);  // pass 0 of a partly kept symbol: see KeyGen
    }
    if (shifts[0] < 0) shifts[0] = 0;  // not used by the key-generating pass
    h2d(ctx, ghist.p, gh.data(), gh.size());
    KeyGen kg;
    memset(&kg, 0, sizeof(kg));
    kg.text = text_d; kg.gt_end = gt_end_bits_d; kg.m = m;
    kg.last = last; kg.code_last_high = lut.code_last_high - dn;
    kg.dshift = 64 - 8 * S; kg.kshift = 64 - kbits; kg.vbits = packed ? pbits : 0;
    for (int c = 0; c < 256; ++c) kg.code[c] = (uint8_t)(lut.code[c] - dn);
    uint64_t key0 = 0;
    for (int j = 0; j < 8; ++j) key0 = (key0 << 8) | (uint64_t)hc[j];
    key0 >>= kg.kshift;
    h2d(ctx, key_of_zero.p, &key0, 1);
    if (packed) {
      vbits = pbits;
      radix_sort_pairs(ctx, keyA.p, nullptr, keyB.p, nullptr, m, shifts, S, &K, nullptr, ghist.p, false, &kg);
      V = nullptr;
    } else {
      radix_sort_pairs(ctx, keyA.p, valA.p, keyB.p, valB.p, m, shifts, S, &K, &V, ghist.p, true, &kg);
    }
    k = kbits / 8;  // whole symbols compared so far
    sync(ctx);      // kg, key0 and gh live on this stack frame
  } else {
    ghist.zero();
    { KTrace kt(ctx, "init_keys"); init_keys_kernel<<<(unsigned)tiles, SC_THREADS, (size_t)(PD_TILE + k + (PD_TILE + k) / 32 + 8) * sizeof(uint16_t) + 8 * 256 * 4, ctx->stream>>>(
        text_d, m, gt_end_bits_d, lut, passes, keyA.p, ghist.p); }
    PSC_LAUNCH_CHECK();
    ctx->launches++;
    PSC_CUDA(cudaMemcpyAsync(key_of_zero.p, keyA.p, 8, cudaMemcpyDeviceToDevice, ctx->stream));
    for (int p = 0; p < passes; ++p) shifts[p] = 8 * p;
    radix_sort_pairs(ctx, keyA.p, valA.p, keyB.p, valB.p, m, shifts, passes, &K, &V, ghist.p, true);
  }
  // pairs: the sorted position array IS the suffix array from here on; later rounds patch the slots of
  // suffixes that were still tied.  The other value buffer and both key buffers become scratch.
  // packed: the positions are still inside the keys; the first grouping pass writes them to valA.
  const bool v_is_a = (vbits > 0) || (V == valA.p);
  res->sa = std::move(v_is_a ? valA : valB);
  DevBuf<uint32_t> &vfree = v_is_a ? valB : valA;
  uint64_t *kbuf[2] = {(K == keyA.p) ? keyB.p : keyA.p, K};  // compaction goes to kbuf[0] first
  DevBuf<uint32_t> valC;                                     // second value buffer, sized once `a` is known
  uint32_t *vbuf[2] = {vfree.p, nullptr};
  int out = 0;  // which (kbuf, vbuf) pair the next compaction writes
  long a = m, h = k;
  const int hpbits = bits_for((uint64_t)(m > 1 ? m - 1 : 1)), lowb = bits_for((uint64_t)m);
  int text_rounds = 0;
  DevBuf<unsigned long long> gp_total(ctx, 1);
  int hshift = -1;  // round 0: SA slot = index
  for (int round = 0;; ++round) {
    long t = div_up(a, GP_TILE);
    gp_total.zero();
    {
      KTrace kt(ctx, "group_pass");
      auto fn = round > 0 ? group_pass_kernel<2> : (vbits > 0 ? group_pass_kernel<1> : group_pass_kernel<0>);
      fn<<<(unsigned)t, GP_THREADS, 0, ctx->stream>>>(K, V, (uint32_t)a, vbits, hshift, res->sa.p, res->isa.p, kbuf[out],
                                                      vbuf[out], key_of_zero.p, i0_d.p, gp_total.p);
    }
    PSC_LAUNCH_CHECK();
    ctx->launches++;
    unsigned long long na = 0;
    d2h(ctx, &na, gp_total.p, 1);
    sync(ctx);
    if (round == 0 && na > 0) {
      valC.alloc(ctx, na);
      vbuf[1] = valC.p;
    }
    ctx->stats.doubling_rounds++;
    if (na == 0) break;
    a = (long)na;
    uint64_t *KY = kbuf[out];
    uint32_t *VY = vbuf[out];
    int nsh = 0;
    const int k1 = (64 - hpbits) / lut.bits;
    // text round: extend the compared prefix by k1 symbols read from the text (no ISA yet).  Always the first
    // choice after round 0; repeated while only a few suffixes (< m/32) are still tied, because a text round
    // costs O(a) and the first doubling round costs the m-element ISA (1.4 ms per 2^27 on top of its own sort)
    if (k1 >= 1 && res->isa.p == nullptr && text_rounds < max_text_rounds && (text_rounds == 0 || a * 32 <= m)) {
      ++text_rounds;
      { KTrace kt(ctx, "second_key_text"); second_key_text_kernel<<<(unsigned)div_up(a, 256), 256, 0, ctx->stream>>>(KY, VY, a, text_d, m, gt_end_bits_d,
                                                                                lut, h, k1); }
      PSC_LAUNCH_CHECK();
      for (int b = 0; b < k1 * lut.bits + hpbits; b += 8) shifts[nsh++] = b;
      hshift = k1 * lut.bits;
      h += k1;
    } else {
      if (res->isa.p == nullptr) {
        // first doubling round: now the full inverse suffix array is needed
        res->isa.alloc(ctx, m);
        if (m >= (1L << 24)) {  // ISA larger than 64 MB: group by 16 MB range first (scratch: the free key buffer)
          KTrace kt(ctx, "isa_scatter");
          const int nb = (int)(((m - 1) >> IP_SHIFT) + 1);
          uint64_t *pairs = kbuf[1 - out];
          DevBuf<unsigned long long> cursor(ctx, IP_MAX_BUCKETS);
          isa_cursor_init_kernel<<<div_up(nb, 256), 256, 0, ctx->stream>>>(cursor.p, nb);
          isa_partition_kernel<<<(unsigned)div_up(m, IP_TILE), IP_THREADS, 0, ctx->stream>>>(res->sa.p, m, nb, cursor.p, pairs);
          PSC_LAUNCH_CHECK();
          isa_from_pairs_kernel<<<(unsigned)div_up(div_up(m, 2), 256), 256, 0, ctx->stream>>>(pairs, m, res->isa.p);
          ctx->launches += 2;
        } else {
          KTrace kt(ctx, "isa_scatter");
          isa_scatter_kernel<<<(unsigned)div_up(m, 256), 256, 0, ctx->stream>>>(res->sa.p, m, res->isa.p);
        }
        PSC_LAUNCH_CHECK();
        { KTrace kt(ctx, "isa_fix"); isa_fix_kernel<<<(unsigned)div_up(a, 256), 256, 0, ctx->stream>>>(KY, VY, a, res->isa.p); }
        PSC_LAUNCH_CHECK();
        ctx->launches += 2;
      }
      { KTrace kt(ctx, "second_key"); second_key_kernel<<<(unsigned)div_up(a, 256), 256, 0, ctx->stream>>>(KY, VY, a, res->isa.p, m, h, lowb); }
      PSC_LAUNCH_CHECK();
      for (int b = 0; b < lowb + hpbits; b += 8) shifts[nsh++] = b;  // head packed right above the rank: 7 passes up to 2^27
      hshift = lowb;
      h *= 2;
    }
    ctx->launches++;
    radix_sort_pairs(ctx, KY, VY, kbuf[1 - out], vbuf[1 - out], a, shifts, nsh, &K, &V);
    out = (K == kbuf[0]) ? 1 : 0;  // next compaction must not overwrite the sorted arrays
  }
  uint32_t i0 = 0;
  d2h(ctx, &i0, i0_d.p, 1);
  sync(ctx);
  res->i0 = i0;
}

// ---- K0: gt_end by matching the block against the text after it -------------------------------
// compute_initial_gt_bitvectors.hpp:61-148 (semantics).  Every position compares at most `cap` symbols
// directly; a match that outlasts them is measured by fingerprints (texthash.cuh) in O(log) steps, so
// the kernel is O(m * cap) in the worst case where the reference's ms-decomposition is O(m).
//   cap <= 0: no fingerprints possible (block and tail not contiguous in one buffer): symbol by symbol
//             under the work guard, as in round 1.
struct GtEndArgs {
  const uint8_t *block; long m;
  const uint8_t *after; long after_len; int after_is_text_end;
  const uint32_t *gt_after; long gt_after_off;
  long off_block, off_after, cap;
};

__device__ __forceinline__ bool gt_end_general(const GtEndArgs &a, const HashView &hv, WorkBudget *wb, long p) {
  // gt_after == nullptr: the match simply continues past the block end into `after` (the block and
  // the text after it are contiguous), so no gt bits of the next block are needed.
  const long lim = a.gt_after ? min(a.m - p, a.after_len) : a.after_len;
  long el;
  if (a.cap <= 0) {
    el = common_prefix_guarded(a.block + p, a.after, lim, wb);
  } else {
    el = common_prefix(a.block + p, a.after, min(lim, a.cap));
    if (el == a.cap && a.cap < lim) {
      if (hv.pre) el = th_lcp(hv, a.off_block + p, a.off_after, a.cap, lim);
      else { wb->need_hash = 1; el = -1; }  // the host builds the fingerprints and runs the kernel again
    }
  }
  if (el < 0) return false;  // result discarded by the host
  if (el == a.after_len && a.after_is_text_end) return true;
  if (a.gt_after && p + el == a.m) return !bit_get(a.gt_after, a.gt_after_off + el - 1);
  return a.block[p + el] > a.after[el];
}

// One thread decides 8 consecutive positions (one byte of the bit array).  Almost every position differs from
// the tail within its first four symbols: those are decided from a 12-byte register window against the first
// word of the tail, without touching memory again; the rest go through gt_end_general.
__global__ void __launch_bounds__(256)
    gt_end_kernel(const __grid_constant__ GtEndArgs a, uint32_t *__restrict__ out_bits, WorkBudget *wb,
                  const __grid_constant__ HashView hv) {
  const long p8 = ((long)blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (p8 >= a.m) return;
  const bool fast = (p8 + 12 <= a.m) && a.after_len >= 4;
  uint32_t w0 = 0, w1 = 0, w2 = 0, a4 = 0;
  if (fast) {
    w0 = load_u32_any(a.block + p8); w1 = load_u32_any(a.block + p8 + 4); w2 = load_u32_any(a.block + p8 + 8);
    a4 = load_u32_any(a.after);
  }
  uint32_t bits = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const long p = p8 + i;
    if (p >= a.m) break;
    bool g, done = false;
    if (fast) {
      const uint32_t w = i < 4 ? __funnelshift_r(w0, w1, 8 * i) : __funnelshift_r(w1, w2, 8 * (i - 4));
      const uint32_t x = w ^ a4;
      if (x) {
        const int sh = (__ffs(x) - 1) & ~7;
        g = ((w >> sh) & 255u) > ((a4 >> sh) & 255u);
        done = true;
      }
    }
    if (!done) g = gt_end_general(a, hv, wb, p);
    bits |= (g ? 1u : 0u) << i;
  }
  reinterpret_cast<uint8_t *>(out_bits)[p8 >> 3] = (uint8_t)bits;
}

void compute_gt_end(psc_context *ctx, const uint8_t *block_d, long m, const uint8_t *after_d,
                    long after_len, int after_is_text_end, const uint32_t *gt_after_bits_d,
                    long gt_after_off, uint32_t *gt_end_bits_d) {
  PhaseTimer timer(ctx, &ctx->stats.ms_gt_end);
  if (m == 0) return;
  const bool contiguous = (after_d == block_d + m) && after_len > 0;
  long cap = (contiguous && hash_domain_ensure(ctx, block_d, after_d + after_len)) ? lcp_cap() : 0;
  HashView hv;  // empty until the tables exist
  if (cap > 0) hash_view(ctx, block_d, after_d + after_len, false, &hv);
  for (int attempt = 0;; ++attempt) {
    WorkBudget *wb = work_budget_arm(ctx);
    GtEndArgs ga{block_d, m, after_d, after_len, after_is_text_end, gt_after_bits_d, gt_after_off,
                 hv.pre ? (long)(block_d - hv.text) : 0, hv.pre ? (long)(after_d - hv.text) : 0, cap};
    { KTrace kt(ctx, "gt_end");
      gt_end_kernel<<<(unsigned)div_up(div_up(m, 8), 256), 256, 0, ctx->stream>>>(ga, gt_end_bits_d, wb, hv); }
    PSC_LAUNCH_CHECK();
    ctx->launches++;
    if (!work_budget_check(ctx, wb, "gt_end (K0)")) return;
    // some match outlasted the cap: build the fingerprints (once per construction) and decide everything again
    if (attempt > 0 || !hash_view(ctx, block_d, after_d + after_len, true, &hv)) {
      set_error("gt_end (K0): fingerprints unavailable");
      throw CudaError{PSC_ECUDA};
    }
  }
}

// ---- K3: BWT and gt_begin ------------------------------------------------------------------------
// inmem_bwt_from_sa.hpp:48-56 and change_gt_reference_point.hpp:62-118
__global__ void __launch_bounds__(256)
    bwt_kernel(const uint8_t *__restrict__ text, const uint32_t *__restrict__ sa, long m,
               uint8_t *__restrict__ bwt) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  uint32_t s = sa[i];
  bwt[i] = s ? text[s - 1] : (uint8_t)0;
}

// The same for blocks that do not fit L2.  The direct gather runs at the random-access rate of the HBM (57 B of
// DRAM traffic per symbol, profiles/r1_summary.md).  Tried first: a stable split of the pairs (sa[i], i) by text
// window followed by a gather -- it trades 2^28 random reads for 2^28 single-byte stores and is no faster (count
// 0.46 + scan 0.92 + split 1.35 + gather 1.69 ms per 256 MiB block against 4.3 ms).  What works is to keep the
// OUTPUT order: one pass over the SA per text window of 2^26 positions (64 MiB, L2-resident), each pass serving
// only the entries whose symbol lies in its window.  Reads of the text hit L2, the stores of a warp fall into
// the same sectors (byte-masked), and the price is reading the SA once per window (4 B x windows per symbol).
__global__ void __launch_bounds__(256)
    bwt_window_kernel(const uint8_t *__restrict__ text, const uint32_t *__restrict__ sa, long m, int shift,
                      uint32_t window, uint8_t *__restrict__ bwt) {
  const long i = ((long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i >= m) return;
  uint32_t s[4];
  if (i + 4 <= m) {
    const uint4 v = *reinterpret_cast<const uint4 *>(sa + i);
    s[0] = v.x; s[1] = v.y; s[2] = v.z; s[3] = v.w;
  } else {
    for (int k = 0; k < 4; ++k) s[k] = i + k < m ? sa[i + k] : 0xffffffffu;
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    if (i + k >= m) break;
    // the symbol before suffix s[k]; the first suffix of the block gets the placeholder 0 (window 0)
    const uint32_t q = s[k] ? s[k] - 1 : 0u;
    if ((q >> shift) == window) bwt[i + k] = s[k] ? text[q] : (uint8_t)0;
  }
}

void bwt_from_sa(psc_context *ctx, const uint8_t *text_d, const uint32_t *sa_d, long m,
                 uint8_t *bwt_d) {
  PhaseTimer timer(ctx, &ctx->stats.ms_bwt);
  if (m == 0) return;
  static int shift = -1;  // PSC_BWT_SHIFT: log2 of the window (tests: small windows on small texts)
  if (shift < 0) { const char *e = getenv("PSC_BWT_SHIFT"); shift = e ? std::max(4, std::min(31, atoi(e))) : 26; }
  const long nw = ((m - 1) >> shift) + 1;
  if (nw > 1 && nw <= 64) {
    KTrace kt(ctx, "bwt(windows)");
    for (long w = 0; w < nw; ++w) {
      bwt_window_kernel<<<(unsigned)div_up(div_up(m, 4), 256), 256, 0, ctx->stream>>>(text_d, sa_d, m, shift, (uint32_t)w, bwt_d);
      PSC_LAUNCH_CHECK();
    }
    ctx->launches += nw;
    return;
  }
  KTrace kt(ctx, "bwt(gather)");
  bwt_kernel<<<(unsigned)div_up(m, 256), 256, 0, ctx->stream>>>(text_d, sa_d, m, bwt_d);
  PSC_LAUNCH_CHECK();
  ctx->launches++;
}

// bit bit_off+q <-> block position q+1: rank(q+1) > i0.  Scatter of single bits from the SA side
// (the bit array of a block is at most a few hundred MB and mostly L2-resident).
__global__ void __launch_bounds__(256)
    gt_begin_kernel(const uint32_t *__restrict__ sa, long m, uint32_t i0, uint32_t *__restrict__ bits,
                    long bit_off) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x + (long)i0 + 1;
  if (i >= m) return;
  uint32_t p = sa[i];
  if (p == 0) return;
  long b = bit_off + (long)p - 1;
  atomicOr(&bits[b >> 5], 1u << (b & 31));
}

void gt_begin_from_sa(psc_context *ctx, const uint32_t *sa_d, long m, uint32_t i0, uint32_t *bits_d,
                      long bit_off) {
  PhaseTimer timer(ctx, &ctx->stats.ms_bwt);
  long cnt = m - 1 - (long)i0;
  if (cnt <= 0) return;
  KTrace kt(ctx, "gt_begin");
  gt_begin_kernel<<<(unsigned)div_up(cnt, 256), 256, 0, ctx->stream>>>(sa_d, m, i0, bits_d, bit_off);
  PSC_LAUNCH_CHECK();
  ctx->launches++;
}

// ---- bit-vector layout conversion (bitvector.hpp:61-67, multifile.hpp:57-68) -----------------------
__global__ void bits_fwd_to_rev_kernel(const uint32_t *__restrict__ fwd, long nbits,
                                       uint8_t *__restrict__ rev, long fwd_off) {
  long b = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= (nbits + 7) / 8) return;
  uint32_t byte = 0;
  for (int k = 0; k < 8; ++k) {
    long r = b * 8 + k;
    if (r < nbits && bit_get(fwd, fwd_off + nbits - 1 - r)) byte |= 1u << k;
  }
  rev[b] = (uint8_t)byte;
}
__global__ void bits_rev_to_fwd_kernel(const uint8_t *__restrict__ rev, long nbits,
                                       uint32_t *__restrict__ fwd) {
  long w = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (w >= (nbits + 31) / 32) return;
  uint32_t word = 0;
  for (int k = 0; k < 32; ++k) {
    long q = w * 32 + k;
    if (q < nbits) {
      long r = nbits - 1 - q;
      if ((rev[r >> 3] >> (r & 7)) & 1) word |= 1u << k;
    }
  }
  fwd[w] = word;
}
void bits_forward_to_rev(psc_context *ctx, const uint32_t *fwd_d, long nbits, uint8_t *rev_d, long fwd_off) {
  long nb = (nbits + 7) / 8;
  if (!nb) return;
  bits_fwd_to_rev_kernel<<<(unsigned)div_up(nb, 256), 256, 0, ctx->stream>>>(fwd_d, nbits, rev_d, fwd_off);
  PSC_LAUNCH_CHECK();
  ctx->launches++;
}
void bits_rev_to_forward(psc_context *ctx, const uint8_t *rev_d, long nbits, uint32_t *fwd_d) {
  long nw = (nbits + 31) / 32;
  if (!nw) return;
  bits_rev_to_fwd_kernel<<<(unsigned)div_up(nw, 256), 256, 0, ctx->stream>>>(rev_d, nbits, fwd_d);
  PSC_LAUNCH_CHECK();
  ctx->launches++;
}

}  // namespace psc
