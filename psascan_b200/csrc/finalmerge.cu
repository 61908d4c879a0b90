// finalmerge.cu -- stage 4 as ONE N-way merge by output range (replaces the level-after-level cascade
// over the whole text that round 1 ran after every block).
//
// The reference merges all (half-)block suffix arrays at the end, guided by their gap arrays
// (/root/reference/include/merge.hpp:54-180).  Here the blocks' SAs and gap arrays are kept until the
// end as well, and the merged sequence is produced chunk by chunk of OUTPUT slots:
//   M_k = merge of blocks k..B-1;  in M_k block k's u-th suffix sits at slot u + E_k[u+1]
//   (E_k = exclusive scan of gap_k) and the entries of M_{k+1} fill the other slots in order.
//   A range [a, b) of M_k therefore consists of block-k suffixes [u(a), u(b)) and the range
//   [a - u(a), b - u(b)) of M_{k+1}, where u(q) = number of block suffixes in slots < q
//   = the largest u with u + E_k[u] <= q.  Going down the levels, a chunk of final output slots
//   needs ONE contiguous slice of every block's SA and gap array ("split", B small searches per
//   chunk boundary), and is then merged bottom-up from those slices alone ("cascade").
// What this buys: (1) nothing of size n is double-buffered -- the pipeline keeps 8 bytes per suffix
// (SA + gap) instead of two merged planes; (2) over several GPUs every rank merges its own output
// range after ONE exchange of slices instead of an all-to-all per block (dist.cu); (3) chunks can be
// streamed to the host while the next one is merged, and the slices can come from host memory.
#include <algorithm>
#include <vector>
#include "scan.cuh"
#include "finalmerge.cuh"

namespace psc {

// ---- split: u(q) for a batch of slots, against one slice of one block's gap array ------------------
// The slice holds gap entries [u_off, u_off + cnt) of gap[0..m]; E[u_off] = e_base and carry[j] is
// the sum of the slice's first j * COARSE_TILE entries (carry[tiles] = the slice total).  P(u) :=
// u + E[u] <= q is true for u = 0 and false from some u on; the slice answers a query only if the
// last true u lies inside it (last_slice: the slice ends with entry m), else it writes 0, so that the
// answers of all slices of a block can simply be summed (dist.cu: one all-reduce per level).
__global__ void __launch_bounds__(64)
    merge_split_kernel(const uint32_t *__restrict__ gap, long cnt, long u_off, uint64_t e_base,
                       const uint64_t *__restrict__ carry, long tiles, int last_slice,
                       const uint64_t *__restrict__ q, int nq, uint64_t *__restrict__ u_out,
                       uint64_t *__restrict__ e_out) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nq) return;
  const uint64_t want = q[j];
  uint64_t ru = 0, re = 0;
  const bool p_first = (uint64_t)u_off + e_base <= want;
  const bool p_end = (uint64_t)(u_off + cnt) + e_base + carry[tiles] <= want;  // P at the first entry after the slice
  if (cnt > 0 && p_first && (last_slice || !p_end)) {
    long lo = 0, hi = tiles - 1;  // largest tile whose first entry satisfies P
    while (lo < hi) {
      const long mid = (lo + hi + 1) >> 1;
      if ((uint64_t)(u_off + mid * COARSE_TILE) + e_base + carry[mid] <= want) lo = mid; else hi = mid - 1;
    }
    long k = lo * COARSE_TILE;  // index inside the slice
    uint64_t e = e_base + carry[lo];
    const long stop = min(cnt, k + COARSE_TILE) - 1;
    while (k < stop) {
      const uint64_t e1 = e + gap[k];
      if ((uint64_t)(u_off + k + 1) + e1 > want) break;
      e = e1;
      ++k;
    }
    ru = (uint64_t)(u_off + k);
    re = e;
  }
  u_out[j] = ru;
  e_out[j] = re;
}

// slots of the next level: q' = q - u(q)
__global__ void merge_split_next_kernel(const uint64_t *__restrict__ q, const uint64_t *__restrict__ u, int nq,
                                        uint64_t *__restrict__ q_next) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < nq) q_next[j] = q[j] - u[j];
}

void merge_split_slice(psc_context *ctx, const uint32_t *gap_slice_d, long cnt, long u_off, uint64_t e_base,
                       const uint64_t *carry_d, bool last_slice, const uint64_t *q_d, int nq, uint64_t *u_d,
                       uint64_t *e_d) {
  merge_split_kernel<<<div_up(nq, 64), 64, 0, ctx->stream>>>(gap_slice_d, cnt, u_off, e_base, carry_d,
                                                             coarse_carry_tiles(cnt), last_slice ? 1 : 0, q_d, nq, u_d, e_d);
  PSC_LAUNCH_CHECK();
  ctx->launches++;
}

void merge_split_next(psc_context *ctx, const uint64_t *q_d, const uint64_t *u_d, int nq, uint64_t *q_next_d) {
  merge_split_next_kernel<<<div_up(nq, 64), 64, 0, ctx->stream>>>(q_d, u_d, nq, q_next_d);
  PSC_LAUNCH_CHECK();
  ctx->launches++;
}

// ---- one chunk of output slots from its slices ------------------------------------------------------------------
// Round 2's first version merged level by level into a growing array of (lo, hi) planes: every level rewrote what
// the levels below had produced, 4.5 x the chunk in total for 8 blocks (36 of 62 ms on the bench workload).  Now a
// level only contributes a BITMAP over its slots (1 = suffix of this block, at slot u + E'[u+1]) with a popcount
// directory, and one pass over the output slots walks down the levels: at level k a slot either is the r-th suffix
// of block k (r = ones before it) or continues as slot (zeros before it) of the level below.  Neighbouring slots
// take neighbouring paths, so the bitmaps are read through L1/L2; each suffix array entry is read once and each
// 40-bit value written once.
struct DescentLevel {
  const uint2 *bd;   // per 32 slots: x = bitmap word, y = ones before the word (counted over the bitmaps of ALL levels
                     // of the chunk); nullptr: bottom level, every slot is a suffix of this block
  uint32_t rebase;   // ones of the levels above (their suffixes), subtracted from y
  const uint32_t *sa;
  uint64_t beg;
};
constexpr int DESCENT_MAX_LEVELS = 64;  // per kernel launch; more blocks than that go through device memory
struct DescentLevels { DescentLevel lv[DESCENT_MAX_LEVELS]; int n; };

// Bits of one level from its gap slice, fused with the scan: tile t (SCAN tile of 2048 counters) knows the sum of
// the counters before it (agg, scanned), every thread the inclusive sums of its 8 counters, hence the slot
// u + E'[u+1] of every suffix u.  Slots grow with u, so the slots of a tile form a range no other tile writes
// into except for its first and last word: the range is built in shared memory and stored, the two boundary words
// (and ranges too long for shared memory) go through atomicOr.
constexpr int LB_ITEMS = 8, LB_TILE = SC_THREADS * LB_ITEMS, LB_WORDS = 1024;  // 64 Ki slots per tile in shared memory
static_assert(LB_TILE == COARSE_TILE, "tiles of the level-bits kernel are the scan's tiles");

__global__ void __launch_bounds__(SC_THREADS)
    level_bits_kernel(const uint32_t *__restrict__ gap, long cb, const uint64_t *__restrict__ agg /* scanned tile sums */,
                      uint64_t skip, unsigned long long *__restrict__ bits) {
  __shared__ uint64_t sm[SC_THREADS / 32 + 1];
  __shared__ unsigned long long shw[LB_WORDS];
  __shared__ uint64_t range[2];
  const long base = (long)blockIdx.x * LB_TILE + (long)threadIdx.x * LB_ITEMS;
  uint32_t v[LB_ITEMS];
  uint64_t s = 0;
#pragma unroll
  for (int k = 0; k < LB_ITEMS; ++k) { v[k] = base + k < cb ? gap[base + k] : 0u; s += v[k]; }
  uint64_t tot;
  uint64_t incl = block_excl_sum<uint64_t>(s, sm, &tot) + agg[blockIdx.x] - skip;  // E'[base] (skip leaves after entry 0)
  // slot of suffix u = u + E'[u + 1] = u + (sum of counters 0..u) - skip
  const long tile_first = (long)blockIdx.x * LB_TILE, tile_last = min(cb, tile_first + LB_TILE) - 1;
  if (threadIdx.x == 0) range[0] = ~0ull;
  __syncthreads();
  uint64_t slot[LB_ITEMS];
#pragma unroll
  for (int k = 0; k < LB_ITEMS; ++k) {
    incl += v[k];
    slot[k] = (uint64_t)(base + k) + incl;
    if (base + k == tile_first) range[0] = slot[k];
    if (base + k == tile_last) range[1] = slot[k];
  }
  __syncthreads();
  const uint64_t w0 = range[0] >> 6, w1 = range[1] >> 6;
  if (w1 - w0 + 1 > (uint64_t)LB_WORDS) {  // a very sparse stretch: straight to global memory
#pragma unroll
    for (int k = 0; k < LB_ITEMS; ++k)
      if (base + k < cb) atomicOr(&bits[slot[k] >> 6], 1ull << (slot[k] & 63));
    return;
  }
  const int nwords = (int)(w1 - w0 + 1);
  for (int w = threadIdx.x; w < nwords; w += SC_THREADS) shw[w] = 0ull;
  __syncthreads();
#pragma unroll
  for (int k = 0; k < LB_ITEMS; ++k)
    if (base + k < cb) atomicOr(&shw[(slot[k] >> 6) - w0], 1ull << (slot[k] & 63));
  __syncthreads();
  for (int w = threadIdx.x; w < nwords; w += SC_THREADS) {
    if (w == 0 || w == nwords - 1) { if (shw[w]) atomicOr(&bits[w0 + w], shw[w]); }
    else bits[w0 + w] = shw[w];
  }
}

// {word, ones before it} side by side: one 8-byte load per level and slot.  Positions inside a chunk fit 32 bits
// (chunks have at most 2^27 slots), which keeps the walk at about ten instructions per level: the pass is bound
// by instruction issue, not by memory (64-bit words and directories measured 25 ms per 2^31 slots).
__global__ void __launch_bounds__(256)
    word_popc_kernel(const uint32_t *__restrict__ bits, long nwords, uint32_t *__restrict__ cnt) {
  const long w = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (w < nwords) cnt[w] = (uint32_t)__popc(bits[w]);
}
__global__ void __launch_bounds__(256)
    interleave_kernel(const uint32_t *__restrict__ bits, const uint64_t *__restrict__ dir, long nwords,
                      uint2 *__restrict__ bd) {
  const long w = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (w < nwords) bd[w] = make_uint2(bits[w], (uint32_t)dir[w]);
}

__device__ __forceinline__ uint64_t descend(const DescentLevel *__restrict__ lv, int n, uint32_t pos) {
  for (int k = 0; k < n; ++k) {
    const uint2 *bd = lv[k].bd;
    if (!bd) return lv[k].beg + lv[k].sa[pos];
    const uint2 e = bd[pos >> 5];
    const uint32_t b = pos & 31u;
    const uint32_t ones = e.y - lv[k].rebase + (uint32_t)__popc(e.x & ((1u << b) - 1u));
    if ((e.x >> b) & 1u) return lv[k].beg + lv[k].sa[ones];
    pos -= ones;
  }
  return 0;  // not reached: the last level has no bitmap
}

// 4 slots = 20 bytes = 5 aligned words per thread
__global__ void __launch_bounds__(256)
    descent_kernel(const __grid_constant__ DescentLevels dl, const DescentLevel *__restrict__ lv_d, long total,
                   uint32_t *out_words, uint8_t *out_bytes) {
  const long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
  const long i = g * 4;
  if (i >= total) return;
  const DescentLevel *lv = lv_d ? lv_d : dl.lv;
  if (i + 4 <= total) {
    uint64_t v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) v[j] = descend(lv, dl.n, (uint32_t)(i + j));
    const uint32_t l0 = (uint32_t)v[0], l1 = (uint32_t)v[1], l2 = (uint32_t)v[2], l3 = (uint32_t)v[3];
    const uint32_t h0 = (uint32_t)(v[0] >> 32) & 255, h1 = (uint32_t)(v[1] >> 32) & 255, h2 = (uint32_t)(v[2] >> 32) & 255,
                   h3 = (uint32_t)(v[3] >> 32) & 255;
    uint32_t *o = out_words + g * 5;
    o[0] = l0;
    o[1] = h0 | (l1 << 8);
    o[2] = (l1 >> 24) | (h1 << 8) | (l2 << 16);
    o[3] = (l2 >> 16) | (h2 << 16) | (l3 << 24);
    o[4] = (l3 >> 8) | (h3 << 24);
  } else {
    for (long k = i; k < total; ++k) {
      const uint64_t v = descend(lv, dl.n, (uint32_t)k);
      uint8_t *o = out_bytes + k * 5;
      for (int b = 0; b < 5; ++b) o[b] = (uint8_t)(v >> (8 * b));
    }
  }
}

void MergeScratch::reserve(psc_context *ctx, long entries, bool) {
  if (entries <= cap) return;
  cap = entries;
  E.alloc(ctx, entries + 2);
}

void cascade_chunk(psc_context *ctx, const LevelSlice *lv, int n_levels, long total, MergeScratch &scr,
                   uint8_t *out5_d) {
  if (total <= 0) return;
  if (total > scr.cap) { set_error("cascade_chunk: chunk of %ld entries exceeds the scratch (%ld)", total, scr.cap); throw CudaError{PSC_EARG}; }
  // levels that hold suffixes of the chunk, left to right; L = slots of the level inside the chunk
  std::vector<int> live;
  long have = 0;
  for (int k = 0; k < n_levels; ++k) if (lv[k].cb > 0) { live.push_back(k); have += lv[k].cb; }
  if (have != total) { set_error("cascade_chunk: slices hold %ld entries, the chunk has %ld slots", have, total); throw CudaError{PSC_EARG}; }
  const int nl = (int)live.size();
  std::vector<DescentLevel> host(nl);
  // one allocation for all bitmaps and directories: words of level j = L_j / 64 + 1
  std::vector<long> L(nl), woff(nl + 1, 0);
  {
    long slots = total;
    for (int j = 0; j < nl; ++j) { L[j] = slots; woff[j + 1] = woff[j] + (j + 1 < nl ? slots / 64 + 1 : 0); slots -= lv[live[j]].cb; }
  }
  const long words = std::max(1L, woff[nl]);
  // `words` 64-bit words are built (level_bits_kernel), read as 2 * words 32-bit words afterwards
  const long w32 = 2 * words;
  DevBuf<uint64_t> bits(ctx, words), dir(ctx, w32);
  DevBuf<uint2> bd(ctx, w32);
  DevBuf<uint32_t> cnt(ctx, w32);
  PSC_CUDA(cudaMemsetAsync(bits.p, 0, (size_t)words * 8, ctx->stream));
  for (int j = 0; j < nl; ++j) {
    const LevelSlice &s = lv[live[j]];
    host[j].sa = s.sa;
    host[j].beg = (uint64_t)s.beg;
    host[j].bd = nullptr;
    host[j].rebase = 0;
    if (j + 1 == nl) break;  // bottom level
    // tile sums of the gap slice, scanned; then the bits of every tile (the scanned counters themselves are never stored)
    KTrace kt(ctx, "merge_bits");
    const long tiles = coarse_carry_tiles(s.cb);
    coarse_carries_u32(ctx, s.gap, s.cb, scr.E.p);  // E[t] = sum of the counters before tile t
    level_bits_kernel<<<(unsigned)tiles, SC_THREADS, 0, ctx->stream>>>(s.gap, s.cb, scr.E.p, s.skip,
                                                                        reinterpret_cast<unsigned long long *>(bits.p + woff[j]));
    PSC_LAUNCH_CHECK();
    ctx->launches++;
    host[j].bd = bd.p + 2 * woff[j];
  }
  if (woff[nl] > 0) {  // directories of all levels at once: per-word popcounts, one scan, interleaved with the words
    KTrace kt(ctx, "merge_dir");
    const uint32_t *b32 = reinterpret_cast<const uint32_t *>(bits.p);
    word_popc_kernel<<<(unsigned)div_up(w32, 256), 256, 0, ctx->stream>>>(b32, w32, cnt.p);
    PSC_LAUNCH_CHECK();
    exclusive_scan_u32_to_u64(ctx, cnt.p, dir.p, w32, nullptr);
    interleave_kernel<<<(unsigned)div_up(w32, 256), 256, 0, ctx->stream>>>(b32, dir.p, w32, bd.p);
    PSC_LAUNCH_CHECK();
    ctx->launches += 2;
  }
  // the scan runs over all levels: level j's directory is offset by the ones of the levels before it, which is the
  // number of suffixes of those levels -- known on the host
  DescentLevels dl;
  memset(&dl, 0, sizeof(dl));
  dl.n = nl;
  DevBuf<DescentLevel> lv_d;
  uint64_t ones_before = 0;
  for (int j = 0; j < nl; ++j) {
    if (host[j].bd) host[j].rebase = (uint32_t)ones_before;
    ones_before += (uint64_t)lv[live[j]].cb;
  }
  if (nl <= DESCENT_MAX_LEVELS) {
    for (int j = 0; j < nl; ++j) dl.lv[j] = host[j];
  } else {
    lv_d.alloc(ctx, nl);
    h2d(ctx, lv_d.p, host.data(), (size_t)nl);
    sync(ctx);
  }
  KTrace kt(ctx, "merge_descent");
  descent_kernel<<<(unsigned)div_up(div_up(total, 4), 256), 256, 0, ctx->stream>>>(dl, lv_d.p, total,
                                                                                     reinterpret_cast<uint32_t *>(out5_d), out5_d);
  PSC_LAUNCH_CHECK();
  ctx->launches++;
}

// ---- one GPU, everything resident -------------------------------------------------------------------
long merge_chunk_entries() {
  static long c = 0;
  if (!c) {
    const char *e = getenv("PSC_MERGE_CHUNK");
    c = e ? std::max(4L, atol(e) / 4 * 4) : (1L << 27);
  }
  return c;
}

void MergePlan::resize(int levels, int boundaries) {
  B = levels;
  nb = boundaries;
  q.assign((size_t)B * nb, 0); u.assign((size_t)B * nb, 0); e.assign((size_t)B * nb, 0);
}

LevelSlice MergePlan::slice(int k, int i, const uint32_t *sa_at_ua, const uint32_t *gap_at_ua, long beg) const {
  LevelSlice s;
  const uint64_t ua = U(k, i), ub = U(k, i + 1);
  s.sa = sa_at_ua;
  s.gap = gap_at_ua;
  s.cb = (long)(ub - ua);
  s.skip = (k == B - 1) ? 0 : Q(k, i) - (ua + Ev(k, i));
  s.beg = beg;
  return s;
}

void final_merge_resident(psc_context *ctx, const std::vector<MergePiece> &pc, long n, uint8_t *out5_d,
                          const ChunkSink *sink) {
  PhaseTimer timer(ctx, &ctx->stats.ms_merge);
  const int B = (int)pc.size();
  const long C = merge_chunk_entries();
  const int nb = (int)div_up(n, C) + 1;
  MergePlan plan;
  plan.resize(B, nb);
  for (int i = 0; i < nb; ++i) plan.q[i] = (uint64_t)std::min(n, (long)i * C);
  {
    KTrace kt(ctx, "merge_split");
    DevBuf<uint64_t> tab(ctx, (size_t)3 * B * nb);
    uint64_t *Qd = tab.p, *Ud = tab.p + (size_t)B * nb, *Ed = tab.p + (size_t)2 * B * nb;
    DevBuf<uint64_t> sums(ctx, B);  // sum of every block's gap counters
    h2d(ctx, Qd, plan.q.data(), nb);
    for (int k = 0; k < B; ++k) {
      uint64_t *qk = Qd + (size_t)k * nb, *uk = Ud + (size_t)k * nb, *ek = Ed + (size_t)k * nb;
      if (k == B - 1) {  // the rightmost block: every slot is a block suffix
        PSC_CUDA(cudaMemcpyAsync(uk, qk, (size_t)nb * 8, cudaMemcpyDeviceToDevice, ctx->stream));
        PSC_CUDA(cudaMemsetAsync(ek, 0, (size_t)nb * 8, ctx->stream));
        break;
      }
      const long cnt = pc[k].m + 1;
      DevBuf<uint64_t> carry(ctx, coarse_carry_tiles(cnt) + 1);
      coarse_carries_u32(ctx, pc[k].gap, cnt, carry.p);
      PSC_CUDA(cudaMemcpyAsync(sums.p + k, carry.p + coarse_carry_tiles(cnt), 8, cudaMemcpyDeviceToDevice, ctx->stream));
      merge_split_slice(ctx, pc[k].gap, cnt, 0, 0, carry.p, true, qk, nb, uk, ek);
      merge_split_next(ctx, qk, uk, nb, qk + nb);
    }
    d2h(ctx, plan.q.data(), Qd, (size_t)B * nb);
    d2h(ctx, plan.u.data(), Ud, (size_t)B * nb);
    d2h(ctx, plan.e.data(), Ed, (size_t)B * nb);
    std::vector<uint64_t> hs(B, 0);
    d2h(ctx, hs.data(), sums.p, (size_t)(B - 1));
    sync(ctx);
    // uint32 counters (the reference: 1 byte + excess list, gap_array.hpp:63, 79-88) must add up to the tail length
    for (int k = 0; k + 1 < B; ++k) {
      const uint64_t tail = (uint64_t)(n - (pc[k].beg + pc[k].m));
      if (hs[k] != tail) {
        set_error("gap counters of block %d add up to %llu, its tail has %llu suffixes: a 32-bit counter overflowed", k,
                  (unsigned long long)hs[k], (unsigned long long)tail);
        throw CudaError{PSC_ELIMIT};
      }
    }
  }
  MergeScratch scr;
  scr.reserve(ctx, std::min(n, C) + 16, need_hi_plane(n));
  std::vector<LevelSlice> lv(B);
  for (int i = 0; i + 1 < nb; ++i) {
    const long a = (long)plan.Q(0, i), b = (long)plan.Q(0, i + 1);
    for (int k = 0; k < B; ++k) {
      const uint64_t ua = plan.U(k, i);
      lv[k] = plan.slice(k, i, pc[k].sa + ua, pc[k].gap ? pc[k].gap + ua : nullptr, pc[k].beg);
    }
    uint8_t *dst = sink ? sink->buffer(ctx, i, b - a) : out5_d + 5 * a;
    cascade_chunk(ctx, lv.data(), B, b - a, scr, dst);
    if (sink) sink->done(ctx, i, a, b - a);
  }
}

}  // namespace psc
