// merge.cu -- stage 4: gap-guided merge as exclusive scan + scatter, and the 40-bit writer.
//
// Replaces the single-threaded N-way merge of /root/reference/include/merge.hpp:54-180.  The
// host cascades right to left: `right` = merged suffixes of everything after the block, and one
// level places the block's u-th suffix at u + S(u) (S = inclusive scan of the block's gap
// array) while the entries of `right` fill the other slots in order (same rule as
// inmem_psascan_src/parallel_merge.hpp:116-154).  The running merged SA is kept as two planes
// (low 32 bits, high 8 bits); pack_sa5 interleaves them into the reference's uint40 layout
// (types/uint40.hpp:42-46: uint32 low, uint8 high).
#include "scan.cuh"

namespace psc {

constexpr int MG_ITEMS = 32;
constexpr int MG_TILE = SC_THREADS * MG_ITEMS;  // outputs per block

// E[0..m] exclusive scan of gap[0..m];  A(u) = u + E[u] = output slot where run u starts
// (run u = gap[u] entries of `right`, then the block's u-th suffix at slot A(u+1)-1 when u < m).
// Per tile of MG_TILE output slots: u0 = the run the tile starts in (one binary search per tile,
// all tiles searched in parallel by merge_tile_start_kernel so the dependent loads overlap), the
// slots taken by block suffixes are marked in a shared bitmap (they belong to u0, u0+1, ... in
// order), and a prefix popcount turns every slot into either "block suffix u0+k" or "entry
// q-(u0+k) of the right side".  O(1) work per slot, all global accesses coalesced.
__global__ void __launch_bounds__(256)
    merge_tile_start_kernel(const uint64_t *__restrict__ E, long m, long ntiles, uint64_t q_base,
                            uint32_t *__restrict__ tile_u0) {
  long tile = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (tile >= ntiles) return;
  const uint64_t q0 = q_base + (uint64_t)tile * MG_TILE;
  long lo = 0, hi = m;  // largest u in [0, m] with u + E[u] <= q0
  while (lo < hi) {
    long mid = (lo + hi + 1) >> 1;
    if ((uint64_t)mid + E[mid] <= q0) lo = mid; else hi = mid - 1;
  }
  tile_u0[tile] = (uint32_t)lo;
}

// HI = false: every position is below 2^32 (text of at most 4 GiB), the high-byte plane does not exist.
template <bool HI>
__global__ void __launch_bounds__(SC_THREADS)
    merge_level_kernel(const uint32_t *__restrict__ sa, const uint64_t *__restrict__ E,
                       const uint32_t *__restrict__ tile_u0, long m, uint64_t beg, const uint32_t *__restrict__ right_lo,
                       const uint8_t *__restrict__ right_hi, long q_base, long total, long r_base,
                       uint32_t *__restrict__ out_lo, uint8_t *__restrict__ out_hi) {
  // output slots [q_base, total) of the level go to out[0 ..); right_lo/hi[0] is entry r_base of the right side
  constexpr int WORDS = MG_TILE / 32;
  __shared__ uint32_t flags[WORDS];
  __shared__ uint32_t before[WORDS];
  const long q0 = q_base + (long)blockIdx.x * MG_TILE;
  if (threadIdx.x < WORDS) flags[threadIdx.x] = 0;
  const long u0 = (long)tile_u0[blockIdx.x];
  __syncthreads();
  // block suffix u sits at slot u + E[u+1]; only u0 .. u0+MG_TILE-1 can fall into this tile
  for (long e = threadIdx.x; e < MG_TILE; e += SC_THREADS) {
    long u = u0 + e;
    if (u >= m) break;
    uint64_t p = (uint64_t)u + E[u + 1];
    if (p >= (uint64_t)q0 + MG_TILE) break;
    if (p >= (uint64_t)q0) atomicOr(&flags[(p - q0) >> 5], 1u << ((p - q0) & 31));
  }
  __syncthreads();
  if (threadIdx.x < 32) {  // exclusive prefix popcount over the WORDS flag words
    uint32_t run = 0;
    for (int w0 = 0; w0 < WORDS; w0 += 32) {
      uint32_t c = __popc(flags[w0 + threadIdx.x]), incl = c;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (threadIdx.x >= o) incl += t; }
      before[w0 + threadIdx.x] = run + incl - c;
      run += __shfl_sync(0xffffffffu, incl, 31);
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < MG_ITEMS; ++k) {
    int i = threadIdx.x + k * SC_THREADS;
    long q = q0 + i;
    if (q >= total) break;
    uint32_t word = flags[i >> 5];
    uint32_t nb = before[i >> 5] + __popc(word & ((1u << (i & 31)) - 1u));
    uint64_t v;
    if ((word >> (i & 31)) & 1u) {
      v = beg + sa[u0 + nb];
    } else {
      uint64_t r = (uint64_t)q - (uint64_t)(u0 + nb) - (uint64_t)r_base;  // block suffixes before slot q: u0 + nb
      v = (uint64_t)right_lo[r];
      if (HI) v |= (uint64_t)right_hi[r] << 32;
    }
    out_lo[q - q_base] = (uint32_t)v;
    if (HI) out_hi[q - q_base] = (uint8_t)(v >> 32);
  }
}

__global__ void __launch_bounds__(256)
    sa_to_planes_kernel(const uint32_t *__restrict__ sa, long m, uint64_t beg,
                        uint32_t *__restrict__ lo, uint8_t *__restrict__ hi) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  uint64_t v = beg + sa[i];
  lo[i] = (uint32_t)v;
  if (hi) hi[i] = (uint8_t)(v >> 32);
}

// number of block suffixes in output slots below q[j] (= the run slot q[j] lies in), for a few slots
__global__ void merge_bounds_kernel(const uint64_t *__restrict__ E, long m, const uint64_t *__restrict__ q, int nq,
                                    uint64_t *__restrict__ out) {
  int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nq) return;
  long lo = 0, hi = m;  // largest u in [0, m] with u + E[u] <= q[j]
  while (lo < hi) {
    long mid = (lo + hi + 1) >> 1;
    if ((uint64_t)mid + E[mid] <= q[j]) lo = mid; else hi = mid - 1;
  }
  out[j] = (uint64_t)lo;
}

void merge_gap_scan(psc_context *ctx, const uint32_t *gap_d, long m, uint64_t *E_d) {
  KTrace kt(ctx, "gap_scan");
  exclusive_scan_u32_to_u64(ctx, gap_d, E_d, m + 1, nullptr);
}

void merge_bounds(psc_context *ctx, const uint64_t *E_d, long m, const uint64_t *slots, int nq, uint64_t *before) {
  DevBuf<uint64_t> q(ctx, nq), out(ctx, nq);
  h2d(ctx, q.p, slots, nq);
  merge_bounds_kernel<<<div_up(nq, 64), 64, 0, ctx->stream>>>(E_d, m, q.p, nq, out.p);
  PSC_LAUNCH_CHECK();
  ctx->launches++;
  d2h(ctx, before, out.p, nq);
  sync(ctx);
}

// (the callers time the merge phase as a whole: a PhaseTimer synchronises the stream)
void sa_slice_to_planes(psc_context *ctx, const uint32_t *sa_d, long count, long beg, uint32_t *out_lo, uint8_t *out_hi) {
  if (count <= 0) return;
  { KTrace kt(ctx, "sa_to_planes"); sa_to_planes_kernel<<<(unsigned)div_up(count, 256), 256, 0, ctx->stream>>>(sa_d, count, (uint64_t)beg, out_lo, out_hi); }
  PSC_LAUNCH_CHECK();
  ctx->launches++;
}

// output slots [q_beg, q_end) of one level; right_lo/hi[0] = entry r_base of the right side
void merge_level_range(psc_context *ctx, const uint32_t *sa_d, const uint64_t *E_d, long m, long beg,
                       const uint32_t *right_lo, const uint8_t *right_hi, long r_base, long q_beg, long q_end,
                       uint32_t *out_lo, uint8_t *out_hi) {
  if (q_end <= q_beg) return;
  const long ntiles = div_up(q_end - q_beg, MG_TILE);
  DevBuf<uint32_t> tile_u0(ctx, ntiles);
  { KTrace kt(ctx, "merge_tile_start"); merge_tile_start_kernel<<<(unsigned)div_up(ntiles, 256), 256, 0, ctx->stream>>>(E_d, m, ntiles, (uint64_t)q_beg, tile_u0.p); }
  PSC_LAUNCH_CHECK();
  ctx->launches++;
  {
    KTrace kt(ctx, "merge_level");
    if (out_hi)
      merge_level_kernel<true><<<(unsigned)ntiles, SC_THREADS, 0, ctx->stream>>>(
          sa_d, E_d, tile_u0.p, m, (uint64_t)beg, right_lo, right_hi, q_beg, q_end, r_base, out_lo, out_hi);
    else
      merge_level_kernel<false><<<(unsigned)ntiles, SC_THREADS, 0, ctx->stream>>>(
          sa_d, E_d, tile_u0.p, m, (uint64_t)beg, right_lo, right_hi, q_beg, q_end, r_base, out_lo, out_hi);
  }
  PSC_LAUNCH_CHECK();
  ctx->launches++;
}

void merge_level(psc_context *ctx, const uint32_t *sa_d, const uint32_t *gap_d, long m, long beg,
                 const uint32_t *right_lo, const uint8_t *right_hi, long right_len,
                 uint32_t *out_lo, uint8_t *out_hi) {
  long total = m + right_len;
  if (total == 0) return;
  PhaseTimer timer(ctx, &ctx->stats.ms_merge);
  if (right_len == 0) { sa_slice_to_planes(ctx, sa_d, m, beg, out_lo, out_hi); return; }
  DevBuf<uint64_t> E(ctx, m + 1);
  merge_gap_scan(ctx, gap_d, m, E.p);
  merge_level_range(ctx, sa_d, E.p, m, beg, right_lo, right_hi, 0, 0, total, out_lo, out_hi);
}

// ---- 40-bit writer: 4 entries = 20 bytes = 5 aligned words per thread --------------------------
__global__ void __launch_bounds__(256)
    pack_sa5_kernel(const uint32_t *__restrict__ lo, const uint8_t *__restrict__ hi, long n,
                    uint32_t *out_words, uint8_t *out_bytes) {
  long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
  long i = g * 4;
  if (i >= n) return;
  if (i + 4 <= n) {
    uint4 l = *reinterpret_cast<const uint4 *>(lo + i);
    uint32_t h = hi ? *reinterpret_cast<const uint32_t *>(hi + i) : 0u;
    uint32_t h0 = h & 255, h1 = (h >> 8) & 255, h2 = (h >> 16) & 255, h3 = h >> 24;
    uint32_t *o = out_words + g * 5;
    o[0] = l.x;
    o[1] = h0 | (l.y << 8);
    o[2] = (l.y >> 24) | (h1 << 8) | (l.z << 16);
    o[3] = (l.z >> 16) | (h2 << 16) | (l.w << 24);
    o[4] = (l.w >> 8) | (h3 << 24);
  } else {
    for (long k = i; k < n; ++k) {
      uint32_t l = lo[k];
      uint8_t *o = out_bytes + k * 5;
      o[0] = (uint8_t)l; o[1] = (uint8_t)(l >> 8); o[2] = (uint8_t)(l >> 16); o[3] = (uint8_t)(l >> 24);
      o[4] = hi ? hi[k] : (uint8_t)0;
    }
  }
}

void pack_sa5(psc_context *ctx, const uint32_t *lo, const uint8_t *hi, long n, uint8_t *out5) {
  if (n <= 0) return;
  { KTrace kt(ctx, "pack_sa5"); pack_sa5_kernel<<<(unsigned)div_up(div_up(n, 4), 256), 256, 0, ctx->stream>>>(
      lo, hi, n, reinterpret_cast<uint32_t *>(out5), out5); }
  PSC_LAUNCH_CHECK();
  ctx->launches++;
}

// single piece: beg + sa[i] straight to the 40-bit layout (no planes in between)
__global__ void __launch_bounds__(256)
    pack_sa32_kernel(const uint32_t *__restrict__ sa, long n, uint64_t beg, uint32_t *out_words,
                     uint8_t *out_bytes) {
  long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
  long i = g * 4;
  if (i >= n) return;
  if (i + 4 <= n) {
    uint4 s = *reinterpret_cast<const uint4 *>(sa + i);
    uint64_t v0 = beg + s.x, v1 = beg + s.y, v2 = beg + s.z, v3 = beg + s.w;
    uint32_t l0 = (uint32_t)v0, l1 = (uint32_t)v1, l2 = (uint32_t)v2, l3 = (uint32_t)v3;
    uint32_t h0 = (uint32_t)(v0 >> 32) & 255, h1 = (uint32_t)(v1 >> 32) & 255, h2 = (uint32_t)(v2 >> 32) & 255,
             h3 = (uint32_t)(v3 >> 32) & 255;
    uint32_t *o = out_words + g * 5;
    o[0] = l0;
    o[1] = h0 | (l1 << 8);
    o[2] = (l1 >> 24) | (h1 << 8) | (l2 << 16);
    o[3] = (l2 >> 16) | (h2 << 16) | (l3 << 24);
    o[4] = (l3 >> 8) | (h3 << 24);
  } else {
    for (long k = i; k < n; ++k) {
      uint64_t v = beg + sa[k];
      uint8_t *o = out_bytes + k * 5;
      for (int b = 0; b < 5; ++b) o[b] = (uint8_t)(v >> (8 * b));
    }
  }
}

static void pack_sa5_from_sa_launch(psc_context *ctx, const uint32_t *sa, long n, long beg, uint8_t *out5);
void pack_sa5_from_sa(psc_context *ctx, const uint32_t *sa, long n, long beg, uint8_t *out5, bool timed) {
  if (n <= 0) return;
  if (!timed) { pack_sa5_from_sa_launch(ctx, sa, n, beg, out5); return; }
  PhaseTimer timer(ctx, &ctx->stats.ms_merge);
  pack_sa5_from_sa_launch(ctx, sa, n, beg, out5);
}
static void pack_sa5_from_sa_launch(psc_context *ctx, const uint32_t *sa, long n, long beg, uint8_t *out5) {
  { KTrace kt(ctx, "pack_sa32"); pack_sa32_kernel<<<(unsigned)div_up(div_up(n, 4), 256), 256, 0, ctx->stream>>>(
      sa, n, (uint64_t)beg, reinterpret_cast<uint32_t *>(out5), out5); }
  PSC_LAUNCH_CHECK();
  ctx->launches++;
}

__global__ void __launch_bounds__(256)
    unpack_sa5_kernel(const uint8_t *__restrict__ in5, long n, uint32_t *__restrict__ lo,
                      uint8_t *__restrict__ hi) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint8_t *p = in5 + i * 5;
  lo[i] = (uint32_t)p[0] | ((uint32_t)p[1] << 8) | ((uint32_t)p[2] << 16) | ((uint32_t)p[3] << 24);
  hi[i] = p[4];
}

void unpack_sa5(psc_context *ctx, const uint8_t *in5, long n, uint32_t *lo, uint8_t *hi) {
  if (n <= 0) return;
  unpack_sa5_kernel<<<(unsigned)div_up(n, 256), 256, 0, ctx->stream>>>(in5, n, lo, hi);
  PSC_LAUNCH_CHECK();
  ctx->launches++;
}

}  // namespace psc
