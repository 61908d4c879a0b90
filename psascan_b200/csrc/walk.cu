// walk.cu -- stage 2 (initial ranks) and stage 3b (the LF walk / gap computation).
//
// K5 replaces em_compute_initial_ranks() (/root/reference/include/em_compute_initial_ranks.hpp:
// 222-319, 513-561): one thread per walker start binary-searches the block SA; a block suffix
// that runs into the block end is decided by the tail's gt bit exactly like lcp_compare()
// (:54-76), so every rank is exact and independent (no right-to-left tie chain, :285-310).
//
// K6 replaces parallel_stream() + gap_parallel_updater (stream.hpp:147-158, update.hpp:86-96,
// compute_gap.hpp:77-85): one walker per thread, thousands per SM, each doing the dependent
// chain i <- C[c] + occ(c, i) (+1) with one 32-byte sector read per step and a RED on the
// uint32 gap counter (optionally warp-aggregated through match.any).
#include <algorithm>
#include <vector>
#include "rank.cuh"
#include "texthash.cuh"

namespace psc {

// ---- K5 ---------------------------------------------------------------------------------------
// Every comparison of the binary search looks at most `cap` symbols directly and finishes by fingerprints
// (texthash.cuh) when they all agree; cap <= 0: symbol by symbol under the work guard.
__global__ void __launch_bounds__(128)
    initial_ranks_kernel(const uint8_t *__restrict__ block, const uint32_t *__restrict__ sa, long m,
                         const uint8_t *__restrict__ after, long after_len, int after_is_text_end,
                         const uint32_t *__restrict__ gt_after, long gt_after_off, long nstarts,
                         const uint64_t *__restrict__ starts, uint64_t *__restrict__ out, WorkBudget *wb,
                         const __grid_constant__ HashView hv, long off_block, long off_after, long cap) {
  long k = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nstarts) return;
  long st = (long)starts[k];
  long plen = after_len - st;
  if (plen <= 0 && after_is_text_end) { out[k] = 0; return; }  // :175-178 empty suffix
  const uint8_t *pat = after + st;
  long left = 0, right = m;
  while (left != right) {
    long mid = (left + right) >> 1;
    long s = sa[mid];
    long slen = m - s;
    long lim = slen < plen ? slen : plen;
    long lcp;
    if (cap <= 0) {
      lcp = common_prefix_guarded(block + s, pat, lim, wb);
    } else {
      lcp = common_prefix(block + s, pat, min(lim, cap));
      if (lcp == cap && cap < lim) {
        if (hv.pre) lcp = th_lcp(hv, off_block + s, off_after + st, cap, lim);
        else { wb->need_hash = 1; lcp = -1; }
      }
    }
    if (lcp < 0) { out[k] = 0; return; }  // guard tripped or fingerprints needed: the host discards the result
    bool pat_greater;
    if (lcp == slen) pat_greater = bit_get(gt_after, gt_after_off + st + slen - 1);  // :66-68
    else if (lcp == plen) pat_greater = false;                                       // :69-71
    else pat_greater = pat[lcp] > block[s + lcp];                                    // :72-75
    if (pat_greater) left = mid + 1; else right = mid;
  }
  out[k] = (uint64_t)left;
}

void initial_ranks(psc_context *ctx, const uint8_t *block_d, const uint32_t *sa_d, long m,
                   const uint8_t *after_d, long after_len, int after_is_text_end,
                   const uint32_t *gt_after_bits_d, long gt_after_off, long nstarts,
                   const uint64_t *starts_d, uint64_t *out_d) {
  PhaseTimer timer(ctx, &ctx->stats.ms_init_ranks);
  if (nstarts <= 0) return;
  const bool contiguous = (after_d == block_d + m) && after_len > 0;
  long cap = (contiguous && hash_domain_ensure(ctx, block_d, after_d + after_len)) ? lcp_cap() : 0;
  HashView hv;
  if (cap > 0) hash_view(ctx, block_d, after_d + after_len, false, &hv);
  for (int attempt = 0;; ++attempt) {
    WorkBudget *wb = work_budget_arm(ctx);
    const long off_block = hv.pre ? (long)(block_d - hv.text) : 0, off_after = hv.pre ? (long)(after_d - hv.text) : 0;
    initial_ranks_kernel<<<(unsigned)div_up(nstarts, 128), 128, 0, ctx->stream>>>(
        block_d, sa_d, m, after_d, after_len, after_is_text_end, gt_after_bits_d, gt_after_off, nstarts,
        starts_d, out_d, wb, hv, off_block, off_after, cap);
    PSC_LAUNCH_CHECK();
    ctx->launches++;
    if (!work_budget_check(ctx, wb, "initial ranks (K5)")) return;
    if (attempt > 0 || !hash_view(ctx, block_d, after_d + after_len, true, &hv)) {
      set_error("initial ranks (K5): fingerprints unavailable");
      throw CudaError{PSC_ECUDA};
    }
  }
}

// ---- K6 ---------------------------------------------------------------------------------------
struct WalkTables {
  uint32_t C[256];     // number of block symbols smaller than c  (compute_gap.hpp:77-85)
  int16_t code[256];   // dense code in the rank structure, -1 = symbol absent from the BWT
};

// How the ranks reach the gap array.  Measured on B200 (1 GiB Markov text, 128 MiB blocks): the
// walk alone runs at 38.7 G steps/s = the random 32-byte sector rate of HBM; a RED per step on a
// 512 MB gap array (one more random sector read + write-back) drops it to 14.6 G steps/s.  So for
// gap arrays larger than L2 the ranks are streamed out, partitioned by their high bits into
// buckets whose counter range (16 MB) stays in L2, and only then counted: 16 B of sequential
// traffic per step instead of 64 B of random traffic.  (The reference partitions its gap buffers
// by value range for another reason -- lock-free CPU updaters, stream.hpp:160-232.)
//   MODE 0: RED on gap[i]   1: warp-aggregated RED (match.any)   2: emit ranks + bucket counts
//   MODE 3: the partition folded into the walk.  Every warp keeps a 32-entry buffer per bucket in shared memory;
//   the ranks of a step go to their buckets (lanes of one bucket found with match.any), a buffer that fills up
//   leaves as one 128-byte PAGE (pages are handed out 16 at a time per warp from one global counter) and a
//   descriptor (bucket << 25 | page) is written for it.  Afterwards only the descriptors are grouped by bucket
//   (1/32 of the elements) and the pages are counted bucket by bucket.  The bookkeeping of a step runs between
//   the issue of the next step's sector load and its use, where the walker would wait anyway.
constexpr int WK_MAX_BUCKETS = 512;
constexpr int WK_PAGE_BUCKETS = 72;   // MODE 3: at most this many buckets (9.3 KB of buffers per warp)
constexpr int WK_PAGE_BATCH = 16;     // pages a warp takes from the global counter at once
constexpr int WK_PAGE_ID_BITS = 25;

struct PagedOut {
  uint32_t *pages;                  // [cap_pages][32]
  uint32_t *desc;                   // [cap_pages]
  unsigned long long *next_page;    // global page counter
  uint32_t cap_pages;
};

// the sector of (code, i): issued early, counted later
__device__ __forceinline__ void rank_occ_issue(const RankView &rv, int code, uint32_t i, uint32_t (&w)[8]) {
  const uint32_t s = i / RK_BITS;
  const uint32_t *p = rv.sectors + ((size_t)code * (size_t)rv.n_sectors + s) * 8;
  asm volatile("ld.global.nc.L1::no_allocate.L2::64B.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7])
               : "l"(p));
}
__device__ __forceinline__ uint32_t rank_occ_finish(uint32_t i, const uint32_t (&w)[8]) {
  const uint32_t r = i % RK_BITS;
  uint32_t cnt = w[0];
#pragma unroll
  for (int k = 0; k < 7; ++k) {
    int nb = (int)r - 32 * k;
    nb = nb < 0 ? 0 : (nb > 32 ? 32 : nb);
    const uint32_t mask = nb == 32 ? 0xffffffffu : ((1u << nb) - 1u);
    cnt += __popc(w[k + 1] & mask);
  }
  return cnt;
}

// MODE 3 bookkeeping of one warp step: `v` = this lane's rank (valid lanes only)
struct WarpPager {
  uint32_t *buf;   // [n_buckets][32] of this warp
  uint32_t *cnt;   // [n_buckets]
  uint32_t next = 0, end = 0;  // the warp's batch of pages (uniform over the lanes)
  __device__ __forceinline__ uint32_t take_page(const PagedOut &po, int lane) {
    if (next == end) {
      unsigned long long p = 0;
      if (lane == 0) p = atomicAdd(po.next_page, (unsigned long long)WK_PAGE_BATCH);
      next = (uint32_t)__shfl_sync(0xffffffffu, p, 0);
      end = next + WK_PAGE_BATCH;
    }
    return next++;
  }
  __device__ __forceinline__ void put(const PagedOut &po, uint32_t *sB, bool valid, uint32_t v, int bshift, int lane) {
    const uint32_t b = valid ? (v >> bshift) : (0x80000000u | (uint32_t)lane);
    const uint32_t peers = __match_any_sync(0xffffffffu, b);
    const int lead = __ffs(peers) - 1;
    uint32_t have = 0;
    if (valid && lane == lead) have = cnt[b];
    have = __shfl_sync(0xffffffffu, have, lead);
    const uint32_t slot = have + (uint32_t)__popc(peers & ((1u << lane) - 1u));
    const uint32_t total = have + (uint32_t)__popc(peers);
    if (valid && slot < 32) buf[b * 32 + slot] = v;
    __syncwarp();
    uint32_t fullm = __ballot_sync(0xffffffffu, valid && lane == lead && total >= 32);
    while (fullm) {
      const int L = __ffs(fullm) - 1;
      fullm &= fullm - 1;
      const uint32_t bb = __shfl_sync(0xffffffffu, b, L);
      const uint32_t page = take_page(po, lane);
      if (page < po.cap_pages) {
        po.pages[(size_t)page * 32 + lane] = buf[bb * 32 + lane];
        if (lane == 0) { po.desc[page] = (bb << WK_PAGE_ID_BITS) | page; atomicAdd(&sB[bb], 1u); }
      }
      __syncwarp();
      if (valid && b == bb && slot >= 32) buf[bb * 32 + slot - 32] = v;  // the entries that did not fit
    }
    if (valid && lane == lead) cnt[b] = total >= 32 ? total - 32 : total;
    __syncwarp();
  }
  // end of the kernel: partly filled buffers leave as padded pages, unused pages of the batch get a descriptor
  // of the extra bucket `n_buckets`
  __device__ __forceinline__ void drain(const PagedOut &po, uint32_t *sB, int n_buckets, int lane) {
    for (int b = 0; b < n_buckets; ++b) {
      const uint32_t c = cnt[b];
      if (c == 0) continue;
      const uint32_t page = take_page(po, lane);
      if (page < po.cap_pages) {
        po.pages[(size_t)page * 32 + lane] = (uint32_t)lane < c ? buf[b * 32 + lane] : 0xffffffffu;
        if (lane == 0) { po.desc[page] = ((uint32_t)b << WK_PAGE_ID_BITS) | page; atomicAdd(&sB[b], 1u); }
      }
    }
    for (uint32_t page = next + lane; page < end; page += 32)
      if (page < po.cap_pages) { po.desc[page] = ((uint32_t)n_buckets << WK_PAGE_ID_BITS) | page; atomicAdd(&sB[n_buckets], 1u); }
  }
};

template <int MODE>
__global__ void __launch_bounds__(256)
    lf_walk_kernel(RankView rv, uint32_t i0, int last, const WalkTables *__restrict__ tables,
                   const uint8_t *__restrict__ tail, long nwalk, const uint64_t *__restrict__ walk_beg,
                   const uint64_t *__restrict__ start_rank, const uint32_t *__restrict__ gt_in,
                   long gt_in_off, uint32_t *__restrict__ gt_out, long gt_out_off,
                   uint32_t *__restrict__ gap, uint32_t *__restrict__ emit, long emit_base, int bshift,
                   int n_buckets, unsigned long long *__restrict__ bucket_count, PagedOut po) {
  extern __shared__ uint32_t dyn_smem[];  // MODE 3: per warp n_buckets * 33 words
  __shared__ uint32_t sC[256];
  __shared__ int16_t sCode[256];
  __shared__ uint32_t sB[MODE == 2 ? WK_MAX_BUCKETS : (MODE == 3 ? WK_PAGE_BUCKETS + 1 : 1)];
  sC[threadIdx.x] = tables->C[threadIdx.x];
  sCode[threadIdx.x] = tables->code[threadIdx.x];
  if (MODE == 2) for (int b = threadIdx.x; b < n_buckets; b += 256) sB[b] = 0;
  const int lane = threadIdx.x & 31;
  WarpPager pager;
  if (MODE == 3) {
    for (int b = threadIdx.x; b <= n_buckets; b += 256) sB[b] = 0;
    pager.buf = dyn_smem + (size_t)(threadIdx.x >> 5) * n_buckets * 33;
    pager.cnt = pager.buf + n_buckets * 32;
    for (int b = lane; b < n_buckets; b += 32) pager.cnt[b] = 0;
  }
  __syncthreads();
  long w = (long)blockIdx.x * blockDim.x + threadIdx.x;
  long s = 0, e = 0;
  uint32_t i = 0;
  if (w < nwalk) { s = (long)walk_beg[w]; e = (long)walk_beg[w + 1]; i = (uint32_t)start_rank[w]; }
  long steps = e - s, max_steps = steps;
#pragma unroll
  for (int o = 16; o; o >>= 1) max_steps = max(max_steps, __shfl_xor_sync(0xffffffffu, max_steps, o));

  long j = e;                      // suffix tail+j is the one whose rank `i` currently holds
  uintptr_t cur_chunk = ~(uintptr_t)0;
  uint4 tb = make_uint4(0, 0, 0, 0);
  long cur_in_w = -1, cur_out_w = -1;
  uint32_t in_word = 0, out_word = 0;
  uint32_t prev = 0;
  bool prev_valid = false;
  for (long k = 0; k < max_steps; ++k) {
    const bool active = k < steps;
    if (MODE == 3) {
      // same step as below, with the previous step's rank put away while the sector is on its way
      uint32_t w8[8];
      int c = 0, code = -1;
      uint32_t gin = 0;
      if (active) {
        uintptr_t a = (uintptr_t)(tail + j - 1);
        uintptr_t ch = a & ~(uintptr_t)15;
        if (ch != cur_chunk) { tb = *reinterpret_cast<const uint4 *>(ch); cur_chunk = ch; }
        uint32_t b = (uint32_t)(a & 15);
        uint32_t word = (b & 8) ? ((b & 4) ? tb.w : tb.z) : ((b & 4) ? tb.y : tb.x);
        c = (int)((word >> ((b & 3) * 8)) & 255u);
        long ob = gt_out_off + j - 1;
        if (gt_out) {
          if ((ob >> 5) != cur_out_w) {
            if (cur_out_w >= 0 && out_word) atomicOr(&gt_out[cur_out_w], out_word);
            cur_out_w = ob >> 5;
            out_word = 0;
          }
          if (i > i0) out_word |= 1u << (ob & 31);
        }
        long ib = gt_in_off + j - 1;
        if ((ib >> 5) != cur_in_w) { cur_in_w = ib >> 5; in_word = gt_in ? gt_in[cur_in_w] : 0u; }
        gin = (in_word >> (ib & 31)) & 1u;
        code = sCode[c];
        if (code >= 0) rank_occ_issue(rv, code, i, w8);
      }
      pager.put(po, sB, prev_valid, prev, bshift, lane);
      if (active) {
        const uint32_t occ = code >= 0 ? rank_occ_finish(i, w8) : 0u;
        i = sC[c] + occ + ((c == last) ? gin : 0u);
        --j;
      }
      prev = i;
      prev_valid = active;
      continue;
    }
    if (active) {
      // symbol tail[j-1]
      uintptr_t a = (uintptr_t)(tail + j - 1);
      uintptr_t ch = a & ~(uintptr_t)15;
      if (ch != cur_chunk) { tb = *reinterpret_cast<const uint4 *>(ch); cur_chunk = ch; }
      uint32_t b = (uint32_t)(a & 15);
      uint32_t word = (b & 8) ? ((b & 4) ? tb.w : tb.z) : ((b & 4) ? tb.y : tb.x);
      int c = (int)((word >> ((b & 3) * 8)) & 255u);
      // gt bits of suffix j: new one out (stream.hpp:150), old one in (:151)
      long ob = gt_out_off + j - 1;
      if (gt_out) {
        if ((ob >> 5) != cur_out_w) {
          if (cur_out_w >= 0 && out_word) atomicOr(&gt_out[cur_out_w], out_word);
          cur_out_w = ob >> 5;
          out_word = 0;
        }
        if (i > i0) out_word |= 1u << (ob & 31);
      }
      long ib = gt_in_off + j - 1;
      if ((ib >> 5) != cur_in_w) { cur_in_w = ib >> 5; in_word = gt_in ? gt_in[cur_in_w] : 0u; }
      uint32_t gin = (in_word >> (ib & 31)) & 1u;
      // LF step (:153-155; no delta: the placeholder has no bit in any plane)
      int code = sCode[c];
      uint32_t occ = code >= 0 ? rank_occ(rv, code, i) : 0u;
      i = sC[c] + occ + ((c == last) ? gin : 0u);
      --j;
    }
    if (MODE == 2) {
      if (active) {
        emit[j - emit_base] = i;  // the rank of suffix tail+j, at the slot of position j
        atomicAdd(&sB[i >> bshift], 1u);
      }
    } else if (MODE == 1) {
      uint32_t peers = __match_any_sync(0xffffffffu, active ? i : (0xffffffffu - lane));
      if (active && lane == __ffs(peers) - 1) atomicAdd(&gap[i], (uint32_t)__popc(peers));
    } else {
      if (active) atomicAdd(&gap[i], 1u);
    }
  }
  if (gt_out && cur_out_w >= 0 && out_word) atomicOr(&gt_out[cur_out_w], out_word);
  if (MODE == 3) {
    pager.put(po, sB, prev_valid, prev, bshift, lane);
    pager.drain(po, sB, n_buckets, lane);
    __syncthreads();
    for (int b = threadIdx.x; b <= n_buckets; b += 256)
      if (sB[b]) atomicAdd(&bucket_count[b], (unsigned long long)sB[b]);
  }
  if (MODE == 2) {
    __syncthreads();
    for (int b = threadIdx.x; b < n_buckets; b += 256)
      if (sB[b]) atomicAdd(&bucket_count[b], (unsigned long long)sB[b]);
  }
}

// pages grouped by bucket (descriptors in bucket order): one warp per page, padding entries skipped
__global__ void __launch_bounds__(256)
    page_accumulate_kernel(const uint32_t *__restrict__ desc, long n_pages, const uint32_t *__restrict__ pages,
                           uint32_t *__restrict__ gap) {
  const long w = ((long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  if (w >= n_pages) return;
  const uint32_t page = desc[w] & ((1u << WK_PAGE_ID_BITS) - 1u);
  const uint32_t v = pages[(size_t)page * 32 + (threadIdx.x & 31)];
  if (v != 0xffffffffu) atomicAdd(&gap[v], 1u);
}

// ranks -> buckets (order inside a bucket is irrelevant: they are only counted afterwards)
constexpr int PT_THREADS = 512;
constexpr int PT_ITEMS = 16;
constexpr int PT_TILE = PT_THREADS * PT_ITEMS;  // 8192 ranks, 32 KB of staging

__global__ void __launch_bounds__(PT_THREADS)
    rank_partition_kernel(const uint32_t *__restrict__ in, long n, int bshift, int n_buckets,
                          unsigned long long *__restrict__ cursor, uint32_t *__restrict__ out) {
  __shared__ uint32_t cnt[WK_MAX_BUCKETS];    // count, then running slot inside the tile
  __shared__ uint32_t toff[WK_MAX_BUCKETS];   // first slot of the bucket inside the tile
  __shared__ unsigned long long gofs[WK_MAX_BUCKETS];
  __shared__ uint32_t stage[PT_TILE];
  __shared__ uint32_t wsum[PT_THREADS / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long base = (long)blockIdx.x * PT_TILE;
  const int valid = (int)min((long)PT_TILE, n - base);
  for (int b = tid; b < n_buckets; b += PT_THREADS) cnt[b] = 0;
  __syncthreads();
  uint32_t v[PT_ITEMS];
#pragma unroll
  for (int k = 0; k < PT_ITEMS; ++k) {
    int e = tid + k * PT_THREADS;
    v[k] = e < valid ? in[base + e] : 0xffffffffu;
    if (e < valid) atomicAdd(&cnt[v[k] >> bshift], 1u);
  }
  __syncthreads();
  // exclusive scan of the bucket counts (one bucket per thread), reserve global space
  static_assert(WK_MAX_BUCKETS <= PT_THREADS, "one bucket per thread");
  {
    uint32_t c = tid < n_buckets ? cnt[tid] : 0u, incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    if (lane == 31) wsum[warp] = incl;
    __syncthreads();
    uint32_t woff = 0;
    for (int w = 0; w < warp; ++w) woff += wsum[w];
    uint32_t ex = woff + incl - c;
    if (tid < n_buckets) {
      toff[tid] = ex; cnt[tid] = ex;
      if (c) gofs[tid] = atomicAdd(&cursor[tid], (unsigned long long)c);
    }
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < PT_ITEMS; ++k) {
    int e = tid + k * PT_THREADS;
    if (e < valid) stage[atomicAdd(&cnt[v[k] >> bshift], 1u)] = v[k];
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < PT_ITEMS; ++k) {
    int e = tid + k * PT_THREADS;
    if (e < valid) {
      uint32_t x = stage[e];
      int b = (int)(x >> bshift);
      out[gofs[b] + (unsigned long long)(e - toff[b])] = x;
    }
  }
}

__global__ void __launch_bounds__(256)
    rank_accumulate_kernel(const uint32_t *__restrict__ in, long n, uint32_t *__restrict__ gap) {
  long i = ((long)blockIdx.x * blockDim.x + threadIdx.x) * 4;
  if (i + 4 <= n) {
    uint4 x = *reinterpret_cast<const uint4 *>(in + i);
    atomicAdd(&gap[x.x], 1u); atomicAdd(&gap[x.y], 1u); atomicAdd(&gap[x.z], 1u); atomicAdd(&gap[x.w], 1u);
  } else {
    for (; i < n; ++i) atomicAdd(&gap[in[i]], 1u);
  }
}

void lf_walk(psc_context *ctx, const psc_rank *r, uint32_t i0, uint8_t last, const uint64_t *Cmap,
             const uint8_t *tail_d, long t, long lo, long nwalk, const uint64_t *walk_beg_d, const uint64_t *start_rank_d, const uint32_t *gt_in_bits_d,
             long gt_in_off, uint32_t *gt_out_bits_d, long gt_out_off, uint32_t *gap_d) {
  PhaseTimer timer(ctx, &ctx->stats.ms_walk);
  if (nwalk <= 0 || t <= 0) return;
  WalkTables tb;
  for (int c = 0; c < 256; ++c) { tb.C[c] = (uint32_t)Cmap[c]; tb.code[c] = r->code_of[c]; }
  DevBuf<WalkTables> tb_d(ctx, 1);
  h2d(ctx, tb_d.p, &tb, 1);
  RankView rv{r->sectors_d, r->n_sectors, (uint32_t)r->m};
  static int mode = -2;
  if (mode == -2) { const char *e = getenv("PSC_WALK_MODE"); mode = e ? atoi(e) : -1; }
  const long hi = lo + t;  // the walkers cover tail positions [lo, hi)
  int use = mode;
  if (use < 0) use = ((size_t)(r->m + 1) * 4 > (size_t)(24 << 20) && hi - lo >= (1 << 20)) ? 2 : 0;
  unsigned blocks = (unsigned)div_up(nwalk, 256);
  // device time of the walk kernel alone: every branch below synchronises after it anyway
  struct KernelClock {
    psc_context *c; cudaEvent_t a = nullptr, b = nullptr;
    explicit KernelClock(psc_context *ctx) : c(ctx) { cudaEventCreate(&a); cudaEventCreate(&b); }
    void start() { cudaEventRecord(a, c->stream); }
    void stop() { cudaEventRecord(b, c->stream); }
    ~KernelClock() {
      float ms = 0;
      if (cudaEventSynchronize(b) == cudaSuccess && cudaEventElapsedTime(&ms, a, b) == cudaSuccess) c->stats.ms_walk_kernel += ms;
      else cudaGetLastError();
      cudaEventDestroy(a); cudaEventDestroy(b);
    }
  } kclock(ctx);
  if (use == 2 && mode < 0 && hi - lo <= (1L << 29)) {
    // partition folded into the walk when the gap array splits into few enough L2-sized buckets
    int bs = 22;
    while (bs < 24 && ((r->m + 1) >> bs) + 1 > WK_PAGE_BUCKETS) ++bs;
    if (((r->m + 1) >> bs) + 1 <= WK_PAGE_BUCKETS) use = 3;
  }
  if (use == 3) {
    const long T = hi - lo;
    int bshift = 22;
    while (((r->m + 1) >> bshift) + 1 > WK_PAGE_BUCKETS) ++bshift;
    const int nb = (int)(((r->m + 1) >> bshift) + 1);
    const long n_warps = (long)blocks * 8;
    // every entry leaves in exactly one page: full pages, one padded page per warp and bucket at the end, and
    // the unused rest of every warp's last batch
    const long cap_pages = T / 32 + n_warps * (nb + WK_PAGE_BATCH) + WK_PAGE_BATCH;
    if (cap_pages >= (1L << WK_PAGE_ID_BITS)) { set_error("lf_walk: %ld pages do not fit the descriptor", cap_pages); throw CudaError{PSC_ECUDA}; }
    long scratch = 1L << 24;  // size classes as in mode 2 (see there)
    while (scratch < cap_pages * 32) scratch <<= 1;
    DevBuf<uint32_t> pages(ctx, scratch), desc(ctx, scratch / 32), sorted(ctx, scratch / 32);
    DevBuf<unsigned long long> bc(ctx, WK_MAX_BUCKETS + 1), cur(ctx, WK_MAX_BUCKETS);
    bc.zero();
    PagedOut po{pages.p, desc.p, bc.p + WK_MAX_BUCKETS, (uint32_t)cap_pages};
    const size_t smem = (size_t)8 * nb * 33 * 4;
    PSC_CUDA(cudaFuncSetAttribute(lf_walk_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    kclock.start();
    { KTrace kt(ctx, "lf_walk(paged)"); lf_walk_kernel<3><<<blocks, 256, smem, ctx->stream>>>(rv, i0, last, tb_d.p, tail_d, nwalk, walk_beg_d, start_rank_d,
                                                        gt_in_bits_d, gt_in_off, gt_out_bits_d, gt_out_off, gap_d,
                                                        nullptr, lo, bshift, nb, bc.p, po); }
    kclock.stop();
    PSC_LAUNCH_CHECK();
    unsigned long long h[WK_MAX_BUCKETS + 1], c[WK_MAX_BUCKETS];
    d2h(ctx, h, bc.p, WK_MAX_BUCKETS + 1);
    sync(ctx);
    const long n_taken = (long)h[WK_MAX_BUCKETS];
    if (n_taken > cap_pages) { set_error("lf_walk: %ld pages taken, %ld reserved", n_taken, cap_pages); throw CudaError{PSC_ECUDA}; }
    unsigned long long run = 0;
    for (int b = 0; b <= nb; ++b) { c[b] = run; run += h[b]; }
    if ((long)run != n_taken) { set_error("lf_walk: %llu page descriptors for %ld pages", run, n_taken); throw CudaError{PSC_ECUDA}; }
    h2d(ctx, cur.p, c, nb + 1);
    { KTrace kt(ctx, "page_partition"); rank_partition_kernel<<<(unsigned)div_up(n_taken, PT_TILE), PT_THREADS, 0, ctx->stream>>>(desc.p, n_taken, WK_PAGE_ID_BITS, nb + 1, cur.p, sorted.p); }
    PSC_LAUNCH_CHECK();
    const long n_real = (long)c[nb];  // the descriptors of unused pages come last
    { KTrace kt(ctx, "page_accumulate"); page_accumulate_kernel<<<(unsigned)div_up(n_real * 32, 256), 256, 0, ctx->stream>>>(sorted.p, n_real, pages.p, gap_d); }
    PSC_LAUNCH_CHECK();
    sync(ctx);
    ctx->launches += 3;
  } else if (use == 2) {
    const long T = hi - lo;
    int bshift = 22;  // 4 Mi counters = 16 MB of gap array per bucket
    while (((r->m + 1) >> bshift) + 1 > WK_MAX_BUCKETS) ++bshift;
    const int nb = (int)(((r->m + 1) >> bshift) + 1);
    // scratch sized to the next power of two (at least 2^24): the last chunk of every tail has its own length,
    // and a fresh pair of up to 2 GB buffers per length cost more than the walk on multi-GiB texts (6 GiB of
    // DNA, 11 blocks: 111 GB of cached buffers, trimmed and reallocated under memory pressure)
    long scratch = 1L << 24;
    while (scratch < T) scratch <<= 1;
    DevBuf<uint32_t> emit(ctx, scratch), part(ctx, scratch);
    DevBuf<unsigned long long> bc(ctx, WK_MAX_BUCKETS), cur(ctx, WK_MAX_BUCKETS);
    bc.zero();
    kclock.start();
    { KTrace kt(ctx, "lf_walk(emit)"); lf_walk_kernel<2><<<blocks, 256, 0, ctx->stream>>>(rv, i0, last, tb_d.p, tail_d, nwalk, walk_beg_d, start_rank_d,
                                                        gt_in_bits_d, gt_in_off, gt_out_bits_d, gt_out_off, gap_d,
                                                        emit.p, lo, bshift, nb, bc.p, PagedOut{}); }
    kclock.stop();
    PSC_LAUNCH_CHECK();
    unsigned long long h[WK_MAX_BUCKETS], c[WK_MAX_BUCKETS];
    d2h(ctx, h, bc.p, nb);
    sync(ctx);
    unsigned long long run = 0;
    for (int b = 0; b < nb; ++b) { c[b] = run; run += h[b]; }
    h2d(ctx, cur.p, c, nb);
    { KTrace kt(ctx, "rank_partition"); rank_partition_kernel<<<(unsigned)div_up(T, PT_TILE), PT_THREADS, 0, ctx->stream>>>(emit.p, T, bshift, nb, cur.p,
                                                                                      part.p); }
    PSC_LAUNCH_CHECK();
    { KTrace kt(ctx, "rank_accumulate"); rank_accumulate_kernel<<<(unsigned)div_up(div_up(T, 4), 256), 256, 0, ctx->stream>>>(part.p, T, gap_d); }
    PSC_LAUNCH_CHECK();
    sync(ctx);
    ctx->launches += 3;
  } else {
    kclock.start();
    if (use == 1)
      lf_walk_kernel<1><<<blocks, 256, 0, ctx->stream>>>(rv, i0, last, tb_d.p, tail_d, nwalk, walk_beg_d, start_rank_d,
                                                          gt_in_bits_d, gt_in_off, gt_out_bits_d, gt_out_off, gap_d,
                                                          nullptr, 0, 0, 0, nullptr, PagedOut{});
    else
      { KTrace kt(ctx, "lf_walk(red)"); lf_walk_kernel<0><<<blocks, 256, 0, ctx->stream>>>(rv, i0, last, tb_d.p, tail_d, nwalk, walk_beg_d, start_rank_d,
                                                          gt_in_bits_d, gt_in_off, gt_out_bits_d, gt_out_off, gap_d,
                                                          nullptr, 0, 0, 0, nullptr, PagedOut{}); }
    kclock.stop();
    PSC_LAUNCH_CHECK();
    sync(ctx);  // tb is a stack object
    ctx->launches++;
  }
  ctx->stats.lf_steps += t;
  ctx->stats.walkers = nwalk;
}

// Walkers for tail positions [lo, hi): chunks of at most PSC_WALK_CHUNK positions (2^29: 4 GB of rank-stream
// scratch whatever the tail length), each chunk cut into equal walker slices with boundaries on multiples of
// 32 from the chunk start (compute_gap.hpp:68-69 cuts the tail over threads the same way).  Walker k covers
// [b[k], b[k+1]) and starts from the rank of suffix(after + b[k+1]).
void make_walk_plan(psc_context *ctx, long lo, long hi, WalkPlan *p) {
  static long chunk = 0, target_env = -1;
  if (!chunk) {
    const char *e = getenv("PSC_WALK_CHUNK");
    chunk = e ? std::max(64L, atol(e) / 32 * 32) : (1L << 29);
    const char *w = getenv("PSC_WALKERS");
    target_env = w ? std::max(1L, atol(w)) : 0;
  }
  const long target = target_env > 0 ? target_env : (long)ctx->n_sm * ctx->walk_threads_per_sm;
  p->b.clear();
  p->chunk_at.clear();
  for (long clo = lo; clo < hi; clo += chunk) {
    const long chi = std::min(hi, clo + chunk);
    const long slice = std::max(64L, div_up(div_up(chi - clo, target), 32) * 32);
    const long W = div_up(chi - clo, slice);
    p->chunk_at.push_back((long)p->b.size());
    for (long k = 0; k < W; ++k) p->b.push_back((uint64_t)(clo + k * slice));
  }
  p->chunk_at.push_back((long)p->b.size());
  p->b.push_back((uint64_t)hi);
}

// K6 over a plan: b_d = the plan's boundaries on the device (W + 1), sr_d[k] = start rank of walker k
void walk_with_plan(psc_context *ctx, const psc_rank *rk, long i0, uint8_t last, const uint64_t *Cmap,
                    const uint8_t *after_d, const WalkPlan &p, const uint64_t *b_d, const uint64_t *sr_d,
                    const uint32_t *gt_cur, long gt_off, uint32_t *gt_new, uint32_t *gap_d) {
  for (size_t c = 0; c + 1 < p.chunk_at.size(); ++c) {
    const long k0 = p.chunk_at[c], k1 = p.chunk_at[c + 1];
    if (k1 <= k0) continue;
    const long clo = (long)p.b[k0], chi = (long)p.b[k1];
    lf_walk(ctx, rk, (uint32_t)i0, last, Cmap, after_d, chi - clo, clo, k1 - k0, b_d + k0, sr_d + k0, gt_cur, gt_off,
            gt_new, gt_off, gap_d);
  }
}

// Stages 2+3 for tail positions [lo, hi) of the text after a block: plan, initial ranks by binary search
// over the block SA (K5), then the walk (K6).
void walk_tail_range(psc_context *ctx, const psc_rank *rk, long i0, uint8_t last, const uint64_t *Cmap,
                     const uint8_t *block_d, const uint32_t *sa_d, long m, const uint8_t *after_d, long t,
                     long lo, long hi, const uint32_t *gt_cur, long gt_off, uint32_t *gt_new, uint32_t *gap_d) {
  if (hi <= lo) return;
  WalkPlan p;
  make_walk_plan(ctx, lo, hi, &p);
  const long W = (long)p.b.size() - 1;
  DevBuf<uint64_t> b_d(ctx, W + 1), sr_d(ctx, W);
  h2d(ctx, b_d.p, p.b.data(), p.b.size());
  // walker k starts from the rank of suffix(after + b[k+1]); the one ending at t from suffix(n): 0
  initial_ranks(ctx, block_d, sa_d, m, after_d, t, 1, gt_cur, gt_off, W, b_d.p + 1, sr_d.p);
  walk_with_plan(ctx, rk, i0, last, Cmap, after_d, p, b_d.p, sr_d.p, gt_cur, gt_off, gt_new, gap_d);
  sync(ctx);  // p.b is a host vector the async copy read
}

}  // namespace psc
