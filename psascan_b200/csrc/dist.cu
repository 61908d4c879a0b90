// dist.cu -- the block pipeline over several GPUs, one process per GPU.
//
//   phase 1  block sorts are independent once gt_end is computed by plain text comparison past the
//            block end (no gt bits of the next block needed), so rank r sorts blocks B-1-r, B-1-r-N, ...
//   phase 2  blocks right to left: the owner broadcasts SA, BWT and gt_begin bits of the block;
//            every rank builds the rank structure, walks ITS contiguous share of the tail
//            (compute_gap.hpp:68-69 splits the tail over threads the same way), and the per-GPU
//            gap counts and gt bits are summed over NCCL.  The merge is sharded by OUTPUT range:
//            after every level rank g holds slots [g*L/G, (g+1)*L/G) of the merged sequence (L = its
//            length so far); the entries of the previous level it needs form one contiguous range,
//            known to every rank from the scanned gap array alone, and arrive by one all-to-all.
//            Nothing of size n lives on a GPU except the text: planes and output are n/G per rank.
// Collectives are delegated to the caller through psc_collective_fn (torch.distributed in
// psascan_b200/dist.py, NCCL over NVLink); the library synchronises its stream around them.
#include <algorithm>
#include <condition_variable>
#include <deque>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include "common.cuh"

namespace psc {

__global__ void or_bits_shifted_kernel(uint32_t *dst, long dst_off, const uint32_t *src, long nbits) {
  long w = (long)blockIdx.x * blockDim.x + threadIdx.x;
  long nw = (nbits + 31) / 32;
  if (w >= nw) return;
  uint32_t word = src[w];
  long rem = nbits - w * 32;
  if (rem < 32) word &= (1u << rem) - 1u;
  if (!word) return;
  long pos = dst_off + w * 32;
  int sh = (int)(pos & 31);
  atomicOr(&dst[pos >> 5], word << sh);
  if (sh && (word >> (32 - sh))) atomicOr(&dst[(pos >> 5) + 1], word >> (32 - sh));
}

__global__ void set_bit_kernel(uint32_t *dst, long idx) { atomicOr(&dst[idx >> 5], 1u << (idx & 31)); }

struct OwnedBlock {
  DevBuf<uint32_t> sa;
  DevBuf<uint8_t> bwt;
  DevBuf<uint32_t> gtb;  // bit q <-> block position q+1, q in [0, m-1)
  long i0 = 0;
  int end_bit = 0;       // gt bit of position `end` relative to the block begin
  uint8_t last = 0;      // text[end - 1]
};

struct BlockMeta { long long i0; int end_bit; int last; };

// K0 (by plain comparison past the block end) + block sort + gt_begin bits + BWT of text[beg, end):
// everything a block contributes that does not depend on the blocks to its right.
static std::unique_ptr<OwnedBlock> sort_owned_block(psc_context *ctx, const uint8_t *text_d, long n, long beg,
                                                    long end) {
  const long m = end - beg, t = n - end;
  auto ob = std::make_unique<OwnedBlock>();
  BlockSortResult res;
  DevBuf<uint32_t> gt_end;
  if (t > 0) {
    gt_end.alloc(ctx, m / 32 + 2);
    compute_gt_end(ctx, text_d + beg, m, text_d + end, t, 1, nullptr, 0, gt_end.p);
  }
  sort_block(ctx, text_d + beg, m, t > 0 ? gt_end.p : nullptr, &res);
  res.isa.release();
  ob->i0 = res.i0;
  ob->gtb.alloc(ctx, m / 32 + 2);
  ob->gtb.zero();
  gt_begin_from_sa(ctx, res.sa.p, m, (uint32_t)res.i0, ob->gtb.p, 0);
  uint32_t w0 = 1;  // gt_end[0]: 1 for the last block
  if (t > 0) { d2h(ctx, &w0, gt_end.p, 1); sync(ctx); }
  ob->end_bit = (w0 & 1u) ? 0 : 1;  // !gt_end[beg]  (change_gt_reference_point.hpp)
  if (t > 0) {
    ob->bwt.alloc(ctx, m);
    bwt_from_sa(ctx, text_d + beg, res.sa.p, m, ob->bwt.p);
  }
  PSC_CUDA(cudaMemcpyAsync(&ob->last, text_d + end - 1, 1, cudaMemcpyDeviceToHost, ctx->stream));
  ob->sa = std::move(res.sa);
  sync(ctx);
  return ob;
}

static void coll(psc_context *ctx, psc_collective_fn fn, void *user, int op, void *buf, size_t bytes, int root) {
  sync(ctx);
  if (fn(user, op, buf, bytes, root) != 0) {
    set_error("collective callback failed (op %d, %zu bytes)", op, bytes);
    throw CudaError{PSC_ECUDA};
  }
}

// first output slot of rank g when `len` slots are shared out: equal shares rounded up to a multiple of 4
// (the 40-bit writer packs 4 entries into 5 aligned words)
static long share_begin(long g, long len, int world) {
  long per = (div_up(len, (long)world) + 3) / 4 * 4;
  return std::min(len, per * g);
}
long share_begin_public(long g, long len, int world) { return share_begin(g, len, world); }
static size_t overlap(long a0, long a1, long b0, long b1) {
  long lo = std::max(a0, b0), hi = std::min(a1, b1);
  return hi > lo ? (size_t)(hi - lo) : 0;
}

// gather == true: out_sa5_d[5n] receives the whole SA on every rank (slices broadcast at the end);
// else out_sa5_d receives this rank's slice only: entries [*slice_beg, *slice_beg + *slice_len).
void construct_sa_dist(psc_context *ctx, const uint8_t *text_d, long n, long max_block, uint8_t *out_sa5_d,
                       int rank, int world, psc_collective_fn fn, void *user, bool gather, long *slice_beg,
                       long *slice_len) {
  memset(&ctx->stats, 0, sizeof(ctx->stats));
  ctx->launches = 0;
  if (n == 0) return;
  PhaseTimer total(ctx, &ctx->stats.ms_total);
  max_block = std::min(max_block, (1L << 31) - 4096);
  const long B = div_up(n, max_block);
  ctx->stats.n_blocks = B;
  auto owner = [&](long id) { return (int)((B - 1 - id) % world); };
  auto bounds = [&](long id, long *beg, long *end) { *beg = id * max_block; *end = std::min(n, *beg + max_block); };

  // ---- phase 1: my blocks, independent of each other
  std::vector<std::unique_ptr<OwnedBlock>> mine(B);
  for (long id = B - 1; id >= 0; --id) {
    if (owner(id) != rank) continue;
    long beg, end;
    bounds(id, &beg, &end);
    mine[id] = sort_owned_block(ctx, text_d, n, beg, end);
  }

  // ---- phase 2: right to left
  const long gwords = (n + 1 + 31) / 32 + 2;
  DevBuf<uint32_t> Ga(ctx, gwords), Gb(ctx, gwords);
  Ga.zero(); Gb.zero();
  uint32_t *gt_cur = Ga.p, *gt_new = Gb.p;
  // my share of the merged sequence (double buffered) and what arrives from the other ranks per level
  const bool hi_plane = need_hi_plane(n);  // high bytes only above 4 GiB
  const long cap = share_begin(1, n, world) + 16;
  DevBuf<uint32_t> Mlo[2], Rlo;
  DevBuf<uint8_t> Mhi[2], Rhi;
  for (int k = 0; k < 2; ++k) { Mlo[k].alloc(ctx, cap); if (hi_plane) Mhi[k].alloc(ctx, cap); }
  if (world > 1) { Rlo.alloc(ctx, cap); if (hi_plane) Rhi.alloc(ctx, cap); }
  int cur = 0;
  long right_len = 0;
  for (long id = B - 1; id >= 0; --id) {
    long beg, end;
    bounds(id, &beg, &end);
    const long m = end - beg, t = n - end;
    const int own = owner(id);
    OwnedBlock *ob = mine[id].get();
    // one broadcast per block: [meta | SA | gt_begin bits | BWT] packed into one buffer (every collective
    // costs a stream synchronisation and a host round trip)
    DevBuf<uint8_t> pack;
    uint32_t *sa_p, *gtb_p;
    uint8_t *bwt_p = nullptr;
    BlockMeta meta{0, 0, 0};
    if (world == 1) {
      sa_p = ob->sa.p; gtb_p = ob->gtb.p; bwt_p = ob->bwt.p;
      meta = BlockMeta{ob->i0, ob->end_bit, (int)ob->last};
    } else {
      const size_t sa_off = 256, gtb_off = sa_off + (((size_t)m * 4 + 255) & ~(size_t)255);
      const size_t gtb_bytes = (size_t)(m / 32 + 2) * 4;
      const size_t bwt_off = gtb_off + ((gtb_bytes + 255) & ~(size_t)255);
      const size_t total = bwt_off + (t > 0 ? (size_t)m : 0);
      pack.alloc(ctx, total);
      sa_p = reinterpret_cast<uint32_t *>(pack.p + sa_off);
      gtb_p = reinterpret_cast<uint32_t *>(pack.p + gtb_off);
      if (t > 0) bwt_p = pack.p + bwt_off;
      if (own == rank) {
        meta = BlockMeta{ob->i0, ob->end_bit, (int)ob->last};
        h2d(ctx, reinterpret_cast<BlockMeta *>(pack.p), &meta, 1);
        PSC_CUDA(cudaMemcpyAsync(sa_p, ob->sa.p, (size_t)m * 4, cudaMemcpyDeviceToDevice, ctx->stream));
        PSC_CUDA(cudaMemcpyAsync(gtb_p, ob->gtb.p, gtb_bytes, cudaMemcpyDeviceToDevice, ctx->stream));
        if (t > 0) PSC_CUDA(cudaMemcpyAsync(bwt_p, ob->bwt.p, (size_t)m, cudaMemcpyDeviceToDevice, ctx->stream));
        sync(ctx);          // `meta` is a stack variable
        mine[id].reset();   // the packed copy is what the rest of the level uses
      }
      coll(ctx, fn, user, PSC_COLL_BCAST, pack.p, total, own);
      d2h(ctx, &meta, reinterpret_cast<BlockMeta *>(pack.p), 1);
      sync(ctx);
    }
    const long i0 = (long)meta.i0;
    PSC_CUDA(cudaMemsetAsync(gt_new, 0, gwords * 4, ctx->stream));
    DevBuf<uint32_t> gap;
    if (t > 0) {
      gap.alloc(ctx, m + 1 + (world > 1 ? (n >> 5) + 2 - ((end - 1) >> 5) : 0));
      gap.zero();
      psc_rank *rk = rank_build(ctx, bwt_p, m, i0);
      try {
        uint64_t Cmap[256], run = 0;
        for (int c = 0; c < 256; ++c) { Cmap[c] = run; run += (uint64_t)rk->counts[c] + (c == meta.last ? 1 : 0); }
        // my contiguous share of the tail, boundaries on multiples of 32
        long per = div_up(div_up(t, (long)world), 32) * 32;
        long lo = std::min(t, per * rank), hi = std::min(t, per * (rank + 1));
        if (hi > lo)
          walk_tail_range(ctx, rk, i0, (uint8_t)meta.last, Cmap, text_d + beg, sa_p, m, text_d + end, t, lo, hi, gt_cur,
                          end, id > 0 ? gt_new : nullptr, gap.p);
      } catch (...) { rank_free(rk); throw; }
      rank_free(rk);
      if (world > 1) {
        // gap counts and (disjoint bits of disjoint tail ranges: SUM == OR) the new gt bits in ONE all-reduce:
        // the gt words ride behind the gap array
        const long w0 = (end - 1) >> 5, w1 = (n >> 5) + 1, gw = id > 0 ? w1 - w0 : 0;
        if (gw) PSC_CUDA(cudaMemcpyAsync(gap.p + m + 1, gt_new + w0, (size_t)gw * 4, cudaMemcpyDeviceToDevice, ctx->stream));
        coll(ctx, fn, user, PSC_COLL_ALLREDUCE_SUM_U32, gap.p, (size_t)(m + 1 + gw) * 4, 0);
        if (gw) PSC_CUDA(cudaMemcpyAsync(gt_new + w0, gap.p + m + 1, (size_t)gw * 4, cudaMemcpyDeviceToDevice, ctx->stream));
      }
    }
    if (id > 0) {
      if (m > 1) {
        or_bits_shifted_kernel<<<(unsigned)div_up(div_up(m - 1, 32), 256), 256, 0, ctx->stream>>>(gt_new, beg, gtb_p, m - 1);
        PSC_LAUNCH_CHECK();
      }
      if (meta.end_bit) { set_bit_kernel<<<1, 1, 0, ctx->stream>>>(gt_new, end - 1); PSC_LAUNCH_CHECK(); }
      ctx->launches += 2;
    }
    int nxt = 1 - cur;
    {
      const long R = right_len, L = R + m;
      const long a = share_begin(rank, L, world), b = share_begin(rank + 1, L, world);
      if (R == 0) {
        sa_slice_to_planes(ctx, sa_p + a, b - a, beg, Mlo[nxt].p, Mhi[nxt].p);
      } else {
        DevBuf<uint64_t> E(ctx, m + 1);
        merge_gap_scan(ctx, gap.p, m, E.p);
        if (world == 1) {
          merge_level_range(ctx, sa_p, E.p, m, beg, Mlo[cur].p, Mhi[cur].p, 0, 0, L, Mlo[nxt].p, Mhi[nxt].p);
        } else {
          // output slot Q[g] is preceded by cnt[g] block suffixes, hence by ra[g] = Q[g] - cnt[g] entries of the
          // right side: rank g needs entries [ra[g], ra[g+1]) of the previous level, held in shares P[.]
          std::vector<uint64_t> Q(world + 1), cnt(world + 1);
          std::vector<long> ra(world + 1), P(world + 1);
          for (int g = 0; g <= world; ++g) { Q[g] = (uint64_t)share_begin(g, L, world); P[g] = share_begin(g, R, world); }
          merge_bounds(ctx, E.p, m, Q.data(), world + 1, cnt.data());
          for (int g = 0; g <= world; ++g) ra[g] = (long)(Q[g] - cnt[g]);
          std::vector<size_t> sb(world), rb(world), sb4(world), rb4(world);
          size_t rtot = 0;
          for (int h = 0; h < world; ++h) {
            sb[h] = overlap(P[rank], P[rank + 1], ra[h], ra[h + 1]);
            rb[h] = overlap(P[h], P[h + 1], ra[rank], ra[rank + 1]);
            sb4[h] = sb[h] * 4; rb4[h] = rb[h] * 4;
            rtot += rb[h];
          }
          if ((long)rtot != ra[rank + 1] - ra[rank] || (long)rtot > cap) { set_error("sharded merge: inconsistent exchange plan"); throw CudaError{PSC_EARG}; }
          psc_alltoallv_args x{Mlo[cur].p, Rlo.p, sb4.data(), rb4.data()};
          coll(ctx, fn, user, PSC_COLL_ALLTOALLV, &x, rtot * 4, 0);
          if (hi_plane) {
            psc_alltoallv_args y{Mhi[cur].p, Rhi.p, sb.data(), rb.data()};
            coll(ctx, fn, user, PSC_COLL_ALLTOALLV, &y, rtot, 0);
          }
          merge_level_range(ctx, sa_p, E.p, m, beg, Rlo.p, Rhi.p, ra[rank], a, b, Mlo[nxt].p, Mhi[nxt].p);
        }
      }
    }
    cur = nxt;
    right_len += m;
    std::swap(gt_cur, gt_new);
    sync(ctx);
    mine[id].reset();
  }
  {
    const long a = share_begin(rank, n, world), b = share_begin(rank + 1, n, world);
    if (slice_beg) *slice_beg = a;
    if (slice_len) *slice_len = b - a;
    pack_sa5(ctx, Mlo[cur].p, Mhi[cur].p, b - a, gather ? out_sa5_d + 5 * a : out_sa5_d);
    if (gather && world > 1) {
      for (int h = 0; h < world; ++h) {
        const long ha = share_begin(h, n, world), hb = share_begin(h + 1, n, world);
        if (hb > ha) coll(ctx, fn, user, PSC_COLL_BCAST, out_sa5_d + 5 * ha, (size_t)(hb - ha) * 5, h);
      }
    }
  }
  sync(ctx);
  ctx->stats.kernel_launches = ctx->launches;
}

// ---- one GPU, two streams -------------------------------------------------------------------------
// The same two phases overlapped on ONE device: a sorter thread with its own context (stream, buffer
// cache) sorts blocks right to left and runs at most PIPE_DEPTH blocks ahead; the calling thread
// streams each sorted block through rank build, LF walk and merge.  The walk is bound by the HBM's
// random-access rate; the sort kernels of the next blocks fill the gaps around it (rank build,
// partition, accumulate, merge, host round trips).  While a walk kernel and a radix pass do run side
// by side (512 walkers per SM leave room for a sort CTA) they contend for the same HBM: the walk
// slows from 98 to 129 ms and the passes from 41 to 106 ms, so the gain on 8 x 128 MiB is modest
// (274 -> 264 ms); it costs the memory of the blocks sorted ahead, so api.cu only takes this path
// when HBM allows.
constexpr int PIPE_DEPTH = 2;

struct SortQueue {
  std::mutex mu;
  std::condition_variable cv;
  std::deque<std::unique_ptr<OwnedBlock>> ready;
  long consumed = 0, produced = 0;
  bool failed = false, cancel = false;
  int fail_code = 0;
  std::string fail_msg;
};

static psc_context *aux_context(psc_context *ctx) {
  if (ctx->aux) return ctx->aux;
  psc_context *a = new psc_context();
  a->device = ctx->device;
  a->n_sm = ctx->n_sm;
  a->pool = ctx->pool;
  a->trace = ctx->trace;
  PSC_CUDA(cudaStreamCreateWithFlags(&a->stream, cudaStreamNonBlocking));
  PSC_CUDA(cudaStreamCreateWithFlags(&a->side, cudaStreamNonBlocking));
  memset(&a->stats, 0, sizeof(a->stats));
  ctx->aux = a;
  return a;
}

void construct_sa_pipelined(psc_context *ctx, const uint8_t *text_d, long n, long max_block, uint8_t *out_sa5_d) {
  memset(&ctx->stats, 0, sizeof(ctx->stats));
  ctx->launches = 0;
  if (n == 0) return;
  psc_context *aux = aux_context(ctx);
  memset(&aux->stats, 0, sizeof(aux->stats));
  aux->launches = 0;
  PhaseTimer total(ctx, &ctx->stats.ms_total);
  max_block = std::min(max_block, (1L << 31) - 4096);
  const long B = div_up(n, max_block);
  ctx->stats.n_blocks = B;
  auto bounds = [&](long id, long *beg, long *end) { *beg = id * max_block; *end = std::min(n, *beg + max_block); };

  SortQueue q;
  std::thread sorter([&] {
    try {
      PSC_CUDA(cudaSetDevice(aux->device));
      for (long id = B - 1; id >= 0; --id) {
        {
          std::unique_lock<std::mutex> lk(q.mu);
          q.cv.wait(lk, [&] { return q.cancel || q.produced - q.consumed < PIPE_DEPTH; });
          if (q.cancel) return;
        }
        long beg, end;
        bounds(id, &beg, &end);
        auto ob = sort_owned_block(aux, text_d, n, beg, end);
        std::lock_guard<std::mutex> lk(q.mu);
        q.ready.push_back(std::move(ob));
        q.produced++;
        q.cv.notify_all();
      }
    } catch (const CudaError &e) {
      std::lock_guard<std::mutex> lk(q.mu);
      q.failed = true;
      q.fail_code = e.code;
      q.fail_msg = psc_last_error();
      q.cv.notify_all();
    }
  });
  struct Joiner {  // the sorter must be gone before the queue or any buffer is
    SortQueue &q; std::thread &t;
    ~Joiner() { { std::lock_guard<std::mutex> lk(q.mu); q.cancel = true; q.cv.notify_all(); } if (t.joinable()) t.join(); }
  } joiner{q, sorter};

  const int walkers_before = ctx->walk_threads_per_sm;
  if (const char *e = getenv("PSC_PIPE_WALK_TPS")) ctx->walk_threads_per_sm = atoi(e);
  struct Restore { psc_context *c; int v; ~Restore() { c->walk_threads_per_sm = v; } } restore{ctx, walkers_before};

  const long gwords = (n + 1 + 31) / 32 + 2;
  DevBuf<uint32_t> Ga(ctx, gwords), Gb(ctx, gwords);
  Ga.zero(); Gb.zero();
  uint32_t *gt_cur = Ga.p, *gt_new = Gb.p;
  DevBuf<uint32_t> Mlo[2];
  DevBuf<uint8_t> Mhi[2];
  for (int k = 0; k < 2; ++k) { Mlo[k].alloc(ctx, n); if (need_hi_plane(n)) Mhi[k].alloc(ctx, n);  /* high bytes only above 4 GiB */ }
  int cur = 0;
  long right_len = 0;
  for (long id = B - 1; id >= 0; --id) {
    long beg, end;
    bounds(id, &beg, &end);
    const long m = end - beg, t = n - end;
    std::unique_ptr<OwnedBlock> ob;
    {
      std::unique_lock<std::mutex> lk(q.mu);
      q.cv.wait(lk, [&] { return q.failed || !q.ready.empty(); });
      if (q.failed) { set_error("%s", q.fail_msg.c_str()); throw CudaError{q.fail_code}; }
      ob = std::move(q.ready.front());
      q.ready.pop_front();
    }
    PSC_CUDA(cudaMemsetAsync(gt_new, 0, gwords * 4, ctx->stream));
    DevBuf<uint32_t> gap;
    if (t > 0) {
      gap.alloc(ctx, m + 1);
      gap.zero();
      psc_rank *rk = rank_build(ctx, ob->bwt.p, m, ob->i0);
      try {
        uint64_t Cmap[256], run = 0;
        for (int c = 0; c < 256; ++c) { Cmap[c] = run; run += (uint64_t)rk->counts[c] + (c == ob->last ? 1 : 0); }
        walk_tail_range(ctx, rk, ob->i0, ob->last, Cmap, text_d + beg, ob->sa.p, m, text_d + end, t, 0, t, gt_cur, end,
                        id > 0 ? gt_new : nullptr, gap.p);
      } catch (...) { rank_free(rk); throw; }
      rank_free(rk);
    }
    if (id > 0) {
      if (m > 1) {
        or_bits_shifted_kernel<<<(unsigned)div_up(div_up(m - 1, 32), 256), 256, 0, ctx->stream>>>(gt_new, beg, ob->gtb.p, m - 1);
        PSC_LAUNCH_CHECK();
      }
      if (ob->end_bit) { set_bit_kernel<<<1, 1, 0, ctx->stream>>>(gt_new, end - 1); PSC_LAUNCH_CHECK(); }
      ctx->launches += 2;
    }
    int nxt = 1 - cur;
    merge_level(ctx, ob->sa.p, gap.p, m, beg, Mlo[cur].p, Mhi[cur].p, right_len, Mlo[nxt].p, Mhi[nxt].p);
    cur = nxt;
    right_len += m;
    std::swap(gt_cur, gt_new);
    sync(ctx);   // the block's buffers go back to the sorter's cache: nothing of ours may still read them
    ob.reset();
    {
      std::lock_guard<std::mutex> lk(q.mu);
      q.consumed++;
      q.cv.notify_all();
    }
  }
  pack_sa5(ctx, Mlo[cur].p, Mhi[cur].p, n, out_sa5_d);
  sync(ctx);
  // the sorter's counters belong to the same construction
  const psc_stats &a = aux->stats;
  ctx->stats.ms_gt_end += a.ms_gt_end; ctx->stats.ms_sort += a.ms_sort; ctx->stats.ms_radix += a.ms_radix;
  ctx->stats.ms_bwt += a.ms_bwt; ctx->stats.ms_radix_pass += a.ms_radix_pass;
  ctx->stats.radix_passes += a.radix_passes; ctx->stats.radix_elems += a.radix_elems; ctx->stats.radix_bytes += a.radix_bytes; ctx->stats.radix_keygen_bytes += a.radix_keygen_bytes; ctx->stats.ms_radix_keygen += a.ms_radix_keygen;
  ctx->stats.doubling_rounds += a.doubling_rounds;
  ctx->stats.kernel_launches = ctx->launches + aux->launches;
  for (auto &r : aux->trace_rows) ctx->trace_rows.push_back(r);
  aux->trace_rows.clear();
}

}  // namespace psc

static int dist_entry(psc_context *ctx, const uint8_t *text_d, long n, long max_block, uint8_t *out_sa5_d, int rank,
                      int world, psc_collective_fn fn, void *user, bool gather, long *slice_beg, long *slice_len,
                      const char *who) {
  try {
    if (!ctx || n < 0 || (n > 0 && (!text_d || !out_sa5_d)) || world < 1 || rank < 0 || rank >= world ||
        max_block <= 0 || (world > 1 && !fn)) {
      psc::set_error("bad argument: %s", who);
      return PSC_EARG;
    }
    PSC_CUDA(cudaSetDevice(ctx->device));
    if (slice_beg) *slice_beg = 0;
    if (slice_len) *slice_len = 0;
    psc::construct_sa_dist(ctx, text_d, n, max_block, out_sa5_d, rank, world, fn, user, gather, slice_beg, slice_len);
    psc::trace_report(ctx);
    return PSC_OK;
  } catch (const psc::CudaError &e) {
    return e.code;
  }
}

extern "C" int psc_construct_sa_dist_d(psc_context *ctx, const uint8_t *text_d, long n, long max_block,
                                       uint8_t *out_sa5_d, int rank, int world, psc_collective_fn fn,
                                       void *user) {
  return dist_entry(ctx, text_d, n, max_block, out_sa5_d, rank, world, fn, user, true, nullptr, nullptr,
                    "psc_construct_sa_dist_d");
}

extern "C" int psc_construct_sa_dist_slice_d(psc_context *ctx, const uint8_t *text_d, long n, long max_block,
                                             uint8_t *out_slice_sa5_d, long *slice_beg, long *slice_len, int rank,
                                             int world, psc_collective_fn fn, void *user) {
  if (!slice_beg || !slice_len) { psc::set_error("bad argument: psc_construct_sa_dist_slice_d"); return PSC_EARG; }
  return dist_entry(ctx, text_d, n, max_block, out_slice_sa5_d, rank, world, fn, user, false, slice_beg, slice_len,
                    "psc_construct_sa_dist_slice_d");
}

extern "C" long psc_dist_slice_capacity(long n, int world) {
  return world < 1 || n < 0 ? 0 : psc::share_begin_public(1, n, world);
}
