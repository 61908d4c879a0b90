// dist.cu -- the block pipeline over several GPUs, one rank per GPU, NCCL over NVLink.
//
//   phase 1  block sorts are independent once gt_end is computed by plain text comparison past the
//            block end (no gt bits of the next block needed): rank r sorts blocks B-1-r, B-1-r-N, ...
//            and keeps their SA; BWT, gt_begin bits and i0 wait for phase 2.
//   phase 2  blocks right to left.  The owner broadcasts [gt_begin bits | BWT] (m bytes + m/8 instead of
//            the 5m of round 1) and every rank builds the rank structure.  The owner then computes the
//            start ranks of ALL ranks' walkers (K5 needs the block SA, which never leaves its owner, and
//            the gt bits the previous walk produced) and broadcasts them (a few MB).  Every rank walks ITS
//            share of the tail (compute_gap.hpp:68-69 cuts the tail over threads the same way; shares
//            begin on 32-bit word boundaries of the gt bit array), then
//              ncclReduceScatter(sum, uint32) of the gap counts: every rank keeps a 1/N slice per block,
//              ncclAllGather of the new gt words (disjoint word ranges, no reduction).
//   phase 3  the N-way merge by output range (finalmerge.cu): rank g produces output slots
//            [g n/N, (g+1) n/N).  The split of every boundary over the blocks is a distributed search
//            (each rank searches its gap slices, one small all-reduce per block), then ONE exchange
//            brings every rank the slices of every block's SA and gap array its range needs
//            (8 n/N bytes per rank), and the cascade runs locally.  Nothing of size n lives on a GPU
//            except the text.
// All collectives are enqueued on the context's stream (comm.cu); the host only waits where it needs
// a value (block meta data, split tables).
// PSC_DIST_HELPER=1 (experimental, off by default) moves the package broadcast + rank build of the next block
// and the reduce-scatter of the previous one to a helper thread with its own stream (BlockFeeder below).
#include <algorithm>
#include <condition_variable>
#include <functional>
#include <memory>
#include <mutex>
#include <string>
#include <thread>
#include "comm.cuh"
#include "finalmerge.cuh"
#include "texthash.cuh"

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

// PSC_GAP_RS8=1: gap counters travel as bytes when they are small enough (they almost always are: a rank's counts
// of one block average below one per entry): 4 entries per thread, the largest value seen goes to *vmax.
__global__ void __launch_bounds__(256)
    narrow_u32_to_u8_kernel(const uint32_t *__restrict__ in, long n4, uint32_t *__restrict__ out4, uint32_t *vmax) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  uint32_t mx = 0;
  if (i < n4) {
    const uint4 v = reinterpret_cast<const uint4 *>(in)[i];
    mx = max(max(v.x, v.y), max(v.z, v.w));
    out4[i] = (v.x & 255u) | ((v.y & 255u) << 8) | ((v.z & 255u) << 16) | (v.w << 24);
  }
#pragma unroll
  for (int o = 16; o; o >>= 1) mx = max(mx, __shfl_xor_sync(0xffffffffu, mx, o));
  if ((threadIdx.x & 31) == 0 && mx) atomicMax(vmax, mx);
}
__global__ void __launch_bounds__(256)
    widen_u8_to_u32_kernel(const uint32_t *__restrict__ in4, long n4, uint32_t *__restrict__ out) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n4) return;
  const uint32_t v = in4[i];
  reinterpret_cast<uint4 *>(out)[i] = make_uint4(v & 255u, (v >> 8) & 255u, (v >> 16) & 255u, v >> 24);
}

struct BlockMeta { long long i0; int end_bit; int last; };
constexpr size_t META_BYTES = 256;

// what a block's owner sends to everybody, one contiguous package: [meta | gt_begin bits | BWT]
struct PackLayout {
  size_t gtb_off = META_BYTES, gtb_bytes = 0, bwt_off = 0, total = 0;
  PackLayout(long m, bool has_tail) {
    gtb_bytes = (size_t)(m / 32 + 2) * 4;
    bwt_off = gtb_off + ((gtb_bytes + 255) & ~(size_t)255);
    total = bwt_off + (has_tail ? (size_t)m : 0);
  }
};

struct OwnedBlock {
  DevBuf<uint32_t> sa;
  DevBuf<uint8_t> pack;  // PackLayout: gt bit q <-> block position q+1, q in [0, m-1); BWT only if the block has a tail
  BlockMeta meta{0, 0, 0};  // i0, gt bit of position `end` relative to the block begin, text[end - 1]
};

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
  const PackLayout lay(m, t > 0);
  ob->pack.alloc(ctx, lay.total);
  uint32_t *gtb = reinterpret_cast<uint32_t *>(ob->pack.p + lay.gtb_off);
  PSC_CUDA(cudaMemsetAsync(ob->pack.p, 0, lay.bwt_off, ctx->stream));
  gt_begin_from_sa(ctx, res.sa.p, m, (uint32_t)res.i0, gtb, 0);
  uint32_t w0 = 1;  // gt_end[0]: 1 for the last block
  if (t > 0) { d2h(ctx, &w0, gt_end.p, 1); sync(ctx); }
  if (t > 0) bwt_from_sa(ctx, text_d + beg, res.sa.p, m, ob->pack.p + lay.bwt_off);
  uint8_t last = 0;
  PSC_CUDA(cudaMemcpyAsync(&last, text_d + end - 1, 1, cudaMemcpyDeviceToHost, ctx->stream));
  sync(ctx);
  ob->meta = BlockMeta{(long long)res.i0, (w0 & 1u) ? 0 : 1 /* !gt_end[beg], change_gt_reference_point.hpp */, (int)last};
  h2d(ctx, reinterpret_cast<BlockMeta *>(ob->pack.p), &ob->meta, 1);
  ob->sa = std::move(res.sa);
  sync(ctx);
  return ob;
}

// first output slot of rank g when `len` slots are shared out: equal shares rounded up to a multiple of 4
// (the 40-bit writer packs 4 entries into 5 aligned words)
static long share_begin(long g, long len, int world) {
  long per = (div_up(len, (long)world) + 3) / 4 * 4;
  return std::min(len, per * g);
}
static long overlap_lo(long a0, long b0) { return std::max(a0, b0); }
static long overlap_len(long a0, long a1, long b0, long b1) {
  long lo = std::max(a0, b0), hi = std::min(a1, b1);
  return hi > lo ? hi - lo : 0;
}

// rank g's share of the tail of a block ending at text position `end`: the walkers of [lo, hi) write the gt
// bits with indices [end + lo, end + hi), and shares are cut on 32-bit word boundaries of that bit array so
// that the new words of all ranks are disjoint (all-gather, no reduction)
struct TailShares {
  long wl = 0, per_w = 0;  // first word touched, words per rank
  std::vector<long> lo;    // world + 1 tail-relative boundaries
};
static TailShares tail_shares(long end, long n, int world) {
  TailShares s;
  s.wl = end >> 5;
  const long wh = (n - 1) >> 5;
  s.per_w = div_up(wh - s.wl + 1, (long)world);
  s.lo.resize(world + 1);
  for (int g = 0; g <= world; ++g) {
    long bit = std::min(n, std::max(end, 32 * (s.wl + g * s.per_w)));
    s.lo[g] = bit - end;
  }
  return s;
}

struct DistPieces {  // what a rank keeps of every block until the final merge
  std::vector<DevBuf<uint32_t>> sa;       // whole SA of the blocks this rank owns
  std::vector<DevBuf<uint32_t>> gap;      // this rank's slice of every block's summed gap array
  std::vector<DevBuf<uint64_t>> carry;    // coarse carries of that slice
  std::vector<long> per;                  // slice length (entries per rank) of every block
};

// PSC_TRACE: host time spent waiting (for the helper thread), in the same table as the kernel times
struct HostTrace {
  psc_context *ctx;
  const char *name;
  std::chrono::steady_clock::time_point t0;
  HostTrace(psc_context *c, const char *n) : ctx(c), name(n), t0(std::chrono::steady_clock::now()) {}
  ~HostTrace() {
    if (!ctx->trace) return;
    const double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    for (auto &r : ctx->trace_rows)
      if (r.name == name) { r.count++; r.ms += ms; return; }
    ctx->trace_rows.push_back({name, 1, ms});
  }
};

// What a rank needs of a block before it can walk the block's tail: the owner's package and the rank structure
// built from its BWT.
struct BlockFeed {
  DevBuf<uint8_t> recv;          // ranks that do not own the block: the received package
  const uint8_t *pack = nullptr;  // the package (the owner's own copy, or `recv`)
  BlockMeta meta{0, 0, 0};
  psc_rank *rk = nullptr;
  void clear() { rank_free(rk); rk = nullptr; recv.release(); pack = nullptr; }
};

// The helper thread of a rank: it works on its own context (stream + buffer cache), so what it enqueues runs on
// the device beside the walk the main thread is busy with:
//   pre(k)   the feed of the k-th block (block B-1 is the 0-th): broadcast of the owner's package + rank structure;
//   post(k)  the job the main thread hands over after the walk of the k-th block: summing the gap counters over
//            the ranks, which nothing needs before the final merge.
// Both threads use ONE communicator, and NCCL wants every rank to issue its collectives in the same order, so
// the right to issue them is a token that moves through a fixed sequence of turns:
//   helper: pre(0) pre(1) | for k = 0 .. B-1:  main: start ranks of k | helper: post(k-1) pre(k+1) | main: gt words
//   of k (after the walk) | and last helper: post(B-1).
// The helper's turn of step k falls into the walk of block k, which is where its device work hides; a feed
// occupies one of two slots until the main thread is done with it.  Waits only ever point to earlier turns.
// Without a helper context everything happens in place on the caller's thread.
class BlockFeeder {
 public:
  using Job = std::function<void(psc_context *, Comm &)>;
  BlockFeeder(psc_context *ctx, Comm &cm, psc_context *hctx, long n, long max_block, long B,
              std::vector<std::unique_ptr<OwnedBlock>> &mine)
      : ctx_(ctx), cm_(cm), hctx_(hctx), n_(n), max_block_(max_block), B_(B), mine_(mine) {
    if (hctx_) {
      memset(&hctx_->stats, 0, sizeof(hctx_->stats));
      hctx_->launches = 0;
      th_ = std::thread([this] { run(); });
    }
  }
  ~BlockFeeder() {
    if (th_.joinable()) {
      { std::lock_guard<std::mutex> lk(mu_); stop_ = true; }
      cv_.notify_all();
      th_.join();
      collect();
    }
    // the caller has synchronised its stream after the last use of a feed (done()), or is unwinding
    cudaStreamSynchronize(ctx_->stream);
    for (auto &sl : slot_) sl.feed.clear();
  }
  // turns of the main thread in step k (k-th block from the right): 0 = start ranks, 1 = gt words
  void acquire(long k, int which) { wait_turn(1 + 3 * k + (which ? 2 : 0), false); }
  void release() { pass(); }
  BlockFeed *get(long id) {
    Slot &sl = slot_[(B_ - 1 - id) & 1];
    if (!th_.joinable()) {
      produce(ctx_, id, sl.feed);
      return &sl.feed;
    }
    std::unique_lock<std::mutex> lk(mu_);
    cv_.wait(lk, [&] { return failed_ || (sl.ready && sl.id == id); });
    rethrow();
    return &sl.feed;
  }
  // the caller's stream has finished with block id's feed; `job` (may be empty) is what is left to do for the block
  void done(long id, Job job) {
    Slot &sl = slot_[(B_ - 1 - id) & 1];
    sl.feed.clear();
    static const bool defer = !getenv("PSC_DIST_DEFER") || atoi(getenv("PSC_DIST_DEFER")) != 0;
    if (!th_.joinable() || !defer) {  // the caller holds the turn
      if (job) job(ctx_, cm_);
      if (!th_.joinable()) return;
      job = nullptr;
    }
    std::lock_guard<std::mutex> lk(mu_);
    sl.job = std::move(job);
    sl.ready = false;
    sl.id = -1;
  }
  // all jobs have run and their results are complete in device memory
  void finish() {
    if (!th_.joinable()) return;
    th_.join();
    collect();
    rethrow();
  }

 private:
  struct Slot { BlockFeed feed; long id = -1; bool ready = false; Job job; };
  void rethrow() {
    if (failed_) { set_error("%s", err_.c_str()); throw CudaError{err_code_}; }
  }
  // false: stopped (helper only)
  bool wait_turn(long pos, bool helper) {
    if (!th_.joinable() && !helper) return true;
    std::unique_lock<std::mutex> lk(mu_);
    cv_.wait(lk, [&] { return failed_ || (helper && stop_) || turn_ == pos; });
    if (!helper) rethrow();
    return turn_ == pos && !failed_;
  }
  void pass() {
    if (!th_.joinable()) return;
    { std::lock_guard<std::mutex> lk(mu_); ++turn_; }
    cv_.notify_all();
  }
  void collect() {
    ctx_->stats.ms_rank_build += hctx_->stats.ms_rank_build;
    ctx_->launches += hctx_->launches;
    hctx_->launches = 0;
    hctx_->stats.ms_rank_build = 0;
  }
  void produce(psc_context *c, long id, BlockFeed &f) {
    const long beg = id * max_block_, end = std::min(n_, beg + max_block_), m = end - beg, t = n_ - end;
    const int own = (int)((B_ - 1 - id) % cm_.world);
    const PackLayout lay(m, t > 0);
    uint8_t *p;
    if (own == cm_.rank) {
      p = mine_[id]->pack.p;
      f.meta = mine_[id]->meta;
    } else {
      f.recv.alloc(c, lay.total);
      p = f.recv.p;
    }
    cm_.bcast(c, p, lay.total, own);
    if (own != cm_.rank) { d2h(c, &f.meta, reinterpret_cast<const BlockMeta *>(p), 1); sync(c); }
    f.pack = p;
    if (t > 0) f.rk = rank_build(c, p + lay.bwt_off, m, (long)f.meta.i0);
    sync(c);
  }
  void publish(long k) {
    { std::lock_guard<std::mutex> lk(mu_); slot_[k & 1].id = B_ - 1 - k; slot_[k & 1].ready = true; }
    cv_.notify_all();
  }
  void run_job(long k) {  // handed over by done() before the main thread passed the turn
    Job job;
    { std::lock_guard<std::mutex> lk(mu_); job = std::move(slot_[k & 1].job); slot_[k & 1].job = nullptr; }
    if (job) { job(hctx_, cm_); sync(hctx_); }
  }
  void run() {
    try {
      PSC_CUDA(cudaSetDevice(hctx_->device));
      for (long k = 0; k < 2 && k < B_; ++k) { produce(hctx_, B_ - 1 - k, slot_[k & 1].feed); publish(k); }
      pass();
      for (long k = 0; k < B_; ++k) {
        if (!wait_turn(2 + 3 * k, true)) return;
        if (k >= 1) {
          run_job(k - 1);
          if (k + 1 < B_) { produce(hctx_, B_ - 1 - (k + 1), slot_[(k + 1) & 1].feed); publish(k + 1); }
        }
        pass();
      }
      if (!wait_turn(1 + 3 * B_, true)) return;
      run_job(B_ - 1);
    } catch (const CudaError &e) {
      std::lock_guard<std::mutex> lk(mu_);
      failed_ = true;
      err_code_ = e.code;
      err_ = psc_last_error();  // the message is thread-local: hand it to the consumer
      cv_.notify_all();
    }
  }
  psc_context *ctx_;
  Comm &cm_;
  psc_context *hctx_;
  long n_, max_block_, B_;
  std::vector<std::unique_ptr<OwnedBlock>> &mine_;
  Slot slot_[2];
  std::thread th_;
  std::mutex mu_;
  std::condition_variable cv_;
  long turn_ = 0;
  bool stop_ = false, failed_ = false;
  int err_code_ = PSC_ECUDA;
  std::string err_;
};

// phases 1 and 2; fills `pcs`, totals_d[g * B + id] = sum of rank g's gap slice of block id
static void dist_blocks(psc_context *ctx, Comm &cm, const uint8_t *text_d, long n, long max_block, long B,
                        DistPieces &pcs, DevBuf<uint64_t> &totals_d) {
  const int rank = cm.rank, world = cm.world;
  auto owner = [&](long id) { return (int)((B - 1 - id) % world); };
  auto bounds = [&](long id, long *beg, long *end) { *beg = id * max_block; *end = std::min(n, *beg + max_block); };
  pcs.sa.resize(B); pcs.gap.resize(B); pcs.carry.resize(B); pcs.per.assign(B, 0);

  // ---- phase 1: my blocks, independent of each other
  std::vector<std::unique_ptr<OwnedBlock>> mine(B);
  for (long id = B - 1; id >= 0; --id) {
    if (owner(id) != rank) continue;
    long beg, end;
    bounds(id, &beg, &end);
    mine[id] = sort_owned_block(ctx, text_d, n, beg, end);
  }

  // ---- phase 2: right to left
  const long gwords = (n + 1 + 31) / 32 + 2 + world + 2;
  DevBuf<uint32_t> Ga(ctx, gwords), Gb(ctx, gwords);
  Ga.zero(); Gb.zero();
  uint32_t *gt_cur = Ga.p, *gt_new = Gb.p;
  totals_d.alloc(ctx, (size_t)B * world);
  totals_d.zero();
  DevBuf<uint64_t> my_tot(ctx, B);
  my_tot.zero();
  sync(ctx);  // the helper thread reads what phase 1 wrote
  BlockFeeder feeder(ctx, cm, ctx->helper, n, max_block, B, mine);
  for (long id = B - 1; id >= 0; --id) {
    long beg, end;
    bounds(id, &beg, &end);
    const long m = end - beg, t = n - end;
    const int own = owner(id);
    OwnedBlock *ob = mine[id].get();
    // walkers of every rank (the same plan on every rank: it only depends on t, world and the SM count)
    TailShares sh;
    std::vector<WalkPlan> plans(world);
    std::vector<long> off(world + 1, 0);
    if (t > 0) {
      sh = tail_shares(end, n, world);
      for (int g = 0; g < world; ++g) {
        if (sh.lo[g + 1] > sh.lo[g]) make_walk_plan(ctx, sh.lo[g], sh.lo[g + 1], &plans[g]);
        off[g + 1] = off[g] + (long)plans[g].b.size();
      }
    }
    const long nbnd = off[world];
    DevBuf<uint64_t> b_d(ctx, std::max(1L, nbnd)), sr(ctx, std::max(1L, nbnd));
    const long step = B - 1 - id;
    { HostTrace ht(ctx, "dist_turn_wait(host)"); feeder.acquire(step, 0); }
    if (nbnd > 0) {
      std::vector<uint64_t> all_b((size_t)nbnd);
      for (int g = 0; g < world; ++g) std::copy(plans[g].b.begin(), plans[g].b.end(), all_b.begin() + off[g]);
      h2d(ctx, b_d.p, all_b.data(), all_b.size());
      sync(ctx);  // all_b is a local vector
      // K5 for everybody's walkers: the rank of suffix(end + b) among the block's suffixes, from the owner
      KTrace kt(ctx, "dist_start_ranks");
      if (own == rank)
        initial_ranks(ctx, text_d + beg, ob->sa.p, m, text_d + end, t, 1, gt_cur, end, nbnd, b_d.p, sr.p);
      cm.bcast(ctx, sr.p, (size_t)nbnd * 8, own);
    }
    feeder.release();
    BlockFeed *feed;  // [meta | gt_begin bits | BWT] and the rank structure, prepared ahead
    { HostTrace ht(ctx, "dist_feed_wait(host)"); feed = feeder.get(id); }
    const BlockMeta meta = feed->meta;
    const PackLayout lay(m, t > 0);
    const uint32_t *gtb_p = reinterpret_cast<const uint32_t *>(feed->pack + lay.gtb_off);
    const uint64_t *sr_p = sr.p;
    const long i0 = (long)meta.i0;
    PSC_CUDA(cudaMemsetAsync(gt_new, 0, gwords * 4, ctx->stream));
    BlockFeeder::Job post;
    bool acquired = false;
    if (t > 0) {
      const long per = (div_up(m + 1, (long)world) + 3) / 4 * 4;
      pcs.per[id] = per;
      DevBuf<uint32_t> gap_full(ctx, (size_t)per * world);
      gap_full.zero();
      const psc_rank *rk = feed->rk;
      uint64_t Cmap[256], run = 0;
      for (int c = 0; c < 256; ++c) { Cmap[c] = run; run += (uint64_t)rk->counts[c] + (c == meta.last ? 1 : 0); }
      const WalkPlan &p = plans[rank];
      if (p.b.size() > 1)  // walker k of my plan starts from the rank of boundary k + 1
        walk_with_plan(ctx, rk, i0, (uint8_t)meta.last, Cmap, text_d + end, p, b_d.p + off[rank], sr_p + off[rank] + 1,
                       gt_cur, end, id > 0 ? gt_new : nullptr, gap_full.p);
      // new gt words of disjoint word ranges: the next block's start ranks and walk need them
      { HostTrace ht(ctx, "dist_turn_wait(host)"); feeder.acquire(step, 1); }
      acquired = true;
      if (id > 0 && world > 1) {
        KTrace kt(ctx, "dist_gt_allgather");
        cm.allgather(ctx, gt_new + sh.wl + sh.per_w * rank, gt_new + sh.wl, (size_t)sh.per_w * 4);
      }
      // per-GPU gap counts summed, every rank keeps a slice: nothing needs it before the final merge, so the
      // helper does it while the next block is walked
      auto counts = std::make_shared<DevBuf<uint32_t>>(std::move(gap_full));
      uint64_t *tot_p = my_tot.p + id;
      post = [counts, per, m, id, tot_p, &pcs](psc_context *c, Comm &hc) {  // runs in one of the helper's turns
        const int world = hc.world, rank = hc.rank;
        pcs.gap[id].alloc(c, per);
        bool as_bytes = false;
        if (world > 1 && getenv("PSC_GAP_RS8")) {
          // opt-in: 1 byte per counter over NVLink instead of 4 when no sum can reach 256 (largest local count *
          // ranks < 256).  Fewer bytes, yet measured SLOWER than the uint32 reduce-scatter (4 GPUs: 173.6 against
          // 163.2 ms per 2 GiB construction, tools/dist_ab.py): NCCL sums uint32 at full NVLink speed, bytes not,
          // and the width decision costs a small all-reduce and a host round trip per block.
          DevBuf<uint8_t> narrow(c, (size_t)per * world), slice8(c, per);
          DevBuf<uint32_t> vmax(c, 1);
          vmax.zero();
          const long n4 = per * world / 4;
          narrow_u32_to_u8_kernel<<<(unsigned)div_up(n4, 256), 256, 0, c->stream>>>(counts->p, n4, reinterpret_cast<uint32_t *>(narrow.p), vmax.p);
          PSC_LAUNCH_CHECK();
          hc.allreduce_max_u32(c, vmax.p, 1);
          uint32_t hmax = 0;
          d2h(c, &hmax, vmax.p, 1);
          sync(c);
          if ((uint64_t)hmax * (uint64_t)world < 256) {
            hc.reduce_scatter_u8(c, narrow.p, slice8.p, (size_t)per);
            widen_u8_to_u32_kernel<<<(unsigned)div_up(per / 4, 256), 256, 0, c->stream>>>(reinterpret_cast<const uint32_t *>(slice8.p), per / 4, pcs.gap[id].p);
            PSC_LAUNCH_CHECK();
            c->launches += 2;
            as_bytes = true;
          }
        }
        if (!as_bytes) hc.reduce_scatter_u32(c, counts->p, pcs.gap[id].p, (size_t)per);
        const long cnt = std::max(0L, std::min(per, m + 1 - per * rank));
        pcs.carry[id].alloc(c, coarse_carry_tiles(cnt) + 1);
        coarse_carries_u32(c, pcs.gap[id].p, cnt, pcs.carry[id].p);
        PSC_CUDA(cudaMemcpyAsync(tot_p, pcs.carry[id].p + coarse_carry_tiles(cnt), 8, cudaMemcpyDeviceToDevice, c->stream));
        sync(c);  // the counters go back to the cache of the context they came from
      };
    }
    if (id > 0) {
      if (m > 1) {
        or_bits_shifted_kernel<<<(unsigned)div_up(div_up(m - 1, 32), 256), 256, 0, ctx->stream>>>(gt_new, beg, gtb_p, m - 1);
        PSC_LAUNCH_CHECK();
      }
      if (meta.end_bit) { set_bit_kernel<<<1, 1, 0, ctx->stream>>>(gt_new, end - 1); PSC_LAUNCH_CHECK(); }
      ctx->launches += 2;
    }
    std::swap(gt_cur, gt_new);
    sync(ctx);  // the feed's buffers go back to the helper
    if (!acquired) feeder.acquire(step, 1);
    feeder.done(id, std::move(post));
    feeder.release();
    if (own == rank) {
      pcs.sa[id] = std::move(ob->sa);
      mine[id].reset();
    }
  }
  { HostTrace ht(ctx, "dist_finish_wait(host)"); feeder.finish(); }
  // slice totals of all ranks
  if (world > 1) cm.allgather(ctx, my_tot.p, totals_d.p, (size_t)B * 8);
  else PSC_CUDA(cudaMemcpyAsync(totals_d.p, my_tot.p, (size_t)B * 8, cudaMemcpyDeviceToDevice, ctx->stream));
  sync(ctx);
}

// ---- phase 3: the merge by output range ---------------------------------------------------------------
// Boundaries: rank g's output range [share_begin(g), share_begin(g+1)) cut into chunks of
// merge_chunk_entries(); bd = all boundaries in order, first[g] = index of rank g's first one.
static void dist_merge(psc_context *ctx, Comm &cm, long n, long max_block, long B, DistPieces &pcs,
                       DevBuf<uint64_t> &totals_d, uint8_t *out_slice5_d) {
  PhaseTimer timer(ctx, &ctx->stats.ms_merge);
  const int rank = cm.rank, world = cm.world;
  auto owner = [&](long id) { return (int)((B - 1 - id) % world); };
  auto blen = [&](long id) { return std::min(n, (id + 1) * max_block) - id * max_block; };
  const long C = merge_chunk_entries();
  std::vector<uint64_t> bd;
  std::vector<int> first(world + 1);
  for (int g = 0; g < world; ++g) {
    first[g] = (int)bd.size();
    for (long a = share_begin(g, n, world); a < share_begin(g + 1, n, world); a += C) bd.push_back((uint64_t)a);
  }
  first[world] = (int)bd.size();
  bd.push_back((uint64_t)n);
  const int nb = (int)bd.size();
  std::vector<uint64_t> tot((size_t)B * world);
  d2h(ctx, tot.data(), totals_d.p, tot.size());
  sync(ctx);
  // gap counters are uint32 (the reference keeps 1 byte + an excess list, gap_array.hpp:63, 79-88): the counters
  // of block k must add up to the length of its tail, else one of them has wrapped around (a text in which more
  // than 2^32 - 1 tail suffixes fall between two neighbouring block suffixes)
  for (long k = 0; k + 1 < B; ++k) {
    uint64_t sum = 0;
    for (int h = 0; h < world; ++h) sum += tot[(size_t)h * B + k];
    const uint64_t tail = (uint64_t)(n - std::min(n, (k + 1) * max_block));
    if (sum != tail) {
      set_error("gap counters of block %ld add up to %llu, its tail has %llu suffixes: a 32-bit counter overflowed", k,
                (unsigned long long)sum, (unsigned long long)tail);
      throw CudaError{PSC_ELIMIT};
    }
  }
  // split: level by level, every rank answers from its gap slices, answers are summed
  MergePlan plan;
  plan.resize((int)B, nb);
  {
    KTrace kt(ctx, "merge_split");
    DevBuf<uint64_t> tab(ctx, (size_t)3 * B * nb);
    uint64_t *Qd = tab.p, *UEd = tab.p + (size_t)B * nb;  // U and E of one level are adjacent: one all-reduce
    h2d(ctx, Qd, bd.data(), nb);
    for (long k = 0; k < B; ++k) {
      uint64_t *qk = Qd + (size_t)k * nb, *uk = UEd + (size_t)2 * k * nb, *ek = uk + nb;
      if (k == B - 1) {
        PSC_CUDA(cudaMemcpyAsync(uk, qk, (size_t)nb * 8, cudaMemcpyDeviceToDevice, ctx->stream));
        PSC_CUDA(cudaMemsetAsync(ek, 0, (size_t)nb * 8, ctx->stream));
        break;
      }
      const long m = blen(k), per = pcs.per[k], u_off = per * rank;
      const long cnt = std::max(0L, std::min(per, m + 1 - u_off));
      uint64_t e_base = 0;
      for (int h = 0; h < rank; ++h) e_base += tot[(size_t)h * B + k];
      merge_split_slice(ctx, pcs.gap[k].p, cnt, u_off, e_base, pcs.carry[k].p, cnt > 0 && u_off + cnt == m + 1, qk, nb, uk, ek);
      cm.allreduce_sum_u64(ctx, uk, (size_t)2 * nb);
      merge_split_next(ctx, qk, uk, nb, qk + nb);
    }
    std::vector<uint64_t> ue((size_t)2 * B * nb);
    d2h(ctx, plan.q.data(), Qd, (size_t)B * nb);
    d2h(ctx, ue.data(), UEd, ue.size());
    sync(ctx);
    for (long k = 0; k < B; ++k)
      for (int i = 0; i < nb; ++i) {
        plan.u[(size_t)k * nb + i] = ue[(size_t)2 * k * nb + i];
        plan.e[(size_t)k * nb + i] = ue[(size_t)2 * k * nb + nb + i];
      }
  }
  // exchange: my range needs entries [U(k, first[rank]), U(k, first[rank+1])) of every block's SA and gap array
  auto need_lo = [&](long k, int g) { return (long)plan.U((int)k, first[g]); };
  auto need_hi = [&](long k, int g) { return (long)plan.U((int)k, first[g + 1]); };
  std::vector<size_t> sa_at(B + 1, 0), gap_at(B + 1, 0);  // offsets (entries) of block k's slices in the receive buffers
  for (long k = 0; k < B; ++k) {
    const size_t c = (size_t)(need_hi(k, rank) - need_lo(k, rank));
    sa_at[k + 1] = sa_at[k] + ((c + 3) & ~(size_t)3);
    gap_at[k + 1] = gap_at[k] + ((c + 1 + 3) & ~(size_t)3);
  }
  DevBuf<uint32_t> rsa(ctx, sa_at[B] + 4), rgap(ctx, gap_at[B] + 4);
  {
    KTrace kt(ctx, "merge_exchange");
    cm.group_begin();
    for (long k = 0; k < B; ++k) {
      const int own = owner(k);
      for (int g = 0; g < world; ++g) {  // block SA: from its owner
        const long lo = need_lo(k, g), cnt = need_hi(k, g) - lo;
        if (cnt <= 0) continue;
        if (own == rank && g == rank)
          PSC_CUDA(cudaMemcpyAsync(rsa.p + sa_at[k], pcs.sa[k].p + lo, (size_t)cnt * 4, cudaMemcpyDeviceToDevice, ctx->stream));
        else if (own == rank) cm.send(ctx, pcs.sa[k].p + lo, (size_t)cnt * 4, g);
        else if (g == rank) cm.recv(ctx, rsa.p + sa_at[k], (size_t)cnt * 4, own);
      }
      if (k == B - 1) continue;  // the rightmost block has no gap array
      const long per = pcs.per[k], m = blen(k);
      for (int h = 0; h < world; ++h) {      // holder of a slice of the gap counters
        const long s0 = per * h, s1 = std::min(m + 1, s0 + per);
        for (int g = 0; g < world; ++g) {    // receiver
          if (h != rank && g != rank) continue;
          const long lo = overlap_lo(s0, need_lo(k, g)), cnt = overlap_len(s0, s1, need_lo(k, g), need_hi(k, g));
          if (cnt <= 0) continue;
          uint32_t *dst = rgap.p + gap_at[k] + (lo - need_lo(k, rank));
          if (h == rank && g == rank)
            PSC_CUDA(cudaMemcpyAsync(dst, pcs.gap[k].p + (lo - s0), (size_t)cnt * 4, cudaMemcpyDeviceToDevice, ctx->stream));
          else if (h == rank) cm.send(ctx, pcs.gap[k].p + (lo - s0), (size_t)cnt * 4, g);
          else cm.recv(ctx, dst, (size_t)cnt * 4, h);
        }
      }
    }
    cm.group_end();
  }
  // cascade, chunk by chunk of my range
  MergeScratch scr;
  scr.reserve(ctx, std::min(share_begin(1, n, world), C) + 16, need_hi_plane(n));
  std::vector<LevelSlice> lv(B);
  const long my_beg = share_begin(rank, n, world);
  for (int i = first[rank]; i < first[rank + 1]; ++i) {
    const long a = (long)plan.Q(0, i), b = (long)plan.Q(0, i + 1);
    for (long k = 0; k < B; ++k) {
      const size_t d = (size_t)(plan.U((int)k, i) - (uint64_t)need_lo(k, rank));
      lv[k] = plan.slice((int)k, i, rsa.p + sa_at[k] + d, rgap.p + gap_at[k] + d, k * max_block);
    }
    cascade_chunk(ctx, lv.data(), (int)B, b - a, scr, out_slice5_d + 5 * (a - my_beg));
  }
  sync(ctx);  // the receive buffers are released with this frame
}

void construct_sa_dist(psc_context *ctx, Comm &cm, const uint8_t *text_d, long n, long max_block,
                       uint8_t *out_slice5_d, long *slice_beg, long *slice_len) {
  memset(&ctx->stats, 0, sizeof(ctx->stats));
  ctx->launches = 0;
  cm.n_calls = 0;
  cm.bytes = 0;
  const long a = share_begin(cm.rank, n, cm.world), b = share_begin(cm.rank + 1, n, cm.world);
  *slice_beg = a;
  *slice_len = b - a;
  if (n == 0) return;
  PhaseTimer total(ctx, &ctx->stats.ms_total);
  max_block = std::min(max_block, (1L << 31) - 4096);
  hash_domain_set(ctx, text_d, n);  // K0 compares past the block end here: fingerprints bound it (texthash.cuh)
  struct DomainGuard { psc_context *c; ~DomainGuard() { hash_domain_clear(c); } } domain_guard{ctx};
  const long B = div_up(n, max_block);
  ctx->stats.n_blocks = B;
  DistPieces pcs;
  DevBuf<uint64_t> totals_d;
  dist_blocks(ctx, cm, text_d, n, max_block, B, pcs, totals_d);
  if (B == 1) {  // one block: its owner (rank 0) holds the SA, everybody gets a slice of it
    if (cm.world > 1) {
      DevBuf<uint32_t> part(ctx, std::max(1L, b - a));
      cm.group_begin();
      for (int g = 0; g < cm.world; ++g) {
        const long ga = share_begin(g, n, cm.world), gb = share_begin(g + 1, n, cm.world);
        if (gb <= ga) continue;
        if (cm.rank == 0 && g == 0)
          PSC_CUDA(cudaMemcpyAsync(part.p, pcs.sa[0].p, (size_t)(gb - ga) * 4, cudaMemcpyDeviceToDevice, ctx->stream));
        else if (cm.rank == 0) cm.send(ctx, pcs.sa[0].p + ga, (size_t)(gb - ga) * 4, g);
        else if (g == cm.rank) cm.recv(ctx, part.p, (size_t)(gb - ga) * 4, 0);
      }
      cm.group_end();
      pack_sa5_from_sa(ctx, part.p, b - a, 0, out_slice5_d);  // values are text positions already (the block begins at 0)
      sync(ctx);
    } else {
      pack_sa5_from_sa(ctx, pcs.sa[0].p, n, 0, out_slice5_d);
    }
  } else {
    dist_merge(ctx, cm, n, max_block, B, pcs, totals_d, out_slice5_d);
  }
  sync(ctx);
  ctx->stats.kernel_launches = ctx->launches;
}

}  // namespace psc

using namespace psc;

extern "C" {

int psc_dist_unique_id(uint8_t id[128]) {
  if (!id) { set_error("psc_dist_unique_id: null pointer"); return PSC_EARG; }
  return nccl_unique_id(id);
}

static void dist_teardown(psc_context *ctx, bool abort) {
  if (ctx->comm) {
    if (abort) nccl_comm_abort(ctx->comm); else nccl_comm_destroy(ctx->comm);
    if (!abort) { delete ctx->comm; ctx->comm = nullptr; }
  }
  if (!abort && ctx->helper) { psc_shutdown(ctx->helper); ctx->helper = nullptr; }
}

int psc_dist_init(psc_context *ctx, const uint8_t id[128], int rank, int world) {
  if (!ctx || world < 1 || rank < 0 || rank >= world || (world > 1 && !id)) { set_error("psc_dist_init: bad argument"); return PSC_EARG; }
  cudaSetDevice(ctx->device);
  dist_teardown(ctx, false);
  Comm *c = new Comm();
  int rc = nccl_comm_init(c, id, rank, world);
  if (rc != PSC_OK) { delete c; return rc; }
  // PSC_DIST_HELPER=1: second stream + buffer cache for a thread that prepares the next blocks during a walk
  // (BlockFeeder).  Off by default: measured on 2 GPUs it costs more than it hides.
  psc_context *h = nullptr;
  if (const char *e = getenv("PSC_DIST_HELPER")) {
    if (atoi(e) != 0) {
      rc = psc_init(ctx->device, 0, &h);
      if (rc != PSC_OK) { nccl_comm_destroy(c); delete c; return rc; }
      h->trace = false;
      // lowest priority: a walker CTA that starts late ends late, and the walk lasts as long as its last walker
      int least = 0, greatest = 0;
      cudaDeviceGetStreamPriorityRange(&least, &greatest);
      cudaStream_t low = nullptr;
      if (cudaStreamCreateWithPriority(&low, cudaStreamNonBlocking, least) == cudaSuccess) {
        cudaStreamDestroy(h->stream);
        h->stream = low;
      }
    }
  }
  ctx->comm = c;
  ctx->helper = h;
  return PSC_OK;
}

int psc_dist_finalize(psc_context *ctx) {
  if (!ctx) return PSC_OK;
  if (ctx->comm || ctx->helper) {
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    dist_teardown(ctx, false);
  }
  return PSC_OK;
}

int psc_dist_abort(psc_context *ctx) {
  if (ctx) dist_teardown(ctx, true);
  return PSC_OK;
}

int psc_dist_stats(psc_context *ctx, long *n_collectives, double *bytes) {
  if (!ctx || !ctx->comm) { set_error("psc_dist_stats: psc_dist_init has not been called"); return PSC_EARG; }
  if (n_collectives) *n_collectives = ctx->comm->n_calls;
  if (bytes) *bytes = ctx->comm->bytes;
  return PSC_OK;
}

long psc_dist_slice_capacity(long n, int world) {
  return world < 1 || n < 0 ? 0 : share_begin(1, n, world);
}

int psc_construct_sa_dist_slice_d(psc_context *ctx, const uint8_t *text_d, long n, long max_block,
                                  uint8_t *out_slice_sa5_d, long *slice_beg, long *slice_len) {
  try {
    if (!ctx || n < 0 || (n > 0 && (!text_d || !out_slice_sa5_d)) || max_block <= 0 || !slice_beg || !slice_len) {
      set_error("bad argument: psc_construct_sa_dist_slice_d");
      return PSC_EARG;
    }
    PSC_CUDA(cudaSetDevice(ctx->device));
    Comm self;  // a context without psc_dist_init is a world of one
    construct_sa_dist(ctx, ctx->comm ? *ctx->comm : self, text_d, n, max_block, out_slice_sa5_d, slice_beg, slice_len);
    trace_report(ctx);
    return PSC_OK;
  } catch (const CudaError &e) {
    return e.code;
  }
}

int psc_construct_sa_dist_d(psc_context *ctx, const uint8_t *text_d, long n, long max_block, uint8_t *out_sa5_d) {
  try {
    if (!ctx || n < 0 || (n > 0 && (!text_d || !out_sa5_d)) || max_block <= 0) {
      set_error("bad argument: psc_construct_sa_dist_d");
      return PSC_EARG;
    }
    PSC_CUDA(cudaSetDevice(ctx->device));
    Comm self;
    Comm &cm = ctx->comm ? *ctx->comm : self;
    long a = 0, len = 0;
    construct_sa_dist(ctx, cm, text_d, n, max_block, out_sa5_d + 5 * share_begin(cm.rank, n, cm.world), &a, &len);
    // every rank receives the other slices
    for (int h = 0; h < cm.world; ++h) {
      const long ha = share_begin(h, n, cm.world), hb = share_begin(h + 1, n, cm.world);
      if (hb > ha) cm.bcast(ctx, out_sa5_d + 5 * ha, (size_t)(hb - ha) * 5, h);
    }
    sync(ctx);
    trace_report(ctx);
    return PSC_OK;
  } catch (const CudaError &e) {
    return e.code;
  }
}

}  // extern "C"
