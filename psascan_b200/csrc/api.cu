// api.cu -- the C ABI (include/psascan_cuda.h) and the block pipeline that keeps everything in HBM.
#include <stdarg.h>
#include <algorithm>
#include "common.cuh"
#include "finalmerge.cuh"
#include "texthash.cuh"

namespace psc {

static thread_local char g_err[1024] = "";
void set_error(const char *fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

template <typename F>
static int guarded(F &&f) {
  try {
    f();
    return PSC_OK;
  } catch (const CudaError &e) {
    return e.code;
  } catch (const std::exception &e) {
    set_error("%s", e.what());
    return PSC_ECUDA;
  }
}

static void arg_check(bool ok, const char *what) {
  if (!ok) { set_error("bad argument: %s", what); throw CudaError{PSC_EARG}; }
}

__global__ void or_not_bit_kernel(const uint32_t *src_bits, long src_idx, int src_default,
                                  uint32_t *dst_bits, long dst_idx) {
  // dst[dst_idx] |= !src[src_idx]   (src == null: src bit = src_default)
  bool s = src_bits ? bit_get(src_bits, src_idx) : (src_default != 0);
  if (!s) atomicOr(&dst_bits[dst_idx >> 5], 1u << (dst_idx & 31));
}

// Stages 2+3 for one block against the text after it (all device pointers).
//   gt_cur: bits relative to the block end, bit index = global position - 1 - bit_base
static void gap_of_block(psc_context *ctx, const uint8_t *block_d, long m, const uint32_t *sa_d,
                         long i0, const uint8_t *after_d, long t, const uint32_t *gt_cur, long gt_off,
                         uint32_t *gt_new, uint32_t *gap_d) {
  uint8_t last;
  PSC_CUDA(cudaMemcpyAsync(&last, block_d + m - 1, 1, cudaMemcpyDeviceToHost, ctx->stream));
  DevBuf<uint8_t> bwt(ctx, m);
  bwt_from_sa(ctx, block_d, sa_d, m, bwt.p);
  psc_rank *rk = rank_build(ctx, bwt.p, m, i0);
  bwt.release();
  try {
    uint64_t Cmap[256], run = 0;
    for (int c = 0; c < 256; ++c) { Cmap[c] = run; run += (uint64_t)rk->counts[c] + (c == last ? 1 : 0); }
    walk_tail_range(ctx, rk, i0, last, Cmap, block_d, sa_d, m, after_d, t, 0, t, gt_cur, gt_off, gt_new, gap_d);
  } catch (...) {
    rank_free(rk);
    throw;
  }
  rank_free(rk);
}

// ---- work guard --------------------------------------------------------------------------------
WorkBudget *work_budget_arm(psc_context *ctx) {
  static unsigned long long limit = 0;
  if (!limit) {
    const char *e = getenv("PSC_COMPARE_BUDGET");
    limit = e ? strtoull(e, nullptr, 10) : (1ull << 42);  // matched bytes per kernel (~ seconds on a B200)
    if (!limit) limit = 1;
  }
  WorkBudget *wb = (WorkBudget *)cache_alloc(ctx, sizeof(WorkBudget));
  WorkBudget h{0, limit, 0, 0};
  cudaMemcpyAsync(wb, &h, sizeof(h), cudaMemcpyHostToDevice, ctx->stream);
  cudaStreamSynchronize(ctx->stream);  // h is a stack object
  return wb;
}

int work_budget_check(psc_context *ctx, WorkBudget *wb, const char *what) {
  WorkBudget h;
  cudaError_t e = cudaMemcpyAsync(&h, wb, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(ctx->stream);
  cache_release(ctx, wb);
  if (e != cudaSuccess) { set_error("%s: %s", what, cudaGetErrorString(e)); throw CudaError{PSC_ECUDA}; }
  if (h.exceeded) {
    set_error("%s: more than %llu matched bytes between block suffixes and the text after the block "
              "(degenerate repeats across a block boundary); use one block or a larger max_block", what, h.limit);
    throw CudaError{PSC_ELIMIT};
  }
  return h.need_hash;
}

// ---- buffer cache ------------------------------------------------------------------------------
static size_t size_class(size_t bytes) {
  // 16 classes per octave (<= 6.25 % slack), at least 512-byte steps
  if (bytes < 8192) return (bytes + 511) & ~(size_t)511;
  int lg = 63 - __builtin_clzll((unsigned long long)bytes);
  size_t step = (size_t)1 << (lg - 4);
  return (bytes + step - 1) & ~(step - 1);
}

void cache_trim(psc_context *ctx) {
  std::lock_guard<std::mutex> lk(ctx->cache_mu);
  size_t k = 0;
  for (auto &b : ctx->cache) {
    if (b.used) ctx->cache[k++] = b;
    else cudaFreeAsync(b.p, ctx->stream);
  }
  ctx->cache.resize(k);
  ctx->cache_free_bytes = 0;
}

void *cache_alloc(psc_context *ctx, size_t bytes) {
  const size_t cls = size_class(bytes);
  {
    std::lock_guard<std::mutex> lk(ctx->cache_mu);
    for (auto &b : ctx->cache)
      if (!b.used && b.bytes == cls) { b.used = true; ctx->cache_free_bytes -= cls; return b.p; }
  }
  void *q = nullptr;
  // a pool allocation issued while the stream still has work in flight was measured to take several
  // times longer (3 GiB cold run: 2.6 s against 1.2 s with this synchronisation); misses only happen
  // on the first construction of a given shape
  cudaStreamSynchronize(ctx->stream);
  auto t0 = std::chrono::steady_clock::now();
  cudaError_t e = cudaMallocAsync(&q, cls, ctx->stream);
  if (e == cudaErrorMemoryAllocation) {  // unused size classes (of either context) hold the memory
    cudaGetLastError();
    cache_trim(ctx);
    cudaDeviceSynchronize();
    e = cudaMallocAsync(&q, cls, ctx->stream);
  }
  if (ctx->trace) {
    double ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    ctx->host_alloc_ms += ms;
    ctx->n_allocs++;
    if (ms > 5.0) fprintf(stderr, "[psc trace]   slow cudaMallocAsync: %.1f MB took %.1f ms\n", cls / 1e6, ms);
  }
  if (e != cudaSuccess) {
    cudaGetLastError();
    set_error("device allocation of %zu bytes failed: %s", cls, cudaGetErrorString(e));
    throw CudaError{e == cudaErrorMemoryAllocation ? PSC_ENOMEM : PSC_ECUDA};
  }
  std::lock_guard<std::mutex> lk(ctx->cache_mu);
  ctx->cache.push_back({q, cls, true});
  return q;
}

void cache_release(psc_context *ctx, void *p) {
  std::lock_guard<std::mutex> lk(ctx->cache_mu);
  for (auto &b : ctx->cache)
    if (b.p == p) { b.used = false; ctx->cache_free_bytes += b.bytes; return; }
  cudaFreeAsync(p, ctx->stream);  // not ours (cannot happen)
}

void trace_report(psc_context *ctx) {
  if (!ctx->trace) return;
  const psc_stats &st = ctx->stats;
  fprintf(stderr, "[psc trace] total %.2f ms | sort %.2f (radix passes %.2f) | walk %.2f | merge %.2f | rank %.2f | bwt %.2f | gt_end %.2f\n",
          st.ms_total, st.ms_sort, st.ms_radix_pass, st.ms_walk, st.ms_merge, st.ms_rank_build, st.ms_bwt, st.ms_gt_end);
  for (auto &r : ctx->trace_rows) fprintf(stderr, "[psc trace]   %-18s x%-4ld %9.3f ms\n", r.name, r.count, r.ms);
  fprintf(stderr, "[psc trace]   host time in %ld cudaMallocAsync calls: %.3f ms; %zu cached buffers, %.1f MB unused\n", ctx->n_allocs,
          ctx->host_alloc_ms, ctx->cache.size(), ctx->cache_free_bytes / 1e6);
  ctx->host_alloc_ms = 0;
  ctx->n_allocs = 0;
  ctx->trace_rows.clear();
}

// bytes this context may still use: free HBM + what the pool and the buffer caches hold unused
double hbm_available(psc_context *ctx) {
  size_t free_b = 0, total_b = 0;
  if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) { cudaGetLastError(); return 0; }
  uint64_t reserved = 0, used = 0;
  cudaMemPoolGetAttribute(ctx->pool, cudaMemPoolAttrReservedMemCurrent, &reserved);
  cudaMemPoolGetAttribute(ctx->pool, cudaMemPoolAttrUsedMemCurrent, &used);
  double avail = (double)free_b + (double)(reserved - used) + (double)ctx->cache_free_bytes;
  if (ctx->hbm_budget && (double)ctx->hbm_budget < avail) avail = (double)ctx->hbm_budget;
  return avail * 0.92;
}

// the text of block `id` is complete in HBM (psc_construct_sa copies it block by block on the side stream)
struct BlockGate {
  virtual void wait(psc_context *ctx, long id) const = 0;
  virtual ~BlockGate() {}
};
static void construct_sa_sequential(psc_context *ctx, const uint8_t *text_d, long n, long max_block,
                                    uint8_t *out_sa5_d, const ChunkSink *sink, const BlockGate *gate,
                                    PieceSink *pieces = nullptr);
// Largest block for a text of n symbols.  io_bytes: text and output buffers the caller has NOT allocated yet
// (the public query: 6n); the construction entry points size the blocks after their I/O buffers exist,
// so those are never charged twice.  Nothing is trimmed here: unused cached buffers count as available
// (hbm_available) and are only given back to the driver when an allocation fails.
static long max_block_size_for(psc_context *ctx, long n, double io_bytes) {
  double avail = psc::hbm_available(ctx) - io_bytes;
  if (avail <= 0) return 0;
  // single block: 40 B/symbol of sort state (keys, values, SA)
  if ((double)n * 40.0 <= avail && n <= (1L << 31) - 4096) return n;
  // several blocks: every block's SA and gap array stay until the final merge (8n), gt bits (n/4), the merge
  // scratch of one output chunk (at most 18 B per slot), the rank pages of one walk chunk (8 B x 2^29
  // positions at most, walk.cu); per block symbol 40 B of sort state, or BWT + rank planes
  // (<= 255 * 32/224 B) while streaming
  const double chunk = (double)std::min(n, merge_chunk_entries());
  double rest = avail - 8.3 * (double)n - 18.0 * chunk - 8.0 * (double)std::min(n, 1L << 29);
  if (rest <= 0) return 0;
  long mb = (long)(rest / 40.0);
  mb = std::min(mb, (1L << 31) - 4096);
  mb = std::min(mb, n);
  return mb / 32 * 32;
}

static long choose_max_block(psc_context *ctx, long n, long max_block) {
  if (n <= 0) return 1;
  if (max_block <= 0) max_block = max_block_size_for(ctx, n, 0.0);
  arg_check(max_block > 0, "text does not fit the HBM budget");
  return std::min(max_block, (1L << 31) - 4096);
}

static void construct_sa_device(psc_context *ctx, const uint8_t *text_d, long n, long max_block,
                                uint8_t *out_sa5_d) {
  construct_sa_sequential(ctx, text_d, n, choose_max_block(ctx, n, max_block), out_sa5_d, nullptr, nullptr);
}

// ---- host buffers in and out: copies overlapped with the pipeline on the side stream ------------------
// Text: block by block, right to left (the order the pipeline consumes it), one event per block.
struct H2DGate : BlockGate {
  std::vector<cudaEvent_t> ev;
  void wait(psc_context *ctx, long id) const override {
    PSC_CUDA(cudaStreamWaitEvent(ctx->stream, ev[id], 0));
  }
  ~H2DGate() { for (auto e : ev) if (e) cudaEventDestroy(e); }
};
// Output: finished chunks are packed into one of two device buffers and leave by D2H on the side stream
// while the next chunk is merged; no 5n-byte output buffer exists on the device.
struct HostSink : ChunkSink {
  uint8_t *host_out = nullptr;
  mutable DevBuf<uint8_t> buf[2];
  mutable cudaEvent_t packed[2] = {nullptr, nullptr}, copied[2] = {nullptr, nullptr};
  mutable bool busy[2] = {false, false};
  HostSink(psc_context *ctx, uint8_t *out, long chunk_entries) : host_out(out) {
    for (int k = 0; k < 2; ++k) {
      buf[k].alloc(ctx, (size_t)chunk_entries * 5 + 64);
      PSC_CUDA(cudaEventCreateWithFlags(&packed[k], cudaEventDisableTiming));
      PSC_CUDA(cudaEventCreateWithFlags(&copied[k], cudaEventDisableTiming));
    }
  }
  uint8_t *buffer(psc_context *ctx, int chunk, long) const override {
    const int k = chunk & 1;
    if (busy[k]) PSC_CUDA(cudaStreamWaitEvent(ctx->stream, copied[k], 0));  // its last copy must have left
    return buf[k].p;
  }
  void done(psc_context *ctx, int chunk, long first_entry, long entries) const override {
    const int k = chunk & 1;
    PSC_CUDA(cudaEventRecord(packed[k], ctx->stream));
    PSC_CUDA(cudaStreamWaitEvent(ctx->side, packed[k], 0));
    PSC_CUDA(cudaMemcpyAsync(host_out + 5 * (size_t)first_entry, buf[k].p, 5 * (size_t)entries, cudaMemcpyDeviceToHost, ctx->side));
    PSC_CUDA(cudaEventRecord(copied[k], ctx->side));
    busy[k] = true;
  }
  void finish(psc_context *ctx) const { PSC_CUDA(cudaStreamSynchronize(ctx->side)); }
  ~HostSink() {
    for (int k = 0; k < 2; ++k) { if (packed[k]) cudaEventDestroy(packed[k]); if (copied[k]) cudaEventDestroy(copied[k]); }
  }
};

void construct_sa_with_sink(psc_context *ctx, const uint8_t *text_d, long n, long max_block, PieceSink *pieces) {
  construct_sa_sequential(ctx, text_d, n, max_block, nullptr, nullptr, nullptr, pieces);
}

static void construct_sa_sequential(psc_context *ctx, const uint8_t *text_d, long n, long max_block,
                                    uint8_t *out_sa5_d, const ChunkSink *sink, const BlockGate *gate,
                                    PieceSink *pieces) {
  memset(&ctx->stats, 0, sizeof(ctx->stats));
  ctx->launches = 0;
  if (n == 0) return;
  PhaseTimer total(ctx, &ctx->stats.ms_total);
  arg_check(max_block > 0, "text does not fit the HBM budget");
  max_block = std::min(max_block, (1L << 31) - 4096);
  hash_domain_set(ctx, text_d, n);  // fingerprints for K0 / K5, built only if a comparison outlasts the cap
  struct DomainGuard { psc_context *c; ~DomainGuard() { hash_domain_clear(c); } } domain_guard{ctx};
  long B = div_up(n, max_block);
  ctx->stats.n_blocks = B;
  long gwords = (n + 1 + 31) / 32 + 2;
  DevBuf<uint32_t> Ga, Gb;
  if (B > 1) {
    Ga.alloc(ctx, gwords); Gb.alloc(ctx, gwords);
    Ga.zero(); Gb.zero();
  }
  // every block's SA and gap array stay until the final merge (8 bytes per suffix), finalmerge.cu
  std::vector<DevBuf<uint32_t>> sa(B), gap(B);
  uint32_t *gt_cur = Ga.p, *gt_new = Gb.p;
  for (long id = B - 1; id >= 0; --id) {
    long beg = id * max_block, end = std::min(n, beg + max_block), m = end - beg, t = n - end;
    const uint8_t *block_d = text_d + beg, *after_d = text_d + end;
    if (gate) gate->wait(ctx, id);  // the block's text is on its way from the host (side stream)
    BlockSortResult res;
    {
      DevBuf<uint32_t> gt_end;
      if (t > 0) {
        gt_end.alloc(ctx, m / 32 + 2);
        compute_gt_end(ctx, block_d, m, after_d, t, 1, gt_cur, end, gt_end.p);
      }
      sort_block(ctx, block_d, m, t > 0 ? gt_end.p : nullptr, &res);
      if (id > 0) {
        PSC_CUDA(cudaMemsetAsync(gt_new, 0, gwords * 4, ctx->stream));
        gt_begin_from_sa(ctx, res.sa.p, m, (uint32_t)res.i0, gt_new, beg);
        or_not_bit_kernel<<<1, 1, 0, ctx->stream>>>(t > 0 ? gt_end.p : nullptr, 0, 1, gt_new, end - 1);
        PSC_LAUNCH_CHECK();
        ctx->launches++;
      }
    }
    res.isa.release();
    if (t > 0) {
      gap[id].alloc(ctx, m + 1);
      gap[id].zero();
      gap_of_block(ctx, block_d, m, res.sa.p, res.i0, after_d, t, gt_cur, end, id > 0 ? gt_new : nullptr,
                   gap[id].p);
    }
    if (pieces) {  // external-memory mode: the block leaves the device now
      pieces->take(ctx, id, beg, m, res.sa.p, t > 0 ? gap[id].p : nullptr, id > 0 ? gt_new : nullptr);
      res.sa.release();
      gap[id].release();
    } else {
      sa[id] = std::move(res.sa);
    }
    std::swap(gt_cur, gt_new);
  }
  Ga.release(); Gb.release();
  if (pieces) {
    sync(ctx);
    ctx->stats.kernel_launches = ctx->launches;
    return;
  }
  if (B == 1) {  // one piece: the block SA goes straight to the 40-bit writer
    if (sink) {
      const long C = merge_chunk_entries();
      PhaseTimer timer(ctx, &ctx->stats.ms_merge);
      for (long a = 0, i = 0; a < n; a += C, ++i) {
        const long cnt = std::min(C, n - a);
        pack_sa5_from_sa(ctx, sa[0].p + a, cnt, 0, sink->buffer(ctx, (int)i, cnt), false);
        sink->done(ctx, (int)i, a, cnt);
      }
    } else {
      pack_sa5_from_sa(ctx, sa[0].p, n, 0, out_sa5_d);
    }
  } else {
    std::vector<MergePiece> pc(B);
    for (long id = 0; id < B; ++id) {
      pc[id].beg = id * max_block;
      pc[id].m = std::min(n, pc[id].beg + max_block) - pc[id].beg;
      pc[id].sa = sa[id].p;
      pc[id].gap = gap[id].p;
    }
    final_merge_resident(ctx, pc, n, out_sa5_d, sink);
  }
  sync(ctx);
  ctx->stats.kernel_launches = ctx->launches;
}

}  // namespace psc

using namespace psc;

extern "C" {

const char *psc_last_error(void) { return g_err; }
int psc_version(void) { return 200; }
int psc_device_count(void) {
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess) { cudaGetLastError(); return 0; }
  return count;
}

int psc_init(int device, size_t hbm_budget, psc_context **out) {
  if (!out) { set_error("psc_init: null out pointer"); return PSC_EARG; }
  *out = nullptr;
  return guarded([&] {
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0) {
      cudaGetLastError();
      set_error("no CUDA device: %s (this library has no CPU fallback)",
                e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
      throw CudaError{PSC_ECUDA};
    }
    if (device < 0) PSC_CUDA(cudaGetDevice(&device));
    PSC_CUDA(cudaSetDevice(device));
    if (const char *e = getenv("PSC_L2_FETCH")) {
      size_t got = 0;
      cudaError_t rc = cudaDeviceSetLimit(cudaLimitMaxL2FetchGranularity, (size_t)atoi(e));
      cudaDeviceGetLimit(&got, cudaLimitMaxL2FetchGranularity);
      fprintf(stderr, "[psc] L2 fetch granularity: set %s -> %s, now %zu\n", e, cudaGetErrorName(rc), got);
    }
    psc_context *c = new psc_context();
    c->device = device;
    cudaDeviceProp prop;
    PSC_CUDA(cudaGetDeviceProperties(&prop, device));
    c->n_sm = prop.multiProcessorCount;
    // the pipeline's stream gets the highest priority: the block scheduler then places its CTAs before those of
    // work that only accompanies it (copies on the side stream, the multi-GPU helper's stream, dist.cu)
    int prio_least = 0, prio_greatest = 0;
    PSC_CUDA(cudaDeviceGetStreamPriorityRange(&prio_least, &prio_greatest));
    PSC_CUDA(cudaStreamCreateWithPriority(&c->stream, cudaStreamNonBlocking, prio_greatest));
    PSC_CUDA(cudaStreamCreateWithFlags(&c->side, cudaStreamNonBlocking));
    PSC_CUDA(cudaDeviceGetDefaultMemPool(&c->pool, device));
    uint64_t thr = UINT64_MAX;
    PSC_CUDA(cudaMemPoolSetAttribute(c->pool, cudaMemPoolAttrReleaseThreshold, &thr));
    c->hbm_budget = hbm_budget;
    if (const char *e = getenv("PSC_TRACE")) c->trace = atoi(e) != 0;
    memset(&c->stats, 0, sizeof(c->stats));
    *out = c;
  });
}

int psc_shutdown(psc_context *ctx) {
  if (!ctx) return PSC_OK;
  cudaSetDevice(ctx->device);
  psc_dist_finalize(ctx);  // no-op unless psc_dist_init was called and not undone
  hash_domain_clear(ctx);
  cache_trim(ctx);
  cudaStreamSynchronize(ctx->stream);
  cudaMemPoolTrimTo(ctx->pool, 0);
  cudaStreamDestroy(ctx->stream);
  cudaStreamDestroy(ctx->side);
  delete ctx;
  return PSC_OK;
}

/* external-memory mode (em.cu): block SAs and gap arrays go to disk, so only one block's working set, the merge
 * scratch of one output chunk and the walk's rank stream share the device with the text */
long psc_max_block_size_files(psc_context *ctx, long n) {
  if (!ctx || n <= 0) return 0;
  cudaSetDevice(ctx->device);
  const double chunk = (double)std::min(n, psc::merge_chunk_entries());
  double rest = psc::hbm_available(ctx) - 0.3 * (double)n - 32.0 * chunk - 8.0 * (double)std::min(n, 1L << 29);
  if (rest <= 0) return 0;
  long mb = (long)(rest / 48.0);
  mb = std::min(mb, (1L << 31) - 4096);
  mb = std::min(mb, n);
  return mb < n ? mb / 32 * 32 : mb;
}

long psc_max_block_size(psc_context *ctx, long n) {
  if (!ctx || n <= 0) return 0;
  cudaSetDevice(ctx->device);
  return psc::max_block_size_for(ctx, n, 6.0 * (double)n);
}

int psc_get_stats(psc_context *ctx, psc_stats *out) {
  if (!ctx || !out) { set_error("psc_get_stats: null pointer"); return PSC_EARG; }
  *out = ctx->stats;
  return PSC_OK;
}

int psc_construct_sa_d(psc_context *ctx, const uint8_t *text_d, long n, long max_block,
                       uint8_t *out_sa5_d) {
  return guarded([&] {
    arg_check(ctx && n >= 0 && (n == 0 || (text_d && out_sa5_d)), "psc_construct_sa_d");
    PSC_CUDA(cudaSetDevice(ctx->device));
    construct_sa_device(ctx, text_d, n, max_block, out_sa5_d);
    trace_report(ctx);
  });
}

int psc_construct_sa(psc_context *ctx, const uint8_t *text, long n, long max_block, uint8_t *out_sa5) {
  return guarded([&] {
    arg_check(ctx && n >= 0 && (n == 0 || (text && out_sa5)), "psc_construct_sa");
    if (n == 0) return;
    PSC_CUDA(cudaSetDevice(ctx->device));
    DevBuf<uint8_t> text_d(ctx, n);
    const long C = merge_chunk_entries();
    HostSink sink(ctx, out_sa5, std::min(n, C));
    max_block = choose_max_block(ctx, n, max_block);
    const long B = div_up(n, max_block);
    // the side stream must not start before earlier work on the main stream that used these buffers is done
    cudaEvent_t start;
    PSC_CUDA(cudaEventCreateWithFlags(&start, cudaEventDisableTiming));
    PSC_CUDA(cudaEventRecord(start, ctx->stream));
    PSC_CUDA(cudaStreamWaitEvent(ctx->side, start, 0));
    cudaEventDestroy(start);
    H2DGate gate;
    gate.ev.assign(B, nullptr);
    for (long id = B - 1; id >= 0; --id) {
      const long beg = id * max_block, end = std::min(n, beg + max_block);
      PSC_CUDA(cudaMemcpyAsync(text_d.p + beg, text + beg, (size_t)(end - beg), cudaMemcpyHostToDevice, ctx->side));
      PSC_CUDA(cudaEventCreateWithFlags(&gate.ev[id], cudaEventDisableTiming));
      PSC_CUDA(cudaEventRecord(gate.ev[id], ctx->side));
    }
    try {
      construct_sa_sequential(ctx, text_d.p, n, max_block, nullptr, &sink, &gate);
    } catch (...) {  // the side stream may still be writing into buffers this frame owns
      cudaStreamSynchronize(ctx->side);
      throw;
    }
    sink.finish(ctx);
    trace_report(ctx);
  });
}

int psc_device_alloc(psc_context *ctx, size_t bytes, void **ptr_d) {
  return guarded([&] {
    arg_check(ctx && ptr_d, "psc_device_alloc");
    PSC_CUDA(cudaSetDevice(ctx->device));
    PSC_CUDA(cudaMalloc(ptr_d, bytes + 256));
  });
}
int psc_device_free(psc_context *ctx, void *ptr_d) {
  return guarded([&] { arg_check(ctx != nullptr, "psc_device_free"); PSC_CUDA(cudaSetDevice(ctx->device)); PSC_CUDA(cudaFree(ptr_d)); });
}
int psc_pinned_alloc(psc_context *ctx, size_t bytes, void **ptr) {
  return guarded([&] { arg_check(ctx && ptr, "psc_pinned_alloc"); PSC_CUDA(cudaSetDevice(ctx->device)); PSC_CUDA(cudaMallocHost(ptr, bytes)); });
}
int psc_pinned_free(psc_context *ctx, void *ptr) {
  return guarded([&] { arg_check(ctx != nullptr, "psc_pinned_free"); PSC_CUDA(cudaFreeHost(ptr)); });
}
int psc_copy_h2d(psc_context *ctx, void *dst_d, const void *src, size_t bytes) {
  return guarded([&] {
    arg_check(ctx && (bytes == 0 || (dst_d && src)), "psc_copy_h2d");
    PSC_CUDA(cudaSetDevice(ctx->device));
    PSC_CUDA(cudaMemcpyAsync(dst_d, src, bytes, cudaMemcpyHostToDevice, ctx->stream));
    sync(ctx);
  });
}
int psc_copy_d2h(psc_context *ctx, void *dst, const void *src_d, size_t bytes) {
  return guarded([&] {
    arg_check(ctx && (bytes == 0 || (dst && src_d)), "psc_copy_d2h");
    PSC_CUDA(cudaSetDevice(ctx->device));
    PSC_CUDA(cudaMemcpyAsync(dst, src_d, bytes, cudaMemcpyDeviceToHost, ctx->stream));
    sync(ctx);
  });
}
int psc_device_info(psc_context *ctx, char *name, size_t name_len, size_t *free_bytes, size_t *total_bytes) {
  return guarded([&] {
    arg_check(ctx != nullptr, "psc_device_info");
    PSC_CUDA(cudaSetDevice(ctx->device));
    cudaDeviceProp prop;
    PSC_CUDA(cudaGetDeviceProperties(&prop, ctx->device));
    if (name && name_len) snprintf(name, name_len, "%s", prop.name);
    size_t f = 0, t = 0;
    PSC_CUDA(cudaMemGetInfo(&f, &t));
    if (free_bytes) *free_bytes = f;
    if (total_bytes) *total_bytes = t;
  });
}

}  // extern "C"
