// common.cuh -- context, error handling and device-buffer helpers shared by all kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <chrono>
#include <mutex>
#include <vector>
#include "../../include/psascan_cuda.h"

namespace psc {

void set_error(const char *fmt, ...);

struct CudaError { int code; };

#define PSC_CUDA(expr)                                                                  \
  do {                                                                                  \
    cudaError_t e__ = (expr);                                                           \
    if (e__ != cudaSuccess) {                                                           \
      psc::set_error("%s:%d: %s: %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e__)); \
      throw psc::CudaError{PSC_ECUDA};                                                  \
    }                                                                                   \
  } while (0)

#define PSC_LAUNCH_CHECK() PSC_CUDA(cudaGetLastError())

static inline long div_up(long a, long b) { return (a + b - 1) / b; }
// the merged SA keeps a high-byte plane only for texts above 4 GiB; PSC_FORCE_HI=1 turns it on for small
// texts (tests: the 40-bit paths of merge and writer)
static inline bool need_hi_plane(long n) { return n > (1L << 32) || getenv("PSC_FORCE_HI") != nullptr; }

}  // namespace psc

namespace psc { struct Comm; struct TextHashStore; }

struct psc_context {
  int device = 0;
  psc::TextHashStore *thash = nullptr;  // texthash.cu: fingerprints of the text of the running construction
  psc::Comm *comm = nullptr;  // psc_dist_init (comm.cu): this context is one rank of a multi-GPU construction
  psc_context *helper = nullptr;  // psc_dist_init: second stream + buffer cache for the thread that prepares the
                                  // next blocks during a walk (dist.cu, BlockFeeder)
  int n_sm = 148;
  size_t hbm_budget = 0;
  cudaStream_t stream = nullptr;
  cudaStream_t side = nullptr;
  cudaMemPool_t pool = nullptr;
  psc_stats stats;
  long launches = 0;
  std::vector<std::pair<cudaEvent_t, cudaEvent_t>> ev_free;
  // PSC_TRACE=1: device time per kernel family, printed to stderr after each construction
  struct TraceRow { const char *name; long count; double ms; };
  bool trace = false;
  std::vector<TraceRow> trace_rows;
  double host_alloc_ms = 0;  // host time inside cudaMallocAsync / cudaFreeAsync (trace only)
  long n_allocs = 0;
  // Device buffers released by the pipeline are kept here and handed out again by size class:
  // the driver's pool took 0.2-1 ms of host time per call once block sizes started to vary
  // (measured 76-390 ms per 1 GiB / 8 block construction), all work is on one stream so reuse
  // needs no synchronisation.
  struct CacheBlock { void *p; size_t bytes; bool used; };
  std::vector<CacheBlock> cache;
  size_t cache_free_bytes = 0;
  std::mutex cache_mu;  // buffers of the sorter context are released by the streaming thread
  int walk_threads_per_sm = 512;  // measured on 8 x 128 MiB: 512 -> 98 ms, 1024 -> 111, 2048 -> 104, 256 -> 123
};

namespace psc {

// api.cu: size-class cache in front of cudaMallocAsync (throws CudaError on failure)
void *cache_alloc(psc_context *ctx, size_t bytes);
void cache_release(psc_context *ctx, void *p);
void cache_trim(psc_context *ctx);  // give every unused cached buffer back to the pool

// Stream-ordered device buffer (cudaMallocAsync pool: freed blocks are reused without
// going back to the driver).
template <typename T>
struct DevBuf {
  psc_context *ctx = nullptr;
  T *p = nullptr;
  size_t n = 0;
  DevBuf() {}
  DevBuf(psc_context *c, size_t count) { alloc(c, count); }
  DevBuf(const DevBuf &) = delete;
  DevBuf &operator=(const DevBuf &) = delete;
  DevBuf(DevBuf &&o) noexcept : ctx(o.ctx), p(o.p), n(o.n) { o.p = nullptr; o.n = 0; }
  DevBuf &operator=(DevBuf &&o) noexcept {
    if (this != &o) { release(); ctx = o.ctx; p = o.p; n = o.n; o.p = nullptr; o.n = 0; }
    return *this;
  }
  void alloc(psc_context *c, size_t count) {
    release();
    ctx = c;
    n = count;
    size_t bytes = (count ? count : 1) * sizeof(T) + 256;  // slack: kernels may read whole vectors
    p = (T *)cache_alloc(c, bytes);
  }
  void zero() { PSC_CUDA(cudaMemsetAsync(p, 0, (n ? n : 1) * sizeof(T) + 256, ctx->stream)); }
  void release() {
    if (p) { cache_release(ctx, p); p = nullptr; n = 0; }
  }
  ~DevBuf() { release(); }
  operator T *() const { return p; }
};

// CUDA-event phase timer accumulating into a psc_stats field.
struct PhaseTimer {
  psc_context *ctx;
  double *acc;
  cudaEvent_t a, b;
  PhaseTimer(psc_context *c, double *field) : ctx(c), acc(field) {
    cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a, c->stream);
  }
  ~PhaseTimer() {
    cudaEventRecord(b, ctx->stream);
    cudaEventSynchronize(b);
    float ms = 0;
    if (cudaEventElapsedTime(&ms, a, b) == cudaSuccess) *acc += ms;
    cudaEventDestroy(a); cudaEventDestroy(b);
  }
};

// scoped per-kernel timer, active only under PSC_TRACE=1 (it synchronises the stream)
struct KTrace {
  psc_context *ctx;
  const char *name;
  cudaEvent_t a, b;
  KTrace(psc_context *c, const char *n) : ctx(c), name(n) {
    if (!ctx->trace) return;
    cudaEventCreate(&a); cudaEventCreate(&b);
    cudaEventRecord(a, ctx->stream);
  }
  ~KTrace() {
    if (!ctx->trace) return;
    cudaEventRecord(b, ctx->stream);
    cudaEventSynchronize(b);
    float ms = 0;
    cudaEventElapsedTime(&ms, a, b);
    cudaEventDestroy(a); cudaEventDestroy(b);
    for (auto &r : ctx->trace_rows)
      if (r.name == name) { r.count++; r.ms += ms; return; }
    ctx->trace_rows.push_back({name, 1, (double)ms});
  }
};
void trace_report(psc_context *ctx);
// device-side work guard (see WorkBudget below): allocate + arm, and check after the kernel
struct WorkBudget;
WorkBudget *work_budget_arm(psc_context *ctx);
int work_budget_check(psc_context *ctx, WorkBudget *wb, const char *what);  // returns the need_hash flag

template <typename T>
static inline void h2d(psc_context *ctx, T *dst, const T *src, size_t n) {
  if (n) PSC_CUDA(cudaMemcpyAsync(dst, src, n * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
}
template <typename T>
static inline void d2h(psc_context *ctx, T *dst, const T *src, size_t n) {
  if (n) PSC_CUDA(cudaMemcpyAsync(dst, src, n * sizeof(T), cudaMemcpyDeviceToHost, ctx->stream));
}
static inline void sync(psc_context *ctx) { PSC_CUDA(cudaStreamSynchronize(ctx->stream)); }

#ifdef __CUDACC__
// relaxed device-scope accesses for look-back chains (status word = flag bits | payload)
__device__ __forceinline__ uint64_t ld_relaxed_u64(const uint64_t *p) {
  uint64_t v;
  asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_relaxed_u64(uint64_t *p, uint64_t v) {
  asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ bool bit_get(const uint32_t *bits, long i) {
  return (bits[i >> 5] >> (i & 31)) & 1u;
}
// 4 bytes from any address (little endian): reads the aligned word(s) that hold p[0..3]
__device__ __forceinline__ uint32_t load_u32_any(const uint8_t *p) {
  uintptr_t a = (uintptr_t)p;
  const uint32_t *w = (const uint32_t *)(a & ~(uintptr_t)3);
  uint32_t sh = (uint32_t)(a & 3) * 8;
  uint32_t w0 = w[0];
  if (sh == 0) return w0;
  return __funnelshift_r(w0, w[1], sh);
}
// ORs `nvalid` low bits of `word` into a zero-initialised bit array at bit index `bitpos`.
__device__ __forceinline__ void store_bits_word(uint32_t *bits, long bitpos, uint32_t word, int nvalid) {
  if (nvalid <= 0) return;
  if (nvalid < 32) word &= (1u << nvalid) - 1u;
  int sh = (int)(bitpos & 31);
  long w = bitpos >> 5;
  if (sh == 0 && nvalid == 32) { bits[w] = word; return; }
  if (word << sh) atomicOr(&bits[w], word << sh);
  if (sh && (word >> (32 - sh))) atomicOr(&bits[w + 1], word >> (32 - sh));
}
// Work guard for the two brute-force comparisons of the path (K0 and the initial-rank binary search):
// they are linear in the LCP between a block suffix and the text after the block, which is short on
// every text but degenerate ones (period far below the block length).  Every 4 KiB of matched
// symbols are charged to a device counter; past the limit the kernels bail out and the host reports
// PSC_ELIMIT instead of hanging the GPU for hours.
struct WorkBudget { unsigned long long used, limit; int exceeded; int need_hash; };

// length of the common prefix of a[0..lim) and b[0..lim).  Whole words are only loaded while four symbols
// remain, the rest goes byte by byte: nothing past a + lim or b + lim is ever read beyond the aligned word
// that holds the last compared byte (callers pass texts of exactly n bytes).
__device__ __forceinline__ long common_prefix(const uint8_t *a, const uint8_t *b, long lim) {
  long el = 0;
  while (el + 4 <= lim) {
    uint32_t x = load_u32_any(a + el) ^ load_u32_any(b + el);
    if (x) return el + ((__ffs(x) - 1) >> 3);
    el += 4;
  }
  while (el < lim && a[el] == b[el]) ++el;
  return el;
}

__device__ __forceinline__ long common_prefix_guarded(const uint8_t *a, const uint8_t *b, long lim, WorkBudget *wb) {
  long el = 0;
  while (el < lim) {
    const long stop = min(lim, el + 4096);
    const long got = common_prefix(a + el, b + el, stop - el);
    el += got;
    if (el < stop) return el;
    if (el < lim) {
      if (atomicAdd(&wb->used, 4096ull) > wb->limit || *(volatile int *)&wb->exceeded) { wb->exceeded = 1; return -1; }
    }
  }
  return el;
}
#endif

// ---- device-level stage functions (implemented in the .cu files) ------------------------

// radix.cu: LSD onesweep sort of (key64, val32) pairs on the digit list `shifts` (8-bit
// digits, least significant first).  Result pointers are returned in *keys_out / *vals_out
// (one of the two buffer pairs).  Passes whose digit is constant over all keys are skipped.
// vals_a == nullptr: keys only (the caller keeps its payload in the low key bits).
// kg != nullptr: the first digit pass builds the keys from the text instead of reading keys_a.
struct KeyGen {
  const uint8_t *text;     // m symbols
  const uint32_t *gt_end;  // renaming bits (may be null)
  long m;
  int last, code_last_high;
  int dshift;  // digit of the first pass = (B >> dshift) & 255, B = the 8 renamed codes at the position, MSB first
  int kshift;  // stored key = B >> kshift ...
  int vbits;   // ... << vbits | position when vbits > 0
  uint8_t code[256];  // 8-bit renamed codes
};
void radix_sort_pairs(psc_context *ctx, uint64_t *keys_a, uint32_t *vals_a, uint64_t *keys_b,
                      uint32_t *vals_b, long n, const int *shifts, int n_shifts,
                      uint64_t **keys_out, uint32_t **vals_out,
                      const unsigned long long *pre_hist_d = nullptr /* [8][256] digit counts already known */,
                      bool vals_iota = false /* values are 0..n-1: vals_a is not read */,
                      const KeyGen *kg = nullptr);

// scan.cu
// out[0] = 0, out[j] = in[0] + .. + in[j-1] - sub_after_first for j > 0 (the chunk-local merge drops the
// first `sub_after_first` entries of its first run, see finalmerge.cu)
void exclusive_scan_u32_to_u64(psc_context *ctx, const uint32_t *in, uint64_t *out, long n,
                               uint64_t *total_d /* device, may be null */, uint64_t sub_after_first = 0);
constexpr int COARSE_TILE = 2048;  // = SCAN_TILE (scan.cu)
long coarse_carry_tiles(long n);
void coarse_carries_u32(psc_context *ctx, const uint32_t *in, long n, uint64_t *carry_d /* tiles + 1 */);

// sufsort.cu
struct BlockSortResult {
  DevBuf<uint32_t> sa;    // m
  DevBuf<uint32_t> isa;   // m, only allocated when doubling rounds were needed
  long i0 = 0;
};
// text_d: the block (m symbols, followed in memory by the tail when after_len > 0 is served
// from the same buffer; `after_d` points at the text right after the block).
// gt_end_bits_d: forward bit array over block positions p in [0,m): suffix(p) > suffix(end);
// null when the block is the last one (then every bit is 1).
void sort_block(psc_context *ctx, const uint8_t *text_d, long m, const uint32_t *gt_end_bits_d,
                BlockSortResult *res);
// gt_end by matching the block against the text after it (K0).
//   after_d[after_len]  text following the block; after_is_text_end: 1 if it reaches the text end
//   gt_after_bits_d     forward bits: bit gt_after_off+q (q in [0, after_len)) <-> suffix(end+q+1) > suffix(end)
void compute_gt_end(psc_context *ctx, const uint8_t *block_d, long m, const uint8_t *after_d,
                    long after_len, int after_is_text_end, const uint32_t *gt_after_bits_d,
                    long gt_after_off, uint32_t *gt_end_bits_d);
// K3: BWT (0 at i0) and forward gt_begin bits: bit p-1 (p in [1,m)) <-> rank(p) > i0.
void bwt_from_sa(psc_context *ctx, const uint8_t *text_d, const uint32_t *sa_d, long m,
                 uint8_t *bwt_d);
void gt_begin_from_sa(psc_context *ctx, const uint32_t *sa_d, long m, uint32_t i0,
                      uint32_t *bits_d, long bit_off /* bit bit_off+q <-> position q+1, q in [0,m-1);
                      the array must be zero there */);

// bit-vector layout conversion: forward bits (bit q <-> position lo+q+1) <-> reversed LSB-first
// bytes with hi = lo + nbits (bit r <-> position hi - r).
void bits_forward_to_rev(psc_context *ctx, const uint32_t *fwd_d, long nbits, uint8_t *rev_d,
                         long fwd_off = 0 /* the range starts at bit fwd_off of fwd_d */);
void bits_rev_to_forward(psc_context *ctx, const uint8_t *rev_d, long nbits, uint32_t *fwd_d);

}  // namespace psc

// rank.cu
struct psc_rank {
  long m = 0;
  long i0 = -1;
  int n_codes = 0;           // distinct symbols present (placeholder excluded)
  long n_sectors = 0;        // sectors per symbol
  int16_t code_of[256];      // -1: absent
  long counts[256];
  uint32_t *sectors_d = nullptr;  // [n_codes][n_sectors][8] uint32
  psc_context *ctx = nullptr;
};

namespace psc {
psc_rank *rank_build(psc_context *ctx, const uint8_t *bwt_d, long m, long i0);
void rank_free(psc_rank *r);
void rank_query(psc_context *ctx, const psc_rank *r, long nq, const uint64_t *i_d,
                const uint8_t *c_d, uint64_t *out_d);

// walk.cu
// K5: ranks of patterns among block suffixes.  after_d = text after the block.
void initial_ranks(psc_context *ctx, const uint8_t *block_d, const uint32_t *sa_d, long m,
                   const uint8_t *after_d, long after_len, int after_is_text_end,
                   const uint32_t *gt_after_bits_d, long gt_after_off, long nstarts,
                   const uint64_t *starts_d, uint64_t *out_d);
// K6: LF walk.  Cmap[c] = number of block symbols < c.  gt bits forward over the tail:
// bit q <-> tail position q+1.  The walkers cover tail positions [lo, lo+t).
void lf_walk(psc_context *ctx, const psc_rank *r, uint32_t i0, uint8_t last, const uint64_t *Cmap,
             const uint8_t *tail_d, long t, long lo, long nwalk, const uint64_t *walk_beg_d,
             const uint64_t *start_rank_d, const uint32_t *gt_in_bits_d, long gt_in_off,
             uint32_t *gt_out_bits_d, long gt_out_off, uint32_t *gap_d);

// walkers over a range of tail positions (walk.cu)
struct WalkPlan {
  std::vector<uint64_t> b;     // W + 1 boundaries, tail-relative
  std::vector<long> chunk_at;  // walker index at which every chunk starts, then W
};
void make_walk_plan(psc_context *ctx, long lo, long hi, WalkPlan *p);
void walk_with_plan(psc_context *ctx, const psc_rank *rk, long i0, uint8_t last, const uint64_t *Cmap,
                    const uint8_t *after_d, const WalkPlan &p, const uint64_t *b_d, const uint64_t *sr_d,
                    const uint32_t *gt_cur, long gt_off, uint32_t *gt_new, uint32_t *gap_d);
// K5 + K6 over tail positions [lo, hi), chunked (walk.cu)
void walk_tail_range(psc_context *ctx, const psc_rank *rk, long i0, uint8_t last, const uint64_t *Cmap,
                     const uint8_t *block_d, const uint32_t *sa_d, long m, const uint8_t *after_d, long t,
                     long lo, long hi, const uint32_t *gt_cur, long gt_off, uint32_t *gt_new, uint32_t *gap_d);

// merge.cu: planes lo32/hi8 for the running merged SA.
void merge_level(psc_context *ctx, const uint32_t *sa_d, const uint32_t *gap_d, long m, long beg,
                 const uint32_t *right_lo, const uint8_t *right_hi, long right_len,
                 uint32_t *out_lo, uint8_t *out_hi);
// pieces of one level for callers that own only a range of the output slots (dist.cu): E = exclusive scan
// of the gap array (m + 1 entries); merge_bounds: number of block suffixes in slots below each given slot
void merge_gap_scan(psc_context *ctx, const uint32_t *gap_d, long m, uint64_t *E_d);
void merge_bounds(psc_context *ctx, const uint64_t *E_d, long m, const uint64_t *slots, int nq, uint64_t *before);
void sa_slice_to_planes(psc_context *ctx, const uint32_t *sa_d, long count, long beg, uint32_t *out_lo, uint8_t *out_hi);
void merge_level_range(psc_context *ctx, const uint32_t *sa_d, const uint64_t *E_d, long m, long beg,
                       const uint32_t *right_lo, const uint8_t *right_hi, long r_base, long q_beg, long q_end,
                       uint32_t *out_lo, uint8_t *out_hi);
void pack_sa5(psc_context *ctx, const uint32_t *lo, const uint8_t *hi, long n, uint8_t *out5);
void pack_sa5_from_sa(psc_context *ctx, const uint32_t *sa, long n, long beg, uint8_t *out5, bool timed = true);
void unpack_sa5(psc_context *ctx, const uint8_t *in5, long n, uint32_t *lo, uint8_t *hi);
void gap_to_bitvector(psc_context *ctx, const uint32_t *gap_l, long len_l, long len_r,
                      uint32_t *bv_words);
void merge_bwt(psc_context *ctx, const uint8_t *bwt_l, const uint8_t *bwt_r, long len_l, long len_r,
               long i0_l, long i0_r, uint8_t last_l, const uint32_t *bv_words, uint8_t *bwt_out,
               long *i0_out);
void split_gap(psc_context *ctx, const uint32_t *bv_words, long len_l, long len_r,
               const uint32_t *block_gap, uint64_t *left_gap, uint64_t *right_gap);
}  // namespace psc
