// texthash.cu -- builds the sampled prefix fingerprints of texthash.cuh.
// H[x+1] = H[x] * B + (text[x] + 1) mod 2^61 - 1.  Appending a chunk of 16 symbols is the affine map
// h -> h * B^16 + A(chunk); prefix fingerprints at the chunk boundaries are a scan under composition of such
// maps, (m1, a1) o (m2, a2) = (m1 m2, a1 m2 + a2), whose first component is the power table for free.
// Three kernels (tile aggregates, scan of the aggregates by one block, apply), two reads of the text.
#include <random>
#include "texthash.cuh"

namespace psc {

struct TextHashStore {
  const uint8_t *base = nullptr;
  long n = 0;
  bool built = false;
  DevBuf<uint64_t> pre, pw;
  uint64_t bpow[2][TH_SAMPLE + 1];
};

struct Aff { uint64_t m0, a0, m1, a1; };  // one affine map per base

__device__ __forceinline__ Aff aff_compose(const Aff &x, const Aff &y) {  // x first, then y
  Aff r;
  r.m0 = th_mulmod(x.m0, y.m0); r.a0 = th_addmod(th_mulmod(x.a0, y.m0), y.a0);
  r.m1 = th_mulmod(x.m1, y.m1); r.a1 = th_addmod(th_mulmod(x.a1, y.m1), y.a1);
  return r;
}

constexpr int TH_THREADS = 256, TH_PER_THREAD = 4, TH_TILE = TH_THREADS * TH_PER_THREAD;  // entries per block

struct Bases { uint64_t b0, b1, b0_16, b1_16; };

__device__ __forceinline__ Aff chunk_map(const uint8_t *__restrict__ text, long c, const Bases &bs) {
  Aff r{bs.b0_16, 0, bs.b1_16, 0};
  const uint4 v = *reinterpret_cast<const uint4 *>(text + (c << TH_SAMPLE_LOG));  // the domain base is 16-byte aligned
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int j = 0; j < 16; ++j) {
    const uint64_t s = (uint64_t)((w[j >> 2] >> (8 * (j & 3))) & 255u) + 1;
    r.a0 = th_addmod(th_mulmod(r.a0, bs.b0), s);
    r.a1 = th_addmod(th_mulmod(r.a1, bs.b1), s);
  }
  return r;
}

// inclusive scan of one map per thread over the block (Hillis-Steele in shared memory)
template <int THREADS>
__device__ __forceinline__ Aff block_scan_aff(Aff v, Aff *sh) {
  sh[threadIdx.x] = v;
  __syncthreads();
  for (int d = 1; d < THREADS; d <<= 1) {
    Aff t = sh[threadIdx.x];
    if ((int)threadIdx.x >= d) t = aff_compose(sh[threadIdx.x - d], t);
    __syncthreads();
    sh[threadIdx.x] = t;
    __syncthreads();
  }
  return sh[threadIdx.x];
}

// entry e (0 <= e <= ns) = composition of chunks [0, e).  A thread owns entries [first, first + 4).
template <bool APPLY>
__global__ void __launch_bounds__(TH_THREADS)
    th_tile_kernel(const uint8_t *__restrict__ text, long ns, Bases bs, Aff *__restrict__ agg,
                   uint64_t *__restrict__ pre, uint64_t *__restrict__ pw) {
  __shared__ Aff sh[TH_THREADS];
  const long first = ((long)blockIdx.x * TH_THREADS + threadIdx.x) * TH_PER_THREAD;
  Aff loc{1, 0, 1, 0};
  Aff cm[TH_PER_THREAD];
#pragma unroll
  for (int k = 0; k < TH_PER_THREAD; ++k) {
    const long c = first + k;
    if (c < ns) { cm[k] = chunk_map(text, c, bs); loc = aff_compose(loc, cm[k]); }
  }
  const Aff incl = block_scan_aff<TH_THREADS>(loc, sh);
  if (!APPLY) {
    if (threadIdx.x == TH_THREADS - 1) agg[blockIdx.x] = incl;
    return;
  }
  // exclusive prefix of this thread = tile prefix o (inclusive of the previous thread)
  Aff cur = agg[blockIdx.x];  // after the aggregate scan: composition of all earlier tiles
  if (threadIdx.x > 0) cur = aff_compose(cur, sh[threadIdx.x - 1]);
#pragma unroll
  for (int k = 0; k < TH_PER_THREAD; ++k) {
    const long e = first + k;
    if (e > ns) break;
    pre[e] = cur.a0; pre[ns + 1 + e] = cur.a1;
    pw[e] = cur.m0; pw[ns + 1 + e] = cur.m1;
    if (e < ns) cur = aff_compose(cur, cm[k]);
  }
}

// aggregates -> exclusive prefixes, in place, by one block
__global__ void __launch_bounds__(1024) th_scan_tiles_kernel(Aff *agg, long ntiles) {
  __shared__ Aff sh[1024];
  const long run = (ntiles + 1023) / 1024;
  const long lo = (long)threadIdx.x * run, hi = min(ntiles, lo + run);
  Aff loc{1, 0, 1, 0};
  for (long i = lo; i < hi; ++i) loc = aff_compose(loc, agg[i]);
  block_scan_aff<1024>(loc, sh);
  Aff cur{1, 0, 1, 0};
  if (threadIdx.x > 0) cur = sh[threadIdx.x - 1];
  for (long i = lo; i < hi; ++i) {
    const Aff a = agg[i];
    agg[i] = cur;
    cur = aff_compose(cur, a);
  }
}

static uint64_t host_mulmod(uint64_t a, uint64_t b) {
  unsigned __int128 p = (unsigned __int128)a * b;
  uint64_t r = (uint64_t)(p & TH_P) + (uint64_t)(p >> 61);
  r = (r & TH_P) + (r >> 61);
  return r >= TH_P ? r - TH_P : r;
}

long lcp_cap() {
  static long cap = 0;
  if (!cap) {
    const char *e = getenv("PSC_LCP_CAP");
    cap = e ? std::max(4L, atol(e) / 4 * 4) : 1024;
  }
  return cap;
}

void hash_domain_clear(psc_context *ctx) {
  if (!ctx->thash) return;
  delete ctx->thash;
  ctx->thash = nullptr;
}

void hash_domain_set(psc_context *ctx, const uint8_t *base_d, long n) {
  if (ctx->thash && ctx->thash->base == base_d && ctx->thash->n == n && !ctx->thash->built) return;
  hash_domain_clear(ctx);  // the text behind a pointer may have changed: never reuse built tables
  if (!base_d || n <= 0) return;
  ctx->thash = new TextHashStore();
  ctx->thash->base = base_d;
  ctx->thash->n = n;
}

bool hash_domain_ensure(psc_context *ctx, const uint8_t *lo, const uint8_t *hi) {
  TextHashStore *st = ctx->thash;
  if (st && lo >= st->base && hi <= st->base + st->n && ((uintptr_t)st->base & 15) == 0) return true;
  if (((uintptr_t)lo & 15) != 0 || hi <= lo) return false;
  hash_domain_set(ctx, lo, (long)(hi - lo));
  return true;
}

static void build(psc_context *ctx, TextHashStore *st) {
  KTrace kt(ctx, "text_fingerprints");
  // bases: random per process (PSC_HASH_SEED pins them), in [2^20, P - 2]
  static uint64_t b[2] = {0, 0};
  if (!b[0]) {
    uint64_t seed;
    if (const char *e = getenv("PSC_HASH_SEED")) seed = strtoull(e, nullptr, 10);
    else { std::random_device rd; seed = ((uint64_t)rd() << 32) ^ rd() ^ (uint64_t)std::chrono::steady_clock::now().time_since_epoch().count(); }
    for (int h = 0; h < 2; ++h) {
      seed += 0x9E3779B97F4A7C15ull;
      uint64_t z = seed;
      z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull; z = (z ^ (z >> 27)) * 0x94D049BB133111EBull; z ^= z >> 31;
      b[h] = (1ull << 20) + z % (TH_P - (1ull << 21));
    }
  }
  for (int h = 0; h < 2; ++h) {
    st->bpow[h][0] = 1;
    for (int j = 1; j <= TH_SAMPLE; ++j) st->bpow[h][j] = host_mulmod(st->bpow[h][j - 1], b[h]);
  }
  const long ns = st->n >> TH_SAMPLE_LOG;
  st->pre.alloc(ctx, 2 * (size_t)(ns + 1));
  st->pw.alloc(ctx, 2 * (size_t)(ns + 1));
  const long ntiles = div_up(ns + 1, (long)TH_TILE);
  DevBuf<Aff> agg(ctx, ntiles);
  Bases bs{b[0], b[1], st->bpow[0][TH_SAMPLE], st->bpow[1][TH_SAMPLE]};
  th_tile_kernel<false><<<(unsigned)ntiles, TH_THREADS, 0, ctx->stream>>>(st->base, ns, bs, agg.p, nullptr, nullptr);
  PSC_LAUNCH_CHECK();
  th_scan_tiles_kernel<<<1, 1024, 0, ctx->stream>>>(agg.p, ntiles);
  PSC_LAUNCH_CHECK();
  th_tile_kernel<true><<<(unsigned)ntiles, TH_THREADS, 0, ctx->stream>>>(st->base, ns, bs, agg.p, st->pre.p, st->pw.p);
  PSC_LAUNCH_CHECK();
  ctx->launches += 3;
  st->built = true;
}

bool hash_view(psc_context *ctx, const uint8_t *lo, const uint8_t *hi, bool do_build, HashView *hv) {
  TextHashStore *st = ctx->thash;
  if (!st || lo < st->base || hi > st->base + st->n) return false;
  if (((uintptr_t)st->base & 15) != 0) return false;  // chunk loads are 16-byte vectors
  if (!st->built) {
    if (!do_build) return false;
    build(ctx, st);
  }
  hv->text = st->base;
  hv->n = st->n;
  hv->ns = st->n >> TH_SAMPLE_LOG;
  hv->pre = st->pre.p;
  hv->pw = st->pw.p;
  memcpy(hv->bpow, st->bpow, sizeof(st->bpow));
  return true;
}

}  // namespace psc
