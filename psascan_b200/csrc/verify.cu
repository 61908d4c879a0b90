// verify.cu -- O(n) suffix-array check on the device (the logic of sufcheck(),
// /root/reference/extlib/libdivsufsort/lib/utils.c:161-241): the entries are a permutation of
// [0, n), consecutive suffixes are ordered by their first symbol, and suffixes with the same first
// symbol are ordered like the suffixes that follow them (through the inverse array).  Used by
// bench.py after the timed loop and by the parity tests at sizes no CPU sorter reaches.
#include "common.cuh"
#include "texthash.cuh"

namespace psc {

__device__ __forceinline__ uint64_t load_sa5(const uint8_t *__restrict__ sa5, long i) {
  const uint8_t *p = sa5 + i * 5;
  return (uint64_t)p[0] | ((uint64_t)p[1] << 8) | ((uint64_t)p[2] << 16) | ((uint64_t)p[3] << 24) | ((uint64_t)p[4] << 32);
}

// isa[v] = i + 1 (0 = not seen); viol[0] = entries >= n, viol[1] = values seen twice
__global__ void __launch_bounds__(256)
    verify_scatter_kernel(const uint8_t *__restrict__ sa5, long n, uint32_t *__restrict__ isa,
                          unsigned long long *__restrict__ viol) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t v = load_sa5(sa5, i);
  if (v >= (uint64_t)n) { atomicAdd(&viol[0], 1ull); return; }
  if (atomicExch(&isa[v], (uint32_t)(i + 1)) != 0u) atomicAdd(&viol[1], 1ull);
}

// viol[2] = first symbols out of order, viol[3] = equal first symbols whose successors are out of order
__global__ void __launch_bounds__(256)
    verify_order_kernel(const uint8_t *__restrict__ text, const uint8_t *__restrict__ sa5, long n,
                        const uint32_t *__restrict__ isa, unsigned long long *__restrict__ viol) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i + 1 >= n) return;
  uint64_t a = load_sa5(sa5, i), b = load_sa5(sa5, i + 1);
  if (a >= (uint64_t)n || b >= (uint64_t)n) return;
  uint8_t ca = text[a], cb = text[b];
  if (ca > cb) { atomicAdd(&viol[2], 1ull); return; }
  if (ca == cb) {
    // rank + 1 of the following suffix; the empty suffix (position n) is the smallest: 0
    uint32_t ra = a + 1 < (uint64_t)n ? isa[a + 1] : 0u, rb = b + 1 < (uint64_t)n ? isa[b + 1] : 0u;
    if (!(ra < rb)) atomicAdd(&viol[3], 1ull);
  }
}

}  // namespace psc
using namespace psc;

extern "C" int psc_verify_sa_d(psc_context *ctx, const uint8_t *text_d, long n, const uint8_t *sa5_d,
                               long violations[4]) {
  try {
    if (!ctx || n < 0 || !violations || (n > 0 && (!text_d || !sa5_d)) || n >= (1L << 32) - 1) {
      set_error("psc_verify_sa_d: bad argument (texts of 2^32 - 1 symbols and more are not supported)");
      return PSC_EARG;
    }
    for (int k = 0; k < 4; ++k) violations[k] = 0;
    if (n == 0) return PSC_OK;
    PSC_CUDA(cudaSetDevice(ctx->device));
    DevBuf<uint32_t> isa(ctx, n);
    DevBuf<unsigned long long> viol(ctx, 4);
    isa.zero();
    viol.zero();
    verify_scatter_kernel<<<(unsigned)div_up(n, 256), 256, 0, ctx->stream>>>(sa5_d, n, isa.p, viol.p);
    PSC_LAUNCH_CHECK();
    verify_order_kernel<<<(unsigned)div_up(n, 256), 256, 0, ctx->stream>>>(text_d, sa5_d, n, isa.p, viol.p);
    PSC_LAUNCH_CHECK();
    unsigned long long h[4];
    d2h(ctx, h, viol.p, 4);
    sync(ctx);
    for (int k = 0; k < 4; ++k) violations[k] = (long)h[k];
    return PSC_OK;
  } catch (const CudaError &e) { return e.code; }
}

// ---- check of a SLICE of a suffix array, for arrays spread over several GPUs and texts beyond 2^32 ----------
// Entries [slice_beg, slice_beg + slice_len) as 5-byte values, plus the first entry of the next slice.  Checked:
// every entry < n, and suffix(sa[i]) < suffix(sa[i+1]) for every adjacent pair, by comparing the two suffixes
// directly (at most PSC_LCP_CAP symbols, then fingerprints: texthash.cuh).  n in-range entries in strictly
// increasing suffix order are all different, hence a permutation, hence THE suffix array: no inverse array, no
// exchange between the ranks.  violations[0] = entries >= n, [1] = adjacent pairs out of order.
__global__ void __launch_bounds__(256)
    verify_slice_kernel(const uint8_t *__restrict__ text, long n, const uint8_t *__restrict__ sa5, long len,
                        long next_first, unsigned long long *__restrict__ viol, WorkBudget *wb,
                        const __grid_constant__ HashView hv, long cap) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= len) return;
  const uint64_t a = load_sa5(sa5, i);
  if (a >= (uint64_t)n) { atomicAdd(&viol[0], 1ull); return; }
  long b = i + 1 < len ? (long)load_sa5(sa5, i + 1) : next_first;
  if (b < 0) return;                       // the very last entry of the array
  if (b >= n) return;                      // counted by its own thread (or by the next slice)
  if ((long)a == b) { atomicAdd(&viol[1], 1ull); return; }
  const long lim = n - max((long)a, b);
  long el = common_prefix(text + a, text + b, min(lim, cap));
  if (el == cap && cap < lim) {
    if (hv.pre) el = th_lcp(hv, (long)a, b, cap, lim);
    else { wb->need_hash = 1; return; }
  }
  // one suffix is a prefix of the other: the shorter one (the later position) is the smaller
  const bool ordered = el == lim ? ((long)a > b) : (text[a + el] < text[b + el]);
  if (!ordered) atomicAdd(&viol[1], 1ull);
}

extern "C" int psc_verify_sa_slice_d(psc_context *ctx, const uint8_t *text_d, long n, const uint8_t *slice_sa5_d,
                                     long slice_beg, long slice_len, long next_first, long violations[2]) {
  try {
    if (!ctx || n < 0 || !violations || slice_beg < 0 || slice_len < 0 || slice_beg + slice_len > n ||
        (slice_len > 0 && (!text_d || !slice_sa5_d)) || (slice_beg + slice_len < n && next_first < 0)) {
      set_error("psc_verify_sa_slice_d: bad argument");
      return PSC_EARG;
    }
    violations[0] = violations[1] = 0;
    if (slice_len == 0) return PSC_OK;
    PSC_CUDA(cudaSetDevice(ctx->device));
    hash_domain_set(ctx, text_d, n);
    struct DomainGuard { psc_context *c; ~DomainGuard() { hash_domain_clear(c); } } domain_guard{ctx};
    const long cap = ((uintptr_t)text_d & 15) == 0 ? lcp_cap() : (1L << 62);  // unaligned text: symbol by symbol
    HashView hv;
    DevBuf<unsigned long long> viol(ctx, 2);
    for (int attempt = 0;; ++attempt) {
      viol.zero();
      WorkBudget *wb = work_budget_arm(ctx);
      verify_slice_kernel<<<(unsigned)div_up(slice_len, 256), 256, 0, ctx->stream>>>(
          text_d, n, slice_sa5_d, slice_len, slice_beg + slice_len < n ? next_first : -1, viol.p, wb, hv, cap);
      PSC_LAUNCH_CHECK();
      if (!work_budget_check(ctx, wb, "verify")) break;
      if (attempt > 0 || !hash_view(ctx, text_d, text_d + n, true, &hv)) {
        set_error("psc_verify_sa_slice_d: fingerprints unavailable");
        return PSC_ECUDA;
      }
    }
    unsigned long long h[2];
    d2h(ctx, h, viol.p, 2);
    sync(ctx);
    violations[0] = (long)h[0];
    violations[1] = (long)h[1];
    return PSC_OK;
  } catch (const CudaError &e) { return e.code; }
}

// ---- the same check without any allocation ---------------------------------------------------------------
// For callers that cannot call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid
// it): the inverse-array scratch of the check lives in a file-scope device array of fixed size, which
// every GPU that loads the library holds until the process exits (16 MiB + 256 bytes).  No counterpart
// in the reference.
//   * every call shares the array: calls that overlap on the GPU (other streams, other threads) race;
//   * the array keeps the last call's contents: the part a call uses is cleared first.
constexpr long VERIFY_NOALLOC_MAX = PSC_VERIFY_NOALLOC_MAX_N;
static __device__ __align__(256) uint32_t g_verify_scratch[64 + VERIFY_NOALLOC_MAX];  // [0, 64): counters

namespace {
constexpr int MAX_DEVICES = 64;
std::mutex g_scratch_mu;
uint32_t *g_scratch_addr[MAX_DEVICES] = {nullptr};

// device address of the scratch on the CURRENT device, looked up once per device (in host code the
// array's name is only a stand-in for it)
uint32_t *scratch_on_device(int device) {
  if (device < 0 || device >= MAX_DEVICES) { set_error("psc_verify_sa_noalloc_d: device %d out of range", device); throw CudaError{PSC_EARG}; }
  std::lock_guard<std::mutex> lk(g_scratch_mu);
  if (!g_scratch_addr[device]) {
    void *p = nullptr;
    PSC_CUDA(cudaGetSymbolAddress(&p, g_verify_scratch));
    g_scratch_addr[device] = static_cast<uint32_t *>(p);
  }
  return g_scratch_addr[device];
}
}  // namespace

extern "C" long psc_verify_noalloc_max_n(void) { return VERIFY_NOALLOC_MAX; }

extern "C" int psc_verify_sa_noalloc_d(psc_context *ctx, const uint8_t *text_d, long n, const uint8_t *sa5_d,
                                       long violations[4]) {
  try {
    if (!ctx || n < 0 || !violations || (n > 0 && (!text_d || !sa5_d))) {
      set_error("psc_verify_sa_noalloc_d: bad argument");
      return PSC_EARG;
    }
    if (n > VERIFY_NOALLOC_MAX) {
      set_error("psc_verify_sa_noalloc_d: n = %ld exceeds the static scratch (at most %ld entries); use psc_verify_sa_d",
                n, VERIFY_NOALLOC_MAX);
      return PSC_EARG;
    }
    for (int k = 0; k < 4; ++k) violations[k] = 0;
    if (n == 0) return PSC_OK;
    PSC_CUDA(cudaSetDevice(ctx->device));
    uint32_t *base = scratch_on_device(ctx->device);
    unsigned long long *viol = reinterpret_cast<unsigned long long *>(base);
    uint32_t *isa = base + 64;
    PSC_CUDA(cudaMemsetAsync(base, 0, (size_t)(64 + n) * 4, ctx->stream));  // what the last call left behind
    verify_scatter_kernel<<<(unsigned)div_up(n, 256), 256, 0, ctx->stream>>>(sa5_d, n, isa, viol);
    PSC_LAUNCH_CHECK();
    verify_order_kernel<<<(unsigned)div_up(n, 256), 256, 0, ctx->stream>>>(text_d, sa5_d, n, isa, viol);
    PSC_LAUNCH_CHECK();
    unsigned long long h[4];
    PSC_CUDA(cudaMemcpyAsync(h, viol, sizeof(h), cudaMemcpyDeviceToHost, ctx->stream));
    PSC_CUDA(cudaStreamSynchronize(ctx->stream));
    for (int k = 0; k < 4; ++k) violations[k] = (long)h[k];
    return PSC_OK;
  } catch (const CudaError &e) { return e.code; }
}
