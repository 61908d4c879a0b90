// texthash.cuh -- O(log n) longest-common-prefix queries between two positions of the device text, for the
// two places of the path that compare a block suffix with the text after the block symbol by symbol
// (K0 gt_end, K5 initial ranks).  On ordinary text those comparisons end after a few symbols; on texts with
// long repeats across a block boundary (period << block length, long runs of one byte) they are O(LCP) each
// and O(m * LCP) in total.  The reference bounds them with its ms-decomposition
// (/root/reference/include/inmem_psascan_src/compute_initial_gt_bitvectors.hpp:97-148, srank_aux.hpp:45-65),
// a sequential scan that reuses earlier answers through the period of the matched prefix.  The GPU version
// bounds every single comparison instead: the first PSC_LCP_CAP symbols are compared directly; a comparison
// that survives them is finished by binary search over Karp-Rabin fingerprints (two independent bases modulo
// 2^61 - 1, prefix fingerprints sampled every 16 symbols, built once per construction and only when some
// comparison needs them).  The fingerprints decide only HOW FAR two substrings agree; the symbol after the
// agreed prefix (or the gt bit at the block end) decides the order exactly as before.  A false "equal" needs a
// simultaneous collision of both 61-bit fingerprints: probability < 2^-100 per query with random bases.
#pragma once
#include "common.cuh"

namespace psc {

constexpr uint64_t TH_P = (1ull << 61) - 1;
constexpr int TH_SAMPLE_LOG = 4, TH_SAMPLE = 1 << TH_SAMPLE_LOG;

struct HashView {
  const uint8_t *text = nullptr;  // domain: text[0, n)
  long n = 0, ns = 0;             // ns = n / TH_SAMPLE; tables hold ns + 1 entries per base
  const uint64_t *pre = nullptr;  // pre[h * (ns + 1) + i] = fingerprint of text[0, 16 i) under base h
  const uint64_t *pw = nullptr;   // pw[h * (ns + 1) + i] = base_h ^ (16 i)
  uint64_t bpow[2][TH_SAMPLE + 1];  // base_h ^ j, j = 0..16
};

// host side (texthash.cu).  The domain is the device text of the running construction; tables are built on
// the first request and dropped when the domain changes or is cleared.
void hash_domain_set(psc_context *ctx, const uint8_t *base_d, long n);
void hash_domain_clear(psc_context *ctx);
// true and *hv filled when [lo, hi) lies inside the domain (build: build the tables if they do not exist yet;
// without build, a view is only returned when they already exist)
bool hash_view(psc_context *ctx, const uint8_t *lo, const uint8_t *hi, bool build, HashView *hv);
// a domain that covers [lo, hi) exists afterwards: the current one if it does, else the range itself becomes
// the domain (stage API: block and tail prefix in one buffer); false when lo is not 16-byte aligned
bool hash_domain_ensure(psc_context *ctx, const uint8_t *lo, const uint8_t *hi);
long lcp_cap();  // symbols compared directly before fingerprints take over (PSC_LCP_CAP, default 1024)

#ifdef __CUDACC__
__device__ __forceinline__ uint64_t th_mulmod(uint64_t a, uint64_t b) {  // a, b < 2^61
  const uint64_t lo = a * b, hi = __umul64hi(a, b);
  uint64_t r = (lo & TH_P) + ((lo >> 61) | (hi << 3));
  r = (r & TH_P) + (r >> 61);
  return r >= TH_P ? r - TH_P : r;
}
__device__ __forceinline__ uint64_t th_addmod(uint64_t a, uint64_t b) {
  const uint64_t r = a + b;
  return r >= TH_P ? r - TH_P : r;
}
// fingerprints of text[0, x) under both bases
__device__ __forceinline__ void th_prefix(const HashView &hv, long x, uint64_t &h0, uint64_t &h1) {
  const long i = x >> TH_SAMPLE_LOG;
  h0 = hv.pre[i];
  h1 = hv.pre[hv.ns + 1 + i];
  for (long j = i << TH_SAMPLE_LOG; j < x; ++j) {
    const uint64_t c = (uint64_t)hv.text[j] + 1;
    h0 = th_addmod(th_mulmod(h0, hv.bpow[0][1]), c);
    h1 = th_addmod(th_mulmod(h1, hv.bpow[1][1]), c);
  }
}
__device__ __forceinline__ void th_power(const HashView &hv, long len, uint64_t &p0, uint64_t &p1) {
  const long i = len >> TH_SAMPLE_LOG;
  const int r = (int)(len & (TH_SAMPLE - 1));
  p0 = th_mulmod(hv.pw[i], hv.bpow[0][r]);
  p1 = th_mulmod(hv.pw[hv.ns + 1 + i], hv.bpow[1][r]);
}
// text[a, a + len) == text[b, b + len) by fingerprints; ha*, hb* = th_prefix of a and b
__device__ __forceinline__ bool th_equal(const HashView &hv, long a, long b, long len, uint64_t ha0, uint64_t ha1,
                                         uint64_t hb0, uint64_t hb1) {
  uint64_t xa0, xa1, xb0, xb1, p0, p1;
  th_prefix(hv, a + len, xa0, xa1);
  th_prefix(hv, b + len, xb0, xb1);
  th_power(hv, len, p0, p1);
  // H[a+len] - H[a] p == H[b+len] - H[b] p   <=>   H[a+len] + H[b] p == H[b+len] + H[a] p
  return th_addmod(xa0, th_mulmod(hb0, p0)) == th_addmod(xb0, th_mulmod(ha0, p0)) &&
         th_addmod(xa1, th_mulmod(hb1, p1)) == th_addmod(xb1, th_mulmod(ha1, p1));
}
// Largest len in [known, lim] with text[a, a + len) == text[b, b + len), given that `known` symbols agree.
// a + lim and b + lim must not exceed hv.n.  Galloping, then binary search: O(log(result)) fingerprint tests.
__device__ __forceinline__ long th_lcp(const HashView &hv, long a, long b, long known, long lim) {
  if (known >= lim) return lim;
  uint64_t ha0, ha1, hb0, hb1;
  th_prefix(hv, a, ha0, ha1);
  th_prefix(hv, b, hb0, hb1);
  long good = known, step = known > 64 ? known : 64, bad = -1;
  while (good < lim) {
    const long t = min(lim, good + step);
    if (th_equal(hv, a, b, t, ha0, ha1, hb0, hb1)) { good = t; step <<= 1; }
    else { bad = t; break; }
  }
  if (bad < 0) return lim;
  while (bad - good > 1) {
    const long mid = good + ((bad - good) >> 1);
    if (th_equal(hv, a, b, mid, ha0, ha1, hb0, hb1)) good = mid; else bad = mid;
  }
  return good;
}
#endif

}  // namespace psc
