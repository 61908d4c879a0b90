// verify.cu -- O(n) suffix-array check on the device (the logic of sufcheck(),
// /root/reference/extlib/libdivsufsort/lib/utils.c:161-241): the entries are a permutation of
// [0, n), consecutive suffixes are ordered by their first symbol, and suffixes with the same first
// symbol are ordered like the suffixes that follow them (through the inverse array).  Used by
// bench.py after the timed loop and by the parity tests at sizes no CPU sorter reaches.
#include "common.cuh"

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
