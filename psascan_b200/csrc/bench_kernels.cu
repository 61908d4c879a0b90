// bench_kernels.cu -- micro-benchmarks for the roofline denominators:
// R32 = rate of independent random 32-byte sector reads, and the time of one radix digit pass.
#include "common.cuh"

namespace psc {

__global__ void __launch_bounds__(256)
    random_sector_kernel(const uint32_t *__restrict__ buf, uint64_t n_sectors, int iters,
                         uint32_t *__restrict__ sink) {
  uint64_t x = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 0x9E3779B97F4A7C15ull + 12345;
  uint32_t acc = 0;
  for (int k = 0; k < iters; ++k) {
    x ^= x >> 12; x ^= x << 25; x ^= x >> 27;
    uint64_t s = (x * 0x2545F4914F6CDD1Dull) % n_sectors;
    uint32_t w[8];
    asm volatile("ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]),
                   "=r"(w[7])
                 : "l"(buf + s * 8));
    acc += w[0] ^ w[7];
  }
  if (acc == 0x12345678u) sink[0] = acc;
}

__global__ void fill_random_keys_kernel(uint64_t *keys, uint32_t *vals, long n) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  uint64_t x = (uint64_t)i * 0x9E3779B97F4A7C15ull + 777;
  x ^= x >> 31; x *= 0xBF58476D1CE4E5B9ull; x ^= x >> 29;
  keys[i] = x;
  vals[i] = (uint32_t)i;
}

}  // namespace psc
using namespace psc;

extern "C" {

int psc_bench_random_sectors(psc_context *ctx, size_t bytes, int iters, double *gbs) {
  try {
    if (!ctx || !gbs || bytes < 4096 || iters <= 0) { set_error("psc_bench_random_sectors: bad argument"); return PSC_EARG; }
    PSC_CUDA(cudaSetDevice(ctx->device));
    DevBuf<uint32_t> buf(ctx, bytes / 4), sink(ctx, 1);
    PSC_CUDA(cudaMemsetAsync(buf.p, 1, bytes, ctx->stream));
    unsigned blocks = (unsigned)ctx->n_sm * 8;
    random_sector_kernel<<<blocks, 256, 0, ctx->stream>>>(buf.p, bytes / 32, 8, sink.p);  // warm-up
    PSC_LAUNCH_CHECK();
    double ms = 0;
    {
      PhaseTimer t(ctx, &ms);
      random_sector_kernel<<<blocks, 256, 0, ctx->stream>>>(buf.p, bytes / 32, iters, sink.p);
      PSC_LAUNCH_CHECK();
    }
    *gbs = (double)blocks * 256.0 * iters * 32.0 / (ms * 1e-3) / 1e9;
    return PSC_OK;
  } catch (const CudaError &e) { return e.code; }
}

int psc_bench_radix_pass(psc_context *ctx, long n, int iters, double *ms_per_pass) {
  try {
    if (!ctx || !ms_per_pass || n <= 0 || iters <= 0) { set_error("psc_bench_radix_pass: bad argument"); return PSC_EARG; }
    PSC_CUDA(cudaSetDevice(ctx->device));
    DevBuf<uint64_t> ka(ctx, n), kb(ctx, n);
    DevBuf<uint32_t> va(ctx, n), vb(ctx, n);
    fill_random_keys_kernel<<<(unsigned)div_up(n, 256), 256, 0, ctx->stream>>>(ka.p, va.p, n);
    PSC_LAUNCH_CHECK();
    int shifts[1] = {0};
    uint64_t *ko; uint32_t *vo;
    radix_sort_pairs(ctx, ka.p, va.p, kb.p, vb.p, n, shifts, 1, &ko, &vo);  // warm-up
    double before = ctx->stats.ms_radix_pass;
    for (int k = 0; k < iters; ++k) {
      shifts[0] = 8 * (k % 8);
      radix_sort_pairs(ctx, ka.p, va.p, kb.p, vb.p, n, shifts, 1, &ko, &vo);
    }
    *ms_per_pass = (ctx->stats.ms_radix_pass - before) / iters;
    return PSC_OK;
  } catch (const CudaError &e) { return e.code; }
}

}  // extern "C"
