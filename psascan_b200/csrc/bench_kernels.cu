// bench_kernels.cu -- micro-benchmarks for the roofline denominators:
// R32 = rate of independent random 32-byte sector reads, and the time of one radix digit pass.
#include "common.cuh"

namespace psc {

// V: 0 = one 256-bit load (what rank_occ issues), 1 = two 128-bit loads, 2 = one 32-bit load,
//    3 = 256-bit load with the default cache policy
template <int V>
__global__ void __launch_bounds__(256)
    random_sector_kernel(const uint32_t *__restrict__ buf, uint64_t n_sectors, int iters,
                         uint32_t *__restrict__ sink) {
  uint64_t x = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) * 0x9E3779B97F4A7C15ull + 12345;
  uint32_t acc = 0;
  for (int k = 0; k < iters; ++k) {
    x ^= x >> 12; x ^= x << 25; x ^= x >> 27;
    uint64_t s = (x * 0x2545F4914F6CDD1Dull) % n_sectors;
    uint32_t w[8];
    if (V == 0) {
      asm volatile("ld.global.nc.L1::no_allocate.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                   : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]),
                     "=r"(w[7])
                   : "l"(buf + s * 8));
    } else if (V == 1) {
      uint4 a = __ldg(reinterpret_cast<const uint4 *>(buf + s * 8));
      uint4 b = __ldg(reinterpret_cast<const uint4 *>(buf + s * 8 + 4));
      w[0] = a.x ^ a.y; w[7] = b.w ^ b.z;
    } else if (V == 2) {
      w[0] = __ldg(buf + s * 8); w[7] = 0;
    } else if (V == 3) {
      asm volatile("ld.global.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                   : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]),
                     "=r"(w[7])
                   : "l"(buf + s * 8));
    } else if (V == 4) {
      asm volatile("ld.global.cv.u32 %0, [%1];" : "=r"(w[0]) : "l"(buf + s * 8));
      w[7] = 0;
    } else if (V == 5) {
      asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[7]) : "l"(buf + s * 8));
    } else if (V == 6) {
      w[0] = atomicAdd(const_cast<uint32_t *>(buf) + s * 8, 0u);
      w[7] = 0;
    } else if (V == 7) {
      unsigned long long pol;
      asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
      asm volatile("ld.global.nc.L1::no_allocate.L2::cache_hint.v4.u32 {%0,%1,%2,%3}, [%4], %5;"
                   : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[7]) : "l"(buf + s * 8), "l"(pol));
    } else if (V == 8) {
      asm volatile("ld.global.nc.L1::no_allocate.L2::64B.v4.u32 {%0,%1,%2,%3}, [%4];"
                   : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[7]) : "l"(buf + s * 8));
    } else if (V == 10) {
      asm volatile("ld.global.nc.L1::no_allocate.L2::64B.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                   : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]),
                     "=r"(w[7])
                   : "l"(buf + s * 8));
    } else if (V == 11) {
      asm volatile("ld.global.nc.L2::64B.u32 %0, [%1];" : "=r"(w[0]) : "l"(buf + s * 8));
      w[7] = 0;
    } else if (V == 12) {
      asm volatile("ld.global.nc.L2::128B.u32 %0, [%1];" : "=r"(w[0]) : "l"(buf + s * 8));
      w[7] = 0;
    } else if (V == 13) {
      asm volatile("ld.global.L2::64B.v8.u32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                   : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]),
                     "=r"(w[7])
                   : "l"(buf + s * 8));
    } else if (V == 9) {
      // 32-byte bulk copy global -> shared, one per lane, one mbarrier per warp
      __shared__ __align__(32) uint32_t stage[256 * 8];
      __shared__ __align__(8) unsigned long long bar[8];
      const int lane = threadIdx.x & 31, wp = threadIdx.x >> 5;
      uint32_t bar_a = (uint32_t)__cvta_generic_to_shared(&bar[wp]);
      uint32_t dst_a = (uint32_t)__cvta_generic_to_shared(&stage[threadIdx.x * 8]);
      if (k == 0) {
        if (lane == 0) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar_a));
        __syncwarp();
      }
      if (lane == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar_a), "r"(32 * 32));
      __syncwarp();
      asm volatile("cp.async.bulk.shared::cta.global.mbarrier::complete_tx::bytes [%0], [%1], 32, [%2];"
                   ::"r"(dst_a), "l"(buf + s * 8), "r"(bar_a) : "memory");
      uint32_t done = 0;
      while (!done) {
        asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                     : "=r"(done) : "r"(bar_a), "r"((uint32_t)(k & 1)) : "memory");
      }
      w[0] = stage[threadIdx.x * 8]; w[7] = stage[threadIdx.x * 8 + 7];
      __syncwarp();
    }
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
    {
      size_t lim = 0;
      cudaDeviceGetLimit(&lim, cudaLimitMaxL2FetchGranularity);
      if (getenv("PSC_R32_VARIANT")) fprintf(stderr, "[psc] L2 fetch granularity limit = %zu\n", lim);
    }
    const char *ev = getenv("PSC_R32_VARIANT");
    const int v = ev ? atoi(ev) : 0;
    auto launch = [&](int it) {
      switch (v) {
        case 1: random_sector_kernel<1><<<blocks, 256, 0, ctx->stream>>>(buf.p, bytes / 32, it, sink.p); break;
        case 2: random_sector_kernel<2><<<blocks, 256, 0, ctx->stream>>>(buf.p, bytes / 32, it, sink.p); break;
        case 3: random_sector_kernel<3><<<blocks, 256, 0, ctx->stream>>>(buf.p, bytes / 32, it, sink.p); break;
        case 4: random_sector_kernel<4><<<blocks, 256, 0, ctx->stream>>>(buf.p, bytes / 32, it, sink.p); break;
        case 5: random_sector_kernel<5><<<blocks, 256, 0, ctx->stream>>>(buf.p, bytes / 32, it, sink.p); break;
        case 6: random_sector_kernel<6><<<blocks, 256, 0, ctx->stream>>>(buf.p, bytes / 32, it, sink.p); break;
        case 7: random_sector_kernel<7><<<blocks, 256, 0, ctx->stream>>>(buf.p, bytes / 32, it, sink.p); break;
        case 8: random_sector_kernel<8><<<blocks, 256, 0, ctx->stream>>>(buf.p, bytes / 32, it, sink.p); break;
        case 10: random_sector_kernel<10><<<blocks, 256, 0, ctx->stream>>>(buf.p, bytes / 32, it, sink.p); break;
        case 11: random_sector_kernel<11><<<blocks, 256, 0, ctx->stream>>>(buf.p, bytes / 32, it, sink.p); break;
        case 12: random_sector_kernel<12><<<blocks, 256, 0, ctx->stream>>>(buf.p, bytes / 32, it, sink.p); break;
        case 13: random_sector_kernel<13><<<blocks, 256, 0, ctx->stream>>>(buf.p, bytes / 32, it, sink.p); break;
        case 9: random_sector_kernel<9><<<blocks, 256, 0, ctx->stream>>>(buf.p, bytes / 32, it, sink.p); break;
        default: random_sector_kernel<0><<<blocks, 256, 0, ctx->stream>>>(buf.p, bytes / 32, it, sink.p);
      }
    };
    launch(8);  // warm-up
    PSC_LAUNCH_CHECK();
    double ms = 0;
    {
      PhaseTimer t(ctx, &ms);
      launch(iters);
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
    const bool keys_only = getenv("PSC_BENCH_KEYS_ONLY") != nullptr;  // 16 B per element instead of 24
    uint32_t *v0 = keys_only ? nullptr : va.p, *v1 = keys_only ? nullptr : vb.p;
    radix_sort_pairs(ctx, ka.p, v0, kb.p, v1, n, shifts, 1, &ko, keys_only ? nullptr : &vo);  // warm-up
    double before = ctx->stats.ms_radix_pass;
    for (int k = 0; k < iters; ++k) {
      shifts[0] = 8 * (k % 8);
      radix_sort_pairs(ctx, ka.p, v0, kb.p, v1, n, shifts, 1, &ko, keys_only ? nullptr : &vo);
    }
    *ms_per_pass = (ctx->stats.ms_radix_pass - before) / iters;
    return PSC_OK;
  } catch (const CudaError &e) { return e.code; }
}

}  // extern "C"
