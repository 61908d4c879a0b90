// byproducts.cu -- cheap extras around the path:
//   * the sentinel filter of /root/reference/tools/delete-sentinel-bytes/src/main.cpp:45-104 (erase every byte of
//     value 255, the one byte pSAscan rejects) as a stable stream compaction on the device;
//   * the full-text BWT and inverse-suffix-array samples from the finished 40-bit suffix array (what an FM-index
//     builder wants next; the reference computes block BWTs on the way, inmem_bwt_from_sa.hpp:48-56 and
//     bwt_merge.hpp:66-140, and drops them).
#include "scan.cuh"

namespace psc {

constexpr int FS_THREADS = 256, FS_PER_THREAD = 16, FS_TILE = FS_THREADS * FS_PER_THREAD;

__device__ __forceinline__ int load_keep(const uint8_t *__restrict__ in, long n, long base, uint8_t (&b)[FS_PER_THREAD]) {
  int keep = 0;
#pragma unroll
  for (int k = 0; k < FS_PER_THREAD; ++k) {
    b[k] = base + k < n ? in[base + k] : (uint8_t)255;
    keep += b[k] != 255;
  }
  return keep;
}

__global__ void __launch_bounds__(FS_THREADS)
    filter_count_kernel(const uint8_t *__restrict__ in, long n, uint32_t *__restrict__ tile_keep) {
  __shared__ uint32_t sm[FS_THREADS / 32 + 1];
  uint8_t b[FS_PER_THREAD];
  const long base = (long)blockIdx.x * FS_TILE + (long)threadIdx.x * FS_PER_THREAD;
  uint32_t tot;
  block_excl_sum<uint32_t>((uint32_t)load_keep(in, n, base, b), sm, &tot);
  if (threadIdx.x == 0) tile_keep[blockIdx.x] = tot;
}

__global__ void __launch_bounds__(FS_THREADS)
    filter_scatter_kernel(const uint8_t *__restrict__ in, long n, const uint64_t *__restrict__ tile_off,
                          uint8_t *__restrict__ out) {
  __shared__ uint32_t sm[FS_THREADS / 32 + 1];
  uint8_t b[FS_PER_THREAD];
  const long base = (long)blockIdx.x * FS_TILE + (long)threadIdx.x * FS_PER_THREAD;
  uint32_t tot;
  const uint32_t ex = block_excl_sum<uint32_t>((uint32_t)load_keep(in, n, base, b), sm, &tot);
  uint8_t *o = out + tile_off[blockIdx.x] + ex;
#pragma unroll
  for (int k = 0; k < FS_PER_THREAD; ++k)
    if (b[k] != 255) *o++ = b[k];
}

__device__ __forceinline__ uint64_t sa5_at(const uint8_t *__restrict__ sa5, long i) {
  const uint8_t *p = sa5 + i * 5;
  return (uint64_t)p[0] | ((uint64_t)p[1] << 8) | ((uint64_t)p[2] << 16) | ((uint64_t)p[3] << 24) | ((uint64_t)p[4] << 32);
}

// bwt[i] = text[sa[i] - 1]; the row of suffix 0 gets byte 0 and its index goes to *i0 (the block convention of
// inmem_bwt_from_sa.hpp:52-53 applied to the whole text)
__global__ void __launch_bounds__(256)
    full_bwt_kernel(const uint8_t *__restrict__ text, const uint8_t *__restrict__ sa5, long n,
                    uint8_t *__restrict__ bwt, unsigned long long *__restrict__ i0) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t s = sa5_at(sa5, i);
  if (s == 0) { bwt[i] = 0; *i0 = (unsigned long long)i; }
  else bwt[i] = text[s - 1];
}

// out[p / rate] = i for every i with sa[i] = p, p a multiple of rate
__global__ void __launch_bounds__(256)
    isa_samples_kernel(const uint8_t *__restrict__ sa5, long n, long rate, uint64_t *__restrict__ out) {
  const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const uint64_t s = sa5_at(sa5, i);
  if (s % (uint64_t)rate == 0) out[s / (uint64_t)rate] = (uint64_t)i;
}

}  // namespace psc
using namespace psc;

extern "C" {

int psc_delete_sentinel_bytes_d(psc_context *ctx, const uint8_t *in_d, long n, uint8_t *out_d, long *n_out) {
  try {
    if (!ctx || n < 0 || !n_out || (n > 0 && (!in_d || !out_d))) { set_error("psc_delete_sentinel_bytes_d: bad argument"); return PSC_EARG; }
    *n_out = 0;
    if (n == 0) return PSC_OK;
    PSC_CUDA(cudaSetDevice(ctx->device));
    const long tiles = div_up(n, (long)FS_TILE);
    DevBuf<uint32_t> keep(ctx, tiles);
    DevBuf<uint64_t> off(ctx, tiles), total(ctx, 1);
    filter_count_kernel<<<(unsigned)tiles, FS_THREADS, 0, ctx->stream>>>(in_d, n, keep.p);
    PSC_LAUNCH_CHECK();
    exclusive_scan_u32_to_u64(ctx, keep.p, off.p, tiles, total.p);
    filter_scatter_kernel<<<(unsigned)tiles, FS_THREADS, 0, ctx->stream>>>(in_d, n, off.p, out_d);
    PSC_LAUNCH_CHECK();
    ctx->launches += 2;
    unsigned long long t = 0;
    d2h(ctx, &t, reinterpret_cast<unsigned long long *>(total.p), 1);
    sync(ctx);
    *n_out = (long)t;
    return PSC_OK;
  } catch (const CudaError &e) { return e.code; }
}

int psc_bwt_from_sa5_d(psc_context *ctx, const uint8_t *text_d, long n, const uint8_t *sa5_d, uint8_t *bwt_d, long *i0) {
  try {
    if (!ctx || n < 0 || (n > 0 && (!text_d || !sa5_d || !bwt_d))) { set_error("psc_bwt_from_sa5_d: bad argument"); return PSC_EARG; }
    if (i0) *i0 = -1;
    if (n == 0) return PSC_OK;
    PSC_CUDA(cudaSetDevice(ctx->device));
    DevBuf<unsigned long long> i0_d(ctx, 1);
    i0_d.zero();
    full_bwt_kernel<<<(unsigned)div_up(n, 256), 256, 0, ctx->stream>>>(text_d, sa5_d, n, bwt_d, i0_d.p);
    PSC_LAUNCH_CHECK();
    ctx->launches++;
    unsigned long long h = 0;
    d2h(ctx, &h, i0_d.p, 1);
    sync(ctx);
    if (i0) *i0 = (long)h;
    return PSC_OK;
  } catch (const CudaError &e) { return e.code; }
}

int psc_isa_samples_d(psc_context *ctx, const uint8_t *sa5_d, long n, long rate, uint64_t *samples_d) {
  try {
    if (!ctx || n < 0 || rate <= 0 || (n > 0 && (!sa5_d || !samples_d))) { set_error("psc_isa_samples_d: bad argument"); return PSC_EARG; }
    if (n == 0) return PSC_OK;
    PSC_CUDA(cudaSetDevice(ctx->device));
    isa_samples_kernel<<<(unsigned)div_up(n, 256), 256, 0, ctx->stream>>>(sa5_d, n, rate, samples_d);
    PSC_LAUNCH_CHECK();
    ctx->launches++;
    sync(ctx);
    return PSC_OK;
  } catch (const CudaError &e) { return e.code; }
}

}  // extern "C"
