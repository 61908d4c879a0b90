// glue.cu -- what the reference needs when one block is sorted as two pieces L, R
// (/root/reference/include/partial_sufsort.hpp:441-546): gap -> bitvector
// (gap_array.hpp:273-364), BWT interleave (bwt_merge.hpp:66-140) and the split of the block's gap
// into the two pieces' gaps (compute_left_gap.hpp:55-123, compute_right_gap.hpp:55-122).
// On the GPU all three are scans: prefix sums of the gap array and prefix popcounts of bv.
#include "scan.cuh"

namespace psc {

// bv starts as all ones over [0, len); the u-th L suffix sits at u + S_L(u) and clears its bit.
__global__ void bv_fill_kernel(uint32_t *bv, long len) {
  long w = (long)blockIdx.x * blockDim.x + threadIdx.x;
  long nw = (len + 31) / 32;
  if (w >= nw) return;
  long rem = len - w * 32;
  bv[w] = rem >= 32 ? 0xffffffffu : ((1u << rem) - 1u);
}
__global__ void bv_clear_kernel(uint32_t *bv, const uint64_t *E, long len_l) {
  long u = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= len_l) return;
  uint64_t pos = (uint64_t)u + E[u + 1];
  atomicAnd(&bv[pos >> 5], ~(1u << (pos & 31)));
}

void gap_to_bitvector(psc_context *ctx, const uint32_t *gap_l, long len_l, long len_r, uint32_t *bv) {
  long len = len_l + len_r;
  if (len == 0) return;
  DevBuf<uint64_t> E(ctx, len_l + 2);
  exclusive_scan_u32_to_u64(ctx, gap_l, E.p, len_l + 1, E.p + len_l + 1);
  bv_fill_kernel<<<(unsigned)div_up(div_up(len, 32), 256), 256, 0, ctx->stream>>>(bv, len);
  PSC_LAUNCH_CHECK();
  if (len_l) {
    bv_clear_kernel<<<(unsigned)div_up(len_l, 256), 256, 0, ctx->stream>>>(bv, E.p, len_l);
    PSC_LAUNCH_CHECK();
  }
  ctx->launches += 2;
}

__global__ void popc_words_kernel(const uint32_t *bv, long nw, uint32_t *pc) {
  long w = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (w < nw) pc[w] = __popc(bv[w]);
}

// ones_before[w] = number of 1 bits in words [0, w)
static void ones_prefix(psc_context *ctx, const uint32_t *bv, long len, DevBuf<uint64_t> *ones) {
  long nw = (len + 31) / 32;
  DevBuf<uint32_t> pc(ctx, nw);
  ones->alloc(ctx, nw + 1);
  popc_words_kernel<<<(unsigned)div_up(nw, 256), 256, 0, ctx->stream>>>(bv, nw, pc.p);
  PSC_LAUNCH_CHECK();
  ctx->launches++;
  exclusive_scan_u32_to_u64(ctx, pc.p, ones->p, nw, ones->p + nw);
}

__global__ void merge_bwt_kernel(const uint8_t *__restrict__ bwt_l, const uint8_t *__restrict__ bwt_r,
                                 long len, long i0_l, long i0_r, uint8_t last_l,
                                 const uint32_t *__restrict__ bv, const uint64_t *__restrict__ ones,
                                 uint8_t *__restrict__ out, unsigned long long *__restrict__ i0_out) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= len) return;
  uint32_t word = bv[i >> 5];
  int b = (int)(i & 31);
  long r1 = (long)ones[i >> 5] + __popc(word & ((1u << b) - 1u));
  if ((word >> b) & 1u) {
    out[i] = (r1 == i0_r) ? last_l : bwt_r[r1];  // bwt_merge.hpp:127
  } else {
    long r0 = i - r1;
    out[i] = bwt_l[r0];
    if (r0 == i0_l) *i0_out = (unsigned long long)i;  // :132 select0(i0_l)
  }
}

void merge_bwt(psc_context *ctx, const uint8_t *bwt_l, const uint8_t *bwt_r, long len_l, long len_r,
               long i0_l, long i0_r, uint8_t last_l, const uint32_t *bv, uint8_t *bwt_out,
               long *i0_out) {
  long len = len_l + len_r;
  if (len == 0) return;
  DevBuf<uint64_t> ones;
  ones_prefix(ctx, bv, len, &ones);
  DevBuf<unsigned long long> i0_d(ctx, 1);
  i0_d.zero();
  merge_bwt_kernel<<<(unsigned)div_up(len, 256), 256, 0, ctx->stream>>>(bwt_l, bwt_r, len, i0_l, i0_r, last_l,
                                                                       bv, ones.p, bwt_out, i0_d.p);
  PSC_LAUNCH_CHECK();
  ctx->launches++;
  unsigned long long v = 0;
  d2h(ctx, &v, i0_d.p, 1);
  sync(ctx);
  *i0_out = (long)v;
}

// G = inclusive prefix sums of block_gap (len+1 entries).  For the k-th zero at p:
// HL[k] = G(p) + ones(p) ; for the k-th one at p: HR[k] = G(p).  Sentinels close the arrays.
__global__ void split_marks_kernel(const uint32_t *__restrict__ bv, const uint64_t *__restrict__ ones,
                                   const uint64_t *__restrict__ E /* exclusive, len+2 */, long len,
                                   uint64_t *__restrict__ HL, uint64_t *__restrict__ HR, long len_l,
                                   long len_r) {
  long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i > len) return;
  if (i == len) {  // closing entries: everything up to the end
    uint64_t Gend = E[len + 1];
    HL[len_l] = Gend + (uint64_t)len_r;
    HR[len_r] = Gend;
    return;
  }
  uint32_t word = bv[i >> 5];
  int b = (int)(i & 31);
  uint64_t r1 = ones[i >> 5] + (uint64_t)__popc(word & ((1u << b) - 1u));
  uint64_t G = E[i + 1];  // inclusive sum through block_gap[i]
  if ((word >> b) & 1u) HR[r1] = G;
  else HL[(uint64_t)i - r1] = G + r1;
}
__global__ void adj_diff_kernel(const uint64_t *__restrict__ H, long n, uint64_t *__restrict__ out) {
  long k = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= n) return;
  out[k] = H[k] - (k ? H[k - 1] : 0);
}

void split_gap(psc_context *ctx, const uint32_t *bv, long len_l, long len_r, const uint32_t *block_gap,
               uint64_t *left_gap, uint64_t *right_gap) {
  long len = len_l + len_r;
  DevBuf<uint64_t> E(ctx, len + 2), HL(ctx, len_l + 1), HR(ctx, len_r + 1), ones;
  exclusive_scan_u32_to_u64(ctx, block_gap, E.p, len + 1, E.p + len + 1);
  if (len) ones_prefix(ctx, bv, len, &ones);
  split_marks_kernel<<<(unsigned)div_up(len + 1, 256), 256, 0, ctx->stream>>>(bv, ones.p, E.p, len, HL.p, HR.p,
                                                                             len_l, len_r);
  PSC_LAUNCH_CHECK();
  adj_diff_kernel<<<(unsigned)div_up(len_l + 1, 256), 256, 0, ctx->stream>>>(HL.p, len_l + 1, left_gap);
  PSC_LAUNCH_CHECK();
  adj_diff_kernel<<<(unsigned)div_up(len_r + 1, 256), 256, 0, ctx->stream>>>(HR.p, len_r + 1, right_gap);
  PSC_LAUNCH_CHECK();
  ctx->launches += 3;
}

}  // namespace psc
