// rank.cu -- stage 3a: rank structure over a block BWT, one 32-byte sector per query.
//
// Replaces rank4n<24,20,8> (/root/reference/include/rank.hpp:74-722; query :566-708).  The
// reference needs 3-4 dependent cache misses per query (sblock header, cblock mapping, cblock
// header, trunk word).  Here every symbol that occurs in the BWT owns a bit plane cut into
// 32-byte sectors: word 0 = occurrences before the sector, words 1..7 = 224 presence bits.
// occ(c, i) = one 256-bit load + popcounts.  The placeholder at i0 (inmem_bwt_from_sa.hpp:
// 52-53) gets no bit in any plane, so the reference's `delta` and `--count[0]` corrections
// (stream.hpp:153, compute_gap.hpp:80) are not needed.  Size: n_codes * (m/224+1) * 32 bytes.
#include <algorithm>
#include "rank.cuh"
#include "scan.cuh"

namespace psc {

constexpr int RK_CHUNK_SECTORS = 32;
constexpr int RK_CHUNK = RK_BITS * RK_CHUNK_SECTORS;  // 7168 symbols
constexpr int RK_GROUP = 32;                          // codes per fill block

struct CodeLut { uint8_t code[256]; };

__global__ void __launch_bounds__(256)
    rank_count_kernel(const uint8_t *__restrict__ bwt, long m, long i0, uint32_t *__restrict__ cc) {
  __shared__ uint32_t sh[256];
  sh[threadIdx.x] = 0;
  __syncthreads();
  long beg = (long)blockIdx.x * RK_CHUNK;
  long end = min(beg + RK_CHUNK, m);
  for (long i = beg + threadIdx.x; i < end; i += 256)
    if (i != i0) atomicAdd(&sh[bwt[i]], 1u);
  __syncthreads();
  cc[(size_t)blockIdx.x * 256 + threadIdx.x] = sh[threadIdx.x];
}

// one block per symbol: exclusive scan over the chunks, in place; totals[c] = grand total
__global__ void __launch_bounds__(SC_THREADS)
    rank_scan_kernel(uint32_t *__restrict__ cc, long nchunks, unsigned long long *__restrict__ totals) {
  __shared__ uint64_t sm[SC_THREADS / 32 + 1];
  const int c = blockIdx.x;
  uint64_t carry = 0;
  for (long base = 0; base < nchunks; base += SC_THREADS) {
    long i = base + threadIdx.x;
    uint64_t v = i < nchunks ? cc[(size_t)i * 256 + c] : 0, tot;
    uint64_t ex = block_excl_sum<uint64_t>(v, sm, &tot);
    if (i < nchunks) cc[(size_t)i * 256 + c] = (uint32_t)(carry + ex);
    carry += tot;
  }
  if (threadIdx.x == 0) totals[c] = carry;
}

__global__ void __launch_bounds__(256)
    rank_fill_kernel(const uint8_t *__restrict__ bwt, long m, long i0, CodeLut lut, int n_codes, int code_bits,
                     long n_sectors, const uint32_t *__restrict__ cc, const uint8_t *sym_of_code,
                     uint32_t *__restrict__ sectors) {
  __shared__ uint8_t codes[RK_CHUNK];
  __shared__ uint32_t words[RK_GROUP][RK_CHUNK_SECTORS][7];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long chunk = blockIdx.x;
  const int gbase = blockIdx.y * RK_GROUP;
  const long beg = chunk * RK_CHUNK;
  for (int e = tid; e < RK_CHUNK; e += 256) {
    long i = beg + e;
    codes[e] = (i < m && i != i0) ? lut.code[bwt[i]] : (uint8_t)0xFF;
  }
  for (int e = tid; e < RK_GROUP * RK_CHUNK_SECTORS * 7; e += 256) (&words[0][0][0])[e] = 0;
  __syncthreads();
  for (int s = warp; s < RK_CHUNK_SECTORS; s += 8) {
#pragma unroll
    for (int g = 0; g < 7; ++g) {
      int code = codes[s * RK_BITS + g * 32 + lane];
      // lanes holding the same code: one ballot per code bit (match.any runs on the ADU pipe at
      // about 60 cycles per warp instruction and made this kernel 0.92 ms per 2^27 symbols)
      uint32_t mask = __ballot_sync(0xffffffffu, code != 0xFF);
      if (code == 0xFF) mask = ~mask;
      for (int b = 0; b < code_bits; ++b) {
        const bool bit = (code >> b) & 1;
        const uint32_t bal = __ballot_sync(0xffffffffu, bit);
        mask &= bit ? bal : ~bal;
      }
      int cl = code - gbase;
      if (code != 0xFF && cl >= 0 && cl < RK_GROUP && lane == __ffs(mask) - 1) words[cl][s][g] = mask;
    }
  }
  __syncthreads();
  for (int cl = warp; cl < RK_GROUP; cl += 8) {
    int code = gbase + cl;
    if (code >= n_codes) break;
    uint32_t w[7];
    uint32_t pc = 0;
#pragma unroll
    for (int g = 0; g < 7; ++g) { w[g] = words[cl][lane][g]; pc += __popc(w[g]); }
    uint32_t incl = pc;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
    uint32_t before = cc[(size_t)chunk * 256 + sym_of_code[code]] + incl - pc;
    long s = chunk * RK_CHUNK_SECTORS + lane;
    if (s < n_sectors) {
      uint4 *out = reinterpret_cast<uint4 *>(sectors + ((size_t)code * (size_t)n_sectors + s) * 8);
      out[0] = make_uint4(before, w[0], w[1], w[2]);
      out[1] = make_uint4(w[3], w[4], w[5], w[6]);
    }
  }
}

// Alphabets of more than 32 symbols: one block per chunk of 32 sectors (7168 symbols) writes ALL planes.  The
// lanes holding the same symbol are found ONCE per 32 symbols and kept in registers; the planes are then
// written 32 symbols (codes) at a time through 28 KB of shared memory.  (rank_fill_kernel above, one block per
// chunk AND group of 32 codes, repeats loads and ballots for each group: 3 x 512 MiB of random bytes 217 -> 156 ms;
// for up to 32 symbols the simpler kernel is the faster one, 6.5 against 8.4 ms on 8 x 128 MiB of Markov text.)
__global__ void __launch_bounds__(256)
    rank_fill_all_kernel(const uint8_t *__restrict__ bwt, long m, long i0, CodeLut lut, int n_codes, int code_bits,
                         long n_sectors, const uint32_t *__restrict__ cc, const uint8_t *sym_of_code,
                         uint32_t *__restrict__ sectors) {
  __shared__ uint8_t codes[RK_CHUNK];
  __shared__ uint32_t words[RK_GROUP][RK_CHUNK_SECTORS][7];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const long chunk = blockIdx.x;
  const long beg = chunk * RK_CHUNK;
  for (int e = tid; e < RK_CHUNK; e += 256) {
    long i = beg + e;
    codes[e] = (i < m && i != i0) ? lut.code[bwt[i]] : (uint8_t)0xFF;
  }
  __syncthreads();
  constexpr int PER_WARP = (RK_CHUNK_SECTORS / 8) * 7;  // (sector, word) pairs of a warp
  uint32_t mk[PER_WARP];                                // lanes with my symbol, kept by the lowest of them only
  uint8_t cd[PER_WARP];
#pragma unroll
  for (int q = 0; q < PER_WARP; ++q) {
    const int s = warp + (q / 7) * 8, g = q % 7;
    const int code = codes[s * RK_BITS + g * 32 + lane];
    uint32_t mask = __ballot_sync(0xffffffffu, code != 0xFF);
    if (code == 0xFF) mask = ~mask;
    for (int b = 0; b < code_bits; ++b) {
      const bool bit = (code >> b) & 1;
      const uint32_t bal = __ballot_sync(0xffffffffu, bit);
      mask &= bit ? bal : ~bal;
    }
    cd[q] = (uint8_t)code;
    mk[q] = (code != 0xFF && lane == __ffs(mask) - 1) ? mask : 0u;
  }
  for (int gbase = 0; gbase < n_codes; gbase += RK_GROUP) {
    for (int e = tid; e < RK_GROUP * RK_CHUNK_SECTORS * 7; e += 256) (&words[0][0][0])[e] = 0;
    __syncthreads();
#pragma unroll
    for (int q = 0; q < PER_WARP; ++q) {
      const int cl = (int)cd[q] - gbase;
      if (mk[q] && cl >= 0 && cl < RK_GROUP) words[cl][warp + (q / 7) * 8][q % 7] = mk[q];
    }
    __syncthreads();
    for (int cl = warp; cl < RK_GROUP; cl += 8) {
      int code = gbase + cl;
      if (code >= n_codes) break;
      uint32_t w[7];
      uint32_t pc = 0;
#pragma unroll
      for (int g = 0; g < 7; ++g) { w[g] = words[cl][lane][g]; pc += __popc(w[g]); }
      uint32_t incl = pc;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { uint32_t t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
      uint32_t before = cc[(size_t)chunk * 256 + sym_of_code[code]] + incl - pc;
      long s = chunk * RK_CHUNK_SECTORS + lane;
      if (s < n_sectors) {
        uint4 *out = reinterpret_cast<uint4 *>(sectors + ((size_t)code * (size_t)n_sectors + s) * 8);
        out[0] = make_uint4(before, w[0], w[1], w[2]);
        out[1] = make_uint4(w[3], w[4], w[5], w[6]);
      }
    }
    __syncthreads();
  }
}

psc_rank *rank_build(psc_context *ctx, const uint8_t *bwt_d, long m, long i0) {
  PhaseTimer timer(ctx, &ctx->stats.ms_rank_build);
  psc_rank *r = new psc_rank();
  r->ctx = ctx;
  r->m = m;
  r->i0 = i0;
  r->n_sectors = m / RK_BITS + 1;
  long nchunks = div_up(r->n_sectors, RK_CHUNK_SECTORS);
  DevBuf<uint32_t> cc(ctx, (size_t)nchunks * 256);
  DevBuf<unsigned long long> totals_d(ctx, 256);
  rank_count_kernel<<<(unsigned)nchunks, 256, 0, ctx->stream>>>(bwt_d, m, i0, cc.p);
  PSC_LAUNCH_CHECK();
  rank_scan_kernel<<<256, SC_THREADS, 0, ctx->stream>>>(cc.p, nchunks, totals_d.p);
  PSC_LAUNCH_CHECK();
  unsigned long long totals[256];
  d2h(ctx, totals, totals_d.p, 256);
  sync(ctx);
  CodeLut lut;
  uint8_t sym_of_code[256];
  memset(&lut, 0xFF, sizeof(lut));
  memset(sym_of_code, 0, sizeof(sym_of_code));
  int nc = 0;
  for (int c = 0; c < 256; ++c) {
    r->counts[c] = (long)totals[c];
    r->code_of[c] = -1;
    if (totals[c]) {
      if (c == 255) { delete r; set_error("byte 255 in the BWT"); throw CudaError{PSC_ESENTINEL}; }
      r->code_of[c] = (int16_t)nc;
      lut.code[c] = (uint8_t)nc;
      sym_of_code[nc] = (uint8_t)c;
      ++nc;
    }
  }
  r->n_codes = nc;
  size_t bytes = (size_t)std::max(nc, 1) * (size_t)r->n_sectors * 32 + 256;
  try {
    r->sectors_d = (uint32_t *)cache_alloc(ctx, bytes);
  } catch (...) { delete r; throw; }
  if (nc > 0) {
    DevBuf<uint8_t> soc(ctx, 256);
    h2d(ctx, soc.p, sym_of_code, 256);
    int code_bits = 1;
    while ((1 << code_bits) < nc) ++code_bits;
    if (nc <= RK_GROUP)
      rank_fill_kernel<<<dim3((unsigned)nchunks, 1), 256, 0, ctx->stream>>>(bwt_d, m, i0, lut, nc, code_bits, r->n_sectors,
                                                                             cc.p, soc.p, r->sectors_d);
    else
      rank_fill_all_kernel<<<(unsigned)nchunks, 256, 0, ctx->stream>>>(bwt_d, m, i0, lut, nc, code_bits, r->n_sectors,
                                                                        cc.p, soc.p, r->sectors_d);
    PSC_LAUNCH_CHECK();
    sync(ctx);  // sym_of_code is a stack array
  }
  ctx->launches += 3;
  return r;
}

void rank_free(psc_rank *r) {
  if (!r) return;
  if (r->sectors_d) cache_release(r->ctx, r->sectors_d);
  delete r;
}

struct CodeLut16 { int16_t code[256]; };

__global__ void rank_query_kernel(RankView rv, CodeLut16 lut, long nq, const uint64_t *__restrict__ qi,
                                  const uint8_t *__restrict__ qc, uint64_t *__restrict__ out) {
  long k = (long)blockIdx.x * blockDim.x + threadIdx.x;
  if (k >= nq) return;
  int code = lut.code[qc[k]];
  uint64_t i = qi[k];
  if (i > rv.m) i = rv.m;  // rank.hpp:567-568
  out[k] = code < 0 ? 0 : rank_occ(rv, code, (uint32_t)i);
}

void rank_query(psc_context *ctx, const psc_rank *r, long nq, const uint64_t *i_d, const uint8_t *c_d,
                uint64_t *out_d) {
  if (nq <= 0) return;
  RankView rv{r->sectors_d, r->n_sectors, (uint32_t)r->m};
  CodeLut16 lut;
  for (int c = 0; c < 256; ++c) lut.code[c] = r->code_of[c];
  rank_query_kernel<<<(unsigned)div_up(nq, 256), 256, 0, ctx->stream>>>(rv, lut, nq, i_d, c_d, out_d);
  PSC_LAUNCH_CHECK();
  ctx->launches++;
}

}  // namespace psc
