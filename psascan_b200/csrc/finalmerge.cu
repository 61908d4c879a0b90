// finalmerge.cu -- stage 4 as ONE N-way merge by output range (replaces the level-after-level cascade
// over the whole text that round 1 ran after every block).
//
// The reference merges all (half-)block suffix arrays at the end, guided by their gap arrays
// (/root/reference/include/merge.hpp:54-180).  Here the blocks' SAs and gap arrays are kept until the
// end as well, and the merged sequence is produced chunk by chunk of OUTPUT slots:
//   M_k = merge of blocks k..B-1;  in M_k block k's u-th suffix sits at slot u + E_k[u+1]
//   (E_k = exclusive scan of gap_k) and the entries of M_{k+1} fill the other slots in order.
//   A range [a, b) of M_k therefore consists of block-k suffixes [u(a), u(b)) and the range
//   [a - u(a), b - u(b)) of M_{k+1}, where u(q) = number of block suffixes in slots < q
//   = the largest u with u + E_k[u] <= q.  Going down the levels, a chunk of final output slots
//   needs ONE contiguous slice of every block's SA and gap array ("split", B small searches per
//   chunk boundary), and is then merged bottom-up from those slices alone ("cascade").
// What this buys: (1) nothing of size n is double-buffered -- the pipeline keeps 8 bytes per suffix
// (SA + gap) instead of two merged planes; (2) over several GPUs every rank merges its own output
// range after ONE exchange of slices instead of an all-to-all per block (dist.cu); (3) chunks can be
// streamed to the host while the next one is merged, and the slices can come from host memory.
#include <algorithm>
#include <vector>
#include "scan.cuh"
#include "finalmerge.cuh"

namespace psc {

// ---- split: u(q) for a batch of slots, against one slice of one block's gap array ------------------
// The slice holds gap entries [u_off, u_off + cnt) of gap[0..m]; E[u_off] = e_base and carry[j] is
// the sum of the slice's first j * COARSE_TILE entries (carry[tiles] = the slice total).  P(u) :=
// u + E[u] <= q is true for u = 0 and false from some u on; the slice answers a query only if the
// last true u lies inside it (last_slice: the slice ends with entry m), else it writes 0, so that the
// answers of all slices of a block can simply be summed (dist.cu: one all-reduce per level).
__global__ void __launch_bounds__(64)
    merge_split_kernel(const uint32_t *__restrict__ gap, long cnt, long u_off, uint64_t e_base,
                       const uint64_t *__restrict__ carry, long tiles, int last_slice,
                       const uint64_t *__restrict__ q, int nq, uint64_t *__restrict__ u_out,
                       uint64_t *__restrict__ e_out) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nq) return;
  const uint64_t want = q[j];
  uint64_t ru = 0, re = 0;
  const bool p_first = (uint64_t)u_off + e_base <= want;
  const bool p_end = (uint64_t)(u_off + cnt) + e_base + carry[tiles] <= want;  // P at the first entry after the slice
  if (cnt > 0 && p_first && (last_slice || !p_end)) {
    long lo = 0, hi = tiles - 1;  // largest tile whose first entry satisfies P
    while (lo < hi) {
      const long mid = (lo + hi + 1) >> 1;
      if ((uint64_t)(u_off + mid * COARSE_TILE) + e_base + carry[mid] <= want) lo = mid; else hi = mid - 1;
    }
    long k = lo * COARSE_TILE;  // index inside the slice
    uint64_t e = e_base + carry[lo];
    const long stop = min(cnt, k + COARSE_TILE) - 1;
    while (k < stop) {
      const uint64_t e1 = e + gap[k];
      if ((uint64_t)(u_off + k + 1) + e1 > want) break;
      e = e1;
      ++k;
    }
    ru = (uint64_t)(u_off + k);
    re = e;
  }
  u_out[j] = ru;
  e_out[j] = re;
}

// slots of the next level: q' = q - u(q)
__global__ void merge_split_next_kernel(const uint64_t *__restrict__ q, const uint64_t *__restrict__ u, int nq,
                                        uint64_t *__restrict__ q_next) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j < nq) q_next[j] = q[j] - u[j];
}

void merge_split_slice(psc_context *ctx, const uint32_t *gap_slice_d, long cnt, long u_off, uint64_t e_base,
                       const uint64_t *carry_d, bool last_slice, const uint64_t *q_d, int nq, uint64_t *u_d,
                       uint64_t *e_d) {
  merge_split_kernel<<<div_up(nq, 64), 64, 0, ctx->stream>>>(gap_slice_d, cnt, u_off, e_base, carry_d,
                                                             coarse_carry_tiles(cnt), last_slice ? 1 : 0, q_d, nq, u_d, e_d);
  PSC_LAUNCH_CHECK();
  ctx->launches++;
}

void merge_split_next(psc_context *ctx, const uint64_t *q_d, const uint64_t *u_d, int nq, uint64_t *q_next_d) {
  merge_split_next_kernel<<<div_up(nq, 64), 64, 0, ctx->stream>>>(q_d, u_d, nq, q_next_d);
  PSC_LAUNCH_CHECK();
  ctx->launches++;
}

// ---- cascade: one chunk of output slots from its slices ---------------------------------------------
void MergeScratch::reserve(psc_context *ctx, long entries, bool hi_plane) {
  if (entries <= cap && hi_plane == has_hi) return;
  cap = entries;
  has_hi = hi_plane;
  for (int k = 0; k < 2; ++k) {
    lo[k].alloc(ctx, entries);
    if (hi_plane) hi[k].alloc(ctx, entries); else hi[k].release();
  }
  E.alloc(ctx, entries + 2);
}

void cascade_chunk(psc_context *ctx, const LevelSlice *lv, int n_levels, long total, MergeScratch &scr,
                   uint8_t *out5_d) {
  if (total <= 0) return;
  if (total > scr.cap) { set_error("cascade_chunk: chunk of %ld entries exceeds the scratch (%ld)", total, scr.cap); throw CudaError{PSC_EARG}; }
  int cur = -1;
  long have = 0;  // entries merged so far: the chunk's share of the levels below
  for (int k = n_levels - 1; k >= 0; --k) {
    const long cb = lv[k].cb;
    if (cb == 0) continue;  // no suffix of this block in the chunk: the level is the identity
    const int nxt = cur < 0 ? 0 : 1 - cur;
    if (have == 0) {
      sa_slice_to_planes(ctx, lv[k].sa, cb, lv[k].beg, scr.lo[nxt].p, scr.hi[nxt].p);
    } else {
      // chunk-local scan: the first `skip` right-side entries of the first run belong to the chunk before
      exclusive_scan_u32_to_u64(ctx, lv[k].gap, scr.E.p, cb + 1, nullptr, lv[k].skip);
      merge_level_range(ctx, lv[k].sa, scr.E.p, cb, lv[k].beg, scr.lo[cur].p, scr.hi[cur].p, 0, 0, have + cb,
                        scr.lo[nxt].p, scr.hi[nxt].p);
    }
    cur = nxt;
    have += cb;
  }
  if (have != total) { set_error("cascade_chunk: slices hold %ld entries, the chunk has %ld slots", have, total); throw CudaError{PSC_EARG}; }
  pack_sa5(ctx, scr.lo[cur].p, scr.hi[cur].p, total, out5_d);
}

// ---- one GPU, everything resident -------------------------------------------------------------------
long merge_chunk_entries() {
  static long c = 0;
  if (!c) {
    const char *e = getenv("PSC_MERGE_CHUNK");
    c = e ? std::max(4L, atol(e) / 4 * 4) : (1L << 27);
  }
  return c;
}

void MergePlan::resize(int levels, int boundaries) {
  B = levels;
  nb = boundaries;
  q.assign((size_t)B * nb, 0); u.assign((size_t)B * nb, 0); e.assign((size_t)B * nb, 0);
}

LevelSlice MergePlan::slice(int k, int i, const uint32_t *sa_at_ua, const uint32_t *gap_at_ua, long beg) const {
  LevelSlice s;
  const uint64_t ua = U(k, i), ub = U(k, i + 1);
  s.sa = sa_at_ua;
  s.gap = gap_at_ua;
  s.cb = (long)(ub - ua);
  s.skip = (k == B - 1) ? 0 : Q(k, i) - (ua + Ev(k, i));
  s.beg = beg;
  return s;
}

void final_merge_resident(psc_context *ctx, const std::vector<MergePiece> &pc, long n, uint8_t *out5_d,
                          const ChunkSink *sink) {
  PhaseTimer timer(ctx, &ctx->stats.ms_merge);
  const int B = (int)pc.size();
  const long C = merge_chunk_entries();
  const int nb = (int)div_up(n, C) + 1;
  MergePlan plan;
  plan.resize(B, nb);
  for (int i = 0; i < nb; ++i) plan.q[i] = (uint64_t)std::min(n, (long)i * C);
  {
    KTrace kt(ctx, "merge_split");
    DevBuf<uint64_t> tab(ctx, (size_t)3 * B * nb);
    uint64_t *Qd = tab.p, *Ud = tab.p + (size_t)B * nb, *Ed = tab.p + (size_t)2 * B * nb;
    DevBuf<uint64_t> sums(ctx, B);  // sum of every block's gap counters
    h2d(ctx, Qd, plan.q.data(), nb);
    for (int k = 0; k < B; ++k) {
      uint64_t *qk = Qd + (size_t)k * nb, *uk = Ud + (size_t)k * nb, *ek = Ed + (size_t)k * nb;
      if (k == B - 1) {  // the rightmost block: every slot is a block suffix
        PSC_CUDA(cudaMemcpyAsync(uk, qk, (size_t)nb * 8, cudaMemcpyDeviceToDevice, ctx->stream));
        PSC_CUDA(cudaMemsetAsync(ek, 0, (size_t)nb * 8, ctx->stream));
        break;
      }
      const long cnt = pc[k].m + 1;
      DevBuf<uint64_t> carry(ctx, coarse_carry_tiles(cnt) + 1);
      coarse_carries_u32(ctx, pc[k].gap, cnt, carry.p);
      PSC_CUDA(cudaMemcpyAsync(sums.p + k, carry.p + coarse_carry_tiles(cnt), 8, cudaMemcpyDeviceToDevice, ctx->stream));
      merge_split_slice(ctx, pc[k].gap, cnt, 0, 0, carry.p, true, qk, nb, uk, ek);
      merge_split_next(ctx, qk, uk, nb, qk + nb);
    }
    d2h(ctx, plan.q.data(), Qd, (size_t)B * nb);
    d2h(ctx, plan.u.data(), Ud, (size_t)B * nb);
    d2h(ctx, plan.e.data(), Ed, (size_t)B * nb);
    std::vector<uint64_t> hs(B, 0);
    d2h(ctx, hs.data(), sums.p, (size_t)(B - 1));
    sync(ctx);
    // uint32 counters (the reference: 1 byte + excess list, gap_array.hpp:63, 79-88) must add up to the tail length
    for (int k = 0; k + 1 < B; ++k) {
      const uint64_t tail = (uint64_t)(n - (pc[k].beg + pc[k].m));
      if (hs[k] != tail) {
        set_error("gap counters of block %d add up to %llu, its tail has %llu suffixes: a 32-bit counter overflowed", k,
                  (unsigned long long)hs[k], (unsigned long long)tail);
        throw CudaError{PSC_ELIMIT};
      }
    }
  }
  MergeScratch scr;
  scr.reserve(ctx, std::min(n, C) + 16, need_hi_plane(n));
  std::vector<LevelSlice> lv(B);
  for (int i = 0; i + 1 < nb; ++i) {
    const long a = (long)plan.Q(0, i), b = (long)plan.Q(0, i + 1);
    for (int k = 0; k < B; ++k) {
      const uint64_t ua = plan.U(k, i);
      lv[k] = plan.slice(k, i, pc[k].sa + ua, pc[k].gap ? pc[k].gap + ua : nullptr, pc[k].beg);
    }
    uint8_t *dst = sink ? sink->buffer(ctx, i, b - a) : out5_d + 5 * a;
    cascade_chunk(ctx, lv.data(), B, b - a, scr, dst);
    if (sink) sink->done(ctx, i, a, b - a);
  }
}

}  // namespace psc
