// api_stages.cu -- per-stage C ABI entry points with HOST buffers: each one is the drop-in for one
// reference seam (see include/psascan_cuda.h) and is what the parity tests call.
#include <algorithm>
#include "common.cuh"
#include "texthash.cuh"

namespace psc {
template <typename F>
static int guarded2(F &&f) {
  try { f(); return PSC_OK; }
  catch (const CudaError &e) { return e.code; }
  catch (const std::exception &e) { set_error("%s", e.what()); return PSC_ECUDA; }
}
static void need(bool ok, const char *what) {
  if (!ok) { set_error("bad argument: %s", what); throw CudaError{PSC_EARG}; }
}
// host reversed gt bytes (nbits) -> device forward words (zero padded)
static void upload_gt(psc_context *ctx, const uint8_t *rev, long nbits, DevBuf<uint32_t> *fwd) {
  fwd->alloc(ctx, nbits / 32 + 2);
  fwd->zero();
  if (nbits <= 0 || !rev) return;
  DevBuf<uint8_t> tmp(ctx, (nbits + 7) / 8);
  h2d(ctx, tmp.p, rev, (nbits + 7) / 8);
  bits_rev_to_forward(ctx, tmp.p, nbits, fwd->p);
}
static void download_gt(psc_context *ctx, const uint32_t *fwd_d, long nbits, uint8_t *rev) {
  if (nbits <= 0) return;
  DevBuf<uint8_t> tmp(ctx, (nbits + 7) / 8);
  bits_forward_to_rev(ctx, fwd_d, nbits, tmp.p);
  d2h(ctx, rev, tmp.p, (nbits + 7) / 8);
  sync(ctx);
}
__global__ void or_not_bit2_kernel(const uint32_t *src_bits, int src_default, uint32_t *dst, long idx) {
  bool s = src_bits ? ((src_bits[0] & 1u) != 0) : (src_default != 0);
  if (!s) atomicOr(&dst[idx >> 5], 1u << (idx & 31));
}
}  // namespace psc
using namespace psc;

extern "C" {

int psc_sort_block(psc_context *ctx, const uint8_t *text, long m, const uint8_t *tail_prefix,
                   long tail_prefix_len, const uint8_t *tail_gt_rev, uint32_t *sa_out,
                   uint8_t *bwt_out, uint8_t *gt_begin_rev, long *i0) {
  return guarded2([&] {
    need(ctx && text && m > 0 && sa_out, "psc_sort_block: ctx/text/m/sa_out");
    need(tail_prefix_len == 0 || (tail_prefix && tail_gt_rev), "psc_sort_block: tail prefix without gt bits");
    need(m <= (1L << 31) - 4096, "psc_sort_block: block too long");
    PSC_CUDA(cudaSetDevice(ctx->device));
    long t = tail_prefix_len;
    // fingerprints (K0) live on this call's own buffer: no domain may survive the call
    hash_domain_clear(ctx);
    struct DomainGuard { psc_context *c; ~DomainGuard() { hash_domain_clear(c); } } domain_guard{ctx};
    DevBuf<uint8_t> buf(ctx, m + t);  // block followed by the tail prefix, as in the text
    h2d(ctx, buf.p, text, m);
    if (t) h2d(ctx, buf.p + m, tail_prefix, t);
    DevBuf<uint32_t> gt_after, gt_end;
    if (t) {
      upload_gt(ctx, tail_gt_rev, t, &gt_after);
      gt_end.alloc(ctx, m / 32 + 2);
      // the caller passes min(m, tail) symbols: a match that exhausts them without reaching the
      // block end can only happen when they reach the text end (inmem_psascan.hpp:150-156)
      compute_gt_end(ctx, buf.p, m, buf.p + m, t, t < m ? 1 : 0, gt_after.p, 0, gt_end.p);
    }
    BlockSortResult res;
    sort_block(ctx, buf.p, m, t ? gt_end.p : nullptr, &res);
    d2h(ctx, sa_out, res.sa.p, m);
    if (bwt_out) {
      DevBuf<uint8_t> bwt(ctx, m);
      bwt_from_sa(ctx, buf.p, res.sa.p, m, bwt.p);
      d2h(ctx, bwt_out, bwt.p, m);
      sync(ctx);
    }
    if (gt_begin_rev) {
      DevBuf<uint32_t> g(ctx, m / 32 + 2);
      g.zero();
      gt_begin_from_sa(ctx, res.sa.p, m, (uint32_t)res.i0, g.p, 0);
      or_not_bit2_kernel<<<1, 1, 0, ctx->stream>>>(t ? gt_end.p : nullptr, 1, g.p, m - 1);
      PSC_LAUNCH_CHECK();
      download_gt(ctx, g.p, m, gt_begin_rev);
    }
    sync(ctx);
    if (i0) *i0 = res.i0;
  });
}

int psc_rank_build(psc_context *ctx, const uint8_t *bwt, long m, long i0, psc_rank **rank) {
  return guarded2([&] {
    need(ctx && bwt && m > 0 && rank, "psc_rank_build");
    PSC_CUDA(cudaSetDevice(ctx->device));
    DevBuf<uint8_t> b(ctx, m);
    h2d(ctx, b.p, bwt, m);
    *rank = rank_build(ctx, b.p, m, i0);
    sync(ctx);
  });
}

int psc_rank_counts(const psc_rank *rank, long counts[256]) {
  if (!rank || !counts) { set_error("psc_rank_counts: null pointer"); return PSC_EARG; }
  for (int c = 0; c < 256; ++c) counts[c] = rank->counts[c];
  return PSC_OK;
}

int psc_rank_query(psc_context *ctx, const psc_rank *rank, long nq, const long *i, const uint8_t *c,
                   long *out) {
  return guarded2([&] {
    need(ctx && rank && nq >= 0 && (nq == 0 || (i && c && out)), "psc_rank_query");
    if (nq == 0) return;
    PSC_CUDA(cudaSetDevice(ctx->device));
    std::vector<uint64_t> qi(nq);
    for (long k = 0; k < nq; ++k) qi[k] = i[k] < 0 ? 0 : (uint64_t)i[k];  // rank.hpp:567
    DevBuf<uint64_t> i_d(ctx, nq), o_d(ctx, nq);
    DevBuf<uint8_t> c_d(ctx, nq);
    h2d(ctx, i_d.p, qi.data(), nq);
    h2d(ctx, c_d.p, c, nq);
    rank_query(ctx, rank, nq, i_d.p, c_d.p, o_d.p);
    std::vector<uint64_t> o(nq);
    d2h(ctx, o.data(), o_d.p, nq);
    sync(ctx);
    for (long k = 0; k < nq; ++k) out[k] = (long)o[k];
  });
}

int psc_rank_free(psc_context *ctx, psc_rank *rank) {
  (void)ctx;
  rank_free(rank);
  return PSC_OK;
}

int psc_initial_ranks(psc_context *ctx, const uint8_t *block_text, const uint32_t *block_sa, long m,
                      const uint8_t *after_text, long after_len, const uint8_t *after_gt_rev,
                      long nstarts, const long *starts, long *out) {
  return guarded2([&] {
    need(ctx && block_text && block_sa && m > 0 && after_len >= 0 && nstarts >= 0, "psc_initial_ranks");
    need(after_len == 0 || (after_text && after_gt_rev), "psc_initial_ranks: after_text/gt");
    if (nstarts == 0) return;
    need(starts && out, "psc_initial_ranks: starts/out");
    PSC_CUDA(cudaSetDevice(ctx->device));
    hash_domain_clear(ctx);
    struct DomainGuard { psc_context *c; ~DomainGuard() { hash_domain_clear(c); } } domain_guard{ctx};
    DevBuf<uint8_t> buf(ctx, m + after_len);
    h2d(ctx, buf.p, block_text, m);
    if (after_len) h2d(ctx, buf.p + m, after_text, after_len);
    DevBuf<uint32_t> sa_d(ctx, m), gt;
    h2d(ctx, sa_d.p, block_sa, m);
    upload_gt(ctx, after_gt_rev, after_len, &gt);
    std::vector<uint64_t> st(nstarts);
    for (long k = 0; k < nstarts; ++k) {
      need(starts[k] >= 0 && starts[k] <= after_len, "psc_initial_ranks: start out of range");
      st[k] = (uint64_t)starts[k];
    }
    DevBuf<uint64_t> st_d(ctx, nstarts), o_d(ctx, nstarts);
    h2d(ctx, st_d.p, st.data(), nstarts);
    initial_ranks(ctx, buf.p, sa_d.p, m, buf.p + m, after_len, 1, gt.p, 0, nstarts, st_d.p, o_d.p);
    std::vector<uint64_t> o(nstarts);
    d2h(ctx, o.data(), o_d.p, nstarts);
    sync(ctx);
    for (long k = 0; k < nstarts; ++k) out[k] = (long)o[k];
  });
}

int psc_compute_gap(psc_context *ctx, const psc_rank *rank, long i0, uint8_t last,
                    const long block_counts[256], const uint8_t *tail, long t, long nwalk,
                    const long *walk_beg, const long *start_rank, const uint8_t *gt_in_rev,
                    uint8_t *gt_out_rev, uint32_t *gap) {
  return guarded2([&] {
    need(ctx && rank && block_counts && tail && t > 0 && nwalk > 0 && walk_beg && start_rank && gap,
         "psc_compute_gap");
    need(walk_beg[0] == 0 && walk_beg[nwalk] == t, "psc_compute_gap: walkers must cover [0, t)");
    for (long k = 1; k < nwalk; ++k)
      need(walk_beg[k] % 32 == 0 && walk_beg[k] >= walk_beg[k - 1], "psc_compute_gap: interior boundary");
    PSC_CUDA(cudaSetDevice(ctx->device));
    long m = rank->m;
    uint64_t Cmap[256], run = 0;
    for (int c = 0; c < 256; ++c) { Cmap[c] = run; run += (uint64_t)block_counts[c]; }
    need((long)run == m, "psc_compute_gap: block_counts must sum to the block length");
    DevBuf<uint8_t> tail_d(ctx, t);
    h2d(ctx, tail_d.p, tail, t);
    DevBuf<uint32_t> gin, gout, gap_d(ctx, m + 1);
    upload_gt(ctx, gt_in_rev, gt_in_rev ? t : 0, &gin);
    if (!gt_in_rev) { gin.alloc(ctx, t / 32 + 2); gin.zero(); }
    if (gt_out_rev) { gout.alloc(ctx, t / 32 + 2); gout.zero(); }
    h2d(ctx, gap_d.p, gap, m + 1);
    std::vector<uint64_t> wb(nwalk + 1), sr(nwalk);
    for (long k = 0; k <= nwalk; ++k) wb[k] = (uint64_t)walk_beg[k];
    for (long k = 0; k < nwalk; ++k) sr[k] = (uint64_t)start_rank[k];
    DevBuf<uint64_t> wb_d(ctx, nwalk + 1), sr_d(ctx, nwalk);
    h2d(ctx, wb_d.p, wb.data(), nwalk + 1);
    h2d(ctx, sr_d.p, sr.data(), nwalk);
    lf_walk(ctx, rank, (uint32_t)i0, last, Cmap, tail_d.p, t, 0, nwalk, wb_d.p, sr_d.p, gin.p, 0,
            gt_out_rev ? gout.p : nullptr, 0, gap_d.p);
    d2h(ctx, gap, gap_d.p, m + 1);
    sync(ctx);
    if (gt_out_rev) download_gt(ctx, gout.p, t, gt_out_rev);
  });
}

int psc_merge_level(psc_context *ctx, const uint32_t *sa, const uint32_t *gap, long m, long beg,
                    const uint8_t *right_sa5, long right_len, uint8_t *out_sa5) {
  return guarded2([&] {
    need(ctx && sa && m > 0 && beg >= 0 && right_len >= 0 && out_sa5, "psc_merge_level");
    need(right_len == 0 || (gap && right_sa5), "psc_merge_level: gap/right");
    PSC_CUDA(cudaSetDevice(ctx->device));
    long total = m + right_len;
    DevBuf<uint32_t> sa_d(ctx, m), gap_d(ctx, m + 1), rlo(ctx, right_len), olo(ctx, total);
    DevBuf<uint8_t> rhi(ctx, right_len), ohi(ctx, total), r5(ctx, 5 * (size_t)right_len),
        o5(ctx, 5 * (size_t)total);
    h2d(ctx, sa_d.p, sa, m);
    if (right_len) {
      h2d(ctx, gap_d.p, gap, m + 1);
      h2d(ctx, r5.p, right_sa5, 5 * (size_t)right_len);
      unpack_sa5(ctx, r5.p, right_len, rlo.p, rhi.p);
    }
    merge_level(ctx, sa_d.p, gap_d.p, m, beg, rlo.p, rhi.p, right_len, olo.p, ohi.p);
    pack_sa5(ctx, olo.p, ohi.p, total, o5.p);
    d2h(ctx, out_sa5, o5.p, 5 * (size_t)total);
    sync(ctx);
  });
}

}  // extern "C"

// ---- glue for a block sorted as two pieces -----------------------------------------------------
extern "C" {

int psc_gap_to_bitvector(psc_context *ctx, const uint32_t *gap_l, long len_l, long len_r, uint8_t *bv) {
  return guarded2([&] {
    need(ctx && gap_l && len_l >= 0 && len_r >= 0 && bv, "psc_gap_to_bitvector");
    PSC_CUDA(cudaSetDevice(ctx->device));
    long len = len_l + len_r;
    DevBuf<uint32_t> g(ctx, len_l + 1), b(ctx, len / 32 + 2);
    h2d(ctx, g.p, gap_l, len_l + 1);
    b.zero();
    gap_to_bitvector(ctx, g.p, len_l, len_r, b.p);
    d2h(ctx, bv, reinterpret_cast<const uint8_t *>(b.p), (len + 7) / 8);
    sync(ctx);
  });
}

int psc_merge_bwt(psc_context *ctx, const uint8_t *bwt_l, const uint8_t *bwt_r, long len_l, long len_r,
                  long i0_l, long i0_r, uint8_t last_l, const uint8_t *bv, uint8_t *bwt_out,
                  long *i0_out) {
  return guarded2([&] {
    need(ctx && bwt_l && bwt_r && len_l > 0 && len_r > 0 && bv && bwt_out && i0_out, "psc_merge_bwt");
    PSC_CUDA(cudaSetDevice(ctx->device));
    long len = len_l + len_r;
    DevBuf<uint8_t> l(ctx, len_l), r(ctx, len_r), o(ctx, len);
    DevBuf<uint32_t> b(ctx, len / 32 + 2);
    b.zero();
    h2d(ctx, l.p, bwt_l, len_l);
    h2d(ctx, r.p, bwt_r, len_r);
    h2d(ctx, reinterpret_cast<uint8_t *>(b.p), bv, (len + 7) / 8);
    merge_bwt(ctx, l.p, r.p, len_l, len_r, i0_l, i0_r, last_l, b.p, o.p, i0_out);
    d2h(ctx, bwt_out, o.p, len);
    sync(ctx);
  });
}

int psc_split_gap(psc_context *ctx, const uint8_t *bv, long len_l, long len_r, const uint32_t *block_gap,
                  uint64_t *left_gap, uint64_t *right_gap) {
  return guarded2([&] {
    need(ctx && bv && len_l >= 0 && len_r >= 0 && block_gap && left_gap && right_gap, "psc_split_gap");
    PSC_CUDA(cudaSetDevice(ctx->device));
    long len = len_l + len_r;
    DevBuf<uint32_t> b(ctx, len / 32 + 2), g(ctx, len + 1);
    DevBuf<uint64_t> lg(ctx, len_l + 1), rg(ctx, len_r + 1);
    b.zero();
    h2d(ctx, reinterpret_cast<uint8_t *>(b.p), bv, (len + 7) / 8);
    h2d(ctx, g.p, block_gap, len + 1);
    split_gap(ctx, b.p, len_l, len_r, g.p, lg.p, rg.p);
    d2h(ctx, left_gap, lg.p, len_l + 1);
    d2h(ctx, right_gap, rg.p, len_r + 1);
    sync(ctx);
  });
}

}  // extern "C"
