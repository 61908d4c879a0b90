/*
 * psascan_oracle.h -- CPU restatement of the pSAscan hot path.  TEST INFRASTRUCTURE ONLY.
 *
 * This is the checker, not the product: only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may load this library.  Nothing under
 * psascan_b200/ links, imports or calls it.
 *
 * Parity pin: the reference ships no golden vectors or tests, so the
 * restatement is pinned against (a) outputs of the unmodified reference binary built by
 * oracle/Makefile into oracle/_ref/ (the .sa5 files under tests/golden/, made by tests/golden/make_golden.py)
 * and (b) a brute-force suffix sort; see tests/test_oracle.py.
 *
 * All file:line citations are into /root/reference/.
 *
 * Conventions used by every function here
 *   - text positions are GLOBAL (0..n), block-relative values are called "rel".
 *   - gt flag arrays are plain byte arrays indexed by global position j in [0, n]:
 *       gt_X[j] = 1  iff  suffix(j) > suffix(X)   (X = the reference point).
 *     The reference stores the same bits reversed and LSB-first
 *     (include/bitvector.hpp:61-67, include/io/multifile.hpp:57-68); orc_pack_gt_rev()
 *     converts.
 *   - gap arrays are long[m+1].
 */
#ifndef PSASCAN_ORACLE_H
#define PSASCAN_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

/* src/main.cpp:76-131  parse_number(): digits + optional k/m/g/t[i] suffix. 1 = ok, 0 = reject. */
int orc_parse_number(const char *str, uint64_t *out);

/* include/psascan.hpp:53-91  memory budget -> max block size.
 * returns max_block_size, or -1 ("not enough memory to run"), -2 ("... to start threads"). */
long orc_max_block_size(long ram_use, long max_threads);

/* include/types/uint40.hpp:42-46,104 + include/io/async_stream_writer.hpp:141:
 * 5 bytes per entry, uint32 low little-endian then uint8 high. */
void orc_encode_uint40(const int64_t *sa, long n, uint8_t *out);
void orc_decode_uint40(const uint8_t *in, long n, int64_t *sa);

/* Ground truth by definition: qsort of suffixes with memcmp (tiny inputs). */
void orc_sa_naive(const uint8_t *text, long n, int64_t *sa);

/* Plain suffix sort of an int string ("shorter is smaller"), standing in for the vendored
 * sorter call divsufsort()/libsais() at inmem_psascan_src/initial_partial_sufsort.hpp:172-178,
 * 278-284 (sais_template.hpp:56-63).  O(n log n) prefix doubling with counting sorts. */
void orc_suffix_sort_int(const int32_t *s, long n, int32_t sigma, int64_t *sa);

/* Whole-text SA through orc_suffix_sort_int. */
void orc_sa(const uint8_t *text, long n, int64_t *sa);

/* inmem_psascan_src/compute_initial_gt_bitvectors.hpp:61-148 (semantics; the reference's
 * ms-decomposition speed-up at :121-148 changes cost only).
 * For p in [beg,end): gt_end[p] = suffix(p) > suffix(end).  gt_tail = gt_end flags for
 * positions (end, n] (ignored when end == n). Writes gt_out[beg..end). */
void orc_gt_end(const uint8_t *text, long n, long beg, long end,
                const uint8_t *gt_tail, uint8_t *gt_out);

/* inmem_psascan() contract, inmem_psascan_src/inmem_psascan.hpp:64-80 and SURVEY 3.4:
 * rename (initial_partial_sufsort.hpp:61-80) -> suffix sort -> BWT with 0 placeholder at i0
 * (inmem_bwt_from_sa.hpp:48-56) -> gt_begin (change_gt_reference_point.hpp:62-118).
 *   psa_rel[m]  block-relative SA in whole-text order
 *   bwt[m]      bwt[i] = text[beg+psa_rel[i]-1], 0 at i0          (may be NULL)
 *   gt_begin    flags written for global positions (beg, end]      (may be NULL)
 * returns i0 (index of the block's first suffix in psa_rel).
 * Returns -1 if the block holds byte 255 (initial_partial_sufsort.hpp:70-75). */
long orc_sort_block(const uint8_t *text, long n, long beg, long end,
                    const uint8_t *gt_tail, int64_t *psa_rel, uint8_t *bwt,
                    uint8_t *gt_begin);

/* include/em_compute_initial_ranks.hpp:54-76 (lcp_compare), :321-363 (lcp_compare_2),
 * :164-220 / :453-511: number of suffixes of block [beg,end) smaller than suffix(pat_beg),
 * pat_beg >= cmp_end >= end.  Block suffixes are followed up to cmp_end (= end for overload 1;
 * = tail_begin, i.e. through the "mid block", for overload 2); gt_cmp = flags relative to
 * cmp_end.  The reference caps the pattern at one slice and resolves ties through the next
 * slice's rank (:271-314); the number returned is the same exact rank. */
long orc_initial_rank(const uint8_t *text, long n, long beg, long end,
                      const int64_t *psa_rel, long pat_beg, long cmp_end,
                      const uint8_t *gt_cmp);

/* include/compute_gap.hpp:77-85 (C array) + include/stream.hpp:147-158 (LF step) +
 * include/update.hpp:86-96 (gap[i]++), with occ() by plain counting instead of rank4n
 * (include/rank.hpp:566-708 answers the same query).
 * Streams j = tail_end .. tail_beg+1; start_rank = rank of suffix(tail_end).
 *   gt_in   flags relative to the block END, read for j in (tail_beg, tail_end]
 *   gt_out  flags relative to the block BEGIN, written for j in (tail_beg, tail_end]
 *   gap     long[m+1], incremented
 * returns the rank of suffix(tail_beg). */
long orc_compute_gap(const uint8_t *text, const uint8_t *bwt, long m, long i0,
                     uint8_t last, long tail_beg, long tail_end, long start_rank,
                     const uint8_t *gt_in, uint8_t *gt_out, int64_t *gap);

/* include/gap_array.hpp:273-364  gap (of the left piece wrt the right piece) -> bitvector of
 * length len_l+len_r, 0 = next suffix comes from the left piece, 1 = from the right. */
void orc_gap_to_bv(const int64_t *gap_l, long len_l, long len_r, uint8_t *bv);

/* include/bwt_merge.hpp:66-140  returns block_i0 = select0(bv, i0_l). */
long orc_merge_bwt(const uint8_t *bwt_l, const uint8_t *bwt_r, long len_l, long len_r,
                   long i0_l, long i0_r, uint8_t last_l, const uint8_t *bv, uint8_t *bwt);

/* include/compute_left_gap.hpp:55-123,162-306 and compute_right_gap.hpp:55-122,161-305.
 * block_gap has len_l+len_r+1 entries; left_gap len_l+1, right_gap len_r+1. */
void orc_split_gap(const uint8_t *bv, long len_l, long len_r, const int64_t *block_gap,
                   int64_t *left_gap, int64_t *right_gap);

/* include/merge.hpp:54-180.  pieces sorted by beg; gap[k] has len[k]+1 entries (the last
 * piece's gap is ignored / may be NULL). Output sa[n] of global positions. */
void orc_merge(long n_pieces, const long *beg, const long *len,
               const int64_t *const *psa_rel, const int64_t *const *gap, int64_t *sa);

/* include/partial_sufsort.hpp:68-551,559-584 + include/psascan.hpp:117-125: the whole block
 * pipeline, right to left, then the merge.  half_blocks = 1 follows process_block exactly
 * (left/right half-blocks, steps 1-6); half_blocks = 0 is the collapsed schedule the GPU build
 * uses (one piece per block: stages 1, 2, 3 then the merge; SURVEY.md 3.2 last paragraph).
 * ram_use only decides how the LAST block is halved (partial_sufsort.hpp:88).
 * returns 0, or -1 when the text holds byte 255 in a position where the reference rejects it. */
int orc_psascan(const uint8_t *text, long n, long max_block_size, long ram_use,
                int half_blocks, int64_t *sa);

/* reversed LSB-first bit packing of flags for positions (lo, hi] exactly as a reference gt
 * file registered for [n-hi, n-lo) holds them (include/bitvector.hpp:61-67,
 * include/partial_sufsort.hpp:233,358): bit r-(n-hi) <-> global position n-r. */
void orc_pack_gt_rev(const uint8_t *gt, long n, long lo, long hi, uint8_t *bits);

#ifdef __cplusplus
}
#endif
#endif
