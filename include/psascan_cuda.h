/*
 * psascan_cuda.h -- C ABI of libpsascan_cuda.so, the B200 (sm_100a) implementation of the
 * pSAscan hot path.  One export per seam of the reference (pdinklag/psascan); citations are
 * file:line into the reference tree.  Plain pointers and sizes only.
 *
 * Conventions
 *   - every function returns 0 on success, a negative PSC_E* code on error; the message of
 *     the last error on the calling thread is psc_last_error().  (The reference prints and
 *     exits, e.g. src/utils.cpp:60-63; the host wrapper host/psascan.hpp does that.)
 *   - pointers are HOST pointers unless the name ends in _d (device pointers on the GPU the
 *     context was created on).  The library works on its own non-blocking CUDA streams and
 *     returns after its work is complete: device inputs must be COMPLETE when a call is made
 *     (synchronise the stream that produced them first), device outputs are ready on return.
 *   - gt bit vectors cross this boundary in the reference's layout: reversed, LSB-first
 *     (include/bitvector.hpp:61-67): for a vector covering text positions (lo, hi], bit r
 *     of byte r>>3 (mask 1 << (r & 7)) belongs to position hi - r, r in [0, hi-lo).
 *   - block-relative suffix array entries are uint32 (blocks are < 2^32 - 2^12 symbols).
 *   - there is no CPU fallback: without a usable CUDA device every call fails with
 *     PSC_ECUDA.
 */
#ifndef PSASCAN_CUDA_H
#define PSASCAN_CUDA_H
#include <stddef.h>
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define PSC_OK 0
#define PSC_EARG (-1)      /* bad argument */
#define PSC_ECUDA (-2)     /* CUDA runtime error / no device */
#define PSC_ESENTINEL (-3) /* byte 255 in the text (initial_partial_sufsort.hpp:70-75) */
#define PSC_ENOMEM (-4)    /* HBM budget exceeded */
#define PSC_ELIMIT (-5)    /* work guard: repeats far longer than expected span a block boundary */

typedef struct psc_context psc_context; /* one GPU, its streams and scratch memory */
typedef struct psc_rank psc_rank;       /* device rank structure over one block BWT */

const char *psc_last_error(void);
int psc_version(void);
int psc_device_count(void); /* visible CUDA devices (0 when there is none or no driver) */

/* Device / stream setup; nothing in the reference corresponds (psascan.hpp:53-131 only
 * budgets host RAM).  device < 0: current device.  hbm_budget = 0: all free memory. */
int psc_init(int device, size_t hbm_budget, psc_context **ctx);
int psc_shutdown(psc_context *ctx);
/* Largest block length (symbols) the budget allows for a text of n symbols whose SA is
 * assembled on the device: replaces max_block_size = ram/5.2 (psascan.hpp:91). */
long psc_max_block_size(psc_context *ctx, long n);

/* ---- stage 1: inmem_psascan<saidx_t,12>()  inmem_psascan_src/inmem_psascan.hpp:64-80,
 * called at partial_sufsort.hpp:166-168, 286-288.
 * Sorts the suffixes of text[0..m) as suffixes of the whole text.
 *   tail_prefix[tail_prefix_len]  the text right after the block (min(m, tail) symbols
 *                                 suffice, inmem_psascan.hpp:150-156); NULL/0 for the last block
 *   tail_gt_rev   gt bits of positions (end, end + tail_prefix_len] relative to `end`
 *                 (suffix(j) > suffix(end)), reversed layout with hi = end+tail_prefix_len;
 *                 NULL when tail_prefix_len == 0
 *   text_after    1 if any text follows the block (tail_prefix_len > 0 is implied)
 *   sa_out[m]     block-relative SA                       (required)
 *   bwt_out[m]    BWT with 0 at i0 (inmem_bwt_from_sa.hpp:48-56); may be NULL
 *   gt_begin_rev  m bits, positions (0, m] relative to the block start, reversed layout with
 *                 hi = m (change_gt_reference_point.hpp:62-118); may be NULL
 *   i0            rank of the block's first suffix */
int psc_sort_block(psc_context *ctx, const uint8_t *text, long m,
                   const uint8_t *tail_prefix, long tail_prefix_len,
                   const uint8_t *tail_gt_rev,
                   uint32_t *sa_out, uint8_t *bwt_out, uint8_t *gt_begin_rev, long *i0);

/* ---- stage 3a: rank4n ctor / dtor / m_count   rank.hpp:115-136, 710, 112.
 * i0 < 0: bwt holds no placeholder.  counts[c] = occurrences of c in bwt, the placeholder
 * NOT counted (the reference counts it as a 0 and compute_gap.hpp:80 takes it out again). */
int psc_rank_build(psc_context *ctx, const uint8_t *bwt, long m, long i0, psc_rank **rank);
int psc_rank_counts(const psc_rank *rank, long counts[256]);
/* rank(i, c)  rank.hpp:566-708, batched: out[k] = #c[k] in bwt[0..i[k]) (placeholder excluded) */
int psc_rank_query(psc_context *ctx, const psc_rank *rank, long nq, const long *i,
                   const uint8_t *c, long *out);
int psc_rank_free(psc_context *ctx, psc_rank *rank);

/* ---- stage 2: em_compute_initial_ranks()  em_compute_initial_ranks.hpp:222-319, 513-561.
 * out[k] = number of suffixes of the block smaller than suffix(starts[k]).
 *   block_text[m], block_sa[m]      the block and its SA
 *   after_text[after_len]           the text from the block end on (pattern source)
 *   starts[k]                       pattern k begins at after_text[starts[k]], 0 <= starts[k] <= after_len
 *   text_end_after                  1 if after_text reaches the end of the whole text
 *   after_gt_rev                    gt bits of positions (0, after_len] of after_text relative
 *                                   to its position 0 (= block end), reversed layout hi = after_len */
int psc_initial_ranks(psc_context *ctx, const uint8_t *block_text, const uint32_t *block_sa,
                      long m, const uint8_t *after_text, long after_len,
                      const uint8_t *after_gt_rev, long nstarts, const long *starts, long *out);

/* ---- stage 3b: compute_gap()  compute_gap.hpp:61-157 + stream.hpp:147-158 + update.hpp:86-96.
 * Streams tail[0..t) right to left against the block's rank structure.
 *   i0, last      rank of the block's first suffix; last symbol of the block
 *   block_counts  counts[c] of the block TEXT symbols (= rank counts + last), C array is
 *                 derived inside (compute_gap.hpp:77-85)
 *   nwalk, walk_beg[nwalk+1]  walker k streams tail[walk_beg[k]..walk_beg[k+1]) starting from
 *                 start_rank[k] = rank of suffix(tail + walk_beg[k+1]); interior boundaries
 *                 must be multiples of 32
 *   gt_in_rev     gt bits of tail positions (0, t] relative to the block END, reversed, hi = t
 *   gt_out_rev    out: gt bits of tail positions (0, t] relative to the block BEGIN; may be NULL
 *   gap[m+1]      out: gap[i] = number of tail suffixes with rank i (uint32 counters, ADDED to) */
int psc_compute_gap(psc_context *ctx, const psc_rank *rank, long i0, uint8_t last,
                    const long block_counts[256], const uint8_t *tail, long t,
                    long nwalk, const long *walk_beg, const long *start_rank,
                    const uint8_t *gt_in_rev, uint8_t *gt_out_rev, uint32_t *gap);

/* ---- stage 4: merge()  merge.hpp:54-180 (one level of it; the host cascades right to left).
 * Interleaves a block SA with the already merged suffixes of everything right of it:
 *   out[u + S(u)] = beg + sa[u],  S = inclusive scan of gap;  the right_len entries of
 *   `right` fill the remaining slots in order.  out has m + right_len entries of 5 bytes
 *   (types/uint40.hpp:42-46).  right is 5-byte entries too. */
int psc_merge_level(psc_context *ctx, const uint32_t *sa, const uint32_t *gap, long m, long beg,
                    const uint8_t *right_sa5, long right_len, uint8_t *out_sa5);

/* ---- glue for a block sorted as two pieces: gap_array.hpp:273-364 (gap -> bitvector),
 * bwt_merge.hpp:66-140, compute_left_gap.hpp:162-306, compute_right_gap.hpp:161-305.
 * bv has len_l + len_r bits, LSB-first, FORWARD (bitvector.hpp:61-67 without reversal). */
int psc_gap_to_bitvector(psc_context *ctx, const uint32_t *gap_l, long len_l, long len_r,
                         uint8_t *bv);
int psc_merge_bwt(psc_context *ctx, const uint8_t *bwt_l, const uint8_t *bwt_r, long len_l,
                  long len_r, long i0_l, long i0_r, uint8_t last_l, const uint8_t *bv,
                  uint8_t *bwt_out, long *i0_out);
int psc_split_gap(psc_context *ctx, const uint8_t *bv, long len_l, long len_r,
                  const uint32_t *block_gap, uint64_t *left_gap, uint64_t *right_gap);

/* ---- the whole path: partial_sufsort() + merge()  partial_sufsort.hpp:559-584,
 * merge.hpp:54-55 as called from psascan.hpp:117-125.  text[n] -> out_sa5[5n].
 * max_block = 0: psc_max_block_size().  Text stays resident in HBM between stages. */
int psc_construct_sa(psc_context *ctx, const uint8_t *text, long n, long max_block,
                     uint8_t *out_sa5);
/* same with device buffers (text_d[n], out_sa5_d[5n]); used for the HBM-resident timing */
int psc_construct_sa_d(psc_context *ctx, const uint8_t *text_d, long n, long max_block,
                       uint8_t *out_sa5_d);

/* ---- the same over several GPUs, one rank per GPU; the reference is single-process, so
 * nothing in it corresponds.  Ranks may be processes (torchrun) or threads of one process (host/psascan.hpp);
 * each creates its own context with psc_init(device) and joins with psc_dist_init.  NCCL (libnccl.so.2,
 * opened at run time) carries every collective, enqueued on the library's stream: no callbacks.
 *   psc_dist_unique_id   rank 0 creates the 128-byte id (ncclGetUniqueId) and hands it to the other ranks
 *                        by any means (torch.distributed, MPI, a shared variable)
 *   psc_dist_init        collective: ncclCommInitRank on the context's device; world == 1 needs no id
 *   psc_dist_finalize    destroys the communicator (psc_shutdown does not)
 *   psc_dist_abort       callable from any thread when a rank has failed: the collectives the other ranks
 *                        wait in end with an error (ncclCommAbort) instead of hanging
 * Schedule (csrc/dist.cu): blocks are sorted on separate GPUs (rank r owns blocks B-1-r, B-1-r-world, ...);
 * per block the owner broadcasts the BWT, gt bits and the walkers' start ranks (the block SA stays with its
 * owner), every GPU walks a contiguous share of the tail against the replicated rank structure, the per-GPU
 * gap counts are summed with ncclReduceScatter, new gt bits are all-gathered; the merge (merge.hpp:54-180)
 * is sharded by OUTPUT range after ONE exchange of SA and gap slices.  Every rank passes the whole text.
 *   psc_construct_sa_dist_slice_d  every rank keeps its slice: entries [*slice_beg, *slice_beg + *slice_len)
 *                                  of the SA as 5-byte values in out_slice_sa5_d, which must hold
 *                                  psc_dist_slice_capacity(n, world) entries
 *   psc_construct_sa_dist_d        every rank receives the whole suffix array (slices broadcast at the end)
 * A context that never called psc_dist_init is a world of one. */
int psc_dist_unique_id(uint8_t id[128]);
int psc_dist_init(psc_context *ctx, const uint8_t id[128], int rank, int world);
int psc_dist_finalize(psc_context *ctx);
int psc_dist_abort(psc_context *ctx);
int psc_construct_sa_dist_d(psc_context *ctx, const uint8_t *text_d, long n, long max_block, uint8_t *out_sa5_d);
int psc_construct_sa_dist_slice_d(psc_context *ctx, const uint8_t *text_d, long n, long max_block,
                                  uint8_t *out_slice_sa5_d, long *slice_beg, long *slice_len);
long psc_dist_slice_capacity(long n, int world);
/* collectives issued by the last distributed construction of this rank and the bytes they moved */
int psc_dist_stats(psc_context *ctx, long *n_collectives, double *bytes);

/* per-phase device time (ms, CUDA events) and counters of the last psc_construct_sa* call */
typedef struct psc_stats {
  double ms_total, ms_gt_end, ms_sort, ms_radix, ms_bwt, ms_rank_build, ms_init_ranks, ms_walk,
      ms_merge, ms_h2d, ms_d2h,
      ms_radix_pass; /* device time of the digit-pass kernels alone (events around each launch) */
  long n_blocks, lf_steps, radix_passes, radix_elems, doubling_rounds, walkers, kernel_launches;
  long radix_bytes; /* algorithmic bytes of the digit passes: 12 in + 12 out per (key64, val32) pair, 8 + 8 for a
                     * keys-only pass, 1 in for a pass that builds its keys from the text */
  long radix_keygen_bytes; /* the share of radix_bytes moved by key-building passes ... */
  double ms_radix_keygen;  /* ... and their share of ms_radix_pass */
  double ms_walk_kernel;   /* device time of the LF-walk kernels alone (events around each launch); ms_walk also
                            * holds the grouping and counting of the ranks they emit */
} psc_stats;
int psc_get_stats(psc_context *ctx, psc_stats *out);

/* device / pinned-host memory for callers that stream a file through the GPU (host/psascan.hpp):
 * the host program never links the CUDA runtime itself. */
int psc_device_alloc(psc_context *ctx, size_t bytes, void **ptr_d);
int psc_device_free(psc_context *ctx, void *ptr_d);
int psc_pinned_alloc(psc_context *ctx, size_t bytes, void **ptr);
int psc_pinned_free(psc_context *ctx, void *ptr);
int psc_copy_h2d(psc_context *ctx, void *dst_d, const void *src, size_t bytes);
int psc_copy_d2h(psc_context *ctx, void *dst, const void *src_d, size_t bytes);
int psc_device_info(psc_context *ctx, char *name, size_t name_len, size_t *free_bytes, size_t *total_bytes);

/* O(n) check of a finished suffix array on the device (sufcheck(), extlib/libdivsufsort/lib/utils.c:161-241):
 * violations[0] = entries >= n, [1] = values that occur twice, [2] = neighbours whose first symbols are out of
 * order, [3] = neighbours with equal first symbols whose following suffixes are out of order (inverse array).
 * All four zero <=> out_sa5_d is THE suffix array of text_d.  n < 2^32 - 1 (4 n bytes of scratch). */
int psc_verify_sa_d(psc_context *ctx, const uint8_t *text_d, long n, const uint8_t *sa5_d, long violations[4]);

/* Check of a SLICE of a suffix array: entries [slice_beg, slice_beg + slice_len) in slice_sa5_d and next_first =
 * entry slice_beg + slice_len (ignored for the last slice).  violations[0] = entries >= n, [1] = adjacent pairs
 * whose suffixes are not in strictly increasing order (compared directly; long common prefixes by fingerprints).
 * All slices clean <=> the array is the suffix array.  Works for any n and needs no exchange between the GPUs that
 * hold the slices (psc_construct_sa_dist_slice_d). */
int psc_verify_sa_slice_d(psc_context *ctx, const uint8_t *text_d, long n, const uint8_t *slice_sa5_d, long slice_beg,
                          long slice_len, long next_first, long violations[2]);

/* The same check for callers that must not allocate (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * The scratch (inverse array + counters) is a static device array inside the library, 16 MiB + 256
 * bytes that every GPU loading the library holds until the process exits; the call issues no cudaMalloc /
 * cudaFree / allocating library routine.  n <= PSC_VERIFY_NOALLOC_MAX_N (= psc_verify_noalloc_max_n()), larger
 * inputs are rejected with PSC_EARG.  All calls share the scratch: it is cleared at the start of every call,
 * and calls that overlap on the GPU (other streams or threads) race -- serialise them.  Same arguments and
 * results as psc_verify_sa_d. */
#define PSC_VERIFY_NOALLOC_MAX_N 4194304L
long psc_verify_noalloc_max_n(void);
int psc_verify_sa_noalloc_d(psc_context *ctx, const uint8_t *text_d, long n, const uint8_t *sa5_d,
                            long violations[4]);

/* ---- external-memory mode: partial suffix arrays and gap arrays as FILES in the reference's
 * formats, merged by streaming.  text_d[n] resident on the device; per block the SA goes to
 * `<out_path>.distrfile.<hash>.part<k>` (block-relative int32 items, at most part_bytes per part:
 * io/distributed_file.hpp:51-125, partial_sufsort.hpp:206-208) and the gap array, v-byte coded on the device, to
 * `<gap_prefix>.gap.<hash>` (gap_array.hpp:156-182, read by io/async_vbyte_stream_reader.hpp:145-168); the merge
 * (merge.hpp:54-180) reads every file front to back once and writes the 40-bit suffix array to out_path.  The
 * temporary files are deleted at the end (merge.hpp:178-179) unless the environment has PSC_KEEP_PARTS.
 * max_block <= 0: psc_max_block_size_files().  part_bytes <= 0: 1 GiB.  This is where -g / gap_filename of
 * pSAscan() goes (a second disk for the gap files, src/main.cpp:207). */
int psc_construct_sa_files_d(psc_context *ctx, const uint8_t *text_d, long n, long max_block, const char *out_path,
                             const char *gap_prefix, long part_bytes);
long psc_max_block_size_files(psc_context *ctx, long n);
/* With PSC_KEEP_PARTS the run also leaves every block's gt_begin bits as the reference stores them (reversed, LSB
 * first, raw bytes: bitvector.hpp:61-79; registered for the reversed index range [n - end, n - beg),
 * partial_sufsort.hpp:231-234) and `<out_path>.parts.txt` with one line per block:
 *   beg end sa_base items_per_part n_parts gap_file gt_begin_file
 * so that any single stage can be re-run on one block from the files (tests/test_em.py does that).
 * psc_multifile_read_bits: bits [r_lo, r_hi) of a multifile bit sequence (io/multifile.hpp:45-68: file k holds bits
 * [beg[k], end[k]) from its own bit 0), packed LSB first from bit 0 of out -- the access pattern of
 * io/multifile_bit_stream_reader.hpp:79-93.  Host only. */
int psc_multifile_read_bits(int n_files, const long *beg, const long *end, const char *const *names, long r_lo, long r_hi,
                            uint8_t *out);
/* the v-byte code of the gap files on host buffers (device codec): out must hold 5 * count bytes;
 * psc_vbyte_decode stores at most cap values and returns the number of complete values in *count */
int psc_vbyte_encode(psc_context *ctx, const uint32_t *vals, long count, uint8_t *out, long *nbytes);
int psc_vbyte_decode(psc_context *ctx, const uint8_t *bytes, long nbytes, uint32_t *vals, long cap, long *count);

/* ---- by-products, all on device buffers.
 * psc_delete_sentinel_bytes_d: tools/delete-sentinel-bytes/src/main.cpp:45-104 -- out_d receives the bytes of
 *   in_d[n] that are not 255, in order (out_d must hold n bytes); *n_out = how many.
 * psc_bwt_from_sa5_d: full-text BWT from the finished suffix array, bwt[i] = text[sa[i] - 1]; the row of suffix 0
 *   holds byte 0 and its index is returned in *i0 (the convention of inmem_bwt_from_sa.hpp:48-56).
 * psc_isa_samples_d: samples_d[p / rate] = i for every i with sa[i] = p, p a multiple of rate
 *   (samples_d holds (n - 1) / rate + 1 entries). */
int psc_delete_sentinel_bytes_d(psc_context *ctx, const uint8_t *in_d, long n, uint8_t *out_d, long *n_out);
int psc_bwt_from_sa5_d(psc_context *ctx, const uint8_t *text_d, long n, const uint8_t *sa5_d, uint8_t *bwt_d, long *i0);
int psc_isa_samples_d(psc_context *ctx, const uint8_t *sa5_d, long n, long rate, uint64_t *samples_d);

/* micro-benchmarks used for the roofline denominators (SURVEY.md 8-d "R32") */
int psc_bench_random_sectors(psc_context *ctx, size_t bytes, int iters, double *gbs);
int psc_bench_radix_pass(psc_context *ctx, long n, int iters, double *ms_per_pass);

#ifdef __cplusplus
}
#endif
#endif
