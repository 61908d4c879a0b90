/*
 * psascan_oracle.c -- CPU restatement of the pSAscan hot path.  TEST INFRASTRUCTURE ONLY
 * (see psascan_oracle.h for the rules and the parity pin).  Plain C11, single-threaded,
 * written for clarity: every function restates what the cited reference code computes, not
 * how fast it computes it.  Citations are into /root/reference/.
 */
#define _GNU_SOURCE
#include "psascan_oracle.h"
#include <ctype.h>
#include <stdlib.h>
#include <string.h>

/* ------------------------------------------------------------------------------------- */
/* src/main.cpp:76-131                                                                    */
int orc_parse_number(const char *str_in, uint64_t *ret) {
  char str[64];
  size_t str_len = strlen(str_in);
  if (str_len >= sizeof(str)) return 0;
  memcpy(str, str_in, str_len + 1);
  *ret = 0;
  size_t n_digits = 0;
  while (n_digits < str_len && isdigit((unsigned char)str[n_digits])) {
    *ret = (*ret) * 10 + (uint64_t)(str[n_digits] - '0');
    ++n_digits;
  }
  if (n_digits == 0) return 0;
  size_t suffix_length = str_len - n_digits;
  if (suffix_length > 0) {
    if (suffix_length > 2) return 0;
    for (size_t j = 0; j < suffix_length; ++j)
      str[n_digits + j] = (char)tolower((unsigned char)str[n_digits + j]);
    if (suffix_length == 2 && str[n_digits + 1] != 'i') return 0;
    int binary = (suffix_length == 2);
    switch (str[n_digits]) {
      case 'k': if (binary) *ret <<= 10; else *ret *= 1000ULL; break;
      case 'm': if (binary) *ret <<= 20; else *ret *= 1000000ULL; break;
      case 'g': if (binary) *ret <<= 30; else *ret *= 1000000000ULL; break;
      case 't': if (binary) *ret <<= 40; else *ret *= 1000000000000ULL; break;
      default: return 0;
    }
  }
  return 1;
}

/* include/psascan.hpp:53-91                                                              */
long orc_max_block_size(long ram_use, long max_threads) {
  const long gap_buf_size = 1L << 21;                       /* :55 */
  long n_gap_buffers = 2 * max_threads;                     /* :56 */
  if (ram_use < 6L) return -1;                              /* :57-60 */
  long ram_for_threads = n_gap_buffers * gap_buf_size;      /* :73 */
  if (ram_use / 5.2L < (long double)(1L << 31))             /* :74-76 */
    ram_for_threads += max_threads * gap_buf_size;
  else
    ram_for_threads += (long)(((4.L / 5) * max_threads) * gap_buf_size);
  ram_for_threads += max_threads * gap_buf_size;            /* :77 */
  ram_for_threads += max_threads * (6L << 20);              /* :78 */
  long excl = ram_use - ram_for_threads;                    /* :80 */
  if (excl < 6L) return -2;                                 /* :81-86 */
  long mbs = (long)(excl / 5.2L);                           /* :91 */
  return mbs < 2 ? 2 : mbs;
}

/* include/types/uint40.hpp:42-46                                                         */
void orc_encode_uint40(const int64_t *sa, long n, uint8_t *out) {
  for (long i = 0; i < n; ++i) {
    uint64_t v = (uint64_t)sa[i];
    out[5 * i + 0] = (uint8_t)(v);
    out[5 * i + 1] = (uint8_t)(v >> 8);
    out[5 * i + 2] = (uint8_t)(v >> 16);
    out[5 * i + 3] = (uint8_t)(v >> 24);
    out[5 * i + 4] = (uint8_t)(v >> 32);
  }
}
void orc_decode_uint40(const uint8_t *in, long n, int64_t *sa) {
  for (long i = 0; i < n; ++i) {
    uint64_t v = 0;
    for (int b = 4; b >= 0; --b) v = (v << 8) | in[5 * i + b];
    sa[i] = (int64_t)v;
  }
}

/* ------------------------------------------------------------------------------------- */
static const uint8_t *g_naive_text;
static long g_naive_n;
static int naive_cmp(const void *a, const void *b) {
  long x = *(const int64_t *)a, y = *(const int64_t *)b;
  long lx = g_naive_n - x, ly = g_naive_n - y;
  int c = memcmp(g_naive_text + x, g_naive_text + y, (size_t)(lx < ly ? lx : ly));
  if (c) return c;
  return (lx < ly) ? -1 : (lx > ly);
}
void orc_sa_naive(const uint8_t *text, long n, int64_t *sa) {
  for (long i = 0; i < n; ++i) sa[i] = i;
  g_naive_text = text; g_naive_n = n;
  qsort(sa, (size_t)n, sizeof(int64_t), naive_cmp);
}

/* Prefix doubling with two stable counting sorts per round.  rank 0 is "past the end". */
void orc_suffix_sort_int(const int32_t *s, long n, int32_t sigma, int64_t *sa) {
  if (n <= 0) return;
  long nb = (n + 1 > (long)sigma + 1 ? n + 1 : (long)sigma + 1) + 1;
  int64_t *rk = malloc(sizeof(int64_t) * (size_t)n);
  int64_t *tmp = malloc(sizeof(int64_t) * (size_t)n);
  int64_t *sa2 = malloc(sizeof(int64_t) * (size_t)n);
  long *cnt = malloc(sizeof(long) * (size_t)(nb + 1));
  /* round 0: sort by first symbol */
  memset(cnt, 0, sizeof(long) * (size_t)(nb + 1));
  for (long i = 0; i < n; ++i) cnt[s[i] + 1]++;
  for (long c = 0; c < nb; ++c) cnt[c + 1] += cnt[c];
  for (long i = 0; i < n; ++i) sa[cnt[s[i]]++] = i;
  long classes = 0;
  for (long i = 0; i < n; ++i) {
    if (i == 0 || s[sa[i]] != s[sa[i - 1]]) ++classes;
    rk[sa[i]] = classes;                        /* ranks 1..classes */
  }
  for (long h = 1; classes < n; h <<= 1) {
    /* sort by second key rk[i+h] (0 if past the end): positions in order of sa shifted */
    memset(cnt, 0, sizeof(long) * (size_t)(nb + 1));
    for (long i = 0; i < n; ++i) cnt[(i + h < n ? rk[i + h] : 0) + 1]++;
    for (long c = 0; c < nb; ++c) cnt[c + 1] += cnt[c];
    for (long i = 0; i < n; ++i) sa2[cnt[(i + h < n ? rk[i + h] : 0)]++] = i;
    /* stable sort by first key */
    memset(cnt, 0, sizeof(long) * (size_t)(nb + 1));
    for (long i = 0; i < n; ++i) cnt[rk[i] + 1]++;
    for (long c = 0; c < nb; ++c) cnt[c + 1] += cnt[c];
    for (long i = 0; i < n; ++i) sa[cnt[rk[sa2[i]]]++] = sa2[i];
    classes = 0;
    for (long i = 0; i < n; ++i) {
      long p = sa[i];
      if (i == 0) ++classes;
      else {
        long q = sa[i - 1];
        long p2 = p + h < n ? rk[p + h] : 0, q2 = q + h < n ? rk[q + h] : 0;
        if (rk[p] != rk[q] || p2 != q2) ++classes;
      }
      tmp[p] = classes;
    }
    memcpy(rk, tmp, sizeof(int64_t) * (size_t)n);
  }
  free(rk); free(tmp); free(sa2); free(cnt);
}

void orc_sa(const uint8_t *text, long n, int64_t *sa) {
  int32_t *s = malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
  for (long i = 0; i < n; ++i) s[i] = text[i];
  orc_suffix_sort_int(s, n, 256, sa);
  free(s);
}

/* ------------------------------------------------------------------------------------- */
/* inmem_psascan_src/compute_initial_gt_bitvectors.hpp:61-148, branch end == text_length of
 * the in-memory text (:72-148): match the block against the tail prefix; decide by
 *   el == tail_length                         -> greater                      (:113)
 *   ran into the block end                    -> !gt_tail[end + el]           (:114)
 *   mismatch                                  -> by the symbols               (:115)       */
void orc_gt_end(const uint8_t *text, long n, long beg, long end,
                const uint8_t *gt_tail, uint8_t *gt_out) {
  long tail_length = n - end;
  for (long p = beg; p < end; ++p) {
    long el = 0;
    while (p + el < end && el < tail_length && text[p + el] == text[end + el]) ++el;
    uint8_t g;
    if (el == tail_length) g = 1;
    else if (p + el == end) g = (uint8_t)!gt_tail[end + el];
    else g = (uint8_t)(text[p + el] > text[end + el]);
    gt_out[p] = g;
  }
}

/* inmem_psascan() output contract (see header). */
long orc_sort_block(const uint8_t *text, long n, long beg, long end,
                    const uint8_t *gt_tail, int64_t *psa_rel, uint8_t *bwt,
                    uint8_t *gt_begin) {
  long m = end - beg;
  if (m <= 0) return 0;
  uint8_t *gt_end = malloc((size_t)(n + 2));
  orc_gt_end(text, n, beg, end, gt_tail, gt_end);
  /* rename_block, initial_partial_sufsort.hpp:61-80: every symbol > last, the last symbol
   * itself, and every symbol == last whose following suffix is greater than suffix(end),
   * is incremented; byte 255 cannot be incremented (:70-75). */
  int32_t *s = malloc(sizeof(int32_t) * (size_t)m);
  uint8_t last = text[end - 1];
  int has255 = 0;
  for (long i = 0; i < m; ++i) {
    uint8_t c = text[beg + i];
    if (c == 255) has255 = 1;
    int inc = 0;
    if (c > last) inc = 1;
    else if (c == last) inc = (i == m - 1) ? 1 : gt_end[beg + i + 1];
    s[i] = (int32_t)c + inc;
  }
  if (has255) { free(gt_end); free(s); return -1; }
  orc_suffix_sort_int(s, m, 257, psa_rel);
  free(s);
  /* inmem_bwt_from_sa.hpp:48-56 */
  long i0 = -1;
  for (long i = 0; i < m; ++i) {
    if (psa_rel[i]) { if (bwt) bwt[i] = text[beg + psa_rel[i] - 1]; }
    else { if (bwt) bwt[i] = 0; i0 = i; }
  }
  /* change_gt_reference_point.hpp:62-118: gt_begin[p] = suffix(p) > suffix(beg) for p in
   * (beg, end]; inside the block that is rank(p) > i0, at p == end it is !gt_end[beg]. */
  if (gt_begin) {
    for (long i = 0; i < m; ++i)
      if (psa_rel[i]) gt_begin[beg + psa_rel[i]] = (uint8_t)(i > i0);
    gt_begin[end] = (uint8_t)!gt_end[beg];
  }
  free(gt_end);
  return i0;
}

/* em_compute_initial_ranks.hpp:54-76 lcp_compare() and :321-363 lcp_compare_2():
 * +1 pattern greater, -1 pattern smaller.  The block suffix is followed up to cmp_end
 * (= the block end for overload 1, = tail_begin through the "mid block" for overload 2,
 * :340-358), where gt_cmp (flags relative to cmp_end) decides. */
static int lcp_compare_full(const uint8_t *text, long n, long cmp_end, long suf,
                            long pat_beg, const uint8_t *gt_cmp) {
  long lcp = 0;
  while (suf + lcp < cmp_end && pat_beg + lcp < n && text[suf + lcp] == text[pat_beg + lcp])
    ++lcp;
  if (suf + lcp >= cmp_end)                                     /* :66-68 */
    return gt_cmp[pat_beg + (cmp_end - suf)] ? 1 : -1;
  if (pat_beg + lcp >= n) return -1;                            /* :69-71 */
  return text[pat_beg + lcp] > text[suf + lcp] ? 1 : -1;        /* :72-75 */
}

long orc_initial_rank(const uint8_t *text, long n, long beg, long end,
                      const int64_t *psa_rel, long pat_beg, long cmp_end,
                      const uint8_t *gt_cmp) {
  if (pat_beg == n) return 0;                                   /* :175-178 */
  long left = 0, right = end - beg;
  while (left != right) {
    long mid = (left + right) / 2;
    if (lcp_compare_full(text, n, cmp_end, beg + psa_rel[mid], pat_beg, gt_cmp) > 0)
      left = mid + 1;
    else
      right = mid;
  }
  return left;
}

/* compute_gap.hpp:77-85 + stream.hpp:147-158 + update.hpp:86-96 */
long orc_compute_gap(const uint8_t *text, const uint8_t *bwt, long m, long i0,
                     uint8_t last, long tail_beg, long tail_end, long start_rank,
                     const uint8_t *gt_in, uint8_t *gt_out, int64_t *gap) {
  /* occ(c, i) on a 64-sample grid; m_count like rank.hpp:112 with the placeholder's 0 */
  const long S = 64;
  long nsamp = m / S + 1;
  long *samp = calloc((size_t)nsamp * 256, sizeof(long));
  long run[256];
  memset(run, 0, sizeof(run));
  for (long i = 0; i < m; ++i) {
    if (i % S == 0) memcpy(samp + (i / S) * 256, run, sizeof(run));
    run[bwt[i]]++;
  }
  if (m % S == 0) memcpy(samp + (m / S) * 256, run, sizeof(run));
  long count[256];
  memcpy(count, run, sizeof(run));
  ++count[last];                                /* compute_gap.hpp:79 */
  --count[0];                                   /* :80 */
  for (long j = 0, s = 0, t; j < 256; ++j) { t = count[j]; count[j] = s; s += t; }
  long i = start_rank;
  for (long j = tail_end; j > tail_beg; --j) {
    uint8_t c = text[j - 1];                    /* stream.hpp:148 */
    if (gt_out) gt_out[j] = (uint8_t)(i > i0);  /* :150 */
    int next_gt = gt_in ? gt_in[j] : 0;         /* :151 */
    int delta = (i > i0 && c == 0);             /* :153 */
    long ii = i < 0 ? 0 : (i > m ? m : i);      /* rank.hpp:567-568 */
    long occ = samp[(ii / S) * 256 + c];
    for (long k = (ii / S) * S; k < ii; ++k) occ += (bwt[k] == c);
    i = count[c] + occ - delta;                 /* :154 */
    if (c == last && next_gt) ++i;              /* :155 */
    gap[i]++;                                   /* update.hpp:86-96 */
  }
  free(samp);
  return i;
}

/* gap_array.hpp:185-224,273-364 */
void orc_gap_to_bv(const int64_t *gap_l, long len_l, long len_r, uint8_t *bv) {
  long p = 0;
  for (long u = 0; u <= len_l; ++u) {
    for (long k = 0; k < gap_l[u]; ++k) bv[p++] = 1;
    if (u < len_l) bv[p++] = 0;
  }
  (void)len_r;
}

/* bwt_merge.hpp:48-56,66-140 */
long orc_merge_bwt(const uint8_t *bwt_l, const uint8_t *bwt_r, long len_l, long len_r,
                   long i0_l, long i0_r, uint8_t last_l, const uint8_t *bv, uint8_t *bwt) {
  long lp = 0, rp = 0, block_i0 = -1;
  for (long i = 0; i < len_l + len_r; ++i) {
    if (bv[i]) {
      bwt[i] = (rp == i0_r) ? last_l : bwt_r[rp];   /* :127 select1(i0_r) patched */
      ++rp;
    } else {
      if (lp == i0_l) block_i0 = i;                 /* :132 select0(i0_l) */
      bwt[i] = bwt_l[lp++];
    }
  }
  return block_i0;
}

/* compute_right_gap.hpp:55-122 and compute_left_gap.hpp:55-123 (the +1 per 1-bit) */
void orc_split_gap(const uint8_t *bv, long len_l, long len_r, const int64_t *block_gap,
                   int64_t *left_gap, int64_t *right_gap) {
  long len = len_l + len_r;
  long lsum = 0, rsum = 0, lk = 0, rk = 0;
  for (long j = 0; j <= len; ++j) {
    lsum += block_gap[j];
    rsum += block_gap[j];
    if (j == len) break;
    if (bv[j]) {              /* element j of the merged block is from the right piece */
      right_gap[rk++] = rsum; rsum = 0;
      lsum += 1;              /* a right-piece suffix also lies right of the left piece */
    } else {
      left_gap[lk++] = lsum; lsum = 0;
    }
  }
  left_gap[lk] = lsum;
  right_gap[rk] = rsum;
}

/* merge.hpp:54-180 without the sqrt(N) super-block bookkeeping (:88-107), which only
 * speeds up "leftmost piece whose gap head is zero" (:123-141). */
void orc_merge(long n_pieces, const long *beg, const long *len,
               const int64_t *const *psa_rel, const int64_t *const *gap, int64_t *sa) {
  long n = 0;
  for (long k = 0; k < n_pieces; ++k) n += len[k];
  long *head = malloc(sizeof(long) * (size_t)n_pieces);
  long *ptr = calloc((size_t)n_pieces, sizeof(long));
  for (long k = 0; k + 1 < n_pieces; ++k) head[k] = gap[k][0];     /* :84-85 */
  head[n_pieces - 1] = 0;                                           /* :86 */
  for (long i = 0; i < n; ++i) {
    long j = 0;
    while (head[j] != 0) { head[j]--; ++j; }                        /* :123-141 */
    sa[i] = psa_rel[j][ptr[j]] + beg[j];                            /* :143 */
    ptr[j]++;
    if (j != n_pieces - 1) head[j] = gap[j][ptr[j]];                /* :145 */
  }
  free(head); free(ptr);
}

void orc_pack_gt_rev(const uint8_t *gt, long n, long lo, long hi, uint8_t *bits) {
  long nbits = hi - lo;
  memset(bits, 0, (size_t)((nbits + 7) / 8));
  for (long j = lo + 1; j <= hi; ++j) {
    long r = (n - j) - (n - hi);                 /* multifile.hpp:57-68 */
    if (gt[j]) bits[r >> 3] |= (uint8_t)(1u << (r & 7));   /* bitvector.hpp:61-67 */
  }
}

/* ------------------------------------------------------------------------------------- */
/* partial_sufsort.hpp:68-551 process_block and :559-584 partial_sufsort, in memory.       */
typedef struct { long beg, len; int64_t *psa; int64_t *gap; } piece_t;

static void stream_tail(const uint8_t *text, long n, const uint8_t *bwt, long m, long i0,
                        uint8_t last, long blk_beg, long blk_end, const int64_t *psa,
                        long tail_beg, long tail_end, long rank_after,
                        const uint8_t *gt_in, uint8_t *gt_out, int64_t *gap) {
  /* one streamer (max_threads = 1): start rank = rank of suffix(tail_end)
   * (partial_sufsort.hpp:193-196: result shifted by one, last entry 0 / after-rank) */
  (void)blk_beg; (void)blk_end; (void)psa; (void)n;
  orc_compute_gap(text, bwt, m, i0, last, tail_beg, tail_end, rank_after, gt_in, gt_out, gap);
}

int orc_psascan(const uint8_t *text, long n, long max_block_size, long ram_use,
                int half_blocks, int64_t *sa) {
  if (n <= 0) return 0;
  long n_blocks = (n + max_block_size - 1) / max_block_size;        /* :564 */
  piece_t *pieces = calloc((size_t)(2 * n_blocks), sizeof(piece_t));
  long n_pieces = 0;
  uint8_t *gt_tail = calloc((size_t)(n + 2), 1);   /* relative to the current block's end  */
  uint8_t *gt_new = calloc((size_t)(n + 2), 1);    /* relative to the current block's begin */
  int rc = 0;
  for (long id = n_blocks - 1; id >= 0 && rc == 0; --id) {          /* :568 */
    long beg = max_block_size * id;
    long end = beg + max_block_size < n ? beg + max_block_size : n;
    long m = end - beg;
    int last_block = (end == n);
    long lsize;                                                     /* :86-89 */
    if (!half_blocks) lsize = m;
    else if (!last_block) lsize = m / 2 > 1 ? m / 2 : 1;
    else { long c = ram_use / 10 > 1 ? ram_use / 10 : 1; lsize = m < c ? m : c; }
    long rsize = m - lsize;
    long lbeg = beg, lend = beg + lsize, rbeg = lend, rend = end;
    memset(gt_new, 0, (size_t)(n + 2));
    if (rsize == 0) {
      /* one piece: stage 1, then stage 2+3 against the tail */
      int64_t *psa = malloc(sizeof(int64_t) * (size_t)m);
      uint8_t *bwt = malloc((size_t)m);
      long i0 = orc_sort_block(text, n, beg, end, gt_tail, psa, bwt, gt_new);
      if (i0 < 0) { rc = -1; free(psa); free(bwt); break; }
      int64_t *gap = calloc((size_t)(m + 1), sizeof(int64_t));
      if (!last_block)
        stream_tail(text, n, bwt, m, i0, text[end - 1], beg, end, psa, end, n, 0,
                    gt_tail, gt_new, gap);
      free(bwt);
      pieces[n_pieces++] = (piece_t){beg, m, psa, gap};
    } else {
      /* STEP 1 right half (:125-239) */
      int64_t *rpsa = malloc(sizeof(int64_t) * (size_t)rsize);
      uint8_t *rbwt = malloc((size_t)rsize);
      uint8_t *gt_r = calloc((size_t)(n + 2), 1);      /* relative to rbeg */
      long ri0 = orc_sort_block(text, n, rbeg, rend, gt_tail, rpsa, rbwt, gt_r);
      /* STEP 2 left half (:245-370): its tail gt is the right half's gt_begin, relative to
       * rbeg = lend and known on (rbeg, rend]; the sorter never looks further because it
       * matches at most min(lsize, n - lend) symbols and lsize <= rsize (+ last block: the
       * match stops at the text end first). */
      int64_t *lpsa = malloc(sizeof(int64_t) * (size_t)lsize);
      uint8_t *lbwt = malloc((size_t)lsize);
      long li0 = ri0 >= 0 ? orc_sort_block(text, n, lbeg, lend, gt_r, lpsa, lbwt, gt_new) : -1;
      if (li0 < 0) { rc = -1; free(rpsa); free(rbwt); free(gt_r); free(lpsa); free(lbwt); break; }
      /* STEP 3 gap of the left half wrt the right half (:376-429); also emits gt relative
       * to lbeg for (rbeg, rend] into gt_new. after_block_initial_rank (:306-317) = rank of
       * suffix(rend) among the left half's suffixes, compared through the right half
       * ("mid block") up to the block end, then by the tail gt (overload 2, :513-561). */
      long after_rank = last_block ? 0 : orc_initial_rank(text, n, lbeg, lend, lpsa, rend, end, gt_tail);
      int64_t *lgap = calloc((size_t)(lsize + 1), sizeof(int64_t));
      orc_compute_gap(text, lbwt, lsize, li0, text[lend - 1], rbeg, rend, after_rank, gt_r,
                      gt_new, lgap);
      if (last_block) {
        pieces[n_pieces++] = (piece_t){lbeg, lsize, lpsa, lgap};
        pieces[n_pieces++] = (piece_t){rbeg, rsize, rpsa, calloc((size_t)(rsize + 1), sizeof(int64_t))};
      } else {
        /* STEP 4 block BWT (:441-489) */
        uint8_t *bv = malloc((size_t)(m + 1));
        orc_gap_to_bv(lgap, lsize, rsize, bv);
        uint8_t *bbwt = malloc((size_t)m);
        long bi0 = orc_merge_bwt(lbwt, rbwt, lsize, rsize, li0, ri0, text[lend - 1], bv, bbwt);
        /* STEP 5 gap of the whole block wrt the tail (:500-517), new gt for (end, n] */
        int64_t *bgap = calloc((size_t)(m + 1), sizeof(int64_t));
        orc_compute_gap(text, bbwt, m, bi0, text[end - 1], end, n, 0, gt_tail, gt_new, bgap);
        /* STEP 6 split (:533-546) */
        int64_t *lg = calloc((size_t)(lsize + 1), sizeof(int64_t));
        int64_t *rg = calloc((size_t)(rsize + 1), sizeof(int64_t));
        orc_split_gap(bv, lsize, rsize, bgap, lg, rg);
        /* the left half's final gap = (vs right half, already in lg via the +1 terms) */
        free(lgap); free(bv); free(bbwt); free(bgap);
        pieces[n_pieces++] = (piece_t){lbeg, lsize, lpsa, lg};
        pieces[n_pieces++] = (piece_t){rbeg, rsize, rpsa, rg};
      }
      free(rbwt); free(lbwt); free(gt_r);
    }
    /* newtail_gt_begin_rev becomes tail_gt_begin_rev (:579-580) */
    uint8_t *t = gt_tail; gt_tail = gt_new; gt_new = t;
  }
  if (rc == 0) {
    /* merge.hpp:59 sort pieces by beg */
    for (long a = 0; a < n_pieces; ++a)
      for (long b = a + 1; b < n_pieces; ++b)
        if (pieces[b].beg < pieces[a].beg) { piece_t t = pieces[a]; pieces[a] = pieces[b]; pieces[b] = t; }
    long *pb = malloc(sizeof(long) * (size_t)n_pieces), *pl = malloc(sizeof(long) * (size_t)n_pieces);
    const int64_t **pp = malloc(sizeof(void *) * (size_t)n_pieces);
    const int64_t **pg = malloc(sizeof(void *) * (size_t)n_pieces);
    for (long k = 0; k < n_pieces; ++k) { pb[k] = pieces[k].beg; pl[k] = pieces[k].len; pp[k] = pieces[k].psa; pg[k] = pieces[k].gap; }
    orc_merge(n_pieces, pb, pl, pp, pg, sa);
    free(pb); free(pl); free(pp); free(pg);
  }
  for (long k = 0; k < n_pieces; ++k) { free(pieces[k].psa); free(pieces[k].gap); }
  free(pieces); free(gt_tail); free(gt_new);
  return rc;
}
